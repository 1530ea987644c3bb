"""Drop-in for the reference module `margin_two_tower` (margin_two_tower.py:9-85)."""
from .losses import InfoNCELoss  # noqa: F401  (normalising, one-directional, default temperature 0.07)
from .models import TwoTowerModel  # noqa: F401
