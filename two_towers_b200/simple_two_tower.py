"""Drop-in for the reference module `simple_two_tower` (model + loss names as exported there,
simple_two_tower.py:14-78).  `from two_towers_b200.simple_two_tower import SimpleTwoTower, InfoNCELoss`."""
from .losses import SymmetricInfoNCELoss as InfoNCELoss  # noqa: F401  (symmetric variant, default temperature 0.1)
from .models import SimpleTwoTower  # noqa: F401
