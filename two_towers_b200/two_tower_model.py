"""Stand-in for the module `two_tower_model` that train_two_tower.py:8 imports but the reference does not ship."""
from .models import TokenTwoTowerModel as TwoTowerModel  # noqa: F401
