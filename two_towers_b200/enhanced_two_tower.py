"""Drop-in for the reference module `enhanced_two_tower` (enhanced_two_tower.py:13-133)."""
from .losses import InfoNCELoss, MarginRankingLoss, get_hard_negatives, get_hard_negatives_batch  # noqa: F401
from .models import EnhancedTwoTower, EnhancedTwoTowerModel  # noqa: F401
