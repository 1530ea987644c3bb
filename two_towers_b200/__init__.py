"""two_towers_b200: B200-native (sm_100a) training + retrieval hot path of mateomarin/two_towers.

Python mirrors the reference's nn.Module / loss surface; all arithmetic runs in libtt_b200.so (hand-written
CUDA behind the C ABI of include/tt_b200.h).  There is no CPU or stock-torch fallback."""
from . import functional  # noqa: F401
from .gru import GRU  # noqa: F401
from .losses import (InfoNCELoss, MarginRankingLoss, SymmetricInfoNCELoss, get_hard_negatives,  # noqa: F401
                     get_hard_negatives_batch)
from .models import (EnhancedTwoTower, EnhancedTwoTowerModel, SimpleTwoTower, TokenTwoTowerModel,  # noqa: F401
                     TwoTowerModel)
from .retrieval import encode_corpus, mrr_recall, search_topk, sharded_search_topk  # noqa: F401

__version__ = "0.1.0"
