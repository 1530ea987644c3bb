"""two_towers_b200: B200-native (sm_100a) training + retrieval hot path of mateomarin/two_towers.

Python mirrors the reference's nn.Module / loss surface; all arithmetic runs in libtt_b200.so (hand-written
CUDA behind the C ABI of include/tt_b200.h).  There is no CPU or stock-torch fallback."""
from . import functional  # noqa: F401
from .functional import engine, set_engine, use_engine  # noqa: F401
from .graph import GraphedTowers, GraphedTrainStep  # noqa: F401
from .gru import GRU  # noqa: F401
from .losses import (InfoNCELoss, MarginRankingLoss, SymmetricInfoNCELoss, get_hard_negatives,  # noqa: F401
                     get_hard_negatives_batch)
from .models import (EnhancedTwoTower, EnhancedTwoTowerModel, SimpleTwoTower, TokenTwoTowerModel,  # noqa: F401
                     TwoTowerModel)
from .optim import FusedAdam, FusedAdamW, load_checkpoint, save_checkpoint  # noqa: F401
from .pipeline import PinnedIdRing, Vocabulary, id_batches  # noqa: F401
from .retrieval import (CorpusIndex, SearchService, compute_mrr, encode_corpus, first_relevant_rank, load_doc_embeddings, load_or_compute_doc_embeddings,  # noqa: F401
                        merge_shard_topk, mrr_recall, save_doc_embeddings, search_topk, sharded_search_topk)

__version__ = "0.1.0"
