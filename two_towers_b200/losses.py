"""Loss modules with the reference's names, constructor defaults and forward signatures, computed by the fused
score+cross-entropy kernel (K-E) and the hinge kernel (K-G).  Every loss takes an optional process group: with
one, local rows are scored against the all-gathered (global) batch - see distributed.py.

  SymmetricInfoNCELoss   simple_two_tower.py:68-78  (exported there as InfoNCELoss)
  InfoNCELoss            margin_two_tower.py:70-85 == enhanced_two_tower.py:67-82
  MarginRankingLoss      enhanced_two_tower.py:84-121
  get_hard_negatives     enhanced_two_tower.py:123-133
"""
import torch
import torch.nn as nn

from . import distributed as dist_ops
from . import functional as Fn


def _global(vec, group):
    """(global matrix, row offset of the local block, global batch)."""
    if group is None:
        return vec, 0, vec.shape[0]
    return dist_ops.all_gather_rows(vec, group), dist_ops.rank(group) * vec.shape[0], \
        dist_ops.world_size(group) * vec.shape[0]


class SymmetricInfoNCELoss(nn.Module):
    """(CE(S, arange) + CE(S^T, arange)) / 2 with S = q d^T / temperature; inputs are used as given."""

    def __init__(self, temperature=0.1, group=None):
        super().__init__()
        self.temperature, self.group = temperature, group

    def forward(self, query_vec, doc_vec):
        it = 1.0 / self.temperature
        d_all, off, Bg = _global(doc_vec, self.group)
        q_all, _, _ = _global(query_vec, self.group)
        q2d = Fn.score_ce_sum(query_vec, d_all, off, it)
        d2q = Fn.score_ce_sum(doc_vec, q_all, off, it)
        return dist_ops.sum_over_ranks((q2d + d2q) / (2 * Bg), self.group)


class InfoNCELoss(nn.Module):
    """L2-normalise both sides, then one-directional CE over q d^T / temperature."""

    def __init__(self, temperature=0.07, group=None):
        super().__init__()
        self.temperature, self.group = temperature, group

    def forward(self, query_vec, doc_vec):
        q, d = Fn.l2_normalize(query_vec), Fn.l2_normalize(doc_vec)
        d_all, off, Bg = _global(d, self.group)
        return dist_ops.sum_over_ranks(Fn.score_ce_sum(q, d_all, off, 1.0 / self.temperature) / Bg, self.group)


class MarginRankingLoss(nn.Module):
    def __init__(self, margin=0.2, temperature=0.1, group=None):
        super().__init__()
        self.margin, self.temperature, self.group = margin, temperature, group

    def forward(self, query_vec, pos_doc_vec, neg_doc_vec=None):
        if neg_doc_vec is None:
            d_all, off, Bg = _global(pos_doc_vec, self.group)
            s = Fn.score_ce_sum(query_vec, d_all, off, 1.0 / self.temperature, self.margin)
            return dist_ops.sum_over_ranks(s / Bg, self.group)
        B = query_vec.shape[0]
        Bg = B * (dist_ops.world_size(self.group) if self.group is not None else 1)
        return dist_ops.sum_over_ranks(Fn.hinge_sum(query_vec, pos_doc_vec, neg_doc_vec, self.margin) / Bg, self.group)


def get_hard_negatives(query_vec, doc_vecs, positive_idx, k=5):
    """One query [D] vs docs [N,D]: cosine, positive := -1, top-k indices (int64 [k]); lowest index wins ties."""
    return Fn.rows_topk(query_vec.unsqueeze(0), doc_vecs, k, cosine=True, mask_pos_offset=int(positive_idx))[1][0]


def get_hard_negatives_batch(query_vecs, doc_vecs, k=5, pos_offset=0):
    """Row i == get_hard_negatives(query_vecs[i], doc_vecs, pos_offset + i, k), one kernel for the whole batch."""
    return Fn.rows_topk(query_vecs, doc_vecs, k, cosine=True, mask_pos_offset=int(pos_offset))[1]
