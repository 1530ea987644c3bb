"""TEST INFRASTRUCTURE ONLY - CPU restatement ("port") of the reference hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this file.  The product package
(two_towers_b200/) never does, and fails loudly without its CUDA library.

The reference's arithmetic lives in stock PyTorch (torch is its only numeric
dependency, pinned torch==2.1.0 only in server/python-api/requirements.txt:5);
this port composes the same stock CPU ops in a table-driven way.  It is pinned
against the real reference by tests/golden/*.npz (made by oracle/gen_golden.py,
which imports /root/reference) and, when /root/reference is present, directly by
tests/test_oracle_pinned.py.

Reference lines restated here
  simple_two_tower.py:14-66    SimpleTwoTower           -> build_port("simple", E, H)
  simple_two_tower.py:68-78    InfoNCELoss (symmetric)  -> info_nce_symmetric
  margin_two_tower.py:9-68     TwoTowerModel            -> build_port("margin", E, H)
  margin_two_tower.py:70-85    InfoNCELoss (q->d)       -> info_nce_q2d
  enhanced_two_tower.py:13-65  EnhancedTwoTowerModel    -> build_port("enhanced", E, H)
  enhanced_two_tower.py:84-121 MarginRankingLoss        -> margin_ce_inbatch / hinge_explicit_negatives
  enhanced_two_tower.py:123-133 get_hard_negatives      -> hard_negatives
  simple_two_tower.py:333-340  search_ms_marco          -> search_topk
  simple_two_tower.py:364-373  MRR, compare_models.py:55-74 recall -> mrr_recall
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

# kind -> (GRU hidden multiplier, head layer spec, shared head?, normalise in encode?)
# head spec entries: ("lin", in_mult, out_mult) in units of hidden_dim, "ln", "relu", "drop"
PORT_SPECS = {
    "simple":   dict(gmul=1, shared=False, norm_in_encode=True,
                     head=[("lin", 2, 1), ("ln", 1), ("relu",), ("drop",), ("lin", 1, 0.5)]),
    "margin":   dict(gmul=1, shared=True, norm_in_encode=False,
                     head=[("lin", 2, 1), ("ln", 1), ("relu",), ("drop",)]),
    "enhanced": dict(gmul=2, shared=False, norm_in_encode=False,
                     head=[("lin", 4, 2), ("ln", 2), ("relu",), ("lin", 2, 1)]),
}


def _make_head(spec, H):
    mods = []
    for item in spec:
        if item[0] == "lin":
            mods.append(nn.Linear(int(item[1] * H), int(item[2] * H)))
        elif item[0] == "ln":
            mods.append(nn.LayerNorm(int(item[1] * H)))
        elif item[0] == "relu":
            mods.append(nn.ReLU())
        else:
            mods.append(nn.Dropout(0.1))
    return nn.Sequential(*mods)


class PortTwoTower(nn.Module):
    """state_dict-compatible CPU port of the three reference models."""

    def __init__(self, kind, embedding_dim, hidden_dim):
        super().__init__()
        s = PORT_SPECS[kind]
        self.kind = kind
        enc = lambda: nn.GRU(embedding_dim, hidden_dim * s["gmul"], num_layers=2,
                             batch_first=True, bidirectional=True, dropout=0.1)
        # creation order == reference order, so torch.manual_seed(s) gives the same init
        self.query_encoder = enc()
        self.doc_encoder = enc()
        if s["shared"]:
            self.projection = _make_head(s["head"], hidden_dim)
        else:
            self.query_proj = _make_head(s["head"], hidden_dim)
            self.doc_proj = _make_head(s["head"], hidden_dim)

    def _tower(self, x, enc, head):
        h_n = enc(x)[1]
        v = head(torch.cat((h_n[-2], h_n[-1]), dim=1))
        return F.normalize(v, p=2, dim=1) if PORT_SPECS[self.kind]["norm_in_encode"] else v

    def encode_query(self, x):
        return self._tower(x, self.query_encoder,
                           self.projection if self.kind == "margin" else self.query_proj)

    def encode_doc(self, x):
        return self._tower(x, self.doc_encoder,
                           self.projection if self.kind == "margin" else self.doc_proj)

    def forward(self, q, d):
        qv, dv = self.encode_query(q), self.encode_doc(d)
        if self.kind != "margin":
            return qv, dv
        qv, dv = F.normalize(qv, p=2, dim=1), F.normalize(dv, p=2, dim=1)
        return (qv, dv) if self.training else qv @ dv.t()


def build_port(kind, embedding_dim, hidden_dim):
    return PortTwoTower(kind, embedding_dim, hidden_dim)


def disable_dropout(model):
    """Parity runs: make train() deterministic."""
    for m in model.modules():
        if isinstance(m, nn.Dropout):
            m.p = 0.0
        if isinstance(m, nn.GRU):
            m.dropout = 0.0
    return model


# ----------------------------------------------------------------------------- losses
def _ce_diag(logits):
    return F.cross_entropy(logits, torch.arange(logits.shape[0], device=logits.device))


def info_nce_symmetric(q, d, temperature=0.1):
    s = (q @ d.t()) / temperature
    return (_ce_diag(s) + _ce_diag(s.t())) / 2


def info_nce_q2d(q, d, temperature=0.07):
    q, d = F.normalize(q, p=2, dim=1), F.normalize(d, p=2, dim=1)
    return _ce_diag((q @ d.t()) / temperature)


def margin_ce_inbatch(q, d, margin=0.2, temperature=0.1):
    s = (q @ d.t()) / temperature
    s = s - margin * (1 - torch.eye(q.shape[0], device=q.device))
    return _ce_diag(s)


def hinge_explicit_negatives(q, pos, neg, margin=0.2):
    B = q.shape[0]
    k = neg.shape[0] // B
    pos_sim = F.cosine_similarity(q, pos)
    neg_sim = F.cosine_similarity(q.unsqueeze(1).expand(-1, k, -1), neg.view(B, k, -1), dim=2).mean(dim=1)
    return torch.clamp(margin - pos_sim + neg_sim, min=0).mean()


# ----------------------------------------------------------------------------- mining / retrieval
def stable_topk(scores, k):
    """Descending top-k with LOWEST-INDEX tie-break (torch.topk's tie order is unspecified)."""
    order = torch.sort(scores, dim=-1, descending=True, stable=True).indices[..., :k]
    return torch.gather(scores, -1, order), order


def hard_negatives(query_vec, doc_vecs, positive_idx, k=5):
    with torch.no_grad():
        sim = F.cosine_similarity(query_vec.unsqueeze(0), doc_vecs).clone()
        sim[positive_idx] = -1
        return stable_topk(sim, k)[1]


def hard_negatives_rows(q, docs, pos_offset=0, k=5):
    """Row i = hard_negatives(q[i], docs, pos_offset + i, k)."""
    return torch.stack([hard_negatives(q[i], docs, pos_offset + i, k) for i in range(q.shape[0])])


def search_topk(query_vec, doc_encodings, k=10):
    """query_vec [1,D] (or [Q,D] -> row by row), doc_encodings [N,D] -> (scores, idx)."""
    outs = [stable_topk(F.cosine_similarity(query_vec[i:i + 1], doc_encodings), k)
            for i in range(query_vec.shape[0])]
    return torch.stack([o[0] for o in outs]), torch.stack([o[1] for o in outs])


def mrr_recall(topk_idx, relevant, ks=(1, 3, 10)):
    """topk_idx [Q,k] int, relevant: list of sets (or 1-D tensor of one id per query)."""
    mrr_sum, hits = 0.0, {k: 0 for k in ks}
    Q = topk_idx.shape[0]
    for i in range(Q):
        rel = relevant[i]
        rel = rel if isinstance(rel, (set, frozenset)) else {int(rel)}
        first = next((r for r, doc in enumerate(topk_idx[i].tolist(), 1) if doc in rel), None)
        mrr_sum += 0.0 if first is None else 1.0 / first
        for k in ks:
            hits[k] += int(first is not None and first <= k)
    out = {"MRR": mrr_sum / Q}
    out.update({f"R@{k}": hits[k] / Q for k in ks})
    return out
