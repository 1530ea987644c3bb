"""Drop-in two-tower models: same constructor arguments, forward signatures, attribute names and state_dict
keys as the reference classes, computing through libtt_b200 (no stock-torch path, no CPU fallback).

  SimpleTwoTower         simple_two_tower.py:14-66
  TwoTowerModel (margin) margin_two_tower.py:9-68
  EnhancedTwoTowerModel  enhanced_two_tower.py:13-65   (alias EnhancedTwoTower: validate_enhanced.py:7)
  TokenTwoTowerModel     call sites train_two_tower.py:29,44-45,61-63,83-84 (its source is missing upstream)

Additive surface (not in the reference): `attach_embeddings(table, pad_id)` lets every model take integer token
ids [B,T] instead of pre-embedded floats - the gather then runs on the device - and `lengths=`
gives pack_padded_sequence semantics (default None = reference behaviour).
"""
import torch
import torch.nn as nn

from . import functional as Fn
from .gru import GRU


def _head_modules(seq):
    lin1, ln = seq[0], seq[1]
    lin2 = next((m for m in list(seq)[2:] if isinstance(m, nn.Linear)), None)
    drop = next((m for m in seq if isinstance(m, nn.Dropout)), None)
    return lin1, ln, lin2, drop


class _TwoTowerBase(nn.Module):
    _normalize_in_encode = False

    def __init__(self):
        super().__init__()
        self._emb_table = None
        self._pad_id = -1

    # ---- additive token-id front end
    def attach_embeddings(self, table, pad_id=0):
        """Frozen embedding matrix [V,E]; row `pad_id` reads as zeros (the dataset's zero padding)."""
        self._emb_table = table.detach()
        self._pad_id = int(pad_id)
        return self

    def _apply(self, fn, *a, **k):
        super()._apply(fn, *a, **k)
        if self._emb_table is not None:
            self._emb_table = fn(self._emb_table)
        return self

    # The two towers are independent until the loss, and the persistent recurrence kernels of one tower leave SMs
    # idle (the BPTT kernel runs one CTA per (128-row tile, direction)): the query tower goes on a side stream so the
    # hardware can co-schedule it with the doc tower, forward and (autograd replays each node on its forward stream)
    # backward.  Set `overlap_towers = False` for strictly serial execution on the caller's stream.
    overlap_towers = True
    _side_streams = {}

    def _two_towers(self, fq, q_in, fd, d_in):
        if not (self.overlap_towers and q_in.is_cuda and d_in.is_cuda):
            return fq(q_in), fd(d_in)
        cur = torch.cuda.current_stream(q_in.device)
        key = (q_in.device.index, cur.cuda_stream)
        side = _TwoTowerBase._side_streams.get(key)
        if side is None:
            side = _TwoTowerBase._side_streams[key] = torch.cuda.Stream(q_in.device)
            # gradients of the query tower are produced on the side stream by design; autograd synchronises the
            # accumulation itself and only warns about it
            warn_off = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
            if warn_off is not None:
                warn_off(False)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            q_in.record_stream(side)
            q = fq(q_in)
        d = fd(d_in)
        cur.wait_stream(side)
        q.record_stream(cur)
        return q, d

    def _tower(self, x, encoder, proj, lengths=None):
        if not x.is_cuda:
            raise RuntimeError("two_towers_b200 models run on CUDA only (no CPU fallback); call .to('cuda')")
        table = self._emb_table if not torch.is_floating_point(x) else None
        _, h_n = encoder(x, lengths=lengths, table=table, pad_id=self._pad_id, need_output=False)
        lin1, ln, lin2, drop = _head_modules(proj)
        p = drop.p if (drop is not None and self.training) else 0.0
        return Fn.head(h_n[-2], h_n[-1], lin1.weight, lin1.bias, ln.weight, ln.bias,
                       lin2.weight if lin2 is not None else None, lin2.bias if lin2 is not None else None,
                       normalize=self._normalize_in_encode, dropout_p=p)


class SimpleTwoTower(_TwoTowerBase):
    _normalize_in_encode = True

    def __init__(self, embedding_dim: int, hidden_dim: int):
        super().__init__()
        mk = lambda: GRU(embedding_dim, hidden_dim, num_layers=2, batch_first=True, bidirectional=True, dropout=0.1)
        self.query_encoder = mk()
        self.doc_encoder = mk()
        hd = lambda: nn.Sequential(nn.Linear(hidden_dim * 2, hidden_dim), nn.LayerNorm(hidden_dim), nn.ReLU(),
                                   nn.Dropout(0.1), nn.Linear(hidden_dim, hidden_dim // 2))
        self.query_proj = hd()
        self.doc_proj = hd()

    def encode_query(self, query_emb, lengths=None):
        return self._tower(query_emb, self.query_encoder, self.query_proj, lengths)

    def encode_doc(self, doc_emb, lengths=None):
        return self._tower(doc_emb, self.doc_encoder, self.doc_proj, lengths)

    def forward(self, query_emb, doc_emb):
        return self._two_towers(self.encode_query, query_emb, self.encode_doc, doc_emb)


class TwoTowerModel(_TwoTowerBase):
    """margin_two_tower.TwoTowerModel: shared projection, mode-dependent forward."""

    def __init__(self, embedding_dim: int, hidden_dim: int):
        super().__init__()
        mk = lambda: GRU(embedding_dim, hidden_dim, num_layers=2, batch_first=True, bidirectional=True, dropout=0.1)
        self.query_encoder = mk()
        self.doc_encoder = mk()
        self.projection = nn.Sequential(nn.Linear(hidden_dim * 2, hidden_dim), nn.LayerNorm(hidden_dim), nn.ReLU(),
                                        nn.Dropout(0.1))

    def compute_similarity(self, query_vec, doc_vec):
        query_vec = Fn.l2_normalize(query_vec)
        doc_vec = Fn.l2_normalize(doc_vec)
        if self.training:
            return query_vec, doc_vec
        return Fn.gemm_nt(query_vec, doc_vec)      # eval: [B,B] similarity (margin_two_tower.py:46-48)

    def encode(self, emb, encoder, lengths=None):
        return self._tower(emb, encoder, self.projection, lengths)

    def encode_query(self, query_emb, lengths=None):
        return self.encode(query_emb, self.query_encoder, lengths)

    def encode_doc(self, doc_emb, lengths=None):
        return self.encode(doc_emb, self.doc_encoder, lengths)

    def forward(self, query_emb, doc_emb):
        q, d = self._two_towers(lambda x: self.encode(x, self.query_encoder), query_emb,
                                lambda x: self.encode(x, self.doc_encoder), doc_emb)
        return self.compute_similarity(q, d)


class EnhancedTwoTowerModel(_TwoTowerBase):
    def __init__(self, embedding_dim: int, hidden_dim: int):
        super().__init__()
        mk = lambda: GRU(embedding_dim, hidden_dim * 2, num_layers=2, batch_first=True, bidirectional=True, dropout=0.1)
        self.query_encoder = mk()
        self.doc_encoder = mk()
        hd = lambda: nn.Sequential(nn.Linear(hidden_dim * 4, hidden_dim * 2), nn.LayerNorm(hidden_dim * 2), nn.ReLU(),
                                   nn.Linear(hidden_dim * 2, hidden_dim))
        self.query_proj = hd()
        self.doc_proj = hd()

    def encode_query(self, query_emb, lengths=None):
        return self._tower(query_emb, self.query_encoder, self.query_proj, lengths)

    def encode_doc(self, doc_emb, lengths=None):
        return self._tower(doc_emb, self.doc_encoder, self.doc_proj, lengths)

    def forward(self, query_emb, doc_emb):
        return self._two_towers(self.encode_query, query_emb, self.encode_doc, doc_emb)


EnhancedTwoTower = EnhancedTwoTowerModel


class TokenTwoTowerModel(nn.Module):
    """Token-id model pinned only by its call sites (train_two_tower.py:29,44-45,61-63,83-84): frozen pretrained
    embedding + one uni-directional 1-layer GRU per tower (checkpoint size 1 323 881 B ~ 2 x GRU(300->128),
    SURVEY 8c), output = final hidden state, NOT normalised (the caller normalises, train_two_tower.py:63,84).
    Row `padding_idx` of the table reads as zeros."""

    def __init__(self, vocab_size, embedding_dim, hidden_dim, pretrained_embeddings=None, freeze_embeddings=True,
                 padding_idx=0):
        super().__init__()
        w = torch.zeros(vocab_size, embedding_dim) if pretrained_embeddings is None else \
            torch.as_tensor(pretrained_embeddings, dtype=torch.float32).clone()
        if not freeze_embeddings:
            raise NotImplementedError("trainable embeddings are outside the hot path (the reference freezes Word2Vec)")
        self.register_buffer("embedding_weight", w)
        self.padding_idx = padding_idx
        self.query_encoder = GRU(embedding_dim, hidden_dim, num_layers=1, batch_first=True)
        self.doc_encoder = GRU(embedding_dim, hidden_dim, num_layers=1, batch_first=True)

    def _enc(self, ids, enc, lengths=None):
        if not ids.is_cuda:
            raise RuntimeError("two_towers_b200 models run on CUDA only (no CPU fallback)")
        return enc(ids, lengths=lengths, table=self.embedding_weight, pad_id=self.padding_idx, need_output=False)[1][-1]

    def encode_query(self, query, lengths=None):
        return self._enc(query, self.query_encoder, lengths)

    def encode_doc(self, doc, lengths=None):
        return self._enc(doc, self.doc_encoder, lengths)

    def forward(self, query, pos, neg=None):
        q, p = self.encode_query(query), self.encode_doc(pos)
        return (q, p) if neg is None else (q, p, self.encode_doc(neg))
