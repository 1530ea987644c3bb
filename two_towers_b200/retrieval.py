"""Retrieval scan + MRR@k (reference: search_ms_marco simple_two_tower.py:333-340 == validate_enhanced.py:73-80,
infer_top_k train_two_tower.py:67-90, app.py:94-101; MRR simple_two_tower.py:364-373; recall compare_models.py:55-74).

The reference scores ONE query against the whole corpus with F.cosine_similarity + torch.topk; here a batch of
queries is scored against a corpus shard by the fused exact-fp32 score + streaming top-k kernel (K-H), shards are
merged by K-I after an all-gather of the [Q,k] lists, and MRR / recall come from the rank kernel.
"""
import torch
import torch.distributed as dist

from . import distributed as dist_ops
from . import functional as Fn


class CorpusIndex:
    """doc_encodings [N,D] + the per-corpus data of the scan (row norms, 16-bit hi/lo planes), built once.
    The reference keeps `doc_encodings` and searches it query by query (simple_two_tower.py:318-340, app.py:38-60,
    94-101); passing a CorpusIndex instead of the tensor to search_topk / sharded_search_topk skips the per-call
    conversion of the corpus (results are bit-identical either way)."""

    def __init__(self, doc_encodings, cosine=True, index_base=0):
        from . import _lib as L
        self.docs = L.f32c(doc_encodings)
        self.cosine = bool(cosine)
        self.index_base = int(index_base)
        self.buffer = Fn.corpus_index_build(self.docs, self.cosine)

    def __len__(self):
        return self.docs.shape[0]

    @torch.no_grad()
    def search(self, query_vecs, k=10, mask_pos_offset=-1):
        if query_vecs.dim() == 1:
            query_vecs = query_vecs.unsqueeze(0)
        return Fn.rows_topk_indexed(query_vecs, self.docs, self.buffer, k, cosine=self.cosine,
                                    mask_pos_offset=mask_pos_offset, index_base=self.index_base)


@torch.no_grad()
def search_topk(query_vecs, doc_encodings, k=10, cosine=True, index_base=0):
    """query_vecs [Q,D] (or [D]) vs doc_encodings [N,D] (tensor or CorpusIndex) -> (scores [Q,k], indices int64 [Q,k]),
    best first; ties go to the lowest index.  Scores are the exact fp32 FMA chain in ascending d (the arithmetic of the
    oracle's restatement); indices equal what a stable descending sort of those scores gives - which is cosine_similarity +
    topk of the reference up to torch's unspecified tie order and cuBLAS summation order.  The tensor-engine candidate
    selection in front of the exact re-score accepts a row only when its k-th score clears every non-candidate by more than
    the measured error bound (corpus AND query rounding residuals included); other rows are re-scanned exactly."""
    if isinstance(doc_encodings, CorpusIndex):
        if doc_encodings.cosine != bool(cosine):
            raise ValueError("search_topk: the CorpusIndex was built for the other score mode")
        return doc_encodings.search(query_vecs, k)
    if query_vecs.dim() == 1:
        query_vecs = query_vecs.unsqueeze(0)
    return Fn.rows_topk(query_vecs, doc_encodings, k, cosine=cosine, index_base=index_base)


@torch.no_grad()
def merge_shard_topk(val, idx, group=None):
    """Per-shard top-k lists (global doc ids) of every rank -> the identical global top-k on every rank: all-gather of the
    [Q,k] lists + K-I merge (score desc, lowest global id on ties)."""
    W = dist_ops.world_size(group)
    if W == 1:
        return val, idx
    vals = torch.empty(W, *val.shape, device=val.device, dtype=val.dtype)
    idxs = torch.empty(W, *idx.shape, device=idx.device, dtype=idx.dtype)
    dist.all_gather_into_tensor(vals, val.contiguous(), group=group)
    dist.all_gather_into_tensor(idxs, idx.contiguous(), group=group)
    return Fn.topk_merge(vals, idxs)


def sharded_search_topk(query_vecs, local_docs, index_base, k=10, group=None, cosine=True):
    """Every rank holds a contiguous corpus shard starting at global id `index_base`; queries are replicated.
    Returns the identical global top-k on every rank."""
    if isinstance(local_docs, CorpusIndex) and local_docs.index_base != int(index_base):
        raise ValueError("sharded_search_topk: index_base differs from the CorpusIndex's")
    val, idx = search_topk(query_vecs, local_docs, k, cosine, index_base)
    return merge_shard_topk(val, idx, group)


@torch.no_grad()
def mrr_recall(topk_idx, relevant, ks=(1, 3, 10)):
    """topk_idx int64 [Q,k]; relevant int64 [Q] (one relevant doc id per query, <0 = none).
    Returns {'MRR': mean reciprocal rank of the first relevant hit within k, 'R@n': hit rate at n}."""
    first, rr = Fn.mrr_ranks(topk_idx, relevant)
    out = {"MRR": rr.double().mean().item()}
    for n in ks:
        out[f"R@{n}"] = ((first > 0) & (first <= n)).double().mean().item()
    return out


@torch.no_grad()
def encode_corpus(model, docs, batch_size=4096, which="doc", lengths=None):
    """Batched corpus encoder (the reference encodes one doc at a time: validate_enhanced.py:61-71, app.py:55-62).
    docs: float [N,T,E] or ids [N,T] on the model's device.  Returns [N,D] fp32."""
    enc = model.encode_doc if which == "doc" else model.encode_query
    was_training = model.training
    model.eval()
    outs = []
    for i in range(0, docs.shape[0], batch_size):
        ln = lengths[i:i + batch_size] if lengths is not None else None
        outs.append(enc(docs[i:i + batch_size], ln) if ln is not None else enc(docs[i:i + batch_size]))
    model.train(was_training)
    return torch.cat(outs, 0)


# ------------------------------------------------------------------------------------------- N2: embedding cache
def save_doc_embeddings(path, doc_embeddings):
    """The reference's on-disk cache is the bare [N,D] tensor (`torch.save(doc_embeddings, DOC_EMBEDDINGS_CACHE)`,
    server/python-api/app.py:65)."""
    torch.save(doc_embeddings, path)


def load_doc_embeddings(path, device):
    """`torch.load(DOC_EMBEDDINGS_CACHE, map_location=device)` (app.py:49)."""
    t = torch.load(path, map_location=device)
    if not torch.is_tensor(t) or t.dim() != 2:
        raise ValueError(f"{path}: expected a 2-D tensor of document embeddings")
    return t.float().contiguous()


@torch.no_grad()
def load_or_compute_doc_embeddings(model, doc_ids, cache_path=None, batch_size=4096):
    """app.py:41-67 (`load_or_compute_doc_embeddings`) with the batch-1 encode loop replaced by `encode_corpus`.
    doc_ids: int ids [N,T] (CPU or device; moved batch by batch)."""
    import os
    dev = next(model.parameters()).device
    if cache_path is not None and os.path.exists(cache_path):
        return load_doc_embeddings(cache_path, dev)
    outs = []
    was_training = model.training
    model.eval()
    for i in range(0, doc_ids.shape[0], batch_size):
        outs.append(model.encode_doc(doc_ids[i:i + batch_size].to(dev, non_blocking=True)))
    model.train(was_training)
    emb = torch.cat(outs, 0) if outs else torch.empty(0, 0, device=dev)
    if cache_path is not None:
        save_doc_embeddings(cache_path, emb)
    return emb


# ------------------------------------------------------------------------------------------- a16: full-rank MRR
@torch.no_grad()
def first_relevant_rank(query_vecs, doc_vecs, relevant):
    """Reciprocal rank of the first relevant candidate in the FULL ranking of the candidates
    (validate_margin.compute_mrr :37-54: cosine similarity, stable descending sort - ties keep the candidate order -,
    1/rank of the first hit, 0 without one).  query_vecs [Q,D]; doc_vecs [N,D] (one candidate list shared by the queries)
    or [Q,N,D]; relevant bool [Q,N].  Returns (rr float64 [Q], top3 indices [Q,min(3,N)], top3 scores).
    A helper on stock torch ops (cosine_similarity + stable sort over a few hundred candidates), not one of the library's
    kernels and not a hot path: the reference ranks ~1000 candidates per query here."""
    if query_vecs.dim() == 1:
        query_vecs = query_vecs.unsqueeze(0)
    docs = doc_vecs.unsqueeze(0) if doc_vecs.dim() == 2 else doc_vecs
    sims = torch.nn.functional.cosine_similarity(query_vecs.unsqueeze(1), docs, dim=2)            # [Q,N], as :38-42
    relevant = torch.as_tensor(relevant, dtype=torch.bool, device=sims.device).reshape(sims.shape)
    val, order = torch.sort(sims, dim=1, descending=True, stable=True)
    hit = torch.gather(relevant, 1, order)
    first = torch.argmax(hit.to(torch.int8), dim=1)
    rr = torch.where(hit.any(dim=1), 1.0 / (first.double() + 1.0), torch.zeros_like(first, dtype=torch.float64))
    n3 = min(3, sims.shape[1])
    return rr, order[:, :n3], val[:, :n3]


def compute_mrr(model, query, relevant_docs, candidate_docs, vocab, max_length=30, batch_size=128):
    """Text-level drop-in for validate_margin.compute_mrr(model, query, relevant_docs, candidate_docs, word2vec, device):
    `vocab` replaces word2vec (the model carries the table, `attach_embeddings`), the device is the model's.  Candidates are
    encoded in batches (the reference: 128 at a time, :21-36) with the margin model's text preprocessing.  Returns
    (1/rank of the first candidate that is in relevant_docs or 0.0, [(doc, score)] of the three best), as :45-54."""
    dev = next(model.parameters()).device
    was_training = model.training
    model.eval()
    try:
        with torch.no_grad():
            qv = model.encode_query(vocab.encode_margin_batch([query], max_length).to(dev))
            outs = []
            for i in range(0, len(candidate_docs), batch_size):
                outs.append(model.encode_doc(vocab.encode_margin_batch(candidate_docs[i:i + batch_size], max_length).to(dev)))
    finally:
        model.train(was_training)
    if not outs:
        return 0.0, []
    relevant = torch.tensor([[d in relevant_docs for d in candidate_docs]], dtype=torch.bool)
    rr, idx, val = first_relevant_rank(qv, torch.cat(outs, 0), relevant)
    top = [(candidate_docs[i], float(v)) for i, v in zip(idx[0].tolist(), val[0].tolist())]
    return float(rr[0].item()), top


# ------------------------------------------------------------------------------------------- N3: serving path
class SearchService:
    """The `/search` handler of server/python-api/app.py:85-120 without the web framework: query text -> ids ->
    encode_query -> cosine top-k over the cached document embeddings (K-H with Q = 1) -> the response schema of
    app.py:72-83 (`{"query", "results": [{"text", "score", "is_ground_truth", "rank"}]}`; text cut to 200 chars + "...").
    `query_to_docs` maps a query string to its ground-truth documents (app.py:29-37)."""

    def __init__(self, model, vocab, docs, max_length=30, top_k=3, cache_path=None, query_to_docs=None,
                 margin_preprocessing=False, batch_size=4096):
        self.model, self.vocab, self.docs = model.eval(), vocab, list(docs)
        self.max_length, self.top_k = max_length, top_k
        self.query_to_docs = query_to_docs or {}
        self._enc = (lambda texts: vocab.encode_margin_batch(texts, max_length)) if margin_preprocessing else \
            (lambda texts: vocab.encode_batch(texts, max_length))
        self.device = next(model.parameters()).device
        import os
        if cache_path is not None and os.path.exists(cache_path):
            self.doc_embeddings = load_doc_embeddings(cache_path, self.device)
        else:
            self.doc_embeddings = load_or_compute_doc_embeddings(model, self._enc(self.docs), cache_path, batch_size)
        if self.doc_embeddings.shape[0] != len(self.docs):
            raise ValueError("cached document embeddings do not match the document list")
        self.index = CorpusIndex(self.doc_embeddings, cosine=True)

    @torch.no_grad()
    def search(self, query, top_k=None):
        k = min(top_k or self.top_k, len(self.docs))
        ids = self._enc([query]).to(self.device)
        qv = self.model.encode_query(ids)
        scores, idx = self.index.search(qv, k=k)
        truth = self.query_to_docs.get(query, [])
        results = []
        for r, (i, s) in enumerate(zip(idx[0].tolist(), scores[0].tolist())):
            text = self.docs[i]
            results.append({"text": text[:200] + "..." if len(text) > 200 else text, "score": float(s),
                            "is_ground_truth": text in truth, "rank": r + 1})
        return {"query": query, "results": results}
