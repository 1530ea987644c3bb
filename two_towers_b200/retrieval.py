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


@torch.no_grad()
def search_topk(query_vecs, doc_encodings, k=10, cosine=True, index_base=0):
    """query_vecs [Q,D] (or [D]) vs doc_encodings [N,D] -> (scores [Q,k], indices int64 [Q,k]), best first;
    ties go to the lowest index.  Same per-query result as cosine_similarity + topk in the reference."""
    if query_vecs.dim() == 1:
        query_vecs = query_vecs.unsqueeze(0)
    return Fn.rows_topk(query_vecs, doc_encodings, k, cosine=cosine, index_base=index_base)


@torch.no_grad()
def sharded_search_topk(query_vecs, local_docs, index_base, k=10, group=None, cosine=True):
    """Every rank holds a contiguous corpus shard starting at global id `index_base`; queries are replicated.
    Returns the identical global top-k on every rank."""
    val, idx = search_topk(query_vecs, local_docs, k, cosine, index_base)
    W = dist_ops.world_size(group)
    if W == 1:
        return val, idx
    vals = torch.empty(W, *val.shape, device=val.device, dtype=val.dtype)
    idxs = torch.empty(W, *idx.shape, device=idx.device, dtype=idx.dtype)
    dist.all_gather_into_tensor(vals, val.contiguous(), group=group)
    dist.all_gather_into_tensor(idxs, idx.contiguous(), group=group)
    return Fn.topk_merge(vals, idxs)


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
