"""Data-parallel self-check: the sharded step against the same step run by ONE process on the concatenated batch.

The reference is single-process (enhanced_two_tower.py:90-101 in-batch loss, :123-133 mining); under data parallelism
each rank encodes its slice, all-gathers the tower outputs, scores its rows against the global docs, reduce-scatters
dL/d(docs) and all-reduces the weight gradients (distributed.py).  `dp_parity_check` runs both forms with this package's
kernels on the calling rank's GPU and compares loss, every parameter gradient and the mined hard-negative indices;
with `graphed=True` it also runs two optimizer steps through the graphed step (world 1: `GraphedTrainStep`, the whole step
in one CUDA graph; world > 1: `GraphedTowers`, tower forward / backward graphed and the NCCL calls eager) against two eager
single-process steps.  Used by tests/test_gpu_dp.py (2 GPUs) and by `bench.py` at N > 1, which puts
the verdict into its JSON line (`parity_check`).
"""
import torch
import torch.distributed as dist

from . import distributed as dops
from .graph import GraphedTowers, GraphedTrainStep
from .losses import MarginRankingLoss, get_hard_negatives_batch
from .models import EnhancedTwoTowerModel
from .optim import FusedAdam


def _mk(E, hidden, dev, table, overlap):
    torch.manual_seed(0)
    m = EnhancedTwoTowerModel(E, hidden)
    for mod in m.modules():
        if hasattr(mod, "dropout") and isinstance(getattr(mod, "dropout"), float):
            mod.dropout = 0.0
    m.to(dev).train().attach_embeddings(table, pad_id=0)
    m.overlap_towers = overlap
    return m


def _rel(a, b):
    d = b.double().norm().item()
    return (a.double() - b.double()).norm().item() / (d if d > 0 else 1.0)


def dp_parity_check(group, device, B_loc=96, E=20, hidden=32, Tq=6, Td=11, k=4, overlap=True, graphed=False,
                    tol=1e-3, vocab=500):
    """Returns {'ok': bool, 'loss_rel': .., 'worst_grad': (name, rel), 'indices_equal': bool, ...}; identical on every rank
    up to the rank-local slices.  Collective: every rank of `group` must call it with the same arguments."""
    W, r = dops.world_size(group), dops.rank(group)
    gen = torch.Generator().manual_seed(4242)
    table = (torch.randn(vocab, E, generator=gen) * 0.3).to(device)
    q_all = torch.randint(0, vocab, (W * B_loc, Tq), generator=gen).to(device)
    d_all = torch.randint(0, vocab, (W * B_loc, Td), generator=gen).to(device)
    sl = slice(r * B_loc, (r + 1) * B_loc)
    out = {"world": W, "overlap_towers": overlap, "graphed": graphed}

    # ---- one process, concatenated batch
    sp = _mk(E, hidden, device, table, overlap)
    qv, dv = sp(q_all, d_all)
    loss_sp = MarginRankingLoss()(qv, dv)
    loss_sp.backward()
    hard_sp = get_hard_negatives_batch(qv.detach(), dv.detach(), k=k)

    # ---- sharded
    dp = _mk(E, hidden, device, table, overlap)
    red = dops.GradientAllReducer(None, group, named_params=dp.named_parameters(), bucket_bytes=64 << 10)
    crit = MarginRankingLoss(group=group)
    qv2, dv2 = dp(q_all[sl], d_all[sl])
    loss_dp = crit(qv2, dv2)
    loss_dp.backward()
    red.finish()
    docs_g = dops.all_gather_rows(dv2.detach(), group)
    hard_dp = get_hard_negatives_batch(qv2.detach(), docs_g, k=k, pos_offset=r * B_loc)
    torch.cuda.synchronize(device)
    out["loss_rel"] = abs(loss_dp.item() - loss_sp.item()) / abs(loss_sp.item())
    worst = ("", 0.0)
    scale = max(p.grad.norm().item() for p in sp.parameters())
    ok = out["loss_rel"] < 1e-5
    for (n, a), (_, b) in zip(dp.named_parameters(), sp.named_parameters()):
        err = (a.grad.double() - b.grad.double()).norm().item()
        # per-parameter bound + an absolute floor for analytically (near-)zero gradients (the last bias under CE sums to ~1e-7
        # of the largest gradient; its rounding noise depends on how many ranks' partial sums are added) - the same rule as
        # tests/test_gpu_bench_shapes.py
        if err > tol * b.grad.norm().item() + 1e-5 * scale:
            ok = False
            out.setdefault("failed", []).append((n, err, b.grad.norm().item()))
        if b.grad.norm().item() > 1e-3 * scale and _rel(a.grad, b.grad) > worst[1]:
            worst = (n, _rel(a.grad, b.grad))
    out["worst_grad"] = worst
    out["indices_equal"] = bool(torch.equal(hard_dp, hard_sp[sl]))
    out["emb_equal"] = bool(torch.equal(qv2.detach(), qv.detach()[sl]) and torch.equal(dv2.detach(), dv.detach()[sl]))
    ok = ok and out["indices_equal"]
    red.remove()

    if graphed:
        sp, dp = _mk(E, hidden, device, table, overlap), _mk(E, hidden, device, table, overlap)
        # eps well above the rounding noise of analytically-zero gradients (the last bias under CE): with the default 1e-8
        # Adam turns +-1e-7 noise into +-lr updates and the two summation orders could not be compared
        # (1e-2: the summation-order noise of an 8-rank all-reduce on ~1e-6 gradients stays far below it; the check is about
        # the plumbing - which gradients reach which update - not about Adam's conditioning)
        o_sp = FusedAdam(sp.parameters(), lr=1e-3, eps=1e-2, max_grad_norm=1.0)
        o_dp = FusedAdam(dp.parameters(), lr=1e-3, eps=1e-2, max_grad_norm=1.0, capturable=True)
        red = dops.GradientAllReducer(None, group, named_params=dp.named_parameters(), bucket_bytes=64 << 10)
        crit_dp = MarginRankingLoss(group=group)
        if W == 1:                                   # whole step in one graph
            step = GraphedTrainStep(dp, crit_dp, o_dp, reducer=red, mine_k=k, pos_offset=r * B_loc)
        else:                                        # towers graphed (forward + backward), collectives / reducer / optimizer eager
            towers = GraphedTowers(dp)

            def step(qs, ds):
                o_dp.zero_grad(set_to_none=True)
                a_, b_ = towers(qs, ds)
                l_ = crit_dp(a_, b_)
                l_.backward()
                red.finish()
                o_dp.step()
                h_ = get_hard_negatives_batch(a_.detach(), dops.all_gather_rows(b_.detach(), group), k=k, pos_offset=r * B_loc)
                return l_.detach(), h_
        for it in range(2):
            perm = torch.randperm(W * B_loc, generator=gen).to(device)
            qa, da = q_all[perm], d_all[perm]
            loss_g, hard_g = step(qa[sl].contiguous(), da[sl].contiguous())
            o_sp.zero_grad(set_to_none=True)
            a, b = sp(qa, da)
            l2 = MarginRankingLoss()(a, b)
            l2.backward()
            o_sp.step()
            hs = get_hard_negatives_batch(a.detach(), b.detach(), k=k)
            out.setdefault("graphed_steps", []).append((loss_g.item(), l2.item(), bool(torch.equal(hard_g, hs[sl]))))
            ok = ok and abs(loss_g.item() - l2.item()) < 1e-5 * abs(l2.item()) and bool(torch.equal(hard_g, hs[sl]))
        wr = max(_rel(a_, b_) for a_, b_ in zip(dp.parameters(), sp.parameters()))
        out["graphed_weight_rel"] = wr
        ok = ok and wr < 1e-3            # the north-star bound; what is left is Adam's amplification of gradient rounding on the zero-initialised LayerNorm biases
        red.remove()

    flag = torch.tensor([1.0 if ok else 0.0], device=device)
    if W > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    out["ok"] = bool(flag.item() == 1.0)
    return out
