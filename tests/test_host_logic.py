"""CPU: host-side mirror of the reference interface - state_dict keys/shapes, init parity from the same seed,
constructor / attribute surface - and the data-parallel plumbing under gloo with world_size 2."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import two_towers_b200 as tt
from oracle import ref_port as rp
from two_towers_b200 import distributed as dops

CASES = [("simple", tt.SimpleTwoTower, 300, 16), ("margin", tt.TwoTowerModel, 300, 16),
         ("enhanced", tt.EnhancedTwoTowerModel, 300, 8)]


@pytest.mark.parametrize("kind,cls,E,H", CASES)
def test_state_dict_matches_reference_layout_and_init(kind, cls, E, H):
    torch.manual_seed(7)
    ours = cls(E, H)
    torch.manual_seed(7)
    port = rp.build_port(kind, E, H)
    a, b = ours.state_dict(), port.state_dict()
    assert list(a.keys()) == list(b.keys())
    for k in a:
        assert a[k].shape == b[k].shape, k
        assert torch.equal(a[k], b[k]), f"{k}: same seed must give the reference's init"
    ours.load_state_dict(b, strict=True)
    port.load_state_dict(a, strict=True)


def test_param_counts_match_survey():
    n = lambda m: sum(p.numel() for p in m.parameters())
    assert n(tt.SimpleTwoTower(300, 128)) == 1336192
    assert n(tt.TwoTowerModel(300, 256)) == 4211456
    assert n(tt.EnhancedTwoTowerModel(300, 128)) == 4409088
    assert len(tt.SimpleTwoTower(300, 8).state_dict()) == 44
    assert len(tt.TwoTowerModel(300, 8).state_dict()) == 36


def test_module_surface():
    m = tt.EnhancedTwoTowerModel(embedding_dim=300, hidden_dim=8)
    assert tt.EnhancedTwoTower is tt.EnhancedTwoTowerModel
    for a in ("query_encoder", "doc_encoder", "query_proj", "doc_proj", "encode_query", "encode_doc"):
        assert hasattr(m, a)
    mm = tt.TwoTowerModel(300, 8)
    for a in ("projection", "encode", "compute_similarity", "encode_query", "encode_doc"):
        assert hasattr(mm, a)
    from two_towers_b200 import enhanced_two_tower, margin_two_tower, simple_two_tower, two_tower_model
    assert simple_two_tower.InfoNCELoss().temperature == 0.1
    assert margin_two_tower.InfoNCELoss().temperature == 0.07
    assert enhanced_two_tower.InfoNCELoss().temperature == 0.07
    l = enhanced_two_tower.MarginRankingLoss()
    assert (l.margin, l.temperature) == (0.2, 0.1)
    tok = two_tower_model.TwoTowerModel(vocab_size=50, embedding_dim=12, hidden_dim=6,
                                        pretrained_embeddings=torch.randn(50, 12), freeze_embeddings=True)
    assert sum(p.numel() for p in tok.parameters()) == 2 * (18 * 12 + 18 * 6 + 36)
    # optimizer / clipping plumbing used by the reference loops (simple_two_tower.py:193,239)
    opt = torch.optim.AdamW(m.parameters(), lr=1e-4, weight_decay=0.01)
    assert len(opt.param_groups[0]["params"]) == 44 - 0


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _dp_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        Bl, D = 3, 5
        qg = torch.randn(world * Bl, D); dg = torch.randn(world * Bl, D)
        w = torch.nn.Parameter(torch.randn(D, D))            # a "tower" shared by all ranks
        # single-process oracle on the concatenated batch
        wf = w.detach().clone().requires_grad_(True)
        full = rp.margin_ce_inbatch(qg @ wf, dg @ wf, 0.2, 0.1)
        full.backward()
        # sharded: local rows vs all-gathered docs, local partial = sum(rows)/B_global
        q, d = qg[rank * Bl:(rank + 1) * Bl] @ w, dg[rank * Bl:(rank + 1) * Bl] @ w
        d_all = dops.all_gather_rows(d)
        s = (q @ d_all.t()) / 0.1
        mask = torch.ones_like(s); idx = torch.arange(Bl); mask[idx, idx + rank * Bl] = 0
        s = s - 0.2 * mask
        part = torch.nn.functional.cross_entropy(s, idx + rank * Bl, reduction="sum") / (world * Bl)
        loss = dops.sum_over_ranks(part, dist.group.WORLD)
        red = dops.GradientAllReducer([w], bucket_bytes=1)
        loss.backward()
        red.finish()
        ok = abs(loss.item() - full.item()) < 1e-5 * abs(full.item()) and \
            (w.grad - wf.grad).norm() < 1e-5 * wf.grad.norm()
        # buckets: never straddle a top-level submodule, reverse registration order, p.grad ends up as a view of the bucket
        net = torch.nn.ModuleDict({"a": torch.nn.Linear(D, D), "b": torch.nn.Linear(D, 2)})
        for p_ in net.parameters():
            dist.broadcast(p_.data, 0)
        ref_net = torch.nn.ModuleDict({"a": torch.nn.Linear(D, D), "b": torch.nn.Linear(D, 2)})
        ref_net.load_state_dict(net.state_dict())
        red2 = dops.GradientAllReducer(None, named_params=net.named_parameters(), bucket_bytes=1 << 20)
        ok = ok and [len(b) for b in red2.buckets] == [2, 2] and red2.buckets[0][0] is net["b"].bias
        xs = torch.randn(world, 2, 4, D)                       # [rank][micro-batch]
        with red2.no_sync():
            net["b"](net["a"](xs[rank, 0])).square().sum().backward()
        net["b"](net["a"](xs[rank, 1])).square().sum().backward()
        try:                                                   # a third backward without finish(): must raise, not drop data
            net["b"](net["a"](xs[rank, 1])).square().sum().backward()
            ok = False
        except RuntimeError as e:
            ok = ok and "no_sync" in str(e)
        red2.finish()
        ref_net["b"](ref_net["a"](xs.reshape(-1, D))).square().sum().backward()
        for (n_, p_), q_ in zip(net.named_parameters(), ref_net.parameters()):
            ok = ok and (p_.grad - q_.grad).norm() < 1e-5 * q_.grad.norm() + 1e-7
            ok = ok and p_.grad.untyped_storage().data_ptr() == red2._flat[red2._bucket_of[id(p_)]].untyped_storage().data_ptr()
        red2.remove()
        # overlap=False (TT_DP_OVERLAP=0): nothing is launched from the hooks, finish() reduces every bucket - same gradients;
        # the bucket size comes from TT_DP_BUCKET_MB when the argument is omitted
        os.environ["TT_DP_OVERLAP"], os.environ["TT_DP_BUCKET_MB"] = "0", "0.000001"
        try:
            net.zero_grad(set_to_none=True)
            red3 = dops.GradientAllReducer(None, named_params=net.named_parameters())
            ok = ok and red3.overlap is False and [len(b) for b in red3.buckets] == [1, 1, 1, 1]
            net["b"](net["a"](xs[rank].reshape(-1, D))).square().sum().backward()
            ok = ok and not any(red3._launched)
            red3.finish()
            for p_, q_ in zip(net.parameters(), ref_net.parameters()):
                ok = ok and (p_.grad - q_.grad).norm() < 1e-5 * q_.grad.norm() + 1e-7
            red3.remove()
        finally:
            del os.environ["TT_DP_OVERLAP"], os.environ["TT_DP_BUCKET_MB"]
        lo, hi = dops.shard_bounds(11)
        ok = ok and (lo, hi) == ((0, 6) if rank == 0 else (6, 11))
        ret[rank] = (bool(ok), loss.item(), full.item(), (w.grad - wf.grad).abs().max().item(), lo, hi)
    finally:
        dist.destroy_process_group()


def test_data_parallel_plumbing_gloo_world2():
    world = 2
    ret = mp.get_context("spawn").Manager().dict()
    mp.spawn(_dp_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert all(ret.get(r) and ret[r][0] for r in range(world)), dict(ret)


def test_single_process_collectives_are_identity():
    x = torch.randn(3, 4, requires_grad=True)
    assert dops.all_gather_rows(x) is x
    assert dops.sum_over_ranks(x.sum()) is not None
    assert dops.shard_bounds(10) == (0, 10)


def test_engine_selection_api():
    """set_engine / use_engine (the Python face of tt_set_engine): names and numbers, restore on exit, bad values raise."""
    import two_towers_b200 as tt
    assert tt.engine() == "tcgen05-f16x3-fwd/f16-bwd"
    prev = tt.set_engine(0)
    try:
        assert prev == 2 and tt.engine() == "simt-fp32"
        with tt.use_engine("tcgen05-bf16x3"):
            assert tt.engine() == "tcgen05-bf16x3"
        assert tt.engine() == "simt-fp32"
        with pytest.raises(RuntimeError):
            tt.set_engine(7)
    finally:
        tt.set_engine(prev)
    assert tt.engine() == "tcgen05-f16x3-fwd/f16-bwd"


def test_bench_synthetic_pairs_share_the_query_tokens():
    """bench.py's synthetic (query, relevant doc) batches: each doc contains its query's tokens as a contiguous window, pads
    stay at the end, shapes / dtype as the model's id inputs."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(os.path.dirname(os.path.dirname(__file__)), "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    g = torch.Generator().manual_seed(3)
    q, d = bench.make_pair_ids(64, 32, 128, 50000, g)
    assert q.shape == (64, 32) and d.shape == (64, 128) and q.dtype == torch.int64
    for i in range(64):
        ql, dl = int((q[i] != 0).sum()), int((d[i] != 0).sum())
        assert (q[i, ql:] == 0).all() and (d[i, dl:] == 0).all() and ql >= 8 and dl >= 32
        n = min(ql, dl)
        row, want = d[i, :dl].tolist(), q[i, :n].tolist()
        assert any(row[s:s + n] == want for s in range(dl - n + 1)), i


def test_graph_wrappers_reject_misuse_before_touching_the_gpu():
    """GraphedTrainStep needs a capturable optimizer; GraphedTowers is a training-step wrapper (eval mode raises)."""
    import two_towers_b200 as tt
    m = tt.EnhancedTwoTowerModel(20, 32)
    with pytest.raises(ValueError, match="capturable"):
        tt.GraphedTrainStep(m, tt.MarginRankingLoss(), tt.FusedAdam(m.parameters()))
    step = tt.GraphedTrainStep(m, tt.MarginRankingLoss(), tt.FusedAdam(m.parameters(), capturable=True), mine_k=4)
    assert step.replays == 0 and step.launches_per_replay(torch.zeros(2, 3), torch.zeros(2, 3)) is None
    towers = tt.GraphedTowers(m.eval())
    with pytest.raises(RuntimeError, match="training step"):
        towers(torch.zeros(2, 3, dtype=torch.long), torch.zeros(2, 5, dtype=torch.long))
