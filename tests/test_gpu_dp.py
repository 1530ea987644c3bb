"""GPU x2: the data-parallel step under NCCL against one process on the concatenated batch (loss, every gradient, mined
indices), with the towers on two streams and on one, eager and graphed (world > 1: `GraphedTowers` - tower forward / backward
replayed from CUDA graphs, collectives eager; world 1: the whole step in one graph).  Needs two visible GPUs
(`gpurun --gpus 2`); the single-GPU driver tier skips it - `bench.py` repeats the same check at every N > 1 and reports it
as `parity_check` in its JSON line."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from two_towers_b200.selfcheck import dp_parity_check
        res = []
        for overlap, graphed in ((True, False), (False, False), (True, True)):
            res.append(dp_parity_check(dist.group.WORLD, dev, overlap=overlap, graphed=graphed))
        ret[rank] = res
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_data_parallel_step_matches_single_process_nccl_world2():
    world = 2
    ret = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    print("[dp parity]", dict(ret))
    for r in range(world):
        assert ret.get(r), dict(ret)
        for res in ret[r]:
            assert res["ok"] and res["indices_equal"], res
    print("[dp parity]", ret[0])


def test_dp_selfcheck_single_process_is_trivially_consistent():
    """world = 1: the same code path without collectives (runs on the 1-GPU tier)."""
    from two_towers_b200.selfcheck import dp_parity_check
    res = dp_parity_check(None, torch.device("cuda", 0), graphed=True)
    assert res["ok"] and res["emb_equal"], res
