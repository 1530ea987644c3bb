"""Where does the HOST spend its time while queueing one train step?  Wraps every libtt_b200 entry point (ctypes) and the
torch optimizer step with perf_counter and prints host ms per call site, summed over the measured steps.  A call that
blocks until the GPU catches up shows up with a host time close to the device time of the work queued before it.

  python tools/host_time_probe.py [--batch 4096] [--steps 5]
"""
import argparse
import collections
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import two_towers_b200 as tt  # noqa: E402
from two_towers_b200 import _lib as L  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--serial", action="store_true")
    ap.add_argument("--torch-opt", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = L.lib()
    acc = collections.defaultdict(lambda: [0.0, 0])

    class Timed:
        def __init__(self, name, fn):
            self.name, self.fn = name, fn

        def __call__(self, *args, **kw):
            t0 = time.perf_counter()
            r = self.fn(*args, **kw)
            d = acc[self.name]
            d[0] += time.perf_counter() - t0
            d[1] += 1
            return r

    class Proxy:
        def __getattr__(self, name):
            fn = getattr(lib, name)
            return Timed(name, fn) if name.startswith("tt_") else fn

    proxy = Proxy()
    L.lib = lambda: proxy
    L.workspace = Timed("(allocator) L.workspace", L.workspace)
    from two_towers_b200 import optim as O
    O.build_tables = Timed("(optimizer) table rebuild", O.build_tables)
    real_empty, real_empty_like = torch.empty, torch.empty_like
    torch.empty = Timed("(allocator) torch.empty", real_empty)
    torch.empty_like = Timed("(allocator) torch.empty_like", real_empty_like)
    torch.manual_seed(0)
    V = 200_000
    model = tt.EnhancedTwoTowerModel(bench.E, bench.HID).to(dev).train()
    table = torch.randn(V, bench.E, device=dev) * 0.1
    table[0] = 0
    model.attach_embeddings(table, 0)
    model.overlap_towers = not a.serial
    crit = tt.MarginRankingLoss(0.2, 0.1)
    opt = torch.optim.Adam(model.parameters(), fused=True) if a.torch_opt else tt.FusedAdam(model.parameters())
    gen = torch.Generator().manual_seed(1)
    q = bench.make_ids(a.batch, bench.TQ, V, gen).to(dev)
    d = bench.make_ids(a.batch, bench.TD, V, gen).to(dev)
    phases = collections.defaultdict(float)

    def step():
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        qv, dv = model(q, d)
        t1 = time.perf_counter()
        loss = crit(qv, dv)
        hard = tt.get_hard_negatives_batch(qv.detach(), dv.detach(), k=8)
        t2 = time.perf_counter()
        loss.backward()
        t3 = time.perf_counter()
        opt.step()
        t4 = time.perf_counter()
        for k, v in (("forward", t1 - t0), ("loss+mining", t2 - t1), ("backward", t3 - t2), ("optimizer", t4 - t3)):
            phases[k] += v

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    acc.clear()
    phases.clear()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step()
    host = time.perf_counter() - t0
    e1.record()
    torch.cuda.synchronize()
    print(f"batch {a.batch}: host {host / a.steps * 1e3:.2f} ms/step to queue, device {e0.elapsed_time(e1) / a.steps:.2f} ms/step")
    print("phases (host ms/step):", {k: round(v / a.steps * 1e3, 2) for k, v in phases.items()})
    for name, (sec, n) in sorted(acc.items(), key=lambda kv: -kv[1][0])[:14]:
        print(f"  {name:32s} {sec / a.steps * 1e3:8.3f} ms/step  {n / a.steps:6.1f} calls/step  {sec / n * 1e6:9.1f} us/call")


if __name__ == "__main__":
    main()
