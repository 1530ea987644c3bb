"""Retrieval scan alone (the `eval` leg of bench.py): Q queries x N docs, top-10; prints ms and queries/s.
Under `ncu --metrics gpu__time_duration.sum` it gives the launch list of one scan."""
import sys, torch
sys.path.insert(0, ".")
import two_towers_b200 as tt
Q = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1_100_000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(1)
q = torch.nn.functional.normalize(torch.randn(Q, 128, device=dev, generator=g), dim=1)
docs = torch.nn.functional.normalize(torch.randn(N, 128, device=dev, generator=g), dim=1)
from two_towers_b200 import _lib as L


def timed(fn):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


index = tt.CorpusIndex(docs, cosine=True)
L.lib().tt_debug_sel_passes(3)
ms, ref = timed(lambda: tt.search_topk(q, docs, 10))
print(f"Q={Q} N={N} per-call corpus conversion: {ms:.2f} ms per scan, {Q / ms * 1e3:.0f} queries/s")
for passes, resident, pf in ((13, 0, 0), (13, 1, 0), (12, 1, 0), (3, 1, 0), (2, 1, 0), (2, 1, 4), (3, 1, 0), (2, 1, 0)):
    L.lib().tt_debug_sel_resident(resident)
    L.lib().tt_debug_sel_prefetch(pf)
    L.lib().tt_debug_sel_passes(passes)
    ms, (v, i) = timed(lambda: tt.search_topk(q, index, 10))
    same = torch.equal(i, ref[1]) and torch.equal(v, ref[0])
    print(f"Q={Q} N={N} passes={passes} resident_A={resident} prefetch={pf} prebuilt CorpusIndex: {ms:.2f} ms per scan, "
          f"{Q / ms * 1e3:.0f} queries/s, {2 * Q * N * 128 / ms / 1e9:.1f} TFLOP/s algorithmic, identical={same}")
