"""The retrieval leg of bench.py (planted relevant documents) under the debug switches of the candidate scan."""
import sys, torch
sys.path.insert(0, ".")
import two_towers_b200 as tt
from two_towers_b200 import _lib as L
Q, Nl, D = 8192, 1_100_000, 128
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(4321)
q = torch.nn.functional.normalize(torch.randn(Q, D, device=dev, generator=g), dim=1)
gl = torch.Generator(device=dev).manual_seed(99)
docs = torch.nn.functional.normalize(torch.randn(Nl, D, device=dev, generator=gl), dim=1)
rel = torch.randint(0, Nl, (Q,), device=dev, generator=g)
noisy = torch.nn.functional.normalize(q + 0.5 * torch.randn(Q, D, device=dev, generator=g) / D ** 0.5, dim=1)
docs[rel % Nl] = noisy
index = tt.CorpusIndex(docs, cosine=True)
from two_towers_b200 import functional as Fn
ref = None
for passes, res in ((3, 1), (2, 1), (2, 5), (2, 1), (3, 1)):
    L.lib().tt_debug_sel_passes(passes)
    L.lib().tt_debug_sel_resident(res)
    v, i, fl = Fn.rows_topk_indexed(q, index.docs, index.buffer, 10, cosine=True, return_fallback_rows=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        v, i = tt.search_topk(q, index, 10)
    e1.record()
    torch.cuda.synchronize()
    if ref is None:
        ref = (v.clone(), i.clone())
    print(f"passes={passes} resident={res}: {e0.elapsed_time(e1) / 3:.2f} ms per scan, fallback rows {int(fl.sum())} of {Q}, "
          f"identical={torch.equal(i, ref[1]) and torch.equal(v, ref[0])}", flush=True)
