"""GPU debug aid: tcgen05 bf16x3 engine vs float64 on a ladder of shapes, with error-pattern diagnostics."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from two_towers_b200 import _lib as L

lib = L.lib()
dev = "cuda"

def run(M, N, K, trans=0, splitk=1, bias=True, beta=0.0, reps=0):
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K)
    A = torch.randn(M, K, generator=g); B = torch.randn(N, K, generator=g)
    bi = torch.randn(N, generator=g) if bias else None
    C0 = torch.randn(M, N, generator=g)
    ref = A.double() @ B.double().t() * 0.5 + (bi.double() if bias else 0) + beta * C0.double()
    Ad = (A.t().contiguous() if trans else A).to(dev); Bd = (B.t().contiguous() if trans else B).to(dev)
    C = C0.to(dev).clone()
    ws = L.workspace(lib.tt_gemm_x3_workspace(M, N, K), dev)
    bd = bi.to(dev) if bias else None
    call = lambda: L.check(lib.tt_gemm_x3(L.ptr(Ad), L.ptr(Bd), L.ptr(C), L.ptr(bd), M, N, K, 0.5, beta, trans, splitk,
                                          ws.data_ptr(), ws.numel(), L.stream()), "gemm_x3")
    call(); torch.cuda.synchronize()
    got = C.cpu().double()
    err = (got - ref).norm() / ref.norm()
    msg = f"M={M} N={N} K={K} trans={trans} splitk={splitk} beta={beta}: rel_err {err:.3e}"
    if not (err < 1e-4):
        bad = ((got - ref).abs() > 1e-3 * ref.abs().max())
        rows = bad.any(1).nonzero().flatten(); cols = bad.any(0).nonzero().flatten()
        msg += f"  BAD frac {bad.float().mean():.3f} rows[{rows[:6].tolist()}..{rows[-3:].tolist()}] n={len(rows)} cols[{cols[:6].tolist()}..{cols[-3:].tolist()}] n={len(cols)}"
        msg += f"\n   got[0,:4]={got[0,:4].tolist()} ref[0,:4]={ref[0,:4].tolist()} nan={torch.isnan(got).any().item()}"
    if reps:
        torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        msg += f"  | {ms:.3f} ms/call incl. split, {2*M*N*K/ms/1e9:.1f} TFLOP/s algorithmic"
    print(msg, flush=True)

print("engine", lib.tt_engine().decode())
cases = [(128, 128, 64), (128, 128, 256), (128, 256, 64), (256, 256, 512), (300, 200, 100), (5, 7, 3), (1000, 768, 300),
         (4096, 768, 256)]
for c in cases: run(*c)
run(512, 300, 100, beta=1.0)
for c in [(128, 128, 64), (128, 128, 256), (256, 256, 512), (300, 200, 100), (768, 304, 5000)]: run(*c, trans=1)
run(768, 300, 20000, trans=1, splitk=0)
run(256, 256, 4096, splitk=4)
run(65536, 768, 512, reps=5)
run(768, 512, 65536, trans=1, splitk=0, reps=5)
