"""Device time of the hard-negative mining call (rows_topk through the tensor route) at the 8-GPU shape, 4096 x 32768 x 128,
under different host pacing and score distributions."""
import sys, time, torch
sys.path.insert(0, ".")
import two_towers_b200 as tt
dev = torch.device("cuda")
Q, N, D, K = 4096, 32768, 128, 8
g = torch.Generator(device=dev).manual_seed(0)

def timed(q, docs, n=20, sync_each=False, sleep=0.0):
    for _ in range(3):
        tt.get_hard_negatives_batch(q, docs, k=K, pos_offset=4096)
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); tt.get_hard_negatives_batch(q, docs, k=K, pos_offset=4096); e1.record()
        if sync_each:
            torch.cuda.synchronize()
            if sleep: time.sleep(sleep)
        ts.append((e0, e1))
    torch.cuda.synchronize()
    v = sorted(a.elapsed_time(b) for a, b in ts)
    return v[len(v) // 2], v[-1]

cases = {
    "random unit vectors": (torch.randn(Q, D, device=dev, generator=g), torch.randn(N, D, device=dev, generator=g)),
    "common direction + 1% noise (collapsing model)": (torch.ones(Q, D, device=dev) + 0.01 * torch.randn(Q, D, device=dev, generator=g),
                                                       torch.ones(N, D, device=dev) + 0.01 * torch.randn(N, D, device=dev, generator=g)),
    "common direction + 1e-4 noise": (torch.ones(Q, D, device=dev) + 1e-4 * torch.randn(Q, D, device=dev, generator=g),
                                      torch.ones(N, D, device=dev) + 1e-4 * torch.randn(N, D, device=dev, generator=g)),
    "identical rows (collapsed)": (torch.ones(Q, D, device=dev), torch.ones(N, D, device=dev)),
}
for name, (q, docs) in cases.items():
    a = timed(q, docs)
    b = timed(q, docs, sync_each=True)
    c = timed(q, docs, sync_each=True, sleep=0.02)
    print(f"{name:50s} back-to-back {a[0]:.2f} (max {a[1]:.2f}) | sync each {b[0]:.2f} (max {b[1]:.2f}) | sync + 20 ms idle {c[0]:.2f} (max {c[1]:.2f}) ms")
