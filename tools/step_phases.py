"""Phase timing of the recurrence step kernel (gru_step2_kernel) from in-kernel %globaltimer marks.
Runs one GRU layer forward at the bench shape and prints, for the LAST step launch, per-CTA phase durations."""
import sys, torch
sys.path.insert(0, ".")
import two_towers_b200 as tt
from two_towers_b200 import _lib as L, functional as Fn

dev = torch.device("cuda")
B, T, In, H = 4096, int(sys.argv[1]) if len(sys.argv) > 1 else 16, 512, 256
torch.manual_seed(0)
gru = tt.GRU(In, H, num_layers=1, bidirectional=True).to(dev)
x = torch.randn(B, T, In, device=dev) * 0.1
nct = (H // 64) * (B // 128) * 2
buf = torch.zeros(nct * 8, dtype=torch.int64, device=dev)
lib = L.lib()
with torch.no_grad():
    gru(x, need_output=False)
    torch.cuda.synchronize()
    lib.tt_debug_step_buffer(buf.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); gru(x, need_output=False); e1.record()
    torch.cuda.synchronize()
    lib.tt_debug_step_buffer(None)
t = buf.view(nct, 8).cpu().double()
t0 = t[:, 0].min()
names = ["start", "mma thread ready", "first tile landed", "all MMAs issued", "accumulator ready (epilogue)", "epilogue math done",
         "TMA store drained", "CTA end"]
print(f"layer forward: {e0.elapsed_time(e1) * 1e3 / T:.1f} us per step (T={T}, incl. projection GEMM)")
for i, n in enumerate(names):
    d = t[:, i] - t0
    print(f"{n:32s} mean {d.mean() / 1e3:7.2f} us  min {d.min() / 1e3:7.2f}  max {d.max() / 1e3:7.2f}   (since first CTA start)")
d = t[:, 7] - t[:, 0]
print(f"CTA lifetime mean {d.mean() / 1e3:.2f} us, min {d.min() / 1e3:.2f}, max {d.max() / 1e3:.2f}")
for i in range(1, 8):
    d = t[:, i] - t[:, i - 1]
    print(f"  {names[i - 1]:30s} -> {names[i]:30s} mean {d.mean() / 1e3:6.2f} us  max {d.max() / 1e3:6.2f}")
first_wave = t[:, 0] - t0 < 1000
print("CTAs starting within 1 us of the first:", int(first_wave.sum()), "of", nct)

# ---- persistent kernel (tt_debug_seq_mode 2/4): per-item marks of CTA 0
if t[:, 0].min() > 0 or True:
    it = buf.view(-1, 8).cpu().double()
    it = it[(it[:, 0] > 0)]
    if len(it) > 8:
        base = it[0, 0]
        print("persistent kernel, CTA 0, per item: [flag wait begin, flag ready, last load issued, epi waits, acc ready, epi math done, published] (us since first mark)")
        for i in range(4, min(len(it), 16)):
            print(f"item {i:3d} " + " ".join(f"{(v - base) / 1e3:8.2f}" for v in it[i, :7]))
        per = (it[-1, 6] - it[8, 6]) / (len(it) - 9)
        print(f"steady state: {per / 1e3:.2f} us per item")
