"""Backward time of one bidirectional GRU layer at the bench shape under each BPTT kernel choice (tt_debug_bwd_seq_mode)."""
import sys, torch
sys.path.insert(0, ".")
import two_towers_b200 as tt
from two_towers_b200 import _lib as L
dev = torch.device("cuda")
B, In, H = 4096, 512, 256
lib = L.lib()
torch.manual_seed(0)
gru = tt.GRU(In, H, num_layers=1, bidirectional=True).to(dev)
for T in (128, 32):
    x = torch.randn(B, T, In, device=dev) * 0.1
    for mode in (2, 3, 1, 4):
        lib.tt_debug_bwd_seq_mode(mode)
        ts = []
        for it in range(6):
            o, hn = gru(x, need_output=False)
            loss = hn.square().sum()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); loss.backward(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print(f"T={T} bwd_seq_mode={mode}: backward {min(ts[2:]):.3f} ms (min of 4), {min(ts[2:]) * 1e3 / T:.2f} us/step incl. GEMMs")
lib.tt_debug_bwd_seq_mode(2)
