"""Timeline of the fused BPTT kernel (gru_bwd_seq_kernel), CTA 0: per step marks
[0 step begin, 1 acc ready, 2 pass-1 done, then per chunk c: 3+3c staging free, 4+3c staging written, 5+3c stores issued]."""
import sys, torch
sys.path.insert(0, ".")
import two_towers_b200 as tt
from two_towers_b200 import _lib as L
dev = torch.device("cuda")
B, T, In, H = 4096, 24, 512, 256
torch.manual_seed(0)
gru = tt.GRU(In, H, num_layers=1, bidirectional=True).to(dev)
x = torch.randn(B, T, In, device=dev) * 0.1
buf = torch.zeros(64 * 16, dtype=torch.int64, device=dev)
lib = L.lib()
import os
if os.environ.get("TT_BWD_SEQ_MODE"): lib.tt_debug_bwd_seq_mode(int(os.environ["TT_BWD_SEQ_MODE"]))
def run():
    o, hn = gru(x, need_output=False)
    hn.square().sum().backward()
run(); torch.cuda.synchronize()
lib.tt_debug_step_buffer(buf.data_ptr())
run(); torch.cuda.synchronize()
lib.tt_debug_step_buffer(None)
t = buf.view(64, 16).cpu().double()
base = t[4, 0]
for i in range(4, 12):
    print(f"step {i:2d} " + " ".join(f"{(v - base) / 1e3:7.2f}" for v in t[i, :16]))
print("per step:", (t[20, 0] - t[8, 0]).item() / 12e3, "us")
