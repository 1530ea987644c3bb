"""Precision study (CPU, torch): what relative error do fp16 / bf16 / tf32-like operand roundings of every GRU GEMM
(forward AND backward operands) give on the enhanced model at the bench shape (Hg=256, Tq=32, Td=128)?
fp32 accumulation everywhere; only the MMA operands are rounded.  Motivates the engine's operand format."""
import sys, torch
sys.path.insert(0, ".")
from oracle import ref_port as rp

def rnd(x, mode):
    if mode == "fp32": return x
    if mode == "fp16": return x.half().float()
    if mode == "bf16": return x.bfloat16().float()
    if mode == "bf16x3":
        hi = x.bfloat16().float(); return hi + (x - hi).bfloat16().float()
    if mode == "tf32":
        i = x.view(torch.int32)
        return ((i + 0x1000) & ~0x1FFF).view(torch.float32)
    raise ValueError(mode)

BWD_MODE = [None]
class MM(torch.autograd.Function):           # y = a @ b.T with rounded operands in fwd and bwd
    @staticmethod
    def forward(ctx, a, b, mode, gscale):
        ctx.mode, ctx.gscale = mode, gscale
        ctx.save_for_backward(a, b)
        return rnd(a, mode) @ rnd(b, mode).t()
    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        m, s = BWD_MODE[0] or ctx.mode, ctx.gscale
        gr = rnd(g * s, m) / s
        return gr @ rnd(b, m), gr.t() @ rnd(a, m), None, None

def gru_layer(x, w, mode, gs, store=None):
    B, T, _ = x.shape
    outs = []
    for d in range(2):
        w_ih, w_hh, b_ih, b_hh = w[4*d:4*d+4]
        H = w_hh.shape[1]
        gi = MM.apply(x.reshape(B*T, -1), w_ih, mode, gs).view(B, T, 3*H) + b_ih
        if store: gi = rnd(gi, store)
        h = x.new_zeros(B, H); hs = [None]*T
        for t in (range(T) if d == 0 else reversed(range(T))):
            gh = MM.apply(h, w_hh, mode, gs) + b_hh
            r = torch.sigmoid(gi[:, t, :H] + gh[:, :H]); z = torch.sigmoid(gi[:, t, H:2*H] + gh[:, H:2*H])
            n = torch.tanh(gi[:, t, 2*H:] + r * gh[:, 2*H:])
            h = (1 - z) * n + z * h
            hs[t] = h
        outs.append(torch.stack(hs, 1))
    return torch.cat(outs, 2), (outs[0][:, -1], outs[1][:, 0])

def tower(x, enc, proj, mode, gs, store):
    ws = [getattr(enc, n) for n in enc._flat_weights_names]
    o, _ = gru_layer(x, ws[:8], mode, gs, store)
    _, (hf, hr) = gru_layer(o, ws[8:], mode, gs, store)
    return proj(torch.cat((hf, hr), 1))

def run(mode, gs=1.0, store=None, B=48, H=128, seed=0):
    torch.manual_seed(seed)
    m = rp.disable_dropout(rp.build_port("enhanced", 300, H)).train()
    g = torch.Generator().manual_seed(1)
    q = torch.randn(B, 32, 300, generator=g) * 0.1; d = torch.randn(B, 128, 300, generator=g) * 0.1
    qv = tower(q, m.query_encoder, m.query_proj, mode, gs, store); dv = tower(d, m.doc_encoder, m.doc_proj, mode, gs, store)
    loss = rp.margin_ce_inbatch(qv, dv, 0.2, 0.1)
    loss.backward()
    return qv.detach(), dv.detach(), loss.detach(), {n: p.grad.clone() for n, p in m.named_parameters()}

if __name__ == "__main__":
    torch.set_num_threads(8)
    ref = run("fp32")
    rel = lambda a, b: ((a.double() - b.double()).norm() / b.double().norm()).item()
    for mode, bmode, gs, store in [("bf16x3", "bf16x3", 1.0, None), ("bf16x3", "fp16", 1.0, None), ("bf16x3", "fp16", 4096.0, None), ("bf16x3", "fp16", 4096.0, "fp16"), ("bf16x3", "bf16", 1.0, None)]:
        BWD_MODE[0] = bmode
        o = run(mode, gs, store)
        mode = mode + "/" + bmode
        gerr = {n: rel(o[3][n], ref[3][n]) for n in ref[3]}
        worst = max(gerr, key=gerr.get)
        print(f"{mode:12s} gscale {gs:6.0f} store {store}: q {rel(o[0], ref[0]):.2e} d {rel(o[1], ref[1]):.2e} loss {abs(o[2]-ref[2]).item()/abs(ref[2]).item():.2e} "
              f"grad max {gerr[worst]:.2e} ({worst}) median {sorted(gerr.values())[len(gerr)//2]:.2e}", flush=True)
