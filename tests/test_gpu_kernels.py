"""GPU: kernel-level parity of the C-ABI ops against the numpy restatement (oracle/ref_numpy.py, float64)
and the golden vectors.  Tolerances: 1e-3 relative is the north_star bound; exact-fp32 kernels are held to
2e-5, indices to bit-exactness."""
import numpy as np
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import ref_numpy as rn
from oracle import ref_port as rp
from two_towers_b200 import functional as Fn

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 2e-5


def _t(a):
    return torch.as_tensor(np.asarray(a), dtype=torch.float32, device=DEV)


class Tol(float):
    """forward tolerance (the float itself) + gradient tolerance (.grad) of one engine"""
    grad = 2e-4


ENGINES = {"simt-fp32": (0, 2e-5, 2e-4), "tcgen05-bf16x3": (1, 1e-4, 2e-4), "tcgen05-v2": (2, 1e-4, 1e-3)}


@pytest.fixture(params=list(ENGINES))
def engine(request):
    """Run GRU tests on all three engines.  Forward: exact fp32 to 2e-5, the 3-pass tcgen05 engines to 1e-4 (bound:
    1e-3).  Gradients: 2e-4, and the north-star bound 1e-3 for v2 whose backward GEMMs are single-pass fp16
    (shapes v2 does not cover - hidden size not a multiple of 64 - fall through to the bf16x3 engine)."""
    from two_towers_b200 import _lib as L
    mode, fwd, grad = ENGINES[request.param]
    L.check(L.lib().tt_set_engine(mode), "set_engine")
    t = Tol(fwd)
    t.grad = grad
    yield t
    L.check(L.lib().tt_set_engine(2), "set_engine")


@pytest.mark.parametrize("M,N,K", [(1, 1, 1), (5, 7, 3), (128, 128, 16), (130, 257, 300), (513, 96, 77)])
def test_gemm_nt(M, N, K):
    rng = np.random.default_rng(M * 1000 + N)
    a, b, bias = rng.standard_normal((M, K)), rng.standard_normal((N, K)), rng.standard_normal(N)
    c = Fn.gemm_nt(_t(a), _t(b), _t(bias), alpha=0.5)
    assert rel_err(c, 0.5 * a.astype(np.float32).astype(np.float64) @ b.astype(np.float32).astype(np.float64).T
                   + bias.astype(np.float32)) < TOL


@pytest.mark.parametrize("dtype", [torch.int64, torch.int32])
def test_embed_gather_pad_and_oob_rows_are_zero(dtype):
    g = torch.Generator().manual_seed(0)
    table = torch.randn(50, 300, generator=g)
    ids = torch.randint(0, 50, (4, 9), generator=g)
    ids[0, 3] = 0
    ids[1, 1] = 49
    out = Fn.embed_gather(ids.to(DEV, dtype), table.to(DEV), pad_id=0).cpu()
    ref = table[ids]
    ref[ids == 0] = 0
    assert torch.equal(out, ref)
    out2 = Fn.embed_gather(torch.tensor([[-1, 50, 7]], device=DEV, dtype=dtype), table.to(DEV), pad_id=-1).cpu()
    assert torch.equal(out2[0, 0], torch.zeros(300)) and torch.equal(out2[0, 1], torch.zeros(300))
    assert torch.equal(out2[0, 2], table[7])
    t2 = torch.randn(11, 7)     # E % 4 != 0 path
    i2 = torch.randint(0, 11, (3, 5))
    assert torch.equal(Fn.embed_gather(i2.to(DEV), t2.to(DEV), pad_id=-1).cpu(), t2[i2])


def _gru_params(rng, In, H, L, nd):
    p = {}
    for l in range(L):
        for sfx in ("", "_reverse")[:nd]:
            il = In if l == 0 else nd * H
            p[f"g.weight_ih_l{l}{sfx}"] = rng.uniform(-0.4, 0.4, (3 * H, il))
            p[f"g.weight_hh_l{l}{sfx}"] = rng.uniform(-0.4, 0.4, (3 * H, H))
            p[f"g.bias_ih_l{l}{sfx}"] = rng.uniform(-0.4, 0.4, 3 * H)
            p[f"g.bias_hh_l{l}{sfx}"] = rng.uniform(-0.4, 0.4, 3 * H)
    return {k: v.astype(np.float32) for k, v in p.items()}


def _flat(p, L, nd):
    out = []
    for l in range(L):
        for sfx in ("", "_reverse")[:nd]:
            out += [p[f"g.{n}_l{l}{sfx}"] for n in ("weight_ih", "weight_hh", "bias_ih", "bias_hh")]
    return [_t(w) for w in out]


@pytest.mark.parametrize("B,T,In,H,L,nd,use_len", [
    (3, 5, 7, 4, 1, 1, False), (3, 5, 7, 4, 2, 2, False), (5, 9, 20, 10, 2, 2, True),
    (130, 6, 300, 33, 2, 2, False), (2, 1, 5, 3, 2, 2, False), (4, 7, 12, 128, 1, 2, True),
    (300, 9, 300, 256, 2, 2, False), (70, 5, 44, 72, 2, 2, True), (9, 3, 16, 8, 2, 2, False), (129, 4, 300, 64, 1, 1, False)])
def test_gru_stack_forward(engine, B, T, In, H, L, nd, use_len):
    rng = np.random.default_rng(B * 100 + T)
    p = _gru_params(rng, In, H, L, nd)
    x = rng.standard_normal((B, T, In)).astype(np.float32)
    lens = rng.integers(1, T + 1, B) if use_len else None
    out_ref, hn_ref = rn.gru_stack(x, p, "g", L, nd == 2, lengths=lens)
    out, hn = Fn.gru_stack(_t(x), _flat(p, L, nd), In, H, L, nd == 2,
                           lengths=None if lens is None else torch.as_tensor(lens))
    assert rel_err(out, out_ref) < engine and rel_err(hn, hn_ref) < engine


@pytest.mark.parametrize("B,T,In,H,use_len", [(3, 4, 6, 5, False), (6, 7, 20, 9, True), (131, 5, 40, 16, False),
                                               (140, 6, 300, 72, True), (260, 3, 52, 256, False), (5, 9, 12, 8, False),
                                               (140, 6, 300, 64, True), (70, 5, 44, 128, True), (33, 12, 300, 128, False)])
def test_gru_stack_backward_vs_torch_autograd(engine, B, T, In, H, use_len):
    """Oracle: stock nn.GRU on CPU in float64 (same arithmetic the reference dispatches), incl. d/dx."""
    torch.manual_seed(B)
    ref = torch.nn.GRU(In, H, num_layers=2, batch_first=True, bidirectional=True).double()
    x = torch.randn(B, T, In, dtype=torch.float64, requires_grad=True)
    lens = torch.randint(1, T + 1, (B,)) if use_len else None
    if use_len:
        pk = torch.nn.utils.rnn.pack_padded_sequence(x, lens, batch_first=True, enforce_sorted=False)
        o, hn = ref(pk)
        o, _ = torch.nn.utils.rnn.pad_packed_sequence(o, batch_first=True, total_length=T)
    else:
        o, hn = ref(x)
    go, gh = torch.randn_like(o), torch.randn_like(hn)
    (o * go).sum().add((hn * gh).sum()).backward()
    ws = [p.detach().float().to(DEV).requires_grad_(True) for p in ref.parameters()]
    xg = x.detach().float().to(DEV).requires_grad_(True)
    o2, hn2 = Fn.gru_stack(xg, ws, In, H, 2, True, lengths=lens)
    assert rel_err(o2, o.detach()) < engine and rel_err(hn2, hn.detach()) < engine
    ((o2 * go.float().to(DEV)).sum() + (hn2 * gh.float().to(DEV)).sum()).backward()
    assert rel_err(xg.grad, x.grad) < engine.grad
    for (n, p), w in zip(ref.named_parameters(), ws):
        assert rel_err(w.grad, p.grad) < engine.grad, (n, rel_err(w.grad, p.grad))


@pytest.mark.parametrize("overlap", [1, 0])
def test_gru_three_layer_backward_helper_stream(overlap):
    """3 layers: the weight-gradient GEMMs of layers 2 and 1 run on the library's helper stream under the BPTT of the
    layer below (double-buffered dG plane, buffer reuse at layer 0 waits for layer 2's GEMMs); same gradients with the
    overlap switched off, both equal to nn.GRU in float64."""
    from two_towers_b200 import _lib as L
    lib = L.lib()
    torch.manual_seed(5)
    B, T, In, H, NL = 140, 6, 40, 64, 3
    ref = torch.nn.GRU(In, H, num_layers=NL, batch_first=True, bidirectional=True).double()
    x = torch.randn(B, T, In, dtype=torch.float64, requires_grad=True)
    o, hn = ref(x)
    go, gh = torch.randn_like(o), torch.randn_like(hn)
    (o * go).sum().add((hn * gh).sum()).backward()
    ws = [p.detach().float().to(DEV).requires_grad_(True) for p in ref.parameters()]
    xg = x.detach().float().to(DEV).requires_grad_(True)
    try:
        lib.tt_debug_wgrad_overlap(overlap)
        for _ in range(2):                                   # twice: helper stream / events are reused
            for w in ws:
                w.grad = None
            xg.grad = None
            o2, hn2 = Fn.gru_stack(xg, ws, In, H, NL, True)
            ((o2 * go.float().to(DEV)).sum() + (hn2 * gh.float().to(DEV)).sum()).backward()
            torch.cuda.synchronize()
            assert rel_err(xg.grad, x.grad) < 1e-3
            for (n, p), w in zip(ref.named_parameters(), ws):
                assert rel_err(w.grad, p.grad) < 1e-3, (n, rel_err(w.grad, p.grad))
    finally:
        lib.tt_debug_wgrad_overlap(1)


@pytest.mark.parametrize("B,T,In,H,NL,nseg,use_len", [(140, 48, 40, 64, 2, 4, False), (130, 64, 24, 128, 1, 4, True),
                                                        (70, 33, 300, 256, 2, 2, True), (260, 80, 16, 64, 3, 5, False),
                                                        (140, 48, 40, 64, 2, 1, False)])
def test_gru_backward_time_segments(B, T, In, H, NL, nseg, use_len):
    """Fused BPTT run as `nseg` launches of the persistent kernel (dL/dh handed over through global memory) with the
    weight-gradient GEMMs of each finished segment accumulated on the helper stream: same gradients as nn.GRU in
    float64, for every layer count (1 layer = the helper stream is used although there is no layer below)."""
    from two_towers_b200 import _lib as L
    lib = L.lib()
    torch.manual_seed(B + T)
    ref = torch.nn.GRU(In, H, num_layers=NL, batch_first=True, bidirectional=True).double()
    x = torch.randn(B, T, In, dtype=torch.float64, requires_grad=True)
    lens = torch.randint(1, T + 1, (B,)) if use_len else None
    if use_len:
        pk = torch.nn.utils.rnn.pack_padded_sequence(x, lens, batch_first=True, enforce_sorted=False)
        o, hn = ref(pk)
        o, _ = torch.nn.utils.rnn.pad_packed_sequence(o, batch_first=True, total_length=T)
    else:
        o, hn = ref(x)
    go, gh = torch.randn_like(o), torch.randn_like(hn)
    (o * go).sum().add((hn * gh).sum()).backward()
    ws = [p.detach().float().to(DEV).requires_grad_(True) for p in ref.parameters()]
    xg = x.detach().float().to(DEV).requires_grad_(True)
    try:
        lib.tt_debug_bwd_segments(nseg)
        for _ in range(2):
            for w in ws:
                w.grad = None
            xg.grad = None
            o2, hn2 = Fn.gru_stack(xg, ws, In, H, NL, True, lengths=lens)
            ((o2 * go.float().to(DEV)).sum() + (hn2 * gh.float().to(DEV)).sum()).backward()
            torch.cuda.synchronize()
            assert rel_err(xg.grad, x.grad) < 1e-3
            for (n, p), w in zip(ref.named_parameters(), ws):
                assert rel_err(w.grad, p.grad) < 1e-3, (n, rel_err(w.grad, p.grad))
    finally:
        lib.tt_debug_bwd_segments(1)


@pytest.mark.parametrize("B,T,In,NL,use_len,drop", [(70, 9, 40, 2, True, 0.0), (300, 17, 300, 2, False, 0.0),
                                                     (131, 33, 24, 1, True, 0.0), (200, 12, 64, 3, False, 0.25),
                                                     (500, 11, 32, 2, True, 0.0), (1024, 40, 48, 1, False, 0.0)])
def test_gru_backward_cta_pairs(B, T, In, NL, use_len, drop):
    """H = 256: the BPTT on 2-CTA clusters (hidden units split over the pair, gate-gradient tiles exchanged through
    distributed shared memory; mode 3) gives the gradients of nn.GRU in float64 and - same saved state, same dropout
    stream - the gradients of the one-CTA kernel (mode 1) up to the summation order of the fp32 accumulators."""
    from two_towers_b200 import _lib as L
    lib = L.lib()
    H = 256
    torch.manual_seed(B + T)
    ref = torch.nn.GRU(In, H, num_layers=NL, batch_first=True, bidirectional=True).double()
    x = torch.randn(B, T, In, dtype=torch.float64, requires_grad=True)
    lens = torch.randint(1, T + 1, (B,)) if use_len else None
    go = torch.randn(B, T, 2 * H, dtype=torch.float64)
    gh = torch.randn(NL * 2, B, H, dtype=torch.float64)
    if drop == 0.0:
        if use_len:
            pk = torch.nn.utils.rnn.pack_padded_sequence(x, lens, batch_first=True, enforce_sorted=False)
            o, hn = ref(pk)
            o, _ = torch.nn.utils.rnn.pad_packed_sequence(o, batch_first=True, total_length=T)
        else:
            o, hn = ref(x)
        (o * go).sum().add((hn * gh).sum()).backward()
    grads = {}
    try:
        for mode in (3, 4, 1):
            lib.tt_debug_bwd_seq_mode(mode)
            ws = [p.detach().float().to(DEV).requires_grad_(True) for p in ref.parameters()]
            xg = x.detach().float().to(DEV).requires_grad_(True)
            torch.manual_seed(77)                                     # same Philox seed for both modes
            o2, hn2 = Fn.gru_stack(xg, ws, In, H, NL, True, lengths=lens, dropout_p=drop, training=drop > 0)
            ((o2 * go.float().to(DEV)).sum() + (hn2 * gh.float().to(DEV)).sum()).backward()
            torch.cuda.synchronize()
            grads[mode] = [xg.grad.clone()] + [w.grad.clone() for w in ws]
            if drop == 0.0:
                assert rel_err(xg.grad, x.grad) < 1e-3
                for (n, p), w in zip(ref.named_parameters(), ws):
                    assert rel_err(w.grad, p.grad) < 1e-3, (mode, n, rel_err(w.grad, p.grad))
    finally:
        lib.tt_debug_bwd_seq_mode(2)
    for mode in (3, 4):
        for a, b in zip(grads[mode], grads[1]):
            assert rel_err(a, b) < 2e-4, (mode, rel_err(a, b))


def test_gru_token_ids_equal_pregathered_floats(engine):
    torch.manual_seed(1)
    table = torch.randn(40, 12)
    table[0] = 0
    ids = torch.randint(0, 40, (5, 6))
    rng = np.random.default_rng(0)
    p = _gru_params(rng, 12, 8, 2, 2)
    a = Fn.gru_stack(ids.to(DEV), _flat(p, 2, 2), 12, 8, 2, True, table=table.to(DEV), pad_id=0)
    b = Fn.gru_stack(table[ids].to(DEV), _flat(p, 2, 2), 12, 8, 2, True)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    ws = [w.requires_grad_(True) for w in _flat(p, 2, 2)]
    o, hn = Fn.gru_stack(ids.to(DEV), ws, 12, 8, 2, True, table=table.to(DEV), pad_id=0)
    hn.sum().backward()
    ws2 = [w.detach().clone().requires_grad_(True) for w in ws]
    Fn.gru_stack(table[ids].to(DEV), ws2, 12, 8, 2, True)[1].sum().backward()
    for g1, g2 in zip(ws, ws2):
        assert rel_err(g1.grad, g2.grad) < 1e-5


def test_gru_interlayer_dropout_statistics_and_determinism():
    rng = np.random.default_rng(5)
    p = _gru_params(rng, 6, 64, 2, 2)
    x = _t(rng.standard_normal((64, 4, 6)))
    torch.manual_seed(11)
    a = Fn.gru_stack(x, _flat(p, 2, 2), 6, 64, 2, True, dropout_p=0.5, training=True)[1]
    torch.manual_seed(11)
    b = Fn.gru_stack(x, _flat(p, 2, 2), 6, 64, 2, True, dropout_p=0.5, training=True)[1]
    c = Fn.gru_stack(x, _flat(p, 2, 2), 6, 64, 2, True, dropout_p=0.5, training=False)[1]
    assert torch.equal(a[:2], c[:2])                    # layer 0 is untouched by inter-layer dropout
    assert torch.equal(a, b)                            # same torch seed -> same Philox seed
    assert not torch.equal(a[2:], c[2:])
    y = torch.empty(1 << 20, device=DEV)
    from two_towers_b200 import _lib as L
    L.check(L.lib().tt_dropout(L.ptr(torch.ones_like(y)), L.ptr(y), y.numel(), 0.1, 1234, 0, L.stream()), "dropout")
    keep = (y > 0).float().mean().item()
    assert abs(keep - 0.9) < 2e-3 and abs(y.max().item() - 1 / 0.9) < 1e-6


# ------------------------------------------------------------------------------------------- head / l2norm
@pytest.mark.parametrize("kind,B,In,P1,D2", [("simple", 5, 16, 8, 4), ("margin", 7, 16, 8, 0),
                                              ("enhanced", 130, 24, 12, 6), ("enhanced", 3, 512, 256, 128)])
def test_head_forward_backward(kind, B, In, P1, D2):
    torch.manual_seed(B)
    lin1, ln = torch.nn.Linear(In, P1).double(), torch.nn.LayerNorm(P1).double()
    ln.weight.data.uniform_(0.5, 1.5); ln.bias.data.uniform_(-0.5, 0.5)
    lin2 = torch.nn.Linear(P1, D2).double() if D2 else None
    v = torch.randn(B, In, dtype=torch.float64, requires_grad=True)
    y = torch.relu(ln(lin1(v)))
    if lin2 is not None:
        y = lin2(y)
    if kind == "simple":
        y = torch.nn.functional.normalize(y, p=2, dim=1)
    go = torch.randn_like(y)
    (y * go).sum().backward()
    f = lambda t: t.detach().float().to(DEV).requires_grad_(True)
    vlo, vhi = f(v[:, :In // 2]), f(v[:, In // 2:])
    params = [f(lin1.weight), f(lin1.bias), f(ln.weight), f(ln.bias)]
    p2 = [f(lin2.weight), f(lin2.bias)] if lin2 is not None else [None, None]
    out = Fn.head(vlo, vhi, *params, *p2, normalize=(kind == "simple"))
    assert rel_err(out, y.detach()) < TOL
    (out * go.float().to(DEV)).sum().backward()
    assert rel_err(torch.cat((vlo.grad, vhi.grad), 1), v.grad) < 1e-4
    refs = [lin1.weight, lin1.bias, ln.weight, ln.bias] + ([lin2.weight, lin2.bias] if lin2 is not None else [])
    for a, b in zip(params + [t for t in p2 if t is not None], refs):
        assert rel_err(a.grad, b.grad) < 1e-4


def test_head_dropout_mask_is_consistent_between_forward_and_backward():
    torch.manual_seed(0)
    B, In, P1 = 64, 8, 256
    f = lambda *s: torch.randn(*s, device=DEV).requires_grad_(True)
    vlo, vhi, w1, b1 = f(B, In // 2), f(B, In // 2), f(P1, In), f(P1)
    lw, lb = torch.ones(P1, device=DEV, requires_grad=True), torch.zeros(P1, device=DEV, requires_grad=True)
    out = Fn.head(vlo, vhi, w1, b1, lw, lb, dropout_p=0.25)
    ref = Fn.head(vlo, vhi, w1, b1, lw, lb, dropout_p=0.0).detach()
    kept = out.detach() != 0
    relu_on = ref > 0
    frac = (kept & relu_on).float().sum() / relu_on.float().sum()
    assert abs(frac.item() - 0.75) < 0.03
    assert torch.allclose(out.detach()[kept], ref[kept] / 0.75, rtol=1e-6)
    out.sum().backward()
    # d out / d ln_bias[c] = sum_rows keep*relu_on / 0.75
    assert torch.allclose(lb.grad, (kept & relu_on).float().sum(0) / 0.75, rtol=1e-5)


def test_l2_normalize_fwd_bwd_and_eps():
    x = torch.randn(9, 33, dtype=torch.float64, requires_grad=True)
    y = torch.nn.functional.normalize(x, p=2, dim=1)
    g = torch.randn_like(y)
    (y * g).sum().backward()
    xg = x.detach().float().to(DEV).requires_grad_(True)
    y2 = Fn.l2_normalize(xg)
    assert rel_err(y2, y.detach()) < TOL
    (y2 * g.float().to(DEV)).sum().backward()
    assert rel_err(xg.grad, x.grad) < 1e-5
    z = Fn.l2_normalize(torch.zeros(2, 4, device=DEV))
    assert torch.equal(z.cpu(), torch.zeros(2, 4))


# ------------------------------------------------------------------------------------------- score + CE, hinge
@pytest.mark.parametrize("Bl,Bg,D,off,inv_t,margin", [
    (9, 9, 16, 0, 10.0, 0.0), (9, 9, 16, 0, 10.0, 0.2), (6, 24, 8, 12, 1 / 0.07, 0.0),
    (130, 300, 64, 170, 10.0, 0.2), (257, 257, 128, 0, 10.0, 0.2)])
def test_score_ce_forward_backward(Bl, Bg, D, off, inv_t, margin):
    rng = np.random.default_rng(Bl + Bg)
    x, y = rng.standard_normal((Bl, D)) * 0.7, rng.standard_normal((Bg, D)) * 0.7
    xt = torch.tensor(x, requires_grad=True); yt = torch.tensor(y, requires_grad=True)
    s = (xt @ yt.t()) * inv_t
    mask = torch.ones_like(s); idx = torch.arange(Bl); mask[idx, idx + off] = 0
    loss = torch.nn.functional.cross_entropy(s - margin * mask, idx + off, reduction="sum")
    assert abs(loss.item() - rn.score_ce(x, y, inv_t, margin, off) * Bl) < 1e-6 * abs(loss.item())
    (loss * 0.37).backward()
    xg, yg = _t(x).requires_grad_(True), _t(y).requires_grad_(True)
    l2 = Fn.score_ce_sum(xg, yg, off, inv_t, margin)
    assert abs(l2.item() - loss.item()) < 1e-5 * abs(loss.item())
    (l2 * 0.37).backward()
    assert rel_err(xg.grad, xt.grad) < 1e-4 and rel_err(yg.grad, yt.grad) < 1e-4


def test_score_ce_large_logits_are_stable():
    # margin loss on raw (un-normalised) vectors gives logits ~ |q||d|/0.1: must not overflow
    torch.manual_seed(0)
    q, d = torch.randn(64, 128) * 3, torch.randn(64, 128) * 3
    ref = rp.margin_ce_inbatch(q.double(), d.double(), 0.2, 0.1).item()
    got = (Fn.score_ce_sum(q.to(DEV), d.to(DEV), 0, 10.0, 0.2) / 64).item()
    assert np.isfinite(got) and abs(got - ref) < 1e-4 * abs(ref)


def test_losses_match_golden_vectors():
    g = load_golden("losses_mining")
    import two_towers_b200 as tt
    from two_towers_b200 import simple_two_tower as S
    q, p, neg = (_t(g[k]) for k in ("q", "pos", "neg"))
    cases = [
        ("simple_infonce_t0.1", lambda a, b, n: S.InfoNCELoss(0.1)(Fn.l2_normalize(a), Fn.l2_normalize(b)), ("dq", "dpos")),
        ("margin_infonce_t0.07", lambda a, b, n: tt.InfoNCELoss()(a, b), ("dq", "dpos")),
        ("margin_ce_m0.2_t0.1", lambda a, b, n: tt.MarginRankingLoss()(a, b), ("dq", "dpos")),
        ("margin_ce_m0.5_t0.05", lambda a, b, n: tt.MarginRankingLoss(0.5, 0.05)(a, b), ("dq", "dpos")),
        ("hinge_m0.2", lambda a, b, n: tt.MarginRankingLoss()(a, b, n), ("dq", "dpos", "dneg")),
        ("hinge_m1.0", lambda a, b, n: tt.MarginRankingLoss(margin=1.0)(a, b, n), ("dq", "dpos", "dneg")),
    ]
    for tag, fn, names in cases:
        a, b, n = (t.clone().requires_grad_(True) for t in (q, p, neg))
        loss = fn(a, b, n)
        assert abs(loss.item() - float(g[f"{tag}/loss"])) < 1e-4 * max(1.0, abs(float(g[f"{tag}/loss"]))), tag
        loss.backward()
        for nm, t in zip(names, (a, b, n)):
            assert rel_err(t.grad, g[f"{tag}/{nm}"]) < 1e-3, (tag, nm)


def test_hinge_vs_float64(B=33, k=5, D=40):
    rng = np.random.default_rng(3)
    q, p, n = rng.standard_normal((B, D)), rng.standard_normal((B, D)), rng.standard_normal((B * k, D))
    qt, pt, nt = (torch.tensor(a, requires_grad=True) for a in (q, p, n))
    loss = rp.hinge_explicit_negatives(qt, pt, nt, 0.6)
    loss.backward()
    a, b, c = (_t(v).requires_grad_(True) for v in (q, p, n))
    l2 = Fn.hinge_sum(a, b, c, 0.6) / B
    assert abs(l2.item() - loss.item()) < 1e-5
    l2.backward()
    for t, r in ((a, qt), (b, pt), (c, nt)):
        assert rel_err(t.grad, r.grad) < 1e-4


# ------------------------------------------------------------------------------------------- top-k / merge / MRR
def _seq_fma_scores(x, y):
    """The kernel's defined arithmetic: one fp32 FMA chain over d ascending (exact product in float64,
    rounded to fp32 after each add)."""
    acc = np.zeros((x.shape[0], y.shape[0]), np.float32)
    x64, y64 = x.astype(np.float64), y.astype(np.float64)
    for d in range(x.shape[1]):
        acc = (acc.astype(np.float64) + np.outer(x64[:, d], y64[:, d])).astype(np.float32)
    return acc


@pytest.mark.parametrize("Q,N,D,k", [(9, 40, 16, 5), (130, 1000, 32, 10), (4, 3, 8, 5), (257, 5000, 128, 8), (1, 700, 64, 32)])
def test_rows_topk_dot_is_bit_exact(Q, N, D, k):
    rng = np.random.default_rng(Q + N)
    x, y = rng.standard_normal((Q, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32)
    y[N // 2] = y[0]                      # exact duplicate rows -> exact ties
    if N > 10:
        y[7] = y[3]
    val, idx = Fn.rows_topk(_t(x), _t(y), k, cosine=False)
    s = _seq_fma_scores(x, y)
    rv, ri = rn.stable_topk(s, min(k, N))
    assert np.array_equal(idx.cpu().numpy()[:, :min(k, N)], ri)
    assert np.array_equal(val.cpu().numpy()[:, :min(k, N)], rv)
    if N < k:
        assert (idx.cpu().numpy()[:, N:] == -1).all() and np.isneginf(val.cpu().numpy()[:, N:]).all()


@pytest.mark.parametrize("cosine", [False, True])
def test_rows_topk_tensor_route_equals_exact_scan(cosine):
    """tensor-engine candidate selection + exact re-score (default engine) must reproduce the exact fused scan
    (engine 0) bit for bit, including rows that fail the acceptance test (40 copies of one document -> more ties
    than candidate slots) and a masked positive."""
    rng = np.random.default_rng(77)
    Q, N, D, k = 300, 20000, 128, 8
    x, y = rng.standard_normal((Q, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32)
    y[100:140] = y[50]                                   # 41 identical rows
    x[:20] = y[50] + 0.01 * rng.standard_normal((20, D)).astype(np.float32)   # queries whose top hits are those ties
    from two_towers_b200 import _lib as L
    lib = L.lib()
    try:
        lib.tt_set_engine(0)
        ev, ei = Fn.rows_topk(_t(x), _t(y), k, cosine=cosine, mask_pos_offset=5, index_base=1000)
    finally:
        lib.tt_set_engine(2)
    tv, ti = Fn.rows_topk(_t(x), _t(y), k, cosine=cosine, mask_pos_offset=5, index_base=1000)
    assert torch.equal(ti, ei) and torch.equal(tv, ev)
    try:                                       # query tile streamed with every corpus tile instead of kept in shared memory
        lib.tt_debug_sel_resident(0)
        sv, si = Fn.rows_topk(_t(x), _t(y), k, cosine=cosine, mask_pos_offset=5, index_base=1000)
    finally:
        lib.tt_debug_sel_resident(1)
    assert torch.equal(si, ei) and torch.equal(sv, ev)
    assert ti[0].tolist() == [1050] + list(range(1100, 1107))


@pytest.mark.parametrize("cosine", [False, True])
@pytest.mark.parametrize("passes", [2, 3, 22])
def test_rows_topk_near_duplicates_inside_the_approximation_error(cosine, passes):
    """Clusters of documents whose scores differ by less than the error of the tensor-engine candidate scan (2 passes:
    2^-11, 3 passes: 2^-13): the approximate order inside a cluster is arbitrary, so the result is only right if the
    acceptance test sends those rows to the exact scan (or the exact re-score restores the order).  Must equal the
    exact engine bit for bit, for clusters smaller and larger than the candidate list."""
    from two_towers_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(99)
    Q, N, D, k = 200, 30000, 128, 10
    y = rng.standard_normal((N, D)).astype(np.float32)
    for c, (size, eps) in enumerate(((6, 3e-4), (12, 1e-4), (14, 3e-5), (40, 1e-4), (40, 2e-3))):
        base = y[1000 * c + 17].copy()
        for j in range(size):
            y[1000 * c + 100 + 7 * j] = base + eps * rng.standard_normal(D).astype(np.float32)
    x = rng.standard_normal((Q, D)).astype(np.float32)
    for c in range(5):                                   # queries aimed at each cluster (and small variations of them)
        for r in range(8):
            x[8 * c + r] = y[1000 * c + 17] + 0.05 * r * rng.standard_normal(D).astype(np.float32)
    try:
        lib.tt_set_engine(0)
        ev, ei = Fn.rows_topk(_t(x), _t(y), k, cosine=cosine)
    finally:
        lib.tt_set_engine(2)
    try:
        lib.tt_debug_sel_passes(passes)
        tv, ti = Fn.rows_topk(_t(x), _t(y), k, cosine=cosine)
    finally:
        lib.tt_debug_sel_passes(2)
    assert torch.equal(ti, ei) and torch.equal(tv, ev)
    # the rows aimed at the clusters were rejected by the acceptance test (and re-done exactly: by the sparse re-scan for
    # passes 2 / 3, by the tile-shaped exact scan for 22); rows of plain random data were not
    import two_towers_b200 as tt
    index = tt.CorpusIndex(_t(y), cosine=cosine)
    try:
        lib.tt_debug_sel_passes(passes)
        iv, ii, flags = Fn.rows_topk_indexed(_t(x), index.docs, index.buffer, k, cosine=cosine, return_fallback_rows=True)
    finally:
        lib.tt_debug_sel_passes(2)
    assert torch.equal(ii, ei) and torch.equal(iv, ev)
    flags = flags.cpu().numpy()
    assert flags[:40].sum() >= 8 and flags[40:].sum() <= 2


@pytest.mark.parametrize("scale", [1e-3, 1e-5, 1e-7])
def test_rows_topk_tiny_query_norms(scale):
    """Queries whose components fall into the fp16 subnormal range: their hi + lo planes no longer carry them exactly, the
    acceptance bound must grow by the measured residual (or reject the row) - the result still equals the exact engine."""
    from two_towers_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(21)
    Q, N, D, k = 256, 20000, 128, 10
    x = (scale * rng.standard_normal((Q, D))).astype(np.float32)
    y = rng.standard_normal((N, D)).astype(np.float32)
    try:
        lib.tt_set_engine(0)
        ev, ei = Fn.rows_topk(_t(x), _t(y), k, cosine=True)
    finally:
        lib.tt_set_engine(2)
    tv, ti = Fn.rows_topk(_t(x), _t(y), k, cosine=True)
    assert torch.equal(ti, ei) and torch.equal(tv, ev)


def test_rows_topk_many_rejected_rows_take_the_tile_scan():
    """More rejected rows than the sparse re-scan holds (256): every query aims at a cluster of 40 near-identical
    documents, so all 600 rows fail the acceptance test and the gated tile-shaped exact scan must produce them."""
    from two_towers_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(5)
    Q, N, D, k = 600, 8000, 64, 10
    y = rng.standard_normal((N, D)).astype(np.float32)
    base = y[0].copy()
    for j in range(40):
        y[50 + 3 * j] = base + 1e-5 * rng.standard_normal(D).astype(np.float32)
    x = (base[None, :] + 0.02 * rng.standard_normal((Q, D))).astype(np.float32)
    try:
        lib.tt_set_engine(0)
        ev, ei = Fn.rows_topk(_t(x), _t(y), k, cosine=True)
    finally:
        lib.tt_set_engine(2)
    import two_towers_b200 as tt
    index = tt.CorpusIndex(_t(y), cosine=True)
    tv, ti, flags = Fn.rows_topk_indexed(_t(x), index.docs, index.buffer, k, cosine=True, return_fallback_rows=True)
    assert int(flags.sum()) > 256
    assert torch.equal(ti, ei) and torch.equal(tv, ev)


def test_hard_negatives_match_golden_and_reference_semantics():
    g = load_golden("losses_mining")
    from two_towers_b200 import enhanced_two_tower as E
    q, docs, off = _t(g["q"]), _t(g["docs"]), int(g["hardneg_pos_offset"])
    for k in (5, 8):
        got = E.get_hard_negatives_batch(q, docs, k=k, pos_offset=off).cpu().numpy()
        assert np.array_equal(got, g[f"hardneg_k{k}"])
        one = E.get_hard_negatives(q[2], docs, positive_idx=off + 2, k=k).cpu().numpy()
        assert np.array_equal(one, g[f"hardneg_k{k}"][2]) and one.dtype == np.int64
    val, idx = Fn.rows_topk(q, docs, 10, cosine=True)
    assert np.array_equal(idx.cpu().numpy(), g["search_top10_idx"])
    assert np.allclose(val.cpu().numpy(), g["search_top10_val"], atol=2e-6)


def test_hard_negatives_ties_and_masked_positive():
    torch.manual_seed(0)
    docs = torch.randn(300, 32)
    docs[200] = docs[10]; docs[250] = docs[10]
    q = docs[10:12].clone()                              # query 0 == doc 10 (its positive), ties at 200/250
    idx = Fn.rows_topk(q.to(DEV), docs.to(DEV), 4, cosine=True, mask_pos_offset=10)[1].cpu()
    assert idx[0, 0].item() == 200 and idx[0, 1].item() == 250 and 10 not in idx[0].tolist()
    ref = rp.hard_negatives_rows(q.double(), docs.double(), 10, 4)
    assert torch.equal(idx, ref)


def test_sharded_topk_merge_equals_global_and_index_base():
    rng = np.random.default_rng(9)
    Q, N, D, k, S = 70, 4096, 32, 10, 4
    x, y = rng.standard_normal((Q, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32)
    y[3000] = y[5]; y[1024] = y[5]
    gv, gi = Fn.rows_topk(_t(x), _t(y), k, cosine=True)
    per = N // S
    vals, idxs = zip(*[Fn.rows_topk(_t(x), _t(y[s * per:(s + 1) * per]), k, cosine=True, index_base=s * per) for s in range(S)])
    mv, mi = Fn.topk_merge(torch.stack(vals), torch.stack(idxs))
    assert torch.equal(mi, gi) and torch.equal(mv, gv)
    # shards given in another order still merge to the same answer (lowest GLOBAL id wins ties)
    mv2, mi2 = Fn.topk_merge(torch.stack(vals[::-1]), torch.stack(idxs[::-1]))
    assert torch.equal(mi2, gi)


@pytest.mark.parametrize("cosine", [False, True])
@pytest.mark.parametrize("N", [100, 20000])
def test_corpus_index_gives_identical_results_and_is_reusable(cosine, N):
    """tt_corpus_index_build + tt_rows_topk_indexed == tt_rows_topk bit for bit (tensor route at N = 20000, exact scan at
    N = 100 where the index holds norms only), for several query batches against ONE index, with a masked positive,
    an index base, ties that overflow the candidate slots, and with the exact engine selected after the build."""
    import two_towers_b200 as tt
    from two_towers_b200 import _lib as L
    rng = np.random.default_rng(1234 + N)
    D, k = 128, 8
    y = rng.standard_normal((N, D)).astype(np.float32) * rng.uniform(0.2, 3.0, (N, 1)).astype(np.float32)
    y[60:90] = y[50]
    yt = _t(y)
    index = tt.CorpusIndex(yt, cosine=cosine, index_base=700)
    assert len(index) == N
    for Q, seed in ((1, 0), (300, 1), (130, 2)):
        x = np.random.default_rng(seed).standard_normal((Q, D)).astype(np.float32)
        x[0] = y[50]
        xt = _t(x)
        for off in (-1, 5):
            wv, wi = Fn.rows_topk(xt, yt, k, cosine=cosine, mask_pos_offset=off, index_base=700)
            gv, gi = index.search(xt, k, mask_pos_offset=off)
            assert torch.equal(gi, wi) and torch.equal(gv, wv)
    sv, si = tt.search_topk(xt, index, k=k, cosine=cosine)
    assert torch.equal(si, Fn.rows_topk(xt, yt, k, cosine=cosine, index_base=700)[1])
    with pytest.raises(ValueError):
        tt.search_topk(xt, index, k=k, cosine=not cosine)
    lib = L.lib()
    try:
        lib.tt_set_engine(0)                 # exact scan reading the index's norms
        ev, ei = index.search(xt, k)
    finally:
        lib.tt_set_engine(2)
    assert torch.equal(ei, si) and torch.equal(ev, sv)
    # a buffer that is too small is refused, not overrun
    with pytest.raises(RuntimeError):
        Fn.rows_topk_indexed(xt, yt, index.buffer[:64], k, cosine=cosine)


def test_mrr_ranks_and_recall():
    from two_towers_b200 import retrieval
    top = torch.tensor([[5, 2, 9], [1, 3, 4], [7, 8, 6]], device=DEV)
    rel = torch.tensor([2, 1, 0], device=DEV)
    first, rr = Fn.mrr_ranks(top, rel)
    assert first.tolist() == [2, 1, 0] and rr.tolist() == [0.5, 1.0, 0.0]
    out = retrieval.mrr_recall(top, rel, ks=(1, 3))
    ref = rp.mrr_recall(top.cpu(), rel.cpu(), ks=(1, 3))
    assert all(abs(out[k] - ref[k]) < 1e-12 for k in ref)


# ------------------------------------------------------------------------------------------- tcgen05 engine alone
@pytest.mark.parametrize("M,N,K,trans,splitk", [(128, 128, 64, 0, 1), (300, 200, 100, 0, 1), (5, 7, 3, 0, 1), (1000, 768, 300, 0, 1),
                                                 (300, 200, 100, 1, 1), (768, 304, 5000, 1, 0), (256, 256, 4096, 0, 4), (64, 1536, 520, 0, 1)])
def test_gemm_x3_engine(M, N, K, trans, splitk):
    from two_towers_b200 import _lib as L
    lib = L.lib()
    g = torch.Generator().manual_seed(M + N + K)
    A, B, bias = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g), torch.randn(N, generator=g)
    ref = 0.5 * A.double() @ B.double().t() + bias.double()
    Ad = (A.t().contiguous() if trans else A).to(DEV)
    Bd = (B.t().contiguous() if trans else B).to(DEV)
    C = torch.empty(M, N, device=DEV)
    ws = L.workspace(lib.tt_gemm_x3_workspace(M, N, K), DEV)
    L.check(lib.tt_gemm_x3(L.ptr(Ad), L.ptr(Bd), L.ptr(C), L.ptr(bias.to(DEV)), M, N, K, 0.5, 0.0, trans, splitk,
                           ws.data_ptr(), ws.numel(), L.stream()), "gemm_x3")
    assert rel_err(C, ref) < 3e-5


@pytest.mark.parametrize("flags", [0, 1, 3, 4 | 8, 1 | 4 | 8, 1 | 2 | 4 | 8, 1 | 2 | 8])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (300, 200, 130), (1000, 520, 96), (384, 300, 2100), (768, 256, 8192)])
def test_gemm_operand_formats(M, N, K, flags):
    """bf16 / fp16 planes, 3 passes / 1 pass, K-major / MN-major operands of the tcgen05 GEMM vs float64."""
    from two_towers_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(M + N + K + flags)
    a = rng.standard_normal((M, K)).astype(np.float32)
    b = rng.standard_normal((N, K)).astype(np.float32)
    ref = a.astype(np.float64) @ b.astype(np.float64).T
    ad = _t(a.T.copy() if flags & 4 else a)
    bd = _t(b.T.copy() if flags & 8 else b)
    c = torch.zeros(M, N, device=DEV)
    ws = L.workspace(lib.tt_gemm_x3_workspace(M, N, K), DEV)
    L.check(lib.tt_gemm_test(L.ptr(ad), L.ptr(bd), L.ptr(c), M, N, K, 1.0, 0.0, flags, ws.data_ptr(), ws.numel(),
                             L.stream()), "tt_gemm_test")
    tol = 2e-3 if flags & 2 else 5e-5          # single fp16 pass: 2^-11 per operand; 3 passes: ~fp32
    assert flags & 1 or not flags & 2          # (bf16 single pass is not a supported mode of the pipeline)
    assert rel_err(c, ref) < tol, rel_err(c, ref)


@pytest.mark.parametrize("flags", [0, 1, 1 | 2])
@pytest.mark.parametrize("M,N,K", [(128 * 76, 1024, 300), (128 * 80, 1000, 128), (128 * 150, 768, 520)])
def test_gemm_cta_pairs_bit_identical_to_single_cta(M, N, K, flags):
    """Large K-major problems run on 2-CTA clusters (tcgen05.mma.cta_group::2, 256 x 256 tiles, each CTA stages half of B):
    vs float64 within the operand-format tolerance, and bit-identical to the one-CTA-per-tile kernel (same products, same
    order per output element).  Ragged N (1000) and K (300, 520) tails included."""
    from two_towers_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(M + N + K + flags)
    a = rng.standard_normal((M, K)).astype(np.float32)
    b = rng.standard_normal((N, K)).astype(np.float32)
    ad, bd = _t(a), _t(b)
    ws = L.workspace(lib.tt_gemm_x3_workspace(M, N, K), DEV)
    out = []
    try:
        for pairs in (1, 0):
            lib.tt_debug_gemm_pairs(pairs)
            c = torch.zeros(M, N, device=DEV)
            L.check(lib.tt_gemm_test(L.ptr(ad), L.ptr(bd), L.ptr(c), M, N, K, 1.0, 0.0, flags, ws.data_ptr(), ws.numel(),
                                     L.stream()), "tt_gemm_test")
            torch.cuda.synchronize()
            out.append(c)
    finally:
        lib.tt_debug_gemm_pairs(1)
    ref = torch.from_numpy(a.astype(np.float64) @ b.astype(np.float64).T)
    tol = 2e-3 if flags & 2 else 5e-5
    assert rel_err(out[0], ref) < tol, rel_err(out[0], ref)
    assert torch.equal(out[0], out[1])
