"""Autograd-aware Python wrappers over the C ABI (include/tt_b200.h).  Each Function only allocates buffers,
fills the argument struct and calls the library; there is no torch arithmetic on the hot path."""
import ctypes as C

import torch

from . import _lib as L

def _next_seed():
    """Per-call Philox seed derived from torch's CPU generator so torch.manual_seed() makes runs repeatable."""
    return int(torch.randint(0, 2 ** 62, (1,)).item())


# Dropout under CUDA-graph replay: a captured call would replay the SAME Philox seed forever.  Inside `device_seed(t)` every
# dropout-bearing call also hands the kernels the device scalar `t` (int64, 1 element), which they add to the host seed when
# they RUN - the graphed step bumps it once per replay (graph.GraphedTrainStep).
_SEED_DEV = None


class device_seed:
    def __init__(self, tensor):
        if tensor is not None and (tensor.dtype != torch.int64 or tensor.numel() != 1 or not tensor.is_cuda):
            raise ValueError("device_seed: a 1-element int64 CUDA tensor")
        self.tensor = tensor

    def __enter__(self):
        global _SEED_DEV
        self.prev, _SEED_DEV = _SEED_DEV, self.tensor
        return self.tensor

    def __exit__(self, *exc):
        global _SEED_DEV
        _SEED_DEV = self.prev


def _arr2(tensors):
    a = (C.c_void_p * 2)()
    for i, t in enumerate(tensors):
        a[i] = L.ptr(t)
    return a


# =============================================================================== GRU stack
class GRUConfig:
    def __init__(self, input_size, hidden_size, num_layers, ndir, dropout_p, training, lengths=None,
                 table=None, pad_id=-1, need_grad=True):
        self.input_size, self.hidden_size, self.num_layers, self.ndir = input_size, hidden_size, num_layers, ndir
        self.dropout_p = float(dropout_p) if training else 0.0
        self.lengths, self.table, self.pad_id, self.need_grad = lengths, table, pad_id, need_grad


def _fill_layer(cfg, l, x, ids, weights, out, h_n, gates, seed, ws, B, T):
    H, nd = cfg.hidden_size, cfg.ndir
    a = L.GruLayer()
    a.B, a.T, a.H, a.ndir = B, T, H, nd
    a.In = cfg.input_size if l == 0 else nd * H
    a.save_for_backward = 1 if gates is not None else 0
    a.x = L.ptr(x)
    a.ids = L.ptr(ids)
    a.ids_are_i64 = 1 if (ids is not None and ids.dtype == torch.int64) else 0
    if ids is not None:
        a.table, a.vocab, a.pad_id = L.ptr(cfg.table), cfg.table.shape[0], cfg.pad_id
    a.dropout_p = cfg.dropout_p if l > 0 else 0.0
    a.seed, a.offset = seed, 0
    w = weights[l * nd * 4:(l + 1) * nd * 4]
    a.w_ih = _arr2([w[4 * d + 0] for d in range(nd)])
    a.w_hh = _arr2([w[4 * d + 1] for d in range(nd)])
    a.b_ih = _arr2([w[4 * d + 2] for d in range(nd)])
    a.b_hh = _arr2([w[4 * d + 3] for d in range(nd)])
    a.lengths = L.ptr(cfg.lengths)
    a.out, a.h_n, a.gates = L.ptr(out), L.ptr(h_n), L.ptr(gates)
    if ws is not None:
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
    return a


def _ptr_array(tensors):
    a = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        a[i] = L.ptr(t)
    return a


def _fill_stack(cfg, x, is_ids, weights, warr, out, h_n, seed, saved, B, T, seed_dev=None):
    a = L.GruStack()
    a.seed_dev = seed_dev.data_ptr() if seed_dev is not None else None
    a.B, a.T, a.In, a.H, a.L, a.ndir = B, T, cfg.input_size, cfg.hidden_size, cfg.num_layers, cfg.ndir
    a.save_for_backward = 1 if saved is not None else 0
    a.ids_are_i64 = 1 if (is_ids and x.dtype == torch.int64) else 0
    if is_ids:
        a.ids, a.table, a.vocab, a.pad_id = L.ptr(x), L.ptr(cfg.table), cfg.table.shape[0], cfg.pad_id
    else:
        a.x = L.ptr(x)
    a.dropout_p, a.seed = cfg.dropout_p, seed
    a.weights = C.cast(warr, C.POINTER(C.c_void_p))
    a.lengths, a.out, a.h_n = L.ptr(cfg.lengths), L.ptr(out), L.ptr(h_n)
    if saved is not None:
        a.saved, a.saved_bytes = saved.data_ptr(), saved.numel()
    return a


class GRUStackFunction(torch.autograd.Function):
    """Multi-layer (bi)directional GRU == nn.GRU(batch_first=True) (reference: simple_two_tower.py:17-32,
    margin_two_tower.py:12-28, enhanced_two_tower.py:17-33) through tt_gru_stack_fwd / tt_gru_stack_bwd.
    Inputs: cfg, x ([B,T,In] fp32 or [B,T] ids), then the flat nn.GRU weight list.
    Returns (output of the last layer or an empty tensor when cfg.need_out is False, h_n [L*ndir,B,H])."""

    @staticmethod
    def forward(ctx, cfg, x, *weights):
        lib = L.lib()
        is_ids = not torch.is_floating_point(x)
        if is_ids and cfg.table is None:
            raise RuntimeError("token ids given but the GRU has no embedding table")
        if is_ids and x.dtype not in (torch.int32, torch.int64):
            x = x.long()
        x = x.contiguous() if is_ids else L.f32c(x)
        weights = [L.f32c(w) for w in weights]
        B, T = x.shape[0], x.shape[1]
        H, nd, dev = cfg.hidden_size, cfg.ndir, x.device
        h_n = torch.empty(cfg.num_layers * nd, B, H, device=dev, dtype=torch.float32)
        out = torch.empty(B, T, nd * H, device=dev, dtype=torch.float32) if cfg.need_out else None
        seed = _next_seed() if (cfg.num_layers > 1 and cfg.dropout_p > 0) else 0
        warr = _ptr_array(weights)
        seed_dev = _SEED_DEV if seed != 0 else None
        a = _fill_stack(cfg, x, is_ids, weights, warr, out, h_n, seed, None, B, T, seed_dev)
        saved = None
        if cfg.need_grad:
            a.save_for_backward = 1
            saved = L.workspace(lib.tt_gru_stack_saved_bytes(C.byref(a)), dev)
            a.saved, a.saved_bytes = saved.data_ptr(), saved.numel()
        ws = L.workspace(lib.tt_gru_stack_fwd_workspace(C.byref(a)), dev)
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        L.run(dev, lib.tt_gru_stack_fwd, "tt_gru_stack_fwd", C.byref(a))
        if cfg.need_grad:
            ctx.cfg, ctx.is_ids, ctx.seed, ctx.nw, ctx.seed_dev = cfg, is_ids, seed, len(weights), seed_dev
            ctx.save_for_backward(x, saved, *weights)
        if out is None:
            out = torch.empty(0, device=dev, dtype=torch.float32)
            ctx.mark_non_differentiable(out)
        return out, h_n

    @staticmethod
    def backward(ctx, d_out, d_h_n):
        lib = L.lib()
        cfg = ctx.cfg
        x, saved = ctx.saved_tensors[0], ctx.saved_tensors[1]
        weights = list(ctx.saved_tensors[2:])
        B, T, dev = x.shape[0], x.shape[1], x.device
        d_out = L.f32c(d_out) if (d_out is not None and cfg.need_out) else None
        d_h_n = L.f32c(d_h_n) if d_h_n is not None else None
        grads = [torch.empty_like(w) for w in weights]
        need_dx0 = (not ctx.is_ids) and ctx.needs_input_grad[1]
        d_x = torch.empty_like(x) if need_dx0 else None
        warr, garr = _ptr_array(weights), _ptr_array(grads)
        g = L.GruStackGrad()
        g.fwd = _fill_stack(cfg, x, ctx.is_ids, weights, warr, None, d_h_n if d_h_n is not None else grads[0], ctx.seed,
                            saved, B, T, ctx.seed_dev)
        g.d_out, g.d_h_n, g.d_x = L.ptr(d_out), L.ptr(d_h_n), L.ptr(d_x)
        g.d_weights = C.cast(garr, C.POINTER(C.c_void_p))
        ws = L.workspace(lib.tt_gru_stack_bwd_workspace(C.byref(g)), dev)
        g.workspace, g.workspace_bytes = ws.data_ptr(), ws.numel()
        L.run(dev, lib.tt_gru_stack_bwd, "tt_gru_stack_bwd", C.byref(g))
        return (None, d_x, *grads)


def gru_stack(x, weights, input_size, hidden_size, num_layers=2, bidirectional=True, dropout_p=0.0, training=False,
              lengths=None, table=None, pad_id=-1, need_out=True):
    need_grad = torch.is_grad_enabled() and (any(w.requires_grad for w in weights) or
                                             (torch.is_floating_point(x) and x.requires_grad))
    if lengths is not None:
        lengths = lengths.to(device=x.device, dtype=torch.int32).contiguous()
    if table is not None:
        table = L.f32c(table.detach())
    cfg = GRUConfig(input_size, hidden_size, num_layers, 2 if bidirectional else 1, dropout_p, training, lengths,
                    table, pad_id, need_grad)
    cfg.need_out = bool(need_out)
    return GRUStackFunction.apply(cfg, x, *weights)


# =============================================================================== projection head
class HeadFunction(torch.autograd.Function):
    """cat(v_lo, v_hi) -> Linear -> LayerNorm -> ReLU [-> Dropout] [-> Linear] [-> L2 normalise]  (K-D)."""

    @staticmethod
    def forward(ctx, v_lo, v_hi, w1, b1, ln_w, ln_b, w2, b2, normalize, dropout_p, need_grad):
        lib = L.lib()
        v_lo, v_hi = L.f32c(v_lo), L.f32c(v_hi)
        w1, b1, ln_w, ln_b = L.f32c(w1), L.f32c(b1), L.f32c(ln_w), L.f32c(ln_b)
        has2 = w2 is not None
        if has2:
            w2, b2 = L.f32c(w2), L.f32c(b2)
        B, dev = v_lo.shape[0], v_lo.device
        In, P1 = w1.shape[1], w1.shape[0]
        D2 = w2.shape[0] if has2 else 0
        Dl = D2 if has2 else P1
        a = L.Head()
        a.B, a.In, a.P1, a.D2, a.normalize = B, In, P1, D2, 1 if normalize else 0
        a.dropout_p = float(dropout_p)
        a.seed, a.offset = (_next_seed() if dropout_p > 0 else 0), 0
        seed_dev = _SEED_DEV if dropout_p > 0 else None
        a.seed_dev = seed_dev.data_ptr() if seed_dev is not None else None
        e = lambda *s: torch.empty(*s, device=dev, dtype=torch.float32)
        y1, stats, y2 = e(B, P1), e(B, 2), e(B, P1)
        y3 = e(B, D2) if has2 else None
        out, norm = (e(B, Dl), e(B)) if normalize else (None, None)
        a.v_lo, a.v_hi = L.ptr(v_lo), L.ptr(v_hi)
        a.w1, a.b1, a.ln_w, a.ln_b, a.w2, a.b2 = (L.ptr(t) for t in (w1, b1, ln_w, ln_b, w2, b2))
        a.y1, a.stats, a.y2, a.y3, a.out, a.norm = (L.ptr(t) for t in (y1, stats, y2, y3, out, norm))
        ws = L.workspace(lib.tt_head_fwd_workspace(C.byref(a)), dev)
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        L.run(dev, lib.tt_head_fwd, "tt_head_fwd", C.byref(a))
        result = out if normalize else (y3 if has2 else y2)
        if need_grad:
            ctx.meta = (B, In, P1, D2, bool(normalize), float(dropout_p), int(a.seed), has2)
            tensors = [v_lo, v_hi, w1, b1, ln_w, ln_b, y1, stats, y2]
            tensors += [w2, b2, y3] if has2 else []
            tensors += [out, norm] if normalize else []
            ctx.save_for_backward(*tensors)
        return result

    @staticmethod
    def backward(ctx, d_out):
        lib = L.lib()
        B, In, P1, D2, normalize, dropout_p, seed, has2 = ctx.meta
        s = list(ctx.saved_tensors)
        v_lo, v_hi, w1, b1, ln_w, ln_b, y1, stats, y2 = s[:9]
        s = s[9:]
        w2 = b2 = y3 = out = norm = None
        if has2:
            w2, b2, y3 = s[:3]; s = s[3:]
        if normalize:
            out, norm = s[:2]
        dev = v_lo.device
        g = L.HeadGrad()
        a = g.fwd
        a.B, a.In, a.P1, a.D2, a.normalize, a.dropout_p, a.seed, a.offset = B, In, P1, D2, int(normalize), dropout_p, seed, 0
        a.v_lo, a.v_hi = L.ptr(v_lo), L.ptr(v_hi)
        a.w1, a.b1, a.ln_w, a.ln_b, a.w2, a.b2 = (L.ptr(t) for t in (w1, b1, ln_w, ln_b, w2, b2))
        a.y1, a.stats, a.y2, a.y3, a.out, a.norm = (L.ptr(t) for t in (y1, stats, y2, y3, out, norm))
        d_out = L.f32c(d_out)
        e = torch.empty_like
        d_v_lo, d_v_hi, d_w1, d_b1, d_ln_w, d_ln_b = e(v_lo), e(v_hi), e(w1), e(b1), e(ln_w), e(ln_b)
        d_w2, d_b2 = (e(w2), e(b2)) if has2 else (None, None)
        g.d_out, g.d_v_lo, g.d_v_hi = L.ptr(d_out), L.ptr(d_v_lo), L.ptr(d_v_hi)
        g.d_w1, g.d_b1, g.d_ln_w, g.d_ln_b, g.d_w2, g.d_b2 = (L.ptr(t) for t in (d_w1, d_b1, d_ln_w, d_ln_b, d_w2, d_b2))
        ws = L.workspace(lib.tt_head_bwd_workspace(C.byref(g)), dev)
        g.workspace, g.workspace_bytes = ws.data_ptr(), ws.numel()
        L.run(dev, lib.tt_head_bwd, "tt_head_bwd", C.byref(g))
        return d_v_lo, d_v_hi, d_w1, d_b1, d_ln_w, d_ln_b, d_w2, d_b2, None, None, None


def head(v_lo, v_hi, w1, b1, ln_w, ln_b, w2=None, b2=None, normalize=False, dropout_p=0.0):
    ts = [t for t in (v_lo, v_hi, w1, b1, ln_w, ln_b, w2, b2) if t is not None]
    need_grad = torch.is_grad_enabled() and any(t.requires_grad for t in ts)
    return HeadFunction.apply(v_lo, v_hi, w1, b1, ln_w, ln_b, w2, b2, normalize, dropout_p, need_grad)


# =============================================================================== L2 normalise
class L2NormFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, eps):
        lib = L.lib()
        x = L.f32c(x)
        B, D = x.shape
        y, norm = torch.empty_like(x), torch.empty(B, device=x.device, dtype=torch.float32)
        L.run(x.device, lib.tt_l2norm_fwd, "tt_l2norm_fwd", L.ptr(x), B, D, eps, L.ptr(y), L.ptr(norm))
        ctx.eps = eps
        ctx.save_for_backward(y, norm)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = L.lib()
        y, norm = ctx.saved_tensors
        dy = L.f32c(dy)
        dx = torch.empty_like(y)
        L.run(y.device, lib.tt_l2norm_bwd, "tt_l2norm_bwd", L.ptr(y), L.ptr(norm), L.ptr(dy), y.shape[0], y.shape[1], ctx.eps, L.ptr(dx))
        return dx, None


def l2_normalize(x, eps=1e-12):
    """F.normalize(x, p=2, dim=1, eps) for [B,D]."""
    return L2NormFunction.apply(x, eps)


# =============================================================================== score + CE
class ScoreCEFunction(torch.autograd.Function):
    """sum_i CE(logits_i, pos_offset+i), logits = x y^T * inv_temp - margin*(j != pos)  (K-E)."""

    @staticmethod
    def forward(ctx, x, y, pos_offset, inv_temp, margin):
        lib = L.lib()
        x, y = L.f32c(x), L.f32c(y)
        Bl, D = x.shape
        Bg = y.shape[0]
        dev = x.device
        row_loss = torch.empty(Bl, device=dev, dtype=torch.float32)
        row_lse = torch.empty(Bl, device=dev, dtype=torch.float32)
        ws = L.workspace(lib.tt_score_ce_workspace(Bl, Bg, D), dev)
        L.run(dev, lib.tt_score_ce_fwd, "tt_score_ce_fwd", L.ptr(x), Bl, L.ptr(y), Bg, D, pos_offset, inv_temp, margin, L.ptr(row_loss),
                                    L.ptr(row_lse), ws.data_ptr(), ws.numel())
        ctx.args = (pos_offset, inv_temp, margin)
        ctx.save_for_backward(x, y, row_lse)
        ctx.mark_non_differentiable(row_lse)
        return row_loss.sum(), row_lse

    @staticmethod
    def backward(ctx, g, _unused):
        lib = L.lib()
        x, y, row_lse = ctx.saved_tensors
        pos_offset, inv_temp, margin = ctx.args
        Bl, D = x.shape
        Bg = y.shape[0]
        dx, dy = torch.empty_like(x), torch.empty_like(y)
        ws = L.workspace(lib.tt_score_ce_workspace(Bl, Bg, D), x.device)
        # the upstream gradient stays on the device (float(g) would drain the launch queue at the start of every backward)
        g = L.f32c(g)
        L.run(x.device, lib.tt_score_ce_bwd_dev, "tt_score_ce_bwd_dev", L.ptr(x), Bl, L.ptr(y), Bg, D, pos_offset, inv_temp, margin, L.ptr(row_lse),
                                        1.0, L.ptr(g), L.ptr(dx), L.ptr(dy), ws.data_ptr(), ws.numel())
        return dx, dy, None, None, None


def score_ce_sum(x, y, pos_offset=0, inv_temp=1.0, margin=0.0):
    """Sum over local rows of the in-batch cross-entropy; divide by the GLOBAL batch for the reference's mean."""
    return ScoreCEFunction.apply(x, y, int(pos_offset), float(inv_temp), float(margin))[0]


# =============================================================================== hinge (explicit negatives)
class HingeFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, pos, neg, margin):
        lib = L.lib()
        q, pos, neg = L.f32c(q), L.f32c(pos), L.f32c(neg)
        B, D = q.shape
        k = neg.shape[0] // B
        row = torch.empty(B, device=q.device, dtype=torch.float32)
        L.run(q.device, lib.tt_hinge_fwd, "tt_hinge_fwd", L.ptr(q), L.ptr(pos), L.ptr(neg), B, k, D, margin, L.ptr(row))
        ctx.margin, ctx.k = margin, k
        ctx.save_for_backward(q, pos, neg)
        return row.sum()

    @staticmethod
    def backward(ctx, g):
        lib = L.lib()
        q, pos, neg = ctx.saved_tensors
        B, D = q.shape
        dq, dpos, dneg = torch.empty_like(q), torch.empty_like(pos), torch.empty_like(neg)
        g = L.f32c(g)
        L.run(q.device, lib.tt_hinge_bwd_dev, "tt_hinge_bwd_dev", L.ptr(q), L.ptr(pos), L.ptr(neg), B, ctx.k, D, ctx.margin, 1.0, L.ptr(g), L.ptr(dq),
                                     L.ptr(dpos), L.ptr(dneg))
        return dq, dpos, dneg, None


def hinge_sum(q, pos, neg, margin=0.2):
    return HingeFunction.apply(q, pos, neg, float(margin))


# =============================================================================== top-k / retrieval (no grad)
@torch.no_grad()
def rows_topk(x, y, k, cosine=True, mask_pos_offset=-1, index_base=0):
    """Exact-fp32 row-wise top-k of x y^T (or cosine); lowest-index tie-break.  Returns (values [Q,k], idx int64 [Q,k])."""
    lib = L.lib()
    x, y = L.f32c(x), L.f32c(y)
    Q, D = x.shape
    N = y.shape[0]
    mode = 1 if cosine else 0
    val = torch.empty(Q, k, device=x.device, dtype=torch.float32)
    idx = torch.empty(Q, k, device=x.device, dtype=torch.int64)
    ws = L.workspace(lib.tt_rows_topk_workspace(Q, N, D, mode), x.device)
    L.run(x.device, lib.tt_rows_topk, "tt_rows_topk", L.ptr(x), Q, L.ptr(y), N, D, k, mode, int(mask_pos_offset), int(index_base), L.ptr(val),
                             L.ptr(idx), ws.data_ptr(), ws.numel())
    return val, idx


@torch.no_grad()
def corpus_index_build(y, cosine=True):
    """Per-corpus part of rows_topk (row norms + hi/lo planes), built once: returns an opaque uint8 device buffer."""
    lib = L.lib()
    if not y.is_cuda:
        raise RuntimeError("two_towers_b200: corpus index needs a CUDA tensor (no CPU fallback)")
    N, D = y.shape
    buf = torch.empty(lib.tt_corpus_index_bytes(N, D), dtype=torch.uint8, device=y.device)
    L.run(y.device, lib.tt_corpus_index_build, "tt_corpus_index_build", L.ptr(y), N, D, 1 if cosine else 0, buf.data_ptr(), buf.numel())
    return buf


@torch.no_grad()
def rows_topk_indexed(x, y, index, k, cosine=True, mask_pos_offset=-1, index_base=0, return_fallback_rows=False):
    """rows_topk against a corpus whose index (corpus_index_build(y, cosine)) already exists; bit-identical results."""
    lib = L.lib()
    x = L.f32c(x)
    Q, D = x.shape
    N = y.shape[0]
    mode = 1 if cosine else 0
    val = torch.empty(Q, k, device=x.device, dtype=torch.float32)
    idx = torch.empty(Q, k, device=x.device, dtype=torch.int64)
    ws = L.workspace(lib.tt_rows_topk_indexed_workspace(Q, N, D, mode), x.device)
    L.run(x.device, lib.tt_rows_topk_indexed, "tt_rows_topk_indexed", L.ptr(x), Q, L.ptr(y), N, D, k, mode, int(mask_pos_offset), int(index_base),
                                     L.ptr(val), L.ptr(idx), index.data_ptr(), index.numel(), ws.data_ptr(), ws.numel())
    if return_fallback_rows:
        off = lib.tt_debug_topk_fallback_flags(Q, N, D, mode, k, 1)
        flags = ws[off:off + 4 * Q].view(torch.int32).clone() if off >= 0 else torch.zeros(Q, dtype=torch.int32, device=x.device)
        return val, idx, flags
    return val, idx


@torch.no_grad()
def topk_merge(vals, idx):
    """vals/idx [S,Q,k] per-shard lists -> global (vals, idx) [Q,k]."""
    lib = L.lib()
    vals, idx = L.f32c(vals), idx.contiguous()
    S, Q, k = vals.shape
    ov = torch.empty(Q, k, device=vals.device, dtype=torch.float32)
    oi = torch.empty(Q, k, device=vals.device, dtype=torch.int64)
    L.run(vals.device, lib.tt_topk_merge, "tt_topk_merge", L.ptr(vals), L.ptr(idx), S, Q, k, L.ptr(ov), L.ptr(oi))
    return ov, oi


@torch.no_grad()
def mrr_ranks(topk_idx, relevant):
    """first_rank int32 [Q] (0 = not found), rr fp32 [Q]."""
    lib = L.lib()
    topk_idx, relevant = topk_idx.contiguous(), relevant.to(torch.int64).contiguous()
    Q, k = topk_idx.shape
    fr = torch.empty(Q, device=topk_idx.device, dtype=torch.int32)
    rr = torch.empty(Q, device=topk_idx.device, dtype=torch.float32)
    L.run(topk_idx.device, lib.tt_mrr_ranks, "tt_mrr_ranks", L.ptr(topk_idx), L.ptr(relevant), Q, k, L.ptr(fr), L.ptr(rr))
    return fr, rr


@torch.no_grad()
def embed_gather(ids, table, pad_id=-1):
    lib = L.lib()
    ids = ids.contiguous()
    table = L.f32c(table)
    out = torch.empty(*ids.shape, table.shape[1], device=table.device, dtype=torch.float32)
    L.run(table.device, lib.tt_embed_gather, "tt_embed_gather", L.ptr(ids), 1 if ids.dtype == torch.int64 else 0, ids.numel(), L.ptr(table),
                                table.shape[0], table.shape[1], int(pad_id), L.ptr(out))
    return out


@torch.no_grad()
def gemm_nt(a, b, bias=None, alpha=1.0):
    """C = alpha * a @ b.T + bias (exact fp32 engine) - exported for kernel-level tests."""
    lib = L.lib()
    a, b = L.f32c(a), L.f32c(b)
    M, K = a.shape
    N = b.shape[0]
    c = torch.empty(M, N, device=a.device, dtype=torch.float32)
    L.run(a.device, lib.tt_gemm_nt, "tt_gemm_nt", L.ptr(a), K, L.ptr(b), K, L.ptr(c), N, L.ptr(bias), M, N, K, alpha, 0.0)
    return c


# ------------------------------------------------------------------------------------------- arithmetic engine
_ENGINES = {"simt-fp32": 0, "tcgen05-bf16x3": 1, "tcgen05-f16x3-fwd/f16-bwd": 2}


def engine():
    """Name of the arithmetic engine the library is using (`tt_engine`)."""
    return L.lib().tt_engine().decode()


def set_engine(mode):
    """Process-wide choice of the arithmetic engine (`tt_set_engine`):
      2  default - tcgen05 GEMMs on fp16 hi/lo planes, three passes forward, ONE fp16 pass on loss-scaled gradients backward.
         This is the training-precision deviation from the fp32 reference: gradients agree to ~4e-4..8e-4 relative (north-star
         bound 1e-3), and values beyond the fp16 range (|x| >= 65504) overflow - FusedAdam(skip_nonfinite=True) then skips
         the step instead of corrupting the weights;
      1  tcgen05 bf16 x3 in both directions (~fp32 accuracy, bf16 range, several times slower: per-timestep kernels);
      0  exact fp32 SIMT kernels (the on-device oracle of the parity tests; ~17x slower).
    Accepts the number or the engine's name; returns the previous engine's number."""
    prev = _ENGINES[engine()]
    m = _ENGINES[mode] if isinstance(mode, str) else int(mode)
    L.check(L.lib().tt_set_engine(m), "tt_set_engine")
    return prev


class use_engine:
    """`with use_engine(0): ...` - run a block on another engine (e.g. re-run a step that produced non-finite gradients on the
    exact engine) and restore the previous one."""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        self.prev = set_engine(self.mode)
        return self

    def __exit__(self, *exc):
        set_engine(self.prev)
