"""Fused gradient clipping + Adam / AdamW and reference-format checkpoints.

Reference call sites: `optim.AdamW(model.parameters(), lr=0.0001, weight_decay=0.01)` + `clip_grad_norm_(.., 1.0)`
(simple_two_tower.py:193,239), `torch.optim.Adam(.., lr)` + clip (train_margin.py:64-67), `torch.optim.Adam(..)`
(train_enhanced.py:43); checkpoints `{'epoch', 'model_state_dict', 'optimizer_state_dict', 'val_loss' | 'loss'}`
(simple_two_tower.py:263-268, train_margin.py:134-139) or a bare state_dict (train_enhanced.py:76).

`FusedAdam` / `FusedAdamW` keep torch's optimizer state layout (`step`, `exp_avg`, `exp_avg_sq`; the same param_group
keys), so `optimizer.state_dict()` is interchangeable with the reference's `optimizer_state_dict` in both directions.
The update itself is `tt_opt_step` (csrc/optim.cu): total-norm reduction, clip coefficient and the Adam update as
three launches over ALL parameters with no host synchronisation.  No CPU path: CPU parameters raise.
"""
import ctypes as C
import math

import torch

from . import _lib as L

OPT_CHUNK = 65536
OPT_NORM, OPT_STEP, OPT_CLIP = 1, 2, 4


class OptTensor(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("numel", C.c_int64)]


class OptChunk(C.Structure):
    _fields_ = [("tensor", C.c_int32), ("reserved", C.c_int32), ("offset", C.c_int64)]


class OptStep(C.Structure):
    _fields_ = [("tensors", C.c_void_p), ("chunks", C.c_void_p),
                ("ntensors", C.c_int32), ("nchunks", C.c_int32), ("flags", C.c_int32), ("decoupled_wd", C.c_int32),
                ("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double),
                ("weight_decay", C.c_double), ("max_norm", C.c_double),
                ("bias_correction1", C.c_double), ("bias_correction2_sqrt", C.c_double),
                ("norm_out", C.c_void_p), ("workspace", C.c_void_p), ("workspace_bytes", C.c_uint64)]


def build_tables(entries):
    """entries: list of (param_ptr, grad_ptr, exp_avg_ptr, exp_avg_sq_ptr, numel).  Returns the two host tables as
    uint8 CPU tensors (tt_opt_tensor[n], tt_opt_chunk[m]) and m.  Pure host logic (tested without a GPU)."""
    n = len(entries)
    tens = (OptTensor * max(n, 1))()
    chunks = []
    for i, (p, g, m, v, numel) in enumerate(entries):
        tens[i] = OptTensor(p, g, m, v, numel)
        for off in range(0, numel, OPT_CHUNK):
            chunks.append((i, off))
    ck = (OptChunk * max(len(chunks), 1))()
    for j, (i, off) in enumerate(chunks):
        ck[j] = OptChunk(i, 0, off)
    tb = torch.frombuffer(bytearray(bytes(tens)), dtype=torch.uint8)[: n * C.sizeof(OptTensor)].clone()
    cb = torch.frombuffer(bytearray(bytes(ck)), dtype=torch.uint8)[: len(chunks) * C.sizeof(OptChunk)].clone()
    return tb, cb, len(chunks)


class _Table:
    """Device copies of the two tables.  Uploaded from PINNED memory with non-blocking copies: a pageable cudaMemcpy
    would block the host until everything queued before it (the whole backward pass) has run."""

    def __init__(self, entries, device):
        tb, cb, self.nchunks = build_tables(entries)
        self.ntensors = len(entries)
        self._host = (tb.pin_memory(), cb.pin_memory())          # kept alive with the table
        self.tensors = self._host[0].to(device, non_blocking=True)
        self.chunks = self._host[1].to(device, non_blocking=True)
        self.ws = L.workspace(L.lib().tt_opt_step_workspace(self.nchunks), device)


class FusedAdam(torch.optim.Optimizer):
    """Adam (decoupled=False; weight decay folded into the gradient) or AdamW (decoupled=True) with optional fused
    clip_grad_norm_: `max_grad_norm=1.0` reproduces `clip_grad_norm_(model.parameters(), 1.0); optimizer.step()`
    (the total norm is taken over every parameter of every group, and the gradients are left clipped in place).
    After step(), `last_grad_norm` is a device scalar with the pre-clip total norm (what clip_grad_norm_ returns)."""

    _decoupled = False

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_grad_norm=None):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError("invalid Adam hyper-parameters")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                        foreach=None, capturable=False, differentiable=False, fused=None,
                        decoupled_weight_decay=self._decoupled)
        super().__init__(params, defaults)
        self.max_grad_norm = max_grad_norm
        self.last_grad_norm = None
        self._tables = {}
        self._norm = None

    def _entries(self, group):
        out = []
        for p in group["params"]:
            if p.grad is None:
                continue
            if not p.is_cuda:
                raise RuntimeError("two_towers_b200.FusedAdam needs CUDA parameters (no CPU fallback)")
            if p.dtype != torch.float32 or not p.is_contiguous() or p.grad.dtype != torch.float32 or p.grad.is_sparse:
                raise RuntimeError("FusedAdam: parameters and gradients must be dense contiguous fp32")
            if not p.grad.is_contiguous():
                p.grad = p.grad.contiguous()
            st = self.state[p]
            if len(st) and st["step"].is_cuda:          # state loaded from a fused=True / capturable torch optimizer
                st["step"] = st["step"].detach().float().cpu()
            if len(st) == 0:
                st["step"] = torch.tensor(0.0, dtype=torch.float32)
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            out.append((p, st))
        return out

    def _table(self, key, entries, device):
        sig = tuple((p.data_ptr(), p.grad.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel())
                    for p, st in entries)
        # gradients are fresh allocations after zero_grad(set_to_none=True); the caching allocator hands out a small
        # set of recurring addresses, so a few tables per group are kept instead of rebuilding one every step
        cache = self._tables.setdefault(key, {})
        tab = cache.get(sig)
        if tab is None:
            if len(cache) >= 8:
                cache.pop(next(iter(cache)))
            tab = cache[sig] = _Table(list(sig), device)
        return tab

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = L.lib()
        groups = [(g, self._entries(g)) for g in self.param_groups]
        groups = [(g, e) for g, e in groups if e]
        if not groups:
            return loss
        device = groups[0][1][0][0].device
        if any(p.device != device for _, e in groups for p, _ in e):
            raise RuntimeError("FusedAdam: all parameters must live on one device")
        clip = self.max_grad_norm is not None and self.max_grad_norm > 0
        if self._norm is None or self._norm.device != device:
            self._norm = torch.zeros(2, dtype=torch.float32, device=device)
        stream = L.stream(device)
        with torch.cuda.device(device):
            if clip and len(groups) > 1:           # one norm over every group first
                every = [x for _, e in groups for x in e]
                tab = self._table("all", every, device)
                a = OptStep(tab.tensors.data_ptr(), tab.chunks.data_ptr(), tab.ntensors, tab.nchunks, OPT_NORM, 0,
                            0.0, 0.0, 0.0, 0.0, 0.0, float(self.max_grad_norm), 1.0, 1.0,
                            self._norm.data_ptr(), tab.ws.data_ptr(), tab.ws.numel())
                L.check(lib.tt_opt_step(C.byref(a), stream), "tt_opt_step(norm)")
            for gi, (g, entries) in enumerate(groups):
                steps = {float(st["step"]) for _, st in entries}
                if len(steps) != 1:
                    # parameters that joined later (grad appeared after step 0) would need their own bias correction
                    raise RuntimeError("FusedAdam: parameters of one group must share the step count")
                t = steps.pop() + 1.0
                for _, st in entries:
                    st["step"] += 1
                b1, b2 = g["betas"]
                flags = OPT_STEP | (OPT_CLIP if clip else 0) | (OPT_NORM if clip and len(groups) == 1 else 0)
                tab = self._table(gi, entries, device)
                a = OptStep(tab.tensors.data_ptr(), tab.chunks.data_ptr(), tab.ntensors, tab.nchunks, flags,
                            1 if g.get("decoupled_weight_decay", self._decoupled) else 0,
                            float(g["lr"]), float(b1), float(b2), float(g["eps"]), float(g["weight_decay"]),
                            float(self.max_grad_norm) if clip else 0.0,
                            1.0 - b1 ** t, math.sqrt(1.0 - b2 ** t),
                            self._norm.data_ptr(), tab.ws.data_ptr(), tab.ws.numel())
                L.check(lib.tt_opt_step(C.byref(a), stream), "tt_opt_step")
        self.last_grad_norm = self._norm[0] if clip else None
        return loss


class FusedAdamW(FusedAdam):
    _decoupled = True

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, max_grad_norm=None):
        super().__init__(params, lr, betas, eps, weight_decay, max_grad_norm)


# ------------------------------------------------------------------------------------------- checkpoints
def save_checkpoint(path, model, optimizer=None, epoch=0, **metrics):
    """The reference's checkpoint dictionary (simple_two_tower.py:263-268 uses val_loss=, train_margin.py:134-139 loss=).
    With optimizer=None writes the bare state_dict form of train_enhanced.py:76."""
    if optimizer is None and not metrics and epoch == 0:
        torch.save(model.state_dict(), path)
        return
    ck = {"epoch": epoch, "model_state_dict": model.state_dict()}
    if optimizer is not None:
        ck["optimizer_state_dict"] = optimizer.state_dict()
    ck.update(metrics)
    torch.save(ck, path)


def load_checkpoint(path, model, optimizer=None, map_location=None):
    """Accepts both reference forms (dict with 'model_state_dict' - compare_models.py:15-16, validate_margin.py:76-77,
    app.py:24-25 - or a bare state_dict).  Returns the checkpoint dict (or {} for the bare form)."""
    ck = torch.load(path, map_location=map_location)
    if isinstance(ck, dict) and "model_state_dict" in ck:
        model.load_state_dict(ck["model_state_dict"])
        if optimizer is not None and "optimizer_state_dict" in ck:
            optimizer.load_state_dict(ck["optimizer_state_dict"])
        return ck
    model.load_state_dict(ck)
    return {}
