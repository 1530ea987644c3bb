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
OPT_NORM, OPT_STEP, OPT_CLIP, OPT_SKIP_NONFINITE = 1, 2, 4, 8


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
                ("norm_out", C.c_void_p), ("workspace", C.c_void_p), ("workspace_bytes", C.c_uint64),
                ("step_dev", C.c_void_p), ("consts_dev", C.c_void_p)]


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

    def __init__(self, entries, device, pinned=None):
        tb, cb, self.nchunks = build_tables(entries)
        self.ntensors = len(entries)
        if pinned is None:
            self._host = (tb.pin_memory(), cb.pin_memory())          # kept alive with the table
        else:                        # preallocated pinned staging (stream capture: no cudaHostAlloc while capturing)
            if pinned[0].numel() < tb.numel() or pinned[1].numel() < cb.numel():
                raise RuntimeError("FusedAdam: the parameter set grew between the eager warm-up step and the capture")
            pinned[0][:tb.numel()].copy_(tb)
            pinned[1][:cb.numel()].copy_(cb)
            self._host = (pinned[0][:tb.numel()], pinned[1][:cb.numel()])
        self.tensors = self._host[0].to(device, non_blocking=True)
        self.chunks = self._host[1].to(device, non_blocking=True)
        self.ws = L.workspace(L.lib().tt_opt_step_workspace(self.nchunks), device)


class FusedAdam(torch.optim.Optimizer):
    """Adam (decoupled=False; weight decay folded into the gradient) or AdamW (decoupled=True) with optional fused
    clip_grad_norm_: `max_grad_norm=1.0` reproduces `clip_grad_norm_(model.parameters(), 1.0); optimizer.step()`
    (the total norm is taken over every parameter of every group, and the gradients are left clipped in place).
    After step(), `last_grad_norm` is a device scalar with the pre-clip total norm (what clip_grad_norm_ returns).

    `skip_nonfinite=True` (default): the total gradient norm is always formed, and a step whose norm is inf / NaN leaves
    parameters and moments untouched and counts in `skipped_steps` (device scalar) - the library's BPTT runs its GEMMs
    as ONE fp16 pass on loss-scaled gradients, so an overflow there must not turn into NaN weights (GradScaler semantics).
    `capturable=True`: the step count lives on the device and the bias corrections are derived there, so `step()` can be
    captured in a CUDA graph and replayed (`state[p]['step']` is then a device scalar, as with torch's capturable Adam)."""

    _decoupled = False

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_grad_norm=None,
                 skip_nonfinite=True, capturable=False):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError("invalid Adam hyper-parameters")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                        foreach=None, capturable=bool(capturable), differentiable=False, fused=None,
                        decoupled_weight_decay=self._decoupled)
        super().__init__(params, defaults)
        self.max_grad_norm = max_grad_norm
        self.skip_nonfinite = bool(skip_nonfinite)
        self.capturable = bool(capturable)
        self.last_grad_norm = None
        self._tables = {}
        self._norm = None
        self._dev_step = {}
        self._spare = {}
        self._spare_size = {}
        self._captured_tables = []

    @property
    def skipped_steps(self):
        """Device scalar: steps skipped because the gradient norm was not finite (None before the first step)."""
        return None if self._norm is None else self._norm[2]

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
            if len(st) and st["step"].is_cuda and not self.capturable:   # state loaded from a fused=True / capturable torch optimizer
                st["step"] = st["step"].detach().float().cpu()
            if len(st) == 0:
                st["step"] = torch.tensor(0.0, dtype=torch.float32)
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            out.append((p, st))
        return out

    @torch.no_grad()
    def init_state(self):
        """Allocate the optimizer state of every parameter now (zeros, step 0) instead of lazily inside the first step():
        state created while a CUDA graph is being captured would be re-zeroed by every replay."""
        for gi, g in enumerate(self.param_groups):
            ps = [p for p in g["params"] if p.requires_grad]
            for p in ps:
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            if self.capturable and ps and gi not in self._dev_step:
                dev = ps[0].device
                first = self.state[ps[0]]["step"]
                self._dev_step[gi] = (first.detach().float().to(dev).reshape(()).clone(),
                                      torch.zeros(2, dtype=torch.float32, device=dev))
                for p in ps:
                    self.state[p]["step"] = self._dev_step[gi][0]

    def reserve_capture_staging(self):
        """One more pinned staging pair per table key seen so far: a capture consumes one (cudaHostAlloc is not allowed while
        a stream is capturing, and every captured graph must keep its OWN staging - its upload node re-reads it on replay)."""
        for key, sizes in self._spare_size.items():
            self._spare.setdefault(key, []).append(tuple(torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes))

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
            if torch.cuda.is_current_stream_capturing():
                spare = self._spare.get(key)
                if not spare:
                    raise RuntimeError("FusedAdam(capturable=True): call reserve_capture_staging() after one eager step() and "
                                       "before capturing a CUDA graph (it allocates the pinned memory the captured table "
                                       "upload reads on every replay)")
                tab = cache[sig] = _Table(list(sig), device, pinned=spare.pop())     # owned by this table from now on
                self._captured_tables.append(tab)          # never evicted: the graph's upload node reads its staging on replay
            else:
                tab = cache[sig] = _Table(list(sig), device)
                self._spare_size[key] = tuple(h.numel() for h in tab._host)
        return tab

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = L.lib()
        groups = [(gi, g, self._entries(g)) for gi, g in enumerate(self.param_groups)]
        groups = [(gi, g, e) for gi, g, e in groups if e]
        if not groups:
            return loss
        device = groups[0][2][0][0].device
        if any(p.device != device for _, _, e in groups for p, _ in e):
            raise RuntimeError("FusedAdam: all parameters must live on one device")
        clip = self.max_grad_norm is not None and self.max_grad_norm > 0
        skip = self.skip_nonfinite
        need_norm = clip or skip
        if self._norm is None or self._norm.device != device:
            self._norm = torch.zeros(4, dtype=torch.float32, device=device)
        stream = L.stream(device)
        with torch.cuda.device(device):
            if need_norm and len(groups) > 1:      # one norm over every group first
                every = [x for _, _, e in groups for x in e]
                tab = self._table("all", every, device)
                a = OptStep(tab.tensors.data_ptr(), tab.chunks.data_ptr(), tab.ntensors, tab.nchunks,
                            OPT_NORM | (OPT_SKIP_NONFINITE if skip else 0), 0,
                            0.0, 0.0, 0.0, 0.0, 0.0, float(self.max_grad_norm) if clip else 0.0, 1.0, 1.0,
                            self._norm.data_ptr(), tab.ws.data_ptr(), tab.ws.numel(), None, None)
                L.check(lib.tt_opt_step(C.byref(a), stream), "tt_opt_step(norm)")
            for gi, g, entries in groups:
                b1, b2 = g["betas"]
                step_ptr = consts_ptr = None
                if self.capturable:
                    # ONE device scalar per group, shared by its parameters' state entries; never read on the host here
                    ds = self._dev_step.get(gi)
                    if ds is None:
                        if torch.cuda.is_current_stream_capturing():
                            raise RuntimeError("FusedAdam(capturable=True): call init_state() (or one eager step) before capture")
                        first = entries[0][1]["step"]
                        ds = self._dev_step[gi] = (first.detach().float().to(device).reshape(()).clone(),
                                                   torch.zeros(2, dtype=torch.float32, device=device))
                    for _, st in entries:
                        st["step"] = ds[0]
                    step_ptr, consts_ptr, t = ds[0].data_ptr(), ds[1].data_ptr(), 1.0
                else:
                    steps = {float(st["step"]) for _, st in entries}
                    if len(steps) != 1:
                        # parameters that joined later (grad appeared after step 0) would need their own bias correction
                        raise RuntimeError("FusedAdam: parameters of one group must share the step count")
                    t = steps.pop() + 1.0
                    for _, st in entries:
                        st["step"] += 1
                flags = OPT_STEP | (OPT_CLIP if clip else 0) | (OPT_NORM if need_norm and len(groups) == 1 else 0) | \
                    (OPT_SKIP_NONFINITE if skip else 0)
                tab = self._table(gi, entries, device)
                a = OptStep(tab.tensors.data_ptr(), tab.chunks.data_ptr(), tab.ntensors, tab.nchunks, flags,
                            1 if g.get("decoupled_weight_decay", self._decoupled) else 0,
                            float(g["lr"]), float(b1), float(b2), float(g["eps"]), float(g["weight_decay"]),
                            float(self.max_grad_norm) if clip else 0.0,
                            1.0 - b1 ** t, math.sqrt(1.0 - b2 ** t),
                            self._norm.data_ptr(), tab.ws.data_ptr(), tab.ws.numel(), step_ptr, consts_ptr)
                L.check(lib.tt_opt_step(C.byref(a), stream), "tt_opt_step")
        self.last_grad_norm = self._norm[0] if need_norm else None
        return loss


class FusedAdamW(FusedAdam):
    _decoupled = True

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, max_grad_norm=None,
                 skip_nonfinite=True, capturable=False):
        super().__init__(params, lr, betas, eps, weight_decay, max_grad_norm, skip_nonfinite, capturable)


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
