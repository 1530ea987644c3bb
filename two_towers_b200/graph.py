"""Whole-step CUDA-graph capture.

The reference's train step (simple_two_tower.py:229-243; variants train_margin.py:111-123, train_enhanced.py:54-66) is
`zero_grad -> forward -> loss -> backward -> [clip] -> optimizer.step()`.  Through this package that body is ~180 kernel
launches queued from Python/ctypes - 12-20 ms of host time against a ~20 ms device step, and far more than the device time
for the small configurations (C1: B = 256, T = 32).  `GraphedTrainStep` runs the same body ONCE under stream capture per
input shape and afterwards replays it with a single `cudaGraphLaunch`: the host cost of a step becomes two small
host->device copies plus one launch.

What makes the body capturable (all of it lives in the library / this package, nothing here is torch arithmetic):
  * no host synchronisation anywhere in the step (upstream loss gradient, clip coefficient and loss scale stay on the device);
  * `FusedAdam(capturable=True)`: step count and bias corrections on the device, tables uploaded from preallocated pinned memory;
  * dropout seeds: the kernels add a device scalar to the host seed (`functional.device_seed`), bumped once per replay, so
    every replay draws fresh masks and forward / backward of one replay agree;
  * collectives (`distributed.py`) are NCCL calls on streams forked from the capture stream - they become graph nodes.

Inputs are copied into static device buffers before each replay; the returned loss (and mined hard negatives) are static
device tensors that the NEXT replay overwrites - read or copy them before stepping again.
"""
import torch

from . import _lib as L
from . import functional as Fn
from .losses import get_hard_negatives_batch


class _Captured:
    __slots__ = ("graph", "q", "d", "loss", "hard", "launches")


class GraphedTrainStep:
    """step = GraphedTrainStep(model, loss_fn, optimizer[, reducer, mine_k]);  loss, hard = step(q, d)

    model      a two_towers_b200 model (train mode), called as model(q, d) -> (query_vec, doc_vec)
    loss_fn    called as loss_fn(query_vec, doc_vec)
    optimizer  FusedAdam / FusedAdamW built with capturable=True
    reducer    optional distributed.GradientAllReducer (its finish() runs between backward and the optimizer)
    mine_k     > 0: also returns get_hard_negatives over the in-batch docs for every query (enhanced_two_tower.py:123-133)
    pos_offset / docs_fn: for data parallelism the mining runs over the all-gathered docs (docs_fn(doc_vec) -> global docs)

    The first call with a new (shape, dtype) pair runs `warmup` eager steps on a side stream (allocator and library
    one-time state), RESTORES parameters and optimizer state to what they were, captures the step and replays it - so the
    sequence of parameter values equals the eager loop's, step for step.
    """

    def __init__(self, model, loss_fn, optimizer, reducer=None, mine_k=0, pos_offset=0, docs_fn=None, warmup=2):
        if not getattr(optimizer, "capturable", False):
            raise ValueError("GraphedTrainStep needs FusedAdam / FusedAdamW(capturable=True)")
        self.model, self.loss_fn, self.optimizer, self.reducer = model, loss_fn, optimizer, reducer
        self.mine_k, self.pos_offset, self.docs_fn, self.warmup = int(mine_k), int(pos_offset), docs_fn, max(1, int(warmup))
        self._graphs = {}
        self._seed = None
        self._stream = None
        self.replays = 0

    # ---- the step body: exactly what an eager loop would run
    def _body(self, q, d):
        self.optimizer.zero_grad(set_to_none=True)
        qv, dv = self.model(q, d)
        loss = self.loss_fn(qv, dv)
        loss.backward()
        if self.reducer is not None:
            self.reducer.finish()
        self.optimizer.step()
        hard = None
        if self.mine_k > 0:
            docs = dv.detach() if self.docs_fn is None else self.docs_fn(dv.detach())
            hard = get_hard_negatives_batch(qv.detach(), docs, k=self.mine_k, pos_offset=self.pos_offset)
        return loss.detach(), hard

    def _snapshot(self):
        """Values of every parameter and optimizer-state tensor (the state exists: init_state() ran)."""
        self.optimizer.init_state()
        params = [p for g in self.optimizer.param_groups for p in g["params"]]
        snap = []
        for p in params:
            st = self.optimizer.state.get(p, {})
            snap.append((p.detach().clone(), {k: v.detach().clone() for k, v in st.items() if torch.is_tensor(v)}))
        return snap

    def _restore(self, snap):
        """In place: graphs captured earlier (other input shapes) keep pointing at these very tensors."""
        params = [p for g in self.optimizer.param_groups for p in g["params"]]
        with torch.no_grad():
            for p, (v, st) in zip(params, snap):
                p.copy_(v)
                cur = self.optimizer.state.get(p, {})
                for k, sv in st.items():
                    cur[k].copy_(sv)

    def _capture(self, q, d):
        dev = q.device
        if self._seed is None:
            self._seed = torch.zeros(1, dtype=torch.int64, device=dev)
            self._stream = torch.cuda.Stream(dev)
        c = _Captured()
        c.q, c.d = torch.empty_like(q), torch.empty_like(d)
        c.q.copy_(q)
        c.d.copy_(d)
        cur = torch.cuda.current_stream(dev)
        snap = self._snapshot()
        self._stream.wait_stream(cur)
        with torch.cuda.stream(self._stream), Fn.device_seed(self._seed):
            for _ in range(self.warmup):
                self._body(c.q, c.d)
        cur.wait_stream(self._stream)
        torch.cuda.synchronize(dev)
        self._restore(snap)
        for p in (p for g in self.optimizer.param_groups for p in g["params"]):
            p.grad = None
        torch.cuda.synchronize(dev)
        self.optimizer.reserve_capture_staging()
        lib = L.lib()
        c.graph = torch.cuda.CUDAGraph()
        n0 = lib.tt_kernel_launches()
        with Fn.device_seed(self._seed), torch.cuda.graph(c.graph, stream=self._stream):
            self._seed.add_(1)
            c.loss, c.hard = self._body(c.q, c.d)
        c.launches = int(lib.tt_kernel_launches() - n0)
        return c

    def __call__(self, q, d):
        key = (tuple(q.shape), q.dtype, tuple(d.shape), d.dtype, q.device)
        c = self._graphs.get(key)
        if c is None:
            c = self._graphs[key] = self._capture(q, d)
        c.q.copy_(q, non_blocking=True)
        c.d.copy_(d, non_blocking=True)
        c.graph.replay()
        self.replays += 1
        return c.loss, c.hard

    def launches_per_replay(self, q, d):
        """Library kernel launches inside one replay for this input shape (counted while capturing)."""
        key = (tuple(q.shape), q.dtype, tuple(d.shape), d.dtype, q.device)
        c = self._graphs.get(key)
        return c.launches if c is not None else None


class _TowersFn(torch.autograd.Function):
    """Autograd node of GraphedTowers: forward / backward = replay of the two captured graphs.  The parameters are inputs of
    the node, so their gradients flow through autograd's AccumulateGrad like any others (post-accumulate hooks - the
    gradient all-reducer's - fire as in the eager step)."""

    @staticmethod
    def forward(ctx, run, q, d, *params):
        run.q.copy_(q, non_blocking=True)
        run.d.copy_(d, non_blocking=True)
        run.fwd.replay()
        ctx.run = run
        return run.qv.detach(), run.dv.detach()

    @staticmethod
    def backward(ctx, gq, gd):
        run = ctx.run
        run.gq.copy_(gq)
        run.gd.copy_(gd)
        run.bwd.replay()
        return (None, None, None) + tuple(g.detach() for g in run.grads)


class _TowersRun:
    __slots__ = ("fwd", "bwd", "q", "d", "qv", "dv", "gq", "gd", "grads", "launches")


class GraphedTowers:
    """towers = GraphedTowers(model);  query_vec, doc_vec = towers(q, d)   - a drop-in for `model(q, d)` in a train step.

    Forward and backward of BOTH towers (embedding gather, GRU stacks, projection heads: ~150 of the ~180 kernel launches of
    a step) are captured as two CUDA graphs per input shape; the loss, the collectives, the gradient all-reduce and the
    optimizer stay eager.  This is the data-parallel form of GraphedTrainStep: NCCL collectives captured INSIDE a CUDA graph
    cost ~2.5 ms per step at 8 GPUs (measured: 25.4 vs 22.8 ms, profiles/r2_summary.md), so under data parallelism only the
    collective-free part of the step is graphed.  Dropout draws fresh masks on every call (device seed, bumped here).
    Train mode only; the model must keep returning a (query_vec, doc_vec) tuple.  The outputs and everything the backward
    graph reads are static buffers of the capture: run backward (or drop the outputs) before calling again with the same
    input shape - two forwards in flight for one shape would share them (gradient accumulation over micro-batches works:
    forward, backward, forward, backward)."""

    def __init__(self, model, warmup=2):
        self.model, self.warmup = model, max(1, int(warmup))
        self._runs = {}
        self._seed = None
        self._stream = None

    def _capture(self, q, d):
        dev = q.device
        if self._seed is None:
            self._seed = torch.zeros(1, dtype=torch.int64, device=dev)
            self._stream = torch.cuda.Stream(dev)
        params = [p for p in self.model.parameters() if p.requires_grad]
        r = _TowersRun()
        r.q, r.d = q.clone(), d.clone()
        cur = torch.cuda.current_stream(dev)
        self._stream.wait_stream(cur)
        with torch.cuda.stream(self._stream), Fn.device_seed(self._seed):
            for _ in range(self.warmup):                       # allocator + library one-time state, outside any capture
                qv, dv = self.model(r.q, r.d)
                torch.autograd.grad((qv, dv), params, (torch.ones_like(qv), torch.ones_like(dv)))
        cur.wait_stream(self._stream)
        torch.cuda.synchronize(dev)
        lib = L.lib()
        n0 = lib.tt_kernel_launches()
        r.fwd, r.bwd = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with Fn.device_seed(self._seed):
            with torch.cuda.graph(r.fwd, stream=self._stream):
                r.qv, r.dv = self.model(r.q, r.d)
            r.gq, r.gd = torch.zeros_like(r.qv), torch.zeros_like(r.dv)
            with torch.cuda.graph(r.bwd, pool=r.fwd.pool(), stream=self._stream):
                r.grads = torch.autograd.grad((r.qv, r.dv), params, (r.gq, r.gd))
        r.launches = int(lib.tt_kernel_launches() - n0)
        return r, params

    def __call__(self, q, d):
        if not self.model.training:
            raise RuntimeError("GraphedTowers captures the training step; use the model directly for inference")
        key = (tuple(q.shape), q.dtype, tuple(d.shape), d.dtype, q.device)
        ent = self._runs.get(key)
        if ent is None:
            ent = self._runs[key] = self._capture(q, d)
        run, params = ent
        self._seed.add_(1)
        return _TowersFn.apply(run, q, d, *params)

    def launches_per_replay(self, q, d):
        ent = self._runs.get((tuple(q.shape), q.dtype, tuple(d.shape), d.dtype, q.device))
        return ent[0].launches if ent is not None else None
