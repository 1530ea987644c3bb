"""Data parallelism for the two-tower step (one process per GPU, torch.distributed / NCCL over NVLink).

The reference is single-process; the oracle for everything here is the reference run once on the
concatenated global batch (enhanced_two_tower.py:90-101,123-133).  Only the in-batch score couples rows, so the batch
is sharded and there is exactly one exchange step:

  forward   all-gather of the [B_loc, D] tower outputs -> every rank scores its rows against ALL docs
  backward  reduce-scatter(SUM) of dL/d(gathered) back to the owning rank (autograd of the all-gather)
  grads     bucketed all-reduce(SUM) of the weight gradients on a communication stream, launched from grad hooks:
            the query tower's buckets (its BPTT is 4x shorter) travel under the doc tower's BPTT

Each rank's loss is  sum(local rows) / B_global ; the sum over ranks is the reference's mean over the global
batch, hence SUM (not mean) for the gradient all-reduce.
"""
import os

import torch
import torch.distributed as dist


def world_size(group=None):
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def rank(group=None):
    return dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0


class _AllGatherRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        x = x.contiguous()
        W = dist.get_world_size(group)
        out = torch.empty(W * x.shape[0], *x.shape[1:], dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        group = ctx.group
        W, r = dist.get_world_size(group), dist.get_rank(group)
        g = g.contiguous()
        n = g.shape[0] // W
        if g.is_cuda:
            dx = torch.empty(n, *g.shape[1:], dtype=g.dtype, device=g.device)
            dist.reduce_scatter_tensor(dx, g, op=dist.ReduceOp.SUM, group=group)
        else:                                   # gloo (CPU tests) has no reduce_scatter
            g = g.clone()
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
            dx = g[r * n:(r + 1) * n].clone()
        return dx, None


def all_gather_rows(x, group=None):
    """[B_loc, ...] -> [W*B_loc, ...] in rank order; backward = reduce-scatter(SUM)."""
    if world_size(group) == 1:
        return x
    return _AllGatherRows.apply(x, group)


class _SumOverRanks(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        y = x.detach().clone()
        dist.all_reduce(y, op=dist.ReduceOp.SUM, group=group)
        return y

    @staticmethod
    def backward(ctx, g):
        return g, None


def sum_over_ranks(x, group=None):
    """Value = sum over ranks (the global mean loss), gradient = identity on the local term."""
    if group is None or world_size(group) == 1:
        return x
    return _SumOverRanks.apply(x, group)


class GradientAllReducer:
    """Bucketed SUM all-reduce of parameter gradients, overlapped with the rest of backward.

    * Buckets follow the order autograd finishes parameters (reverse registration: heads, then the encoders) and
      never straddle a top-level submodule, so a bucket is complete as soon as ITS tower's backward has run; the
      default 4 MiB cap splits a tower into a few buckets that pipeline through NCCL.
    * Every bucket owns a persistent flat buffer.  When its last gradient arrives the gradients are packed into it
      with one `_foreach_copy_` and all-reduced in place on a dedicated communication stream; `finish()` then points
      each `p.grad` at its slice of the buffer (no `torch.cat`, no copy back - the optimizer reads the flat buffer).
    * Stream safety: the towers run on different streams (models.overlap_towers), so every hook records an event on
      the stream its gradient was produced on and the communication stream waits for all of a bucket's events
      before it reads a single gradient; `finish()` makes the caller's stream wait for the communication stream.
    * One backward per `finish()`: a hook firing for a bucket that is already in flight (gradient accumulation
      without `no_sync()`) raises instead of silently dropping the second micro-batch.  Inside `no_sync()` hooks
      do nothing and `finish()` reduces whatever has accumulated.
    """

    def __init__(self, params, group=None, bucket_bytes=None, named_params=None, overlap=None):
        # TT_DP_BUCKET_MB / TT_DP_OVERLAP: measurement knobs (bench.py); overlap=False launches every bucket from finish()
        if bucket_bytes is None:
            bucket_bytes = int(float(os.environ.get("TT_DP_BUCKET_MB", "4")) * (1 << 20))
        self.overlap = (os.environ.get("TT_DP_OVERLAP", "1") != "0") if overlap is None else bool(overlap)
        self.group = group
        if named_params is not None:
            named = [(n, p) for n, p in named_params if p.requires_grad]
        else:
            named = [("", p) for p in params if p.requires_grad]
        self.params = [p for _, p in named]
        self.buckets, cur, size, cur_mod = [], [], 0, None
        for name, p in reversed(named):
            mod = name.split(".")[0]
            if cur and (mod != cur_mod or size >= bucket_bytes):
                self.buckets.append(cur)
                cur, size = [], 0
            cur_mod = mod
            cur.append(p)
            size += p.numel() * p.element_size()
        if cur:
            self.buckets.append(cur)
        self._bucket_of = {id(p): i for i, b in enumerate(self.buckets) for p in b}
        self._flat = [None] * len(self.buckets)
        self._views = [None] * len(self.buckets)
        self._reset()
        self._sync = True
        self._comm = None
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]

    def _reset(self):
        self._pending = [len(b) for b in self.buckets]
        self._launched = [False] * len(self.buckets)
        self._events = [[] for _ in self.buckets]
        self._work = []

    def _buffers(self, i):
        ps = self.buckets[i]
        if self._flat[i] is None or self._flat[i].device != ps[0].device:
            n = sum(p.numel() for p in ps)
            self._flat[i] = torch.empty(n, dtype=ps[0].dtype, device=ps[0].device)
            views, off = [], 0
            for p in ps:
                views.append(self._flat[i][off:off + p.numel()].view_as(p))
                off += p.numel()
            self._views[i] = views
        return self._flat[i], self._views[i]

    def _launch(self, i):
        ps = self.buckets[i]
        flat, views = self._buffers(i)
        self._launched[i] = True
        if flat.is_cuda:
            if self._comm is None:
                self._comm = torch.cuda.Stream(flat.device)
            cur = torch.cuda.current_stream(flat.device)
            ev = torch.cuda.Event()
            ev.record(cur)                              # orders the comm stream after the buffer's previous consumers
            self._comm.wait_event(ev)
            for e in self._events[i]:
                self._comm.wait_event(e)
            with torch.cuda.stream(self._comm):
                torch._foreach_copy_(views, [p.grad for p in ps])
                work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            for p in ps:
                p.grad.record_stream(self._comm)
        else:
            torch._foreach_copy_(views, [p.grad for p in ps])
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self._work.append((work, i))

    def _on_grad(self, p):
        if not self._sync or world_size(self.group) == 1:
            return
        i = self._bucket_of[id(p)]
        if self._launched[i]:
            raise RuntimeError("GradientAllReducer: a gradient arrived for a bucket that is already being reduced "
                               "(two backward() calls per finish()); wrap all but the last micro-batch in no_sync()")
        if p.grad.is_cuda:
            e = torch.cuda.Event()
            e.record(torch.cuda.current_stream(p.grad.device))
            self._events[i].append(e)
        self._pending[i] -= 1
        if self._pending[i] == 0 and self.overlap:
            self._launch(i)

    class _NoSync:
        def __init__(self, owner):
            self.owner = owner

        def __enter__(self):
            self.prev, self.owner._sync = self.owner._sync, False

        def __exit__(self, *exc):
            self.owner._sync = self.prev

    def no_sync(self):
        """Gradient accumulation: backward() calls inside this context only accumulate; the last micro-batch runs
        outside of it (or `finish()` reduces what has accumulated)."""
        return GradientAllReducer._NoSync(self)

    def finish(self):
        """Call after loss.backward(): launches what the hooks did not (unused parameters, no_sync accumulation),
        waits for every bucket and points p.grad at the summed gradients."""
        if world_size(self.group) > 1:
            for i, ps in enumerate(self.buckets):
                if not self._launched[i] and any(p.grad is not None for p in ps):
                    for p in ps:
                        if p.grad is None:
                            p.grad = torch.zeros_like(p)
                    # (backward() has already joined every stream it used into the caller's stream, which _launch orders
                    #  the communication stream after)
                    self._launch(i)
            for work, i in self._work:
                work.wait()                               # current stream waits for NCCL's stream
                for p, v in zip(self.buckets[i], self._views[i]):
                    p.grad = v
            if self._comm is not None:
                torch.cuda.current_stream(self._comm.device).wait_stream(self._comm)
        self._reset()

    def remove(self):
        for h in self._hooks:
            h.remove()


def shard_bounds(n, group=None):
    """Contiguous [lo, hi) slice of n items owned by this rank (corpus sharding for the MRR scan)."""
    W, r = world_size(group), rank(group)
    per = (n + W - 1) // W
    return min(r * per, n), min((r + 1) * per, n)
