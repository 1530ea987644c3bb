"""Data parallelism for the two-tower step (one process per GPU, torch.distributed / NCCL over NVLink).

The reference is single-process; the oracle for everything here is the reference run once on the
concatenated global batch.  Only the in-batch score couples rows, so the batch is sharded and there is exactly
one exchange step:

  forward   all-gather of the [B_loc, D] tower outputs -> every rank scores its rows against ALL docs
  backward  reduce-scatter(SUM) of dL/d(gathered) back to the owning rank (autograd of the all-gather)
  grads     bucketed all-reduce(SUM) of the weight gradients, launched from grad hooks so the buckets of the
            layers that finish first (heads, doc tower layer 1) overlap the rest of BPTT

Each rank's loss is  sum(local rows) / B_global ; the sum over ranks is the reference's mean over the global
batch, hence SUM (not mean) for the gradient all-reduce.
"""
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

    Parameters are bucketed in reverse registration order (the order autograd finishes them: heads and the
    last GRU layer first).  A post-accumulate hook counts ready parameters; when a bucket is complete its
    gradients are flattened and all-reduced asynchronously.  `finish()` waits and scatters the results back.
    """

    def __init__(self, params, group=None, bucket_bytes=32 << 20):
        self.group = group
        self.params = [p for p in params if p.requires_grad]
        self.buckets, cur, size = [], [], 0
        for p in reversed(self.params):
            cur.append(p)
            size += p.numel() * p.element_size()
            if size >= bucket_bytes:
                self.buckets.append(cur)
                cur, size = [], 0
        if cur:
            self.buckets.append(cur)
        self._bucket_of = {id(p): i for i, b in enumerate(self.buckets) for p in b}
        self._pending = [len(b) for b in self.buckets]
        self._work = []
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]

    def _launch(self, i):
        ps = self.buckets[i]
        flat = torch.cat([p.grad.reshape(-1) for p in ps])
        work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self._work.append((work, flat, ps))

    def _on_grad(self, p):
        if world_size(self.group) == 1:
            return
        i = self._bucket_of[id(p)]
        self._pending[i] -= 1
        if self._pending[i] == 0:
            self._launch(i)

    def finish(self):
        """Call after loss.backward(): waits for every bucket and writes the summed gradients back."""
        if world_size(self.group) > 1:
            for i, n in enumerate(self._pending):      # buckets the hooks did not complete (unused parameters)
                if n > 0 and any(p.grad is not None for p in self.buckets[i]):
                    for p in self.buckets[i]:
                        if p.grad is None:
                            p.grad = torch.zeros_like(p)
                    self._launch(i)
            for work, flat, ps in self._work:
                work.wait()
                off = 0
                for p in ps:
                    n = p.numel()
                    p.grad.copy_(flat[off:off + n].view_as(p))
                    off += n
        self._work = []
        self._pending = [len(b) for b in self.buckets]

    def remove(self):
        for h in self._hooks:
            h.remove()


def shard_bounds(n, group=None):
    """Contiguous [lo, hi) slice of n items owned by this rank (corpus sharding for the MRR scan)."""
    W, r = world_size(group), rank(group)
    per = (n + W - 1) // W
    return min(r * per, n), min((r + 1) * per, n)
