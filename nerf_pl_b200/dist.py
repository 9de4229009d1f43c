"""Ray-sharded data parallelism: every rank renders its own slice of the rays; the only
exchange is ONE flat all-reduce of the MLP gradients (2 x 595 844 fp32 = 4.77 MB) over NCCL/NVLink in
training, and one gather of the per-ray outputs for full-frame renders.  Works with gloo on CPU for the
host-logic tests."""
import torch
import torch.distributed as dist


def shard_bounds(n_items: int, rank: int, world: int):
    """Contiguous, balanced [lo, hi) slice of n_items for `rank` (first n_items % world ranks get one more)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class GradBucket:
    """One flat fp32 buffer aliasing every parameter's .grad, so the gradient all-reduce is a single
    collective launch (latency-bound at 4.77 MB; NVSwitch makes it ~tens of microseconds).

    NeRF modules are also told where their gradients live (`_nrf_grad_sinks`): the fused render pass then lets its
    backward kernels add into the bucket directly instead of handing zero-initialised temporaries to autograd
    (mlp.backward_rays), as long as the parameters' .grad still are the bucket's views."""

    def __init__(self, modules):
        self.params = [p for m in modules for p in m.parameters() if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(n, dtype=torch.float32, device=self.params[0].device)
        off = 0
        for p in self.params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()
        for m in modules:
            if hasattr(m, "ordered_parameters") and all(p.requires_grad for p in m.parameters()):
                m.__dict__["_nrf_grad_sinks"] = [p.grad for p in m.ordered_parameters()]

    def zero(self):
        self.flat.zero_()

    def allreduce_mean(self, group=None, async_op=False):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        self.flat.div_(dist.get_world_size(group))
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)


def render_sharded(render_fn, rays, group=None, gather_dst=0):
    """Full-frame render: rank r renders rays[lo:hi]; results are gathered on `gather_dst`
    (eval.py:58-86 restated for N ranks).  render_fn(rays_slice) -> dict of per-ray tensors."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_bounds(rays.shape[0], rank, world)
    local = render_fn(rays[lo:hi])
    if world == 1:
        return local
    out = {}
    sizes = [shard_bounds(rays.shape[0], r, world) for r in range(world)]
    pad_to = max(b - a for a, b in sizes)
    for k in sorted(local):
        v = local[k].contiguous()
        if v.shape[0] < pad_to:   # equal-size all_gather (NCCL); the tail is sliced off below
            v = torch.cat([v, v.new_zeros((pad_to - v.shape[0],) + tuple(v.shape[1:]))], 0)
        bufs = [torch.empty_like(v) for _ in range(world)]
        dist.all_gather(bufs, v, group=group)
        if rank == gather_dst:
            out[k] = torch.cat([bufs[r][: b - a] for r, (a, b) in enumerate(sizes)], 0)
    return out if rank == gather_dst else None


class _GatherRows(torch.autograd.Function):
    """all_gather of contiguous row shards with a gradient: forward returns the full (n_total, ...) tensor on every
    rank; backward sums the ranks' gradients of the full tensor and hands each rank its own rows (a reduce-scatter,
    written as all_reduce + slice: the tensors here are a 64 KB depth map)."""

    @staticmethod
    def forward(ctx, local, n_total, group):
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        sizes = [shard_bounds(n_total, r, world) for r in range(world)]
        lo, hi = sizes[rank]
        if local.shape[0] != hi - lo:
            raise RuntimeError(f"rank {rank} holds {local.shape[0]} rows, its shard of {n_total} is {hi - lo}")
        pad_to = max(b - a for a, b in sizes)
        v = local.contiguous()
        if v.shape[0] < pad_to:
            v = torch.cat([v, v.new_zeros((pad_to - v.shape[0],) + tuple(v.shape[1:]))], 0)
        bufs = [torch.empty_like(v) for _ in range(world)]
        dist.all_gather(bufs, v, group=group)
        ctx.group, ctx.span = group, (lo, hi)
        return torch.cat([bufs[r][: b - a] for r, (a, b) in enumerate(sizes)], 0)

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous().clone()
        dist.all_reduce(g, op=dist.ReduceOp.SUM, group=ctx.group)
        lo, hi = ctx.span
        return g[lo:hi], None, None


def gather_rows(local, n_total, group=None):
    """Differentiable all_gather of row shards cut by shard_bounds (identity without a process group)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    return _GatherRows.apply(local, n_total, group)


def render_light_sharded(render_fn, light_rays, group=None):
    """Shadow-mapping variant on N ranks (option ii): the light-view depth map is a dependency shared by
    every rank's camera rays.  Each rank renders only its contiguous slice of the light rays and the per-ray results
    (depth_coarse / depth_fine / ...) are all-gathered WITH gradient, so `efficient_sm` sees the whole light map on every
    rank and, when the light render carries gradients (train_efficient_sm.py grad_on_light), every rank's shadow loss
    reaches the slice that produced each texel.  Together with the mean all-reduce of the parameter gradients this
    equals the replicated computation (every rank rendering all light rays) averaged over ranks."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n = light_rays.shape[0]
    lo, hi = shard_bounds(n, rank, world)
    local = render_fn(light_rays[lo:hi])
    return {k: gather_rows(v, n, group) for k, v in local.items()}
