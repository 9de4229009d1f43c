"""Ray-sharded data parallelism: every rank renders its own slice of the rays; the only
exchange is ONE flat all-reduce of the MLP gradients (2 x 595 844 fp32 = 4.77 MB) over NCCL/NVLink in
training, and one gather of the per-ray outputs for full-frame renders.  Works with gloo on CPU for the
host-logic tests."""
import ctypes
import os
import socket

import torch
import torch.distributed as dist


def shard_bounds(n_items: int, rank: int, world: int):
    """Contiguous, balanced [lo, hi) slice of n_items for `rank` (first n_items % world ranks get one more)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class _DevMem:
    """torch view of a raw device allocation (CUDA array interface, zero copy)."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class PeerExchange:
    """The flat gradient buffers of all ranks of ONE node mapped into every rank's process (CUDA IPC over NVLink peer
    access), for nrf_peer_allreduce_mean (csrc/peer_allreduce.cu): the training step's only collective as one kernel
    launch per rank that loads / stores the peers' memory directly.  Collective constructor (every rank of `group`
    must call it); `ok` is False on every rank alike when any rank cannot take part (ranks on several hosts, more than
    8 ranks, no CUDA, IPC refused), and the caller keeps NCCL."""

    def __init__(self, n_floats, device, group=None):
        from . import _lib
        self.ok, self.n, self.device = False, int(n_floats), device
        self.own, self.mapped, self.flat = [], [], None
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        usable = (device.type == "cuda" and 2 <= world <= 8 and "nccl" in str(dist.get_backend(group))
                  and os.environ.get("NRF_PEER_ALLREDUCE", "1") != "0")
        hosts = [None] * world
        dist.all_gather_object(hosts, (socket.gethostname(), bool(usable)), group=group)
        if len({h for h, _ in hosts}) != 1 or not all(u for _, u in hosts):
            return
        lib = _lib.load()
        err = ""
        handles = None
        try:
            with torch.cuda.device(device):
                buf, flg = ctypes.c_void_p(), ctypes.c_void_p()
                _lib.call("nrf_peer_alloc", self.n * 4, ctypes.byref(buf))
                self.own.append(buf.value)
                _lib.call("nrf_peer_alloc", 256, ctypes.byref(flg))      # 8 flag words, then seq (+64) and counter (+128)
                self.own.append(flg.value)
                hb, hf = ctypes.create_string_buffer(64), ctypes.create_string_buffer(64)
                _lib.call("nrf_peer_export", buf.value, hb)
                _lib.call("nrf_peer_export", flg.value, hf)
                handles = (hb.raw, hf.raw)
        except Exception as e:      # noqa: BLE001  (every rank must reach the collectives below)
            err = f"{type(e).__name__}: {e}"
        everyone = [None] * world
        dist.all_gather_object(everyone, (handles, err), group=group)
        good = all(h is not None for h, _ in everyone)
        comm = _lib.PeerComm()
        if good:
            try:
                with torch.cuda.device(device):
                    for q, (h, _) in enumerate(everyone):
                        if q == rank:
                            comm.bufs[q], comm.flags[q] = self.own[0], self.own[1]
                            continue
                        pb, pf = ctypes.c_void_p(), ctypes.c_void_p()
                        _lib.call("nrf_peer_import", h[0], ctypes.byref(pb))
                        self.mapped.append(pb.value)
                        _lib.call("nrf_peer_import", h[1], ctypes.byref(pf))
                        self.mapped.append(pf.value)
                        comm.bufs[q], comm.flags[q] = pb.value, pf.value
            except Exception as e:      # noqa: BLE001
                good, err = False, f"{type(e).__name__}: {e}"
        flags = [None] * world
        dist.all_gather_object(flags, (bool(good), err), group=group)
        if not all(g for g, _ in flags):
            self.error = "; ".join(e for _, e in flags if e)
            self.close()
            return
        comm.rank, comm.world, comm.n = rank, world, self.n
        comm.seq, comm.counter = self.own[1] + 64, self.own[1] + 128
        self.comm = comm
        self.flat = torch.as_tensor(_DevMem(self.own[0], self.n), device=device)
        self.ok = True

    def allreduce_mean(self):
        from . import _lib
        from .ops import _stream
        with torch.cuda.device(self.device):
            _lib.call("nrf_peer_allreduce_mean", ctypes.byref(self.comm), _stream())

    def close(self):
        """Unmap the peers' buffers and free this rank's (after a barrier: nobody may still be reading them)."""
        from . import _lib
        if not (self.own or self.mapped):
            return
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        for ptr in self.mapped:
            _lib.call("nrf_peer_close", ptr)
        self.mapped = []
        if self.ok and dist.is_initialized():
            dist.barrier()
        for ptr in self.own:
            _lib.call("nrf_peer_free", ptr)
        self.own, self.ok, self.flat = [], False, None


class GradBucket:
    """One flat fp32 buffer aliasing every parameter's .grad, so the gradient exchange of the ray-sharded training step
    is a single launch: `nrf_peer_allreduce_mean` over NVLink peer memory when all ranks share a node (PeerExchange
    above; `peer=False` or NRF_PEER_ALLREDUCE=0 keeps NCCL), one NCCL all-reduce otherwise (gloo on CPU for the
    host-logic tests).

    NeRF modules are also told where their gradients live (`_nrf_grad_sinks`): the fused render pass then lets its
    backward kernels add into the bucket directly instead of handing zero-initialised temporaries to autograd
    (mlp.backward_rays), as long as the parameters' .grad still are the bucket's views."""

    def __init__(self, modules, peer=True, group=None):
        self.params = [p for m in modules for p in m.parameters() if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        self.peer = None
        dev = self.params[0].device
        if peer and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1 and dev.type == "cuda":
            px = PeerExchange(n, dev, group)      # collective: every rank decides alike
            if px.ok:
                self.peer = px
        self.flat = self.peer.flat if self.peer is not None else torch.zeros(n, dtype=torch.float32, device=dev)
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
        if self.peer is not None:
            self.peer.allreduce_mean()
            return None
        self.flat.div_(dist.get_world_size(group))
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)

    def close(self):
        """Release the peer mapping (call before torch.distributed.destroy_process_group())."""
        if self.peer is not None:
            for p in self.params:
                p.grad = None
            self.flat = None
            self.peer.close()
            self.peer = None


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
