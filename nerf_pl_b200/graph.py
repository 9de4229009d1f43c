"""CUDA-graph capture of a whole training step ("capture launch-bound inner loops in CUDA graphs").

One nerf_pl step is ~20 kernel launches driven by a few hundred microseconds of Python (render_rays, autograd, the
optimizer); the GPU idles at the start of every step while the host catches up after the loss read-back.  A
`GraphedStep` records the step once - sampling, both MLP passes, compositing, loss, backward, gradient all-reduce,
optimizer step, operand re-pack - and replays it with one launch.  Everything on the path is capture-safe: kernels
take torch's current stream, random numbers come from torch's graph-aware CUDA generator (fresh values every replay),
`optim.FusedAdam(capturable=True)` keeps its step count and learning rate on the device, and
`NeRF.packed_weights()` always records the re-pack while capturing."""
import torch


class GraphedStep:
    """fn(*inputs) -> tensor or tuple of tensors; inputs are tensors of fixed shape/dtype.

    g = GraphedStep(step_fn, (rays, rgbs))     # warm-up runs + capture (fn really executes: parameters train)
    loss = g(rays_host_pinned, rgbs_host_pinned)   # copies into the static inputs (H2D allowed), one graph launch

    The returned tensors are the graph's static outputs: read them (e.g. `.item()`) before the next call."""

    def __init__(self, fn, example_inputs, warmup=3, optimizers=()):
        if not torch.cuda.is_available():
            raise RuntimeError("GraphedStep needs a CUDA device (no CPU path)")
        self.fn = fn
        self.optimizers = list(optimizers)
        self.static_in = [x.detach().clone() for x in example_inputs]
        if any(not x.is_cuda for x in self.static_in):
            raise RuntimeError("example inputs must be CUDA tensors")
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):              # warm-up off the default stream: allocator, lazy inits, caches
            for _ in range(max(1, warmup)):
                fn(*self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        for o in self.optimizers:
            if hasattr(o, "sync_lr"):
                o.sync_lr()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.static_out = fn(*self.static_in)
        self.replays = 0

    def __call__(self, *inputs):
        if len(inputs) != len(self.static_in):
            raise ValueError(f"expected {len(self.static_in)} inputs, got {len(inputs)}")
        for o in self.optimizers:
            if hasattr(o, "sync_lr"):
                o.sync_lr()
        for dst, src in zip(self.static_in, inputs):
            if src is not dst:
                if src.shape != dst.shape or src.dtype != dst.dtype:
                    raise ValueError("GraphedStep inputs must keep the captured shape and dtype")
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        self.replays += 1
        return self.static_out

    def close(self):
        """Release the captured graph and its memory pool.  Call before torch.distributed.destroy_process_group():
        NCCL waits for graphs that captured its collectives to be destroyed."""
        torch.cuda.synchronize()
        self.static_out = None
        if self.graph is not None:
            self.graph.reset()
            self.graph = None
