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
        self._ring = None
        self._inflight = []

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
        # the captured optimizer step changed the parameters on the device without any host code running: bump their
        # version counters so that caches keyed on them (NeRF.packed_weights) re-pack on the next EAGER render
        for o in self.optimizers:
            for group in o.param_groups:
                torch.autograd.graph.increment_version(group["params"])
        return self.static_out

    def submit(self, *inputs, depth=2):
        """Pipelined form of __call__ for training loops that read the loss one step late: queues the pinned-host ->
        device copies of this step's inputs, the graph launch and an asynchronous read-back of the outputs into a
        pinned slot, and returns at once with a handle; `handle.result()` waits for THAT step only and returns host
        copies of its outputs.  With `depth` steps in flight the GPU never idles while Python turns around after a
        read-back (the ~0.2 ms per step that the blocking form loses).  The caller must not overwrite a host input
        buffer before the step that reads it has been collected."""
        if self._ring is None or len(self._ring) != depth:
            outs = self.static_out if isinstance(self.static_out, (tuple, list)) else (self.static_out,)
            self._ring = [([torch.empty(o.shape, dtype=o.dtype, pin_memory=True) for o in outs], torch.cuda.Event())
                          for _ in range(depth)]
            self._inflight = [False] * depth
        slot = self.replays % depth
        host_out, done = self._ring[slot]
        if self._inflight[slot]:
            done.synchronize()          # the slot's previous result has not been collected: do not overwrite in flight
        self(*inputs)
        outs = self.static_out if isinstance(self.static_out, (tuple, list)) else (self.static_out,)
        for h, o in zip(host_out, outs):
            h.copy_(o.detach(), non_blocking=True)
        done.record()
        self._inflight[slot] = True
        return _Pending(self, slot, isinstance(self.static_out, (tuple, list)))

    def close(self):
        """Release the captured graph and its memory pool.  Call before torch.distributed.destroy_process_group():
        NCCL waits for graphs that captured its collectives to be destroyed."""
        torch.cuda.synchronize()
        self.static_out = None
        if self.graph is not None:
            self.graph.reset()
            self.graph = None


class _Pending:
    """Handle returned by GraphedStep.submit."""

    def __init__(self, owner, slot, many):
        self.owner, self.slot, self.many = owner, slot, many

    def result(self):
        host_out, done = self.owner._ring[self.slot]
        done.synchronize()
        self.owner._inflight[self.slot] = False
        vals = [h.clone() for h in host_out]
        return tuple(vals) if self.many else vals[0]
