"""Fused optimizer steps over flat buffers (reference utils/__init__.py:10-30: adam, sgd, radam, ranger).

`FusedAdam` re-homes every parameter of the given modules into ONE flat fp32 buffer (p.data becomes a view) and shares
the flat gradient buffer of a `dist.GradBucket`, so a step is one elementwise launch over 2 x 595 844 values instead of
torch's per-tensor (or foreach) kernels, and `zero_grad()` is one fill.  It is a torch.optim.Optimizer, so the reference's
schedulers (`MultiStepLR`, `CosineAnnealingLR`, `GradualWarmupScheduler`, utils/__init__.py:32-49) drive `param_groups`
as usual.  The bf16 operand images are re-packed lazily by NeRF.packed_weights() when the parameters' version changes."""
import torch

from . import _lib
from .dist import GradBucket
from .ops import _p, _stream


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, modules, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, bucket=None,
                 capturable=False):
        modules = list(modules)
        self.bucket = bucket if bucket is not None else GradBucket(modules)
        params = self.bucket.params
        if any(not p.is_cuda or p.dtype != torch.float32 for p in params):
            raise RuntimeError("FusedAdam needs fp32 CUDA parameters (no CPU path)")
        n = self.bucket.flat.numel()
        self.flat_param = torch.empty(n, dtype=torch.float32, device=self.bucket.flat.device)
        off = 0
        with torch.no_grad():
            for p in params:
                view = self.flat_param[off:off + p.numel()].view_as(p)
                view.copy_(p)
                p.data = view
                off += p.numel()
        self.exp_avg = torch.zeros_like(self.flat_param)
        self.exp_avg_sq = torch.zeros_like(self.flat_param)
        self._steps = 0
        # capturable: step count and lr live on the device so a captured step (graph.GraphedStep) replays correctly
        self._state = None
        if capturable:
            assert _lib.load().nrf_adam_state_bytes() == 24
            self._state = torch.zeros(3, dtype=torch.int64, device=self.flat_param.device)   # {step, lr (f64), scratch}
            self._state_lr = None
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    def zero_grad(self, set_to_none=False):
        self.bucket.zero()

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        g = self.param_groups[0]
        self._steps += 1
        with torch.cuda.device(self.flat_param.device):
            if self._state is not None:
                if not torch.cuda.is_current_stream_capturing():
                    self.sync_lr()
                _lib.call("nrf_adam_step_dev", _p(self.flat_param), _p(self.bucket.flat), _p(self.exp_avg),
                          _p(self.exp_avg_sq), self.flat_param.numel(), float(g["betas"][0]), float(g["betas"][1]),
                          float(g["eps"]), float(g["weight_decay"]), _p(self._state), _stream())
            else:
                _lib.call("nrf_adam_step", _p(self.flat_param), _p(self.bucket.flat), _p(self.exp_avg),
                          _p(self.exp_avg_sq), self.flat_param.numel(), float(g["lr"]), float(g["betas"][0]),
                          float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]), self._steps, _stream())
        # parameter version counters drive the weight re-pack cache (NeRF.packed_weights); the kernel wrote through
        # the flat buffer, so bump them here (one C++ call, no launches)
        torch.autograd.graph.increment_version(self.bucket.params)
        return loss

    def sync_lr(self):
        """capturable mode: push param_groups[0]['lr'] to the device if a scheduler changed it (call before a replay)."""
        lr = float(self.param_groups[0]["lr"])
        if self._state is not None and lr != self._state_lr:
            self._state[1:2].view(torch.float64).copy_(torch.tensor([lr], dtype=torch.float64))
            self._state_lr = lr

    @property
    def steps(self):
        return int(self._state[0].item()) if self._state is not None else self._steps

    def state_dict(self):
        return {"steps": self.steps, "flat_param": self.flat_param, "exp_avg": self.exp_avg,
                "exp_avg_sq": self.exp_avg_sq, "param_groups": [{k: v for k, v in g.items() if k != "params"}
                                                                  for g in self.param_groups]}

    def load_state_dict(self, sd):
        self._steps = int(sd["steps"])
        if self._state is not None:
            self._state[0:1].fill_(self._steps)
        with torch.no_grad():
            self.flat_param.copy_(sd["flat_param"])
            self.exp_avg.copy_(sd["exp_avg"])
            self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        for g, s in zip(self.param_groups, sd["param_groups"]):
            g.update(s)
        torch.autograd.graph.increment_version(self.bucket.params)


class _FlatOptimizer(torch.optim.Optimizer):
    """Shared plumbing of the non-Adam fused optimizers: parameters re-homed into one flat fp32 buffer beside the flat
    gradient buffer, one elementwise launch per step, version counters bumped for the operand re-pack cache."""

    def __init__(self, modules, defaults, bucket=None):
        modules = list(modules)
        self.bucket = bucket if bucket is not None else GradBucket(modules)
        params = self.bucket.params
        if any(not p.is_cuda or p.dtype != torch.float32 for p in params):
            raise RuntimeError(f"{type(self).__name__} needs fp32 CUDA parameters (no CPU path)")
        self.flat_param = torch.empty_like(self.bucket.flat)
        off = 0
        with torch.no_grad():
            for p in params:
                view = self.flat_param[off:off + p.numel()].view_as(p)
                view.copy_(p)
                p.data = view
                off += p.numel()
        self._steps = 0
        self._buffers = {}
        super().__init__(params, defaults)

    def _buf(self, name, init=None):
        if name not in self._buffers:
            self._buffers[name] = torch.zeros_like(self.flat_param) if init is None else init.clone()
        return self._buffers[name]

    def zero_grad(self, set_to_none=False):
        self.bucket.zero()

    @property
    def steps(self):
        return self._steps

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self._steps += 1
        with torch.cuda.device(self.flat_param.device):
            self._launch(self.param_groups[0])
        torch.autograd.graph.increment_version(self.bucket.params)
        return loss

    def state_dict(self):
        return {"steps": self._steps, "flat_param": self.flat_param, "buffers": dict(self._buffers),
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd):
        self._steps = int(sd["steps"])
        with torch.no_grad():
            self.flat_param.copy_(sd["flat_param"])
            for k, v in sd["buffers"].items():
                self._buf(k).copy_(v)
        for g, s in zip(self.param_groups, sd["param_groups"]):
            g.update(s)
        torch.autograd.graph.increment_version(self.bucket.params)


class FusedSGD(_FlatOptimizer):
    """torch.optim.SGD as utils/__init__.py:14-16 builds it (lr, momentum, weight_decay), one launch (nrf_sgd_step)."""

    def __init__(self, modules, lr=5e-4, momentum=0.0, weight_decay=0.0, bucket=None):
        super().__init__(modules, dict(lr=lr, momentum=momentum, weight_decay=weight_decay), bucket)

    def _launch(self, g):
        buf = self._buf("momentum_buffer") if g["momentum"] != 0 else None
        _lib.call("nrf_sgd_step", _p(self.flat_param), _p(self.bucket.flat), _p(buf), self.flat_param.numel(),
                  float(g["lr"]), float(g["momentum"]), float(g["weight_decay"]), int(self._steps == 1), _stream())


class FusedRAdam(_FlatOptimizer):
    """utils/optimizers.py:6-95 (`--optimizer radam`), one launch (nrf_radam_step)."""

    def __init__(self, modules, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, degenerated_to_sgd=True,
                 bucket=None):
        if not 0.0 <= lr:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if not 0.0 <= eps:
            raise ValueError("Invalid epsilon value: {}".format(eps))
        for i in (0, 1):
            if not 0.0 <= betas[i] < 1.0:
                raise ValueError("Invalid beta parameter at index {}: {}".format(i, betas[i]))
        self.degenerated_to_sgd = degenerated_to_sgd
        super().__init__(modules, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay), bucket)

    def _launch(self, g):
        _lib.call("nrf_radam_step", _p(self.flat_param), _p(self.bucket.flat), _p(self._buf("exp_avg")),
                  _p(self._buf("exp_avg_sq")), None, self.flat_param.numel(), float(g["lr"]), float(g["betas"][0]),
                  float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]), self._steps, 0, 5.0,
                  int(bool(self.degenerated_to_sgd)), 0.0, 1, _stream())


class FusedRanger(_FlatOptimizer):
    """utils/optimizers.py:266-404 (`--optimizer ranger`): RAdam + lookahead, one launch (nrf_radam_step)."""

    def __init__(self, modules, lr=1e-3, alpha=0.5, k=6, N_sma_threshhold=5, betas=(.95, 0.999), eps=1e-5,
                 weight_decay=0, bucket=None):
        if not 0.0 <= alpha <= 1.0:
            raise ValueError(f'Invalid slow update rate: {alpha}')
        if not 1 <= k:
            raise ValueError(f'Invalid lookahead steps: {k}')
        if not lr > 0:
            raise ValueError(f'Invalid Learning Rate: {lr}')
        if not eps > 0:
            raise ValueError(f'Invalid eps: {eps}')
        self.N_sma_threshhold, self.alpha, self.k = N_sma_threshhold, alpha, k
        super().__init__(modules, dict(lr=lr, alpha=alpha, k=k, step_counter=0, betas=betas,
                                       N_sma_threshhold=N_sma_threshhold, eps=eps, weight_decay=weight_decay), bucket)

    def _launch(self, g):
        slow = self._buf("slow_buffer", init=self.flat_param)       # seeded with the parameters on the first step (:347-348)
        _lib.call("nrf_radam_step", _p(self.flat_param), _p(self.bucket.flat), _p(self._buf("exp_avg")),
                  _p(self._buf("exp_avg_sq")), _p(slow), self.flat_param.numel(), float(g["lr"]), float(g["betas"][0]),
                  float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]), self._steps, 1,
                  float(self.N_sma_threshhold), 1, float(self.alpha), int(g["k"]), _stream())


def get_optimizer(hparams, models):
    """utils/__init__.py:10-30: the same four choices, each as one fused launch over the flat buffers."""
    eps = 1e-8
    wd = getattr(hparams, "weight_decay", 0.0)
    name = getattr(hparams, "optimizer", "adam")
    if name == 'sgd':
        return FusedSGD(models, lr=hparams.lr, momentum=getattr(hparams, "momentum", 0.9), weight_decay=wd)
    if name == 'adam':
        return FusedAdam(models, lr=hparams.lr, eps=eps, weight_decay=wd)
    if name == 'radam':
        return FusedRAdam(models, lr=hparams.lr, eps=eps, weight_decay=wd)
    if name == 'ranger':
        return FusedRanger(models, lr=hparams.lr, eps=eps, weight_decay=wd)
    raise ValueError('optimizer not recognized!')
