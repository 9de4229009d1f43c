"""Drop-in for losses.py (reference :4-80): MSELoss / SMMSELoss as ONE fused, deterministic reduction launch forward and
ONE launch backward (the reference runs ~6 elementwise/reduction kernels each way per term); OpactiyLoss keeps the
reference's torch formulation (data-dependent index sets; not on the benchmarked path)."""
import torch
from torch import nn

from . import _lib
from .ops import _dev, _p, _stream

_ws = {}


def _workspace(dev):
    ws = _ws.get(dev)
    if ws is None:
        ws = torch.zeros(_lib.load().nrf_mse_workspace_bytes(), dtype=torch.uint8, device=dev)   # zeroed once; reusable
        _ws[dev] = ws
    return ws


def mse_pair(pred_a, pred_b, target, mask=None):
    """mean((pred_a - target)^2) [+ mean((pred_b - target)^2)] as a 0-dim tensor, no autograd."""
    a = _dev(pred_a, "pred_a")
    t = _dev(target, "target")
    b = _dev(pred_b, "pred_b") if pred_b is not None else None
    if a.shape != t.shape or (b is not None and b.shape != t.shape):
        raise RuntimeError(f"prediction / target shapes differ: {tuple(a.shape)} vs {tuple(t.shape)}")
    m = None
    if mask is not None:
        m = mask.to(device=a.device, dtype=torch.bool)
        if m.shape != a.shape:
            m = m.reshape(m.shape + (1,) * (a.dim() - m.dim())).expand_as(a)
        m = m.contiguous().view(torch.uint8)
    loss = torch.empty((), device=a.device, dtype=torch.float32)
    with torch.cuda.device(a.device):
        _lib.call("nrf_mse_pair", _p(a), _p(b), _p(t), _p(m), a.numel(), _p(loss), _p(_workspace(a.device)), _stream())
    return loss


class _FusedMSE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred_a, pred_b, target):
        ctx.has_b = pred_b is not None
        ctx.save_for_backward(pred_a, pred_b if ctx.has_b else pred_a, target)
        return mse_pair(pred_a, pred_b, target)

    @staticmethod
    def backward(ctx, g):
        a, b, t = ctx.saved_tensors
        a, t = a.contiguous(), t.contiguous()
        b = b.contiguous() if ctx.has_b else None
        need_a, need_b = ctx.needs_input_grad[0], ctx.has_b and ctx.needs_input_grad[1]
        ga = torch.empty_like(a) if need_a else None
        gb = torch.empty_like(b) if need_b else None
        if need_a or need_b:
            g = g.to(torch.float32).contiguous()
            with torch.cuda.device(a.device):
                _lib.call("nrf_mse_pair_bwd", _p(a) if need_a else None, _p(b) if need_b else None, _p(t), a.numel(), _p(g),
                          _p(ga), _p(gb), _stream())
        return ga, gb, None


def fused_mse(pred_a, pred_b, target):
    """Differentiable mean((pred_a - target)^2) [+ mean((pred_b - target)^2)]."""
    return _FusedMSE.apply(pred_a, pred_b, target)


class MSELoss(nn.Module):
    """reference losses.py:4-14"""

    def forward(self, inputs, targets):
        return fused_mse(inputs['rgb_coarse'], inputs.get('rgb_fine'), targets)


class SMMSELoss(nn.Module):
    """reference losses.py:16-26"""

    def forward(self, inputs, targets):
        return fused_mse(inputs['sm_coarse'], inputs.get('sm_fine'), targets)


class OpactiyLoss(nn.Module):
    """reference losses.py:28-78 (name as spelled there)"""

    def __init__(self, coeff=2000.0, sm_thres=0.4):
        super().__init__()
        self.coeff = coeff
        self.sm_thres = sm_thres
        self.loss = nn.L1Loss(reduction='mean')

    def get_sm_pixels(self, targets):
        targets = (targets[:, 0] + targets[:, 1] + targets[:, 2]) / 3.
        return torch.where(targets > self.sm_thres)[0], torch.where(targets <= self.sm_thres)[0]

    def forward(self, inputs, targets):
        sm_idx, non_sm_idx = self.get_sm_pixels(targets)
        if len(sm_idx) == 0 or len(non_sm_idx) == 0:
            return 0.
        loss = 0.
        for key in ('opacity_coarse', 'opacity_fine'):
            if key in inputs:
                op = inputs[key]
                loss = loss + (self.coeff - self.loss(torch.mean(op[non_sm_idx]), torch.mean(op[sm_idx])))
        return loss


loss_dict = {'mse': MSELoss, 'sm': SMMSELoss, 'opacity': OpactiyLoss}
