"""MLP back ends behind one interface: fp32 FFMA (parity) and bf16 tcgen05 (performance).

`forward_rays` evaluates the NeRF MLP on the sample points o+d*z of a batch of rays with the
positional encodings computed inside the kernels; `backward_rays` returns parameter gradients.
`RenderPass` fuses MLP + compositing into one autograd node (models/rendering.py:117-198)."""
from ctypes import byref

import torch

from . import _lib, ops
from .ops import _dev, _p, _stream


def pack_weights(spec, params, out=None):
    """bf16 UMMA operand images of the parameters (nrf_mlp_tc_pack).  Callers that own the parameters
    (NeRF.packed_weights) cache the result per parameter version; there is no global cache, because device
    addresses are recycled between models."""
    nbytes = _lib.load().nrf_mlp_tc_packed_bytes()
    buf = out if out is not None else torch.empty(nbytes, dtype=torch.uint8, device=params[0].device)
    _lib.call("nrf_mlp_tc_pack", byref(spec.params_struct(params)), _p(buf), _stream())
    return buf


def _use_tc(spec, precision):
    if precision == "fp32":
        return False
    if not spec.tc_capable:
        raise RuntimeError("the bf16 tensor-core MLP supports D=8, W=256, in_xyz=63, in_dir=27, skips=[4] only; "
                           "call nerf_pl_b200.set_mlp_precision('fp32') for other shapes")
    return True


def forward_rays(spec, params, rays, z, sigma_only, need_grad, precision, xyz_freqs=10, dir_freqs=4, packed=None):
    """Returns (sigma (N,S), rgb (N,S,3) or None, saved) — `saved` feeds backward_rays."""
    lib = _lib.load()
    N, S = z.shape
    dev = z.device
    sigma = torch.empty(N, S, device=dev, dtype=torch.float32)
    rgb = None if sigma_only else torch.empty(N, S, 3, device=dev, dtype=torch.float32)
    if N == 0:
        return sigma, rgb, None
    if _use_tc(spec, precision):
        if packed is None:
            packed = pack_weights(spec, params)
        nb = lib.nrf_mlp_tc_stash_bytes(N * S, int(sigma_only)) if need_grad else 0
        stash = torch.empty(nb, dtype=torch.uint8, device=dev) if nb else None
        _lib.call("nrf_mlp_tc_fwd", _p(packed), _p(rays), _p(z), N, S, int(sigma_only), _p(sigma), _p(rgb),
                  _p(stash), nb, _stream())
        return sigma, rgb, ("tc", packed, stash)
    ps = spec.params_struct(params)
    nb = lib.nrf_mlp_fp32_workspace_bytes(byref(ps), N * S, N, int(sigma_only), int(need_grad))
    ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    _lib.call("nrf_mlp_fp32_fwd_rays", byref(ps), _p(rays), _p(z), N, S, xyz_freqs, dir_freqs, int(sigma_only),
              int(need_grad), _p(sigma), _p(rgb), _p(ws), nb, _stream())
    return sigma, rgb, ("fp32", ws if need_grad else None)


def backward_rays(spec, params, rays, z, sigma_only, saved, d_sigma, d_rgb, rgb, grad_sinks=None):
    """Parameter gradients.  The kernels ACCUMULATE (atomic adds from many CTAs), so by default they run on
    zero-filled tensors that are returned to autograd.  With `grad_sinks` (dist.GradBucket: one fp32 tensor per
    parameter, aliasing its .grad) they add straight into the sinks and a list of None is returned: no 24 fills and
    no 24 AccumulateGrad additions per pass."""
    lib = _lib.load()
    N, S = z.shape
    if grad_sinks is not None:
        grads, ret = grad_sinks, [None] * len(params)
    else:
        grads = ret = [torch.zeros_like(p) for p in params]
    if N == 0:
        return ret
    gs = spec.grads_struct(grads)
    if saved[0] == "tc":
        _, packed, stash = saved
        nb = lib.nrf_mlp_tc_bwd_workspace_bytes(N * S, int(sigma_only))
        ws = torch.empty(nb, dtype=torch.uint8, device=z.device)
        _lib.call("nrf_mlp_tc_bwd", _p(packed), byref(gs), _p(rays), _p(z), N, S, int(sigma_only), _p(d_sigma),
                  _p(d_rgb), _p(rgb), _p(stash), stash.numel(), _p(ws), nb, _stream())
    else:
        ws = saved[1]
        ps = spec.params_struct(params)
        _lib.call("nrf_mlp_fp32_bwd", byref(ps), byref(gs), None, N * S, N, S, int(sigma_only), _p(d_sigma),
                  _p(d_rgb), _p(rgb), _p(ws), ws.numel(), _stream())
    return ret


def _live_sinks(cfg, params):
    """cfg.grad_sinks if they still alias the parameters' .grad (the user may have dropped or replaced them)."""
    sinks = getattr(cfg, "grad_sinks", None)
    if sinks is None or len(sinks) != len(params):
        return None
    for p, s in zip(params, sinks):
        g = p.grad
        if g is None or g.data_ptr() != s.data_ptr() or g.shape != s.shape:
            return None
    return sinks


class RenderPass(torch.autograd.Function):
    """One MLP evaluation + volume-rendering integral over a batch of rays.

    forward(rays, z, noise_or_None, cfg, *params) ->
        (rgb (N,3) [empty when cfg.sigma_only], depth, opacity, disp, weights)
    cfg: namespace with spec, sigma_only, white_back, noise_std, precision, xyz_freqs, dir_freqs, need_grad."""

    @staticmethod
    def forward(ctx, rays, z, noise, cfg, *params):
        ctx.set_materialize_grads(False)      # unused outputs (depth, opacity, disp) arrive as None, not as zero fills
        # grad mode is always off inside Function.forward and needs_input_grad ignores torch.no_grad(): the caller
        # decided (cfg.need_grad) whether this pass will be differentiated, i.e. whether the training stash is written
        need_grad = bool(cfg.need_grad) and any(ctx.needs_input_grad[4:])
        sigma, rgb_s, saved = forward_rays(cfg.spec, params, rays, z, cfg.sigma_only, need_grad, cfg.precision,
                                           cfg.xyz_freqs, cfg.dir_freqs, getattr(cfg, "packed", None))
        c = ops.composite(sigma, rgb_s, z, rays, noise, cfg.noise_std, cfg.white_back)
        ctx.cfg, ctx.saved, ctx.n_params = cfg, saved, len(params)
        if need_grad:
            ctx.save_for_backward(rays, z, noise if noise is not None else torch.empty(0), sigma,
                                  rgb_s if rgb_s is not None else torch.empty(0), *params)
        rgb = c["rgb"] if c["rgb"] is not None else torch.empty(0, device=z.device)
        ctx.mark_non_differentiable(c["weights"])
        return rgb, c["depth"], c["opacity"], c["disp"], c["weights"]

    @staticmethod
    def backward(ctx, g_rgb, g_depth, g_opacity, g_disp, _g_weights):
        cfg = ctx.cfg
        rays, z, noise, sigma, rgb_s, *params = ctx.saved_tensors
        noise = noise if noise.numel() else None
        rgb_s = rgb_s if rgb_s.numel() else None
        d_sigma, d_rgb = ops.composite_bwd(sigma, rgb_s, z, rays, noise, cfg.noise_std, cfg.white_back,
                                           g_rgb if rgb_s is not None else None, g_depth, g_opacity, g_disp)
        grads = backward_rays(cfg.spec, params, rays, z, cfg.sigma_only, ctx.saved, d_sigma, d_rgb, rgb_s,
                              _live_sinks(cfg, params))
        ctx.saved = None
        return (None, None, None, None, *grads)


class MlpEmbedded(torch.autograd.Function):
    """NeRF.forward on already-encoded inputs (models/nerf.py:83-124), fp32 kernels."""

    @staticmethod
    def forward(ctx, x, spec, sigma_only, *params):
        if ctx.needs_input_grad[0]:
            raise NotImplementedError("gradients w.r.t. the encoded input of NeRF.forward are not provided "
                                      "(no caller in the reference differentiates through it)")
        lib = _lib.load()
        x = _dev(x, "x")
        rows = x.shape[0]
        want = spec.in_xyz + (0 if sigma_only else spec.in_dir)
        if x.dim() != 2 or x.shape[1] != want:
            raise RuntimeError(f"NeRF.forward expects x of shape (B, {want}), got {tuple(x.shape)}")
        need_grad = any(ctx.needs_input_grad[3:])
        sigma = torch.empty(rows, device=x.device)
        rgb = None if sigma_only else torch.empty(rows, 3, device=x.device)
        ps = spec.params_struct(params)
        nb = lib.nrf_mlp_fp32_workspace_bytes(byref(ps), rows, 0, int(sigma_only), int(need_grad))
        ws = torch.empty(nb, dtype=torch.uint8, device=x.device)
        _lib.call("nrf_mlp_fp32_fwd_embedded", byref(ps), _p(x), rows, int(sigma_only), int(need_grad), _p(sigma),
                  _p(rgb), _p(ws), nb, _stream())
        ctx.spec, ctx.sigma_only = spec, sigma_only
        if need_grad:
            ctx.save_for_backward(x, ws, rgb if rgb is not None else torch.empty(0), *params)
        if sigma_only:
            return sigma.view(rows, 1)
        return torch.cat([rgb, sigma.view(rows, 1)], -1)

    @staticmethod
    def backward(ctx, g):
        x, ws, rgb, *params = ctx.saved_tensors
        spec, sigma_only = ctx.spec, ctx.sigma_only
        g = g.contiguous()
        rows = x.shape[0]
        if sigma_only:
            d_sigma, d_rgb, rgb = g.reshape(rows).contiguous(), None, None
        else:
            d_sigma, d_rgb = g[:, 3].contiguous(), g[:, :3].contiguous()
        grads = [torch.zeros_like(p) for p in params]
        ps, gs = spec.params_struct(params), spec.grads_struct(grads)
        _lib.call("nrf_mlp_fp32_bwd", byref(ps), byref(gs), _p(x), rows, 0, 0, int(sigma_only), _p(d_sigma),
                  _p(d_rgb), _p(rgb), _p(ws), ws.numel(), _stream())
        return (None, None, None, *grads)
