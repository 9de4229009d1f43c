"""Thin custom-op layer: torch tensors in, C-ABI calls (include/nerf_b200.h) out.

PyTorch is used for device memory, streams and autograd bookkeeping only; every numerical
operation of the hot path is a kernel in libnerf_b200.so.  CPU tensors raise — there is no
fallback path by design."""
import os
from ctypes import byref

import torch

from . import _lib

_PRECISION = os.environ.get("NERF_B200_MLP", "bf16")


def set_mlp_precision(mode: str):
    """'bf16' = tcgen05 tensor-core MLP (2e-3 abs end-to-end); 'fp32' = FFMA parity mode (1e-5)."""
    global _PRECISION
    if mode not in ("bf16", "fp32"):
        raise ValueError("precision must be 'bf16' or 'fp32'")
    _PRECISION = mode


def get_mlp_precision() -> str:
    return _PRECISION


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _dev(t: torch.Tensor, name: str, dtype=torch.float32):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: nerf_pl_b200 has no CPU path")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def _p(t):
    return None if t is None else t.data_ptr()


def check_current_device(t: torch.Tensor, name: str):
    """The launchers use torch's CURRENT stream: a tensor on another GPU than the current device would be handed to a
    kernel running on the wrong GPU.  One process per GPU with torch.cuda.set_device(local_rank) (or a
    `with torch.cuda.device(...)` block) is the supported arrangement; anything else raises here."""
    if t.is_cuda and t.device.index != torch.cuda.current_device():
        raise RuntimeError(f"{name} lives on {t.device} but the current CUDA device is cuda:{torch.cuda.current_device()}: "
                           "call torch.cuda.set_device() (one process per GPU) or wrap the call in torch.cuda.device(...)")


# --------------------------------------------------------------------------------------------
# MLP parameter plumbing
# --------------------------------------------------------------------------------------------

class MlpSpec:
    """Shape of a NeRF MLP + the order in which its parameters are passed around."""

    def __init__(self, D, W, in_xyz, in_dir, skips):
        self.D, self.W, self.in_xyz, self.in_dir = int(D), int(W), int(in_xyz), int(in_dir)
        self.skips = tuple(int(s) for s in skips)
        if self.D > _lib.NRF_MAX_D:
            raise ValueError(f"D={D} > {_lib.NRF_MAX_D}")
        self.names = []
        for i in range(self.D):
            self.names += [f"xyz_encoding_{i + 1}.0.weight", f"xyz_encoding_{i + 1}.0.bias"]
        self.names += ["xyz_encoding_final.weight", "xyz_encoding_final.bias", "dir_encoding.0.weight",
                       "dir_encoding.0.bias", "sigma.weight", "sigma.bias", "rgb.0.weight", "rgb.0.bias"]

    @property
    def skip_mask(self):
        m = 0
        for s in self.skips:
            if 0 < s < self.D:
                m |= 1 << s
        return m

    @property
    def tc_capable(self):
        return (self.D, self.W, self.in_xyz, self.in_dir, self.skips) == (8, 256, 63, 27, (4,))

    def fill(self, struct, tensors):
        """Write the data pointers of `tensors` (ordered like self.names) into a C struct."""
        D = self.D
        for i in range(D):
            struct.xyz_w[i] = tensors[2 * i].data_ptr()
            struct.xyz_b[i] = tensors[2 * i + 1].data_ptr()
        rest = tensors[2 * D:]
        (struct.final_w, struct.final_b, struct.dir_w, struct.dir_b,
         struct.sigma_w, struct.sigma_b, struct.rgb_w, struct.rgb_b) = [t.data_ptr() for t in rest]
        return struct

    def params_struct(self, tensors):
        s = _lib.MlpParams()
        s.D, s.W, s.in_xyz, s.in_dir, s.skip_mask = self.D, self.W, self.in_xyz, self.in_dir, self.skip_mask
        return self.fill(s, tensors)

    def grads_struct(self, tensors):
        return self.fill(_lib.MlpGrads(), tensors)


def _check_params(spec, params):
    out = []
    for n, t in zip(spec.names, params):
        out.append(_dev(t, n))
    return out


# --------------------------------------------------------------------------------------------
# Stand-alone operators
# --------------------------------------------------------------------------------------------

_T_STEPS = {}


def sample_coarse(rays, n_samples, use_disp, perturb, u=None, t_steps=None):
    rays = _dev(rays, "rays")
    N = rays.shape[0]
    if t_steps is None:
        key = (n_samples, rays.device)
        if key not in _T_STEPS:                   # constant of the call: one linspace per (Nc, device), not per step
            _T_STEPS[key] = torch.linspace(0, 1, n_samples, device=rays.device)
        t_steps = _T_STEPS[key]
    t_steps = _dev(t_steps, "t_steps")
    if perturb > 0:
        u = _dev(u, "u")
        assert u.shape == (N, n_samples)
    z = torch.empty(N, n_samples, device=rays.device, dtype=torch.float32)
    _lib.call("nrf_sample_coarse", _p(rays), _p(t_steps), N, n_samples, int(bool(use_disp)), float(perturb),
              _p(u) if perturb > 0 else None, _p(z), _stream())
    return z


def sample_pdf(rays, weights_coarse, n_importance, u, jitter, eps=1e-5, return_index=False, return_cdf=False):
    rays, w = _dev(rays, "rays"), _dev(weights_coarse, "weights")
    u, jitter = _dev(u, "u"), _dev(jitter, "jitter")
    N, nc = w.shape
    assert u.shape == (N, n_importance) and jitter.shape == (N, n_importance)
    z = torch.empty(N, n_importance, device=w.device, dtype=torch.float32)
    idx = torch.empty(N, n_importance, device=w.device, dtype=torch.int64) if return_index else None
    cdf = torch.empty(N, nc - 1, device=w.device, dtype=torch.float32) if return_cdf else None
    _lib.call("nrf_sample_pdf", _p(rays), _p(w), N, nc, n_importance, _p(u), _p(jitter), float(eps), _p(z),
              _p(idx), _p(cdf), _stream())
    out = (z,)
    if return_index:
        out += (idx,)
    if return_cdf:
        out += (cdf,)
    return out if len(out) > 1 else z


def merge_sort(z_coarse, z_fine):
    zc, zf = _dev(z_coarse, "z_coarse"), _dev(z_fine, "z_fine")
    N = zc.shape[0]
    out = torch.empty(N, zc.shape[1] + zf.shape[1], device=zc.device, dtype=torch.float32)
    _lib.call("nrf_merge_sort", _p(zc), _p(zf), N, zc.shape[1], zf.shape[1], _p(out), _stream())
    return out


_SEARCHSORTED_DTYPES = {torch.float32: 0, torch.float64: 1, torch.int64: 2, torch.int32: 3, torch.int16: 4, torch.int8: 5,
                        torch.uint8: 6}


def _dev_any(t, name):
    """Contiguous CUDA tensor of any dtype (no CPU path)."""
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (nerf_pl_b200 has no CPU path)")
    return t.contiguous()


def searchsorted(a, v, out=None, side="left"):
    """Drop-in for torchsearchsorted.searchsorted (src/torchsearchsorted/searchsorted.py:20-53)."""
    assert len(a.shape) == 2, "input `a` must be 2-D."
    assert len(v.shape) == 2, "input `v` mus(t be 2-D."
    assert (a.shape[0] == v.shape[0] or a.shape[0] == 1 or v.shape[0] == 1), \
        "`a` and `v` must have the same number of rows or one of them must have only one "
    assert a.device == v.device, "`a` and `v` must be on the same device"
    shape = (max(a.shape[0], v.shape[0]), v.shape[1])
    if out is not None:
        assert out.device == a.device, "`out` must be on the same device as `a`"
        assert out.dtype == torch.long, "out.dtype must be torch.long"
        assert tuple(out.shape) == shape, "If the output tensor is provided, its shape must be correct."
        assert out.is_contiguous()
    else:
        out = torch.empty(shape, device=v.device, dtype=torch.long)
    if a.dtype != v.dtype:
        raise RuntimeError(f"expected `a` and `v` of the same dtype, got {a.dtype} and {v.dtype}")
    code = _SEARCHSORTED_DTYPES.get(a.dtype)
    if code is None:      # the reference's dispatch error (AT_DISPATCH_ALL_TYPES, searchsorted_cuda_kernel.cu:134)
        raise RuntimeError(f'"searchsorted cuda" not implemented for \'{str(a.dtype).replace("torch.", "")}\'')
    a, v = _dev_any(a, "a"), _dev_any(v, "v")
    _lib.call("nrf_searchsorted_typed", _p(a), _p(v), _p(out), a.shape[0], v.shape[0], a.shape[1], v.shape[1],
              1 if side == "left" else 0, code, _stream())
    return out


class _Embed(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, freqs):
        x = _dev(x, "x")
        rows, C = x.shape
        L = freqs.numel()
        out = torch.empty(rows, C * (2 * L + 1), device=x.device, dtype=torch.float32)
        _lib.call("nrf_embed_fwd", _p(x), rows, C, _p(freqs), L, _p(out), _stream())
        ctx.save_for_backward(x, freqs)
        return out

    @staticmethod
    def backward(ctx, g):
        x, freqs = ctx.saved_tensors
        g = _dev(g, "grad")
        gx = torch.empty_like(x)
        _lib.call("nrf_embed_bwd", _p(x), _p(g), x.shape[0], x.shape[1], _p(freqs), freqs.numel(), _p(gx), _stream())
        return gx, None


def embed(x, freqs):
    lead = x.shape[:-1]
    out = _Embed.apply(x.reshape(-1, x.shape[-1]), _dev(freqs, "freqs"))
    return out.reshape(*lead, out.shape[-1])


def composite(sigma, rgb, z, rays, noise, noise_std, white_back, want_disp=True):
    """Forward-only compositing (used by tests and the no-grad paths)."""
    sigma, z, rays = _dev(sigma, "sigma"), _dev(z, "z"), _dev(rays, "rays")
    N, S = sigma.shape
    rgb = None if rgb is None else _dev(rgb, "rgb")
    noise = None if noise is None else _dev(noise, "noise")
    dev = sigma.device
    w = torch.empty(N, S, device=dev)
    rgb_o = torch.empty(N, 3, device=dev) if rgb is not None else None
    depth, opac = torch.empty(N, device=dev), torch.empty(N, device=dev)
    disp = torch.empty(N, device=dev) if want_disp else None
    _lib.call("nrf_composite_fwd", _p(sigma), _p(rgb), _p(z), _p(rays), _p(noise), float(noise_std),
              int(bool(white_back)), N, S, _p(w), _p(rgb_o), _p(depth), _p(opac), _p(disp), _stream())
    return {"weights": w, "rgb": rgb_o, "depth": depth, "opacity": opac, "disp": disp}


def composite_bwd(sigma, rgb, z, rays, noise, noise_std, white_back, g_rgb, g_depth, g_opacity, g_disp, g_weights=None):
    N, S = sigma.shape
    d_sigma = torch.empty_like(sigma)
    d_rgb = torch.empty_like(rgb) if rgb is not None else None
    cont = lambda t: None if t is None else _dev(t, "grad")
    g_rgb, g_depth, g_opacity, g_disp, g_weights = map(cont, (g_rgb, g_depth, g_opacity, g_disp, g_weights))
    _lib.call("nrf_composite_bwd", _p(sigma), _p(rgb), _p(z), _p(rays), _p(noise), float(noise_std),
              int(bool(white_back)), N, S, _p(g_rgb), _p(g_depth), _p(g_opacity), _p(g_disp), _p(g_weights),
              _p(d_sigma), _p(d_rgb), _stream())
    return d_sigma, d_rgb
