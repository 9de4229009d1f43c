"""Shared implementation of the three `render_rays` variants of the reference
(models/rendering.py:84-272, rendering_shadows.py:94-281, rendering_rgb_sm.py:87-278).

Per call: sample_coarse kernel -> fused MLP(+encoding) kernel(s) -> compositing kernel ->
sample_pdf kernel -> per-ray sort kernel -> fused MLP -> compositing.  The reference's two
chunk loops (train.py:53, rendering.py:151) exist to bound activation memory of the unfused
PyTorch graph; here no (N*S, 90) tensor is ever materialised, so `chunk` is accepted and ignored.

Random numbers come from torch's CUDA generator with the reference's DISTRIBUTIONS but not its stream order:
one fused uniform draw and one fused normal draw per call, sliced per use (the reference makes five separate
draws: rendering.py:231, :178, :36, :41, :178), so a seeded run sees different jitter / noise values than the
reference.  Bit-parity checks inject the random tensors through the private `_inject` dict."""
from types import SimpleNamespace

import torch

from .. import ops
from ..mlp import RenderPass


def _is_pow2_bands(emb):
    # checked once per (buffer, version): a device-resident freq_bands would cost a host sync per render otherwise
    # (and a sync is illegal while a CUDA graph is being captured)
    fb = emb.freq_bands
    tag = (fb.data_ptr(), fb._version, int(emb.N_freqs))
    cache = emb.__dict__.get("_nrf_pow2")
    if cache is None or cache[0] != tag:
        ok = bool(torch.equal(fb.cpu().float(), 2.0 ** torch.arange(emb.N_freqs, dtype=torch.float32)))
        emb.__dict__["_nrf_pow2"] = cache = (tag, ok)
    return cache[1]


def _model_cfg(model, embeddings, sigma_only, white_back, noise_std):
    spec = getattr(model, "spec", None)
    if spec is None:
        raise TypeError("render_rays needs nerf_pl_b200.models.nerf.NeRF models (got %r)" % type(model).__name__)
    ex, ed = embeddings[0], embeddings[1]
    if not (_is_pow2_bands(ex) and _is_pow2_bands(ed)):
        raise RuntimeError("the fused renderer supports logscale Embedding frequency bands only")
    if ex.out_channels != spec.in_xyz or (not sigma_only and ed.out_channels != spec.in_dir):
        raise RuntimeError("Embedding out_channels do not match the NeRF in_channels")
    precision = ops.get_mlp_precision()
    packed = model.packed_weights() if (precision == "bf16" and spec.tc_capable) else None
    return SimpleNamespace(spec=spec, sigma_only=bool(sigma_only), white_back=bool(white_back),
                           noise_std=float(noise_std), precision=precision, packed=packed,
                           xyz_freqs=int(ex.N_freqs), dir_freqs=int(ed.N_freqs),
                           grad_sinks=model.__dict__.get("_nrf_grad_sinks"),
                           need_grad=torch.is_grad_enabled() and any(p.requires_grad for p in model.parameters()))


def sample_pdf(rays, weights, N_importance, det=False, eps=1e-5, _u=None, _jitter=None):
    """The fork's sampler (models/rendering.py:14-48): `weights` is (N, n_bins) =
    weights_coarse[:, 1:-1]; `det` is ignored, as in the reference."""
    N = weights.shape[0]
    dev = weights.device
    u = torch.rand(N, N_importance, device=dev) if _u is None else _u
    jit = torch.rand(N, N_importance, device=dev) if _jitter is None else _jitter
    full = torch.nn.functional.pad(weights.detach(), (1, 1))
    return ops.sample_pdf(rays, full, N_importance, u, jit, eps)


def render_rays_impl(variant, models, embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance,
                     chunk, white_back, test_time, _inject=None):
    if not isinstance(rays, torch.Tensor) or rays.dim() != 2 or rays.shape[1] != 8:
        raise RuntimeError("rays must be a (N_rays, 8) tensor [o, d, near, far]")
    rays = ops._dev(rays.detach(), "rays")
    ops.check_current_device(rays, "rays")
    inj = _inject or {}
    N = rays.shape[0]
    dev = rays.device
    depth_only = variant == "depth"
    with_disp = variant in ("depth", "rgb_sm")

    # Random tensors the caller did not inject: ONE uniform draw and ONE normal draw per call, sliced per use
    # (the reference makes five separate draws: rendering.py:231, :178, :36, :41, :178).
    S_fine = N_samples + N_importance
    need_u = [("u_perturb", N_samples)] if perturb > 0 else []
    if N_importance > 0 and "z_fine" not in inj:
        need_u += [("u_fine", N_importance), ("jitter_fine", N_importance)]
    need_n = [("noise_coarse", N_samples)] + ([("noise_fine", S_fine)] if N_importance > 0 else []) if noise_std != 0 else []
    drawn = {}
    for need, fn in ((need_u, torch.rand), (need_n, torch.randn)):
        need = [(k, w) for k, w in need if k not in inj]
        if need:
            pad = lambda v: (v + 63) // 64 * 64               # every slice starts 256-byte aligned (vector loads)
            flat = fn(sum(pad(N * w) for _, w in need), device=dev)
            off = 0
            for k, w in need:
                drawn[k] = flat[off:off + N * w].view(N, w)
                off += pad(N * w)

    def rnd(key):
        return inj[key] if key in inj else drawn[key]

    u = rnd("u_perturb") if perturb > 0 else None
    z = ops.sample_coarse(rays, N_samples, use_disp, perturb, u)

    def run(model, z_vals, noise_key, sigma_only):
        S = z_vals.shape[1]
        cfg = _model_cfg(model, embeddings, sigma_only, white_back, noise_std)
        return RenderPass.apply(rays, z_vals, rnd(noise_key) if noise_std != 0 else None, cfg, *model.ordered_parameters())

    rgb_c, depth_c, opac_c, disp_c, w_c = run(models[0], z, "noise_coarse", test_time or depth_only)
    if test_time:
        result = {"opacity_coarse": opac_c}
    else:
        result = {}
        if not depth_only:
            result["rgb_coarse"] = rgb_c
        result["depth_coarse"] = depth_c
        result["opacity_coarse"] = opac_c
        if with_disp:
            result["disp_map_coarse"] = disp_c

    if N_importance > 0:
        if "z_fine" in inj:
            z_fine = inj["z_fine"]
        else:
            z_fine = ops.sample_pdf(rays, w_c, N_importance, rnd("u_fine"), rnd("jitter_fine"))
        z_all = ops.merge_sort(z, z_fine)
        rgb_f, depth_f, opac_f, disp_f, _ = run(models[1], z_all, "noise_fine", depth_only)
        if not depth_only:
            result["rgb_fine"] = rgb_f
        result["depth_fine"] = depth_f
        result["opacity_fine"] = opac_f
        if with_disp:
            result["disp_map_fine"] = disp_f
    return result
