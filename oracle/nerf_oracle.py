"""CPU oracle for the nerf_pl ray-rendering hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, in plain fp32 torch-on-CPU tensor arithmetic, the algorithm of
the reference (ktiwary2/nerf_pl @ 99d40cb).  It is the *checker* for the CUDA path:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it.  Nothing under ``nerf_pl_b200/`` imports it
and the product path has no CPU fallback.

Parity pin: every function here is checked against outputs of the reference's own
``models/*.py`` (imported from /root/reference by ``tests/make_golden.py`` in the
build container, fixtures committed under ``tests/golden/``) by
``tests/test_oracle_golden.py``.  Parity is therefore *pinned* (see DESIGN.md §3).

Unlike the reference, every random tensor is an explicit argument, so that the
oracle and the CUDA kernels can be driven with identical noise.  The order in which
the reference draws them is (rendering.py:231, :178, :36, :41, :178):
``u_perturb (N,Nc)`` -> ``noise_coarse (N,Nc)`` -> ``u_fine (N,Nf)`` ->
``jitter_fine (N,Nf)`` -> ``noise_fine (N,Nc+Nf)``.

All ``file:line`` citations are relative to /root/reference.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

# Synthetic-input generators (random-init parameters in the checkpoint layout, Blender-style rays, the random tensors
# in the reference's draw order) live in a neutral module so that bench.py's GPU arm does not depend on oracle/.
from nerf_pl_b200.synthetic import blender_rays, init_mlp_params, make_noise  # noqa: F401

Tensor = torch.Tensor


# ---------------------------------------------------------------------------
# models/nerf.py
# ---------------------------------------------------------------------------

def embed(x: Tensor, n_freqs: int, logscale: bool = True) -> Tensor:
    """Positional encoding (models/nerf.py:21-38).

    Output channel order: x, then for k in 0..L-1: sin(f_k x), cos(f_k x), with
    f_k = 2**k (logscale, nerf.py:16-17) or linspace(1, 2**(L-1), L) (nerf.py:19).
    """
    if logscale:
        freqs = 2.0 ** torch.linspace(0, n_freqs - 1, n_freqs)
    else:
        freqs = torch.linspace(1, 2 ** (n_freqs - 1), n_freqs)
    pieces = [x]
    for f in freqs:
        fx = f * x
        pieces.append(torch.sin(fx))
        pieces.append(torch.cos(fx))
    return torch.cat(pieces, dim=-1)


def mlp_param_names(D: int = 8) -> List[str]:
    """state_dict keys in the reference order (models/nerf.py:57-81)."""
    names: List[str] = []
    for i in range(D):
        names += [f"xyz_encoding_{i + 1}.0.weight", f"xyz_encoding_{i + 1}.0.bias"]
    names += ["xyz_encoding_final.weight", "xyz_encoding_final.bias",
              "dir_encoding.0.weight", "dir_encoding.0.bias",
              "sigma.weight", "sigma.bias", "rgb.0.weight", "rgb.0.bias"]
    return names


def mlp_forward(p: Dict[str, Tensor], x: Tensor, sigma_only: bool = False,
                D: int = 8, skips: Sequence[int] = (4,), in_xyz: int = 63) -> Tensor:
    """NeRF.forward (models/nerf.py:83-124).  Skip concat puts the encoding FIRST
    (nerf.py:108-109); sigma is raw (nerf.py:112); rgb is sigmoid (nerf.py:79-81)."""
    F = torch.nn.functional
    enc = x[..., :in_xyz]
    h = enc
    for i in range(D):
        if i in skips:
            h = torch.cat([enc, h], dim=-1)
        h = torch.relu(F.linear(h, p[f"xyz_encoding_{i + 1}.0.weight"], p[f"xyz_encoding_{i + 1}.0.bias"]))
    sigma = F.linear(h, p["sigma.weight"], p["sigma.bias"])
    if sigma_only:
        return sigma
    feat = F.linear(h, p["xyz_encoding_final.weight"], p["xyz_encoding_final.bias"])
    d = torch.relu(F.linear(torch.cat([feat, x[..., in_xyz:]], dim=-1),
                            p["dir_encoding.0.weight"], p["dir_encoding.0.bias"]))
    rgb = torch.sigmoid(F.linear(d, p["rgb.0.weight"], p["rgb.0.bias"]))
    return torch.cat([rgb, sigma], dim=-1)


# ---------------------------------------------------------------------------
# models/rendering.py
# ---------------------------------------------------------------------------

def stratified_z(rays: Tensor, n_samples: int, use_disp: bool, perturb: float,
                 u_perturb: Optional[Tensor]) -> Tensor:
    """Coarse depths (models/rendering.py:216-232)."""
    near, far = rays[:, 6:7], rays[:, 7:8]
    t = torch.linspace(0, 1, n_samples, device=rays.device)
    if use_disp:
        z = 1 / (1 / near * (1 - t) + 1 / far * t)
    else:
        z = near * (1 - t) + far * t
    z = z.expand(rays.shape[0], n_samples)
    if perturb > 0:
        mid = 0.5 * (z[:, :-1] + z[:, 1:])
        hi = torch.cat([mid, z[:, -1:]], -1)
        lo = torch.cat([z[:, :1], mid], -1)
        z = lo + (hi - lo) * (perturb * u_perturb)
    return z


def composite(sigmas: Tensor, rgbs: Optional[Tensor], z: Tensor, dirs: Tensor,
              noise: Optional[Tensor], noise_std: float, white_back: bool):
    """Volume-rendering integral (models/rendering.py:170-198; disp from
    rendering_shadows.py:198 / rendering_rgb_sm.py:197).

    Returns dict(weights, opacity, depth, disp[, rgb]).
    """
    delta = z[:, 1:] - z[:, :-1]
    delta = torch.cat([delta, torch.full_like(delta[:, :1], 1e10)], -1)
    delta = delta * torch.norm(dirs.unsqueeze(1), dim=-1)
    raw = sigmas if noise is None else sigmas + noise * noise_std
    alpha = 1 - torch.exp(-delta * torch.relu(raw))
    trans = torch.cumprod(torch.cat([torch.ones_like(alpha[:, :1]), 1 - alpha + 1e-10], -1), -1)[:, :-1]
    w = alpha * trans
    out = {"weights": w, "opacity": w.sum(1), "depth": (w * z).sum(-1), "raw": raw, "rgbs": rgbs}
    out["disp"] = 1.0 / torch.max(1e-10 * torch.ones_like(out["depth"]), out["depth"] / w.sum(-1))
    if rgbs is not None:
        rgb = (w.unsqueeze(-1) * rgbs).sum(-2)
        if white_back:
            rgb = rgb + 1 - out["opacity"].unsqueeze(-1)
        out["rgb"] = rgb
    return out


def pdf_cdf(weights_mid: Tensor, eps: float = 1e-5) -> Tensor:
    """cdf of sample_pdf (models/rendering.py:29-34): (N, n_bins+1), cdf[:,0]=0."""
    w = weights_mid + eps
    pdf = w / torch.sum(w, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    return torch.cat([torch.zeros_like(cdf[:, :1]), cdf], -1)


def searchsorted_rows(a: np.ndarray, v: np.ndarray, side: str = "left") -> np.ndarray:
    """Row-wise searchsorted with 1-row broadcast, int64 result.

    Semantics of torchsearchsorted (src/torchsearchsorted/searchsorted.py:20-53,
    src/cpu/searchsorted_cpu_wrapper.cpp:41-122): result = number of a[row,:] that are
    < v (side='left') or <= v (side='right').  The reference's own known-answer test
    (torchsearchsorted/test/test_searchsorted.py:27-44) defines exactly this via
    numpy.searchsorted, which is what we use."""
    rows = max(a.shape[0], v.shape[0])
    out = np.empty((rows, v.shape[1]), dtype=np.int64)
    for r in range(rows):
        ar = a[0] if a.shape[0] == 1 else a[r]
        vr = v[0] if v.shape[0] == 1 else v[r]
        out[r] = np.searchsorted(ar, vr, side=side)
    return out


def sample_pdf(rays: Tensor, weights_mid: Tensor, n_importance: int, u: Tensor, jitter: Tensor,
               eps: float = 1e-5, return_index: bool = False):
    """The fork's pixel-NeRF-style sampler (models/rendering.py:14-48): bin index by
    upper-bound search in the cdf, second uniform jitter, linear map of
    (index+jitter)/n_bins between near and far.  `det` is ignored by the reference."""
    n_bins = weights_mid.shape[1]
    cdf = pdf_cdf(weights_mid, eps)
    idx = torch.searchsorted(cdf, u.contiguous(), right=True)
    ind = torch.clamp_min(idx.float() - 1.0, 0.0)
    t = (ind + jitter) / n_bins
    near, far = rays[:, -2:-1], rays[:, -1:]
    z = near * (1 - t) + far * t
    if return_index:
        return z, ind.long()
    return z


def render_rays(params: Sequence[Dict[str, Tensor]], rays: Tensor, *, n_samples: int = 64,
                use_disp: bool = False, perturb: float = 0.0, noise_std: float = 1.0,
                n_importance: int = 0, white_back: bool = False, test_time: bool = False,
                variant: str = "rgb", noise: Optional[Dict[str, Tensor]] = None,
                xyz_freqs: int = 10, dir_freqs: int = 4, D: int = 8, skips=(4,),
                return_aux: bool = False) -> Dict[str, Tensor]:
    """render_rays of the three reference variants.

    variant='rgb'      models/rendering.py:84-272        keys rgb_*/depth_*/opacity_*
    variant='depth'    models/rendering_shadows.py:94-281 sigma-only MLP, keys depth_*/opacity_*/disp_map_*
    variant='rgb_sm'   models/rendering_rgb_sm.py:87-278  'rgb' + disp_map_*

    `noise` holds the injected random tensors (see module docstring); an optional
    key 'z_fine' overrides the output of sample_pdf (used to compare end to end with
    bin flips excluded).
    """
    noise = noise or {}
    in_xyz = 3 * (2 * xyz_freqs + 1)
    N = rays.shape[0]
    o, d = rays[:, 0:3], rays[:, 3:6]
    dir_enc = embed(d, dir_freqs)
    depth_only = variant == "depth"

    def run_pass(p, z, nz, weights_only):
        S = z.shape[1]
        pts = (o.unsqueeze(1) + d.unsqueeze(1) * z.unsqueeze(2)).reshape(-1, 3)
        x = embed(pts, xyz_freqs)
        sig_only = weights_only or depth_only
        if not sig_only:
            x = torch.cat([x, torch.repeat_interleave(dir_enc, S, dim=0)], 1)
        raw = mlp_forward(p, x, sigma_only=sig_only, D=D, skips=skips, in_xyz=in_xyz)
        if sig_only:
            sig, rgb = raw.view(N, S), None
        else:
            raw = raw.view(N, S, 4)
            sig, rgb = raw[..., 3], raw[..., :3]
        return composite(sig, rgb, z, d, nz, noise_std, white_back)

    z = stratified_z(rays, n_samples, use_disp, perturb, noise.get("u_perturb"))
    aux = {"z_coarse": z}
    c = run_pass(params[0], z, noise.get("noise_coarse"), weights_only=test_time)
    res: Dict[str, Tensor] = {}
    if test_time:
        res["opacity_coarse"] = c["opacity"]
    else:
        if not depth_only:
            res["rgb_coarse"] = c["rgb"]
        res["depth_coarse"] = c["depth"]
        res["opacity_coarse"] = c["opacity"]
        if variant in ("depth", "rgb_sm"):
            res["disp_map_coarse"] = c["disp"]
    aux["weights_coarse"] = c["weights"]
    aux["raw_coarse"] = c["raw"]
    if c["rgbs"] is not None:
        aux["rgbs_coarse"] = c["rgbs"]
    if n_importance > 0:
        if "z_fine" in noise:
            zf = noise["z_fine"]
        else:
            zf = sample_pdf(rays, c["weights"][:, 1:-1], n_importance,
                            noise["u_fine"], noise["jitter_fine"]).detach()
        aux["z_fine"] = zf
        z2, _ = torch.sort(torch.cat([z, zf], -1), -1)
        aux["z_all"] = z2
        f = run_pass(params[1], z2, noise.get("noise_fine"), weights_only=False)
        if not depth_only:
            res["rgb_fine"] = f["rgb"]
        res["depth_fine"] = f["depth"]
        res["opacity_fine"] = f["opacity"]
        if variant in ("depth", "rgb_sm"):
            res["disp_map_fine"] = f["disp"]
        aux["weights_fine"] = f["weights"]
        aux["raw_fine"] = f["raw"]
        if f["rgbs"] is not None:
            aux["rgbs_fine"] = f["rgbs"]
    if return_aux:
        return res, aux
    return res


# ---------------------------------------------------------------------------
# models/camera.py + models/efficient_shadow_mapping.py
# ---------------------------------------------------------------------------

SHADOW_EPS = 1e-5  # efficient_shadow_mapping.py:8, rendering_shadows.py:357


def camera_matrix(hfov_deg: float, res: Tuple[int, int]) -> Tensor:
    """3x3 planar-pinhole matrix with columns a,b,c (models/camera.py:20-31)."""
    w, h = res
    hf = torch.tensor(hfov_deg) / torch.tensor(180.0) * np.pi
    a = torch.tensor([1.0, 0.0, 0.0])
    b = torch.tensor([0.0, -1.0, 0.0])
    c = torch.tensor([-w / 2.0, h / 2, -w / (2 * torch.tan(hf / 2.0))])
    return torch.stack([a, b, c]).T


def pose_camera(cam: Tensor, c2w: Tensor) -> Tuple[Tensor, Tensor]:
    """(eye_pos, M) for a Blender c2w (3,4) (models/camera.py:69-89, non-transform branch)."""
    return c2w[:, 3].float(), c2w[:, :3].float() @ cam.float()


def transformation_to(cam_M: Tensor, cam_eye: Tensor, light_M: Tensor, light_eye: Tensor):
    """R = ML^-1 MC, Q = ML^-1 (OC - OL) (models/camera.py:121-132)."""
    inv = torch.inverse(light_M)
    return inv @ cam_M, inv @ (cam_eye - light_eye)


def normed_w(M: Tensor, pixel_depth: Tensor) -> Tensor:
    """[i,j,1,depth] -> [i,j,1, depth/(|M [i,j,1]^T| + 1e-5)] (efficient_shadow_mapping.py:47-62)."""
    pix = pixel_depth[:, :3]
    coords = (pix[:, None, :] * M).sum(-1)
    nrm = torch.linalg.norm(coords, dim=1) + SHADOW_EPS
    return torch.cat([pix, (pixel_depth[:, 3] / nrm).view(-1, 1)], 1)


def diff_projections(pix: Tensor, w_cam: Tensor, R: Tensor, Q: Tensor) -> Tensor:
    """K=[ul,vl,wl] (efficient_shadow_mapping.py:65-83)."""
    proj = (pix[:, None, :] * R).sum(-1)
    uvw = w_cam[:, None] * proj + Q
    wl = uvw[:, 2]
    return torch.stack([uvw[:, 0] / wl, uvw[:, 1] / wl, wl], 1)


def projected_depths(res: Tuple[int, int], K: Tensor, w_light: Tensor):
    """clamp / truncate / gather (efficient_shadow_mapping.py:85-100).  Row index is vl."""
    w, h = res
    ul = K[:, 0].clamp(min=0.0).clamp(max=w - 1.0)
    vl = K[:, 1].clamp(min=0.0).clamp(max=h - 1.0)
    return K[:, 2], w_light.view(w, h)[vl.long(), ul.long()]


def shadow_map(wl: Tensor, w_light_b: Tensor, mode: str, delta=1e-2, epsilon=0.0,
               new_min=0.0, new_max=1.0, sigmoid=False) -> Tensor:
    """efficient_shadow_mapping.py:103-130 (+ normalize_min_max :10-11)."""
    diff = wl - w_light_b
    if mode == "shadow_method_1":
        diff = torch.max(diff / delta, torch.tensor(epsilon))
    elif mode == "shadow_method_2":
        # the reference calls normalize_min_max(diff) without forwarding new_min / new_max (:124): they have no effect
        diff = (diff - diff.min()) / (diff.max() - diff.min() + SHADOW_EPS)
        if sigmoid:
            diff = torch.sigmoid(diff)
    else:
        raise ValueError("{} not found".format(mode))
    return torch.stack([diff, diff, diff], 1).clip(0.0, 1.0)


def run_shadow_mapping(res, cam_M, cam_eye, light_M, light_eye, mesh_range_cam, normed_light,
                       mode="shadow_method_1", **kw) -> Tensor:
    """efficient_shadow_mapping.py:19-44 for one same-camera segment."""
    wc = normed_w(cam_M, mesh_range_cam)
    R, Q = transformation_to(cam_M, cam_eye, light_M, light_eye)
    K = diff_projections(wc[:, :3], wc[:, 3], R, Q)
    wl, wlb = projected_depths(res, K, normed_light[:, 3])
    return shadow_map(wl, wlb, mode, **kw)


def camera_segments(eye_pos: Tensor) -> List[Tuple[int, int]]:
    """Consecutive same-eye_pos runs, as found by the Python loop at
    rendering_shadows.py:385-406 (a new segment starts where eye_pos differs from the
    eye_pos of the current segment's first ray)."""
    n = eye_pos.shape[0]
    segs, start = [], 0
    for i in range(n):
        if not torch.equal(eye_pos[start], eye_pos[i]):
            segs.append((start, i))
            start = i
    segs.append((start, n))
    return segs


def efficient_sm(cam_pixels, light_pixels, cam_results, light_results, ppc, light_M, light_eye,
                 image_shape, fine_sampling, light_n_importance, shadow_method,
                 out_prefix: str = "rgb") -> Dict[str, Tensor]:
    """rendering_shadows.py:359-482 (out_prefix='rgb') and its twin
    rendering_rgb_sm.py:356-480 (out_prefix='sm')."""
    eye, cams = ppc["eye_pos"], ppc["camera"]
    segs = camera_segments(eye)

    def run(cam_depth, light_depth):
        mrc = torch.cat([cam_pixels, cam_depth.view(-1, 1)], 1)
        nl = normed_w(light_M, torch.cat([light_pixels, light_depth.view(-1, 1)], 1))
        outs = [run_shadow_mapping(image_shape, cams[s], eye[s], light_M, light_eye, mrc[s:e], nl,
                                   mode=shadow_method) for (s, e) in segs]
        return torch.cat(outs, 0).view(-1, 3) + SHADOW_EPS

    cam_results[f"{out_prefix}_coarse"] = run(cam_results["depth_coarse"], light_results["depth_coarse"])
    if fine_sampling:
        ld = light_results["depth_fine"] if light_n_importance else light_results["depth_coarse"]
        cam_results[f"{out_prefix}_fine"] = run(cam_results["depth_fine"], ld)
    return cam_results


def normalize_min_max(t: Tensor, new_max: float = 1.0, new_min: float = 0.0) -> Tensor:
    """efficient_shadow_mapping.py:10-11"""
    return (t - t.min()) / (t.max() - t.min() + SHADOW_EPS) * (new_max - new_min) + new_min


def projections(cam_M: Tensor, cam_eye: Tensor, light_M: Tensor, light_eye: Tensor, mesh_range_cam: Tensor) -> Tensor:
    """get_projections (efficient_shadow_mapping.py:13-17) for one camera: get_normed_w then get_diff_projections."""
    wc = normed_w(cam_M, mesh_range_cam)
    R, Q = transformation_to(cam_M, cam_eye, light_M, light_eye)
    return diff_projections(wc[:, :3], wc[:, 3], R, Q)


def get_K(cam_pixels: Tensor, cam_results: Dict[str, Tensor], ppc, light_M: Tensor, light_eye: Tensor,
          fine_sampling: bool):
    """rendering_shadows.py:484-546: per same-eye_pos run (each with the camera of its first ray) the projection
    [ul, vl, wl] of every camera ray into the light's image, for depth_coarse and (optionally) depth_fine."""
    eye, cams = ppc["eye_pos"], ppc["camera"]
    segs = camera_segments(eye)

    def run(depth):
        mrc = torch.cat([cam_pixels, depth.view(-1, 1)], 1)
        return torch.cat([projections(cams[s], eye[s], light_M, light_eye, mrc[s:e]) for (s, e) in segs], 0).view(-1, 3)

    coarse = run(cam_results["depth_coarse"])
    return coarse, (run(cam_results["depth_fine"]) if fine_sampling else None)


# ---------------------------------------------------------------------------
# models/shadow_mapping_utils.py: whole-image shadow mapping (train_shadow_mapping.py)
# ---------------------------------------------------------------------------

def image_pixels(res: Tuple[int, int]) -> Tensor:
    """(w,h,3) grid [i+.5, j+.5, 1] with i the column index (shadow_mapping_utils.py:46-58)."""
    w, h = res
    i = (torch.arange(h, dtype=torch.float32) + 0.5).view(1, h).expand(w, h)
    j = (torch.arange(w, dtype=torch.float32) + 0.5).view(w, 1).expand(w, h)
    return torch.stack([i, j, torch.ones(w, h)], -1)


def image_w_mapping(res, M: Tensor, range_img: Tensor) -> Tensor:
    """range (w,h) -> w = range / |M pix| -- no epsilon (shadow_mapping_utils.py:39-66)."""
    coords = (image_pixels(res)[..., None, :] * M).sum(-1)
    return range_img / torch.linalg.norm(coords, dim=2)


def image_projections(res, w_cam: Tensor, cam_M, cam_eye, light_M, light_eye) -> Tensor:
    """K (w,h,3) (shadow_mapping_utils.py:68-104)."""
    R, Q = transformation_to(cam_M, cam_eye, light_M, light_eye)
    proj = (image_pixels(res)[..., None, :] * R).sum(-1)
    uvw = w_cam[..., None] * proj + Q
    wl = uvw[..., 2]
    return torch.stack([uvw[..., 0] / wl, uvw[..., 1] / wl, wl], 2)


def image_depths(res, K: Tensor, w_light: Tensor):
    """(wl, w_light_bounded, ul_, vl_) (shadow_mapping_utils.py:106-147); K[j,i] is the identity on a square grid."""
    w, h = res
    ul_, vl_, wl = K.unbind(2)
    ul = ul_.clamp(min=0.0).clamp(max=w - 1.0).long()
    vl = vl_.clamp(min=0.0).clamp(max=h - 1.0).long()
    return wl, w_light[vl, ul], ul_, vl_


def image_normalize_min_max(t: Tensor, new_max: float = 1.0, new_min: float = 0.0) -> Tensor:
    """shadow_mapping_utils.py:8-9 (no epsilon in the denominator)."""
    return (t - t.min()) / (t.max() - t.min()) * (new_max - new_min) + new_min


def image_shadow_map(wl: Tensor, w_light_b: Tensor, mode: str, delta=1e-2, epsilon=0.0, sigmoid=False) -> Tensor:
    """(w,h,3) (shadow_mapping_utils.py:149-174)."""
    diff = wl - w_light_b
    if mode == "shadow_method_1":
        diff = torch.max(diff / delta, torch.tensor(epsilon))
    elif mode == "shadow_method_2":
        diff = image_normalize_min_max(diff)
        if sigmoid:
            diff = torch.sigmoid(diff)
    else:
        raise ValueError("{} not found".format(mode))
    return torch.stack([diff, diff, diff], 2).clip(0.0, 1.0)


def image_run_shadow_mapping(res, cam_M, cam_eye, light_M, light_eye, range_cam, range_light, mode, **kw) -> Tensor:
    """shadow_mapping_utils.py:11-36"""
    w_cam = image_w_mapping(res, cam_M, range_cam)
    w_light = image_w_mapping(res, light_M, range_light)
    K = image_projections(res, w_cam, cam_M, cam_eye, light_M, light_eye)
    wl, wlb, _, _ = image_depths(res, K, w_light)
    return image_shadow_map(wl, wlb, mode, **kw)


# ---------------------------------------------------------------------------
# Synthetic inputs: blender_rays / make_noise are imported from nerf_pl_b200.synthetic (top of file).
# ---------------------------------------------------------------------------

# --------------------------------------------------------------------------------------------
# The steps either side of the path: ray generation, MSE loss, psnr
# --------------------------------------------------------------------------------------------

def get_ray_directions(H: int, W: int, focal: float) -> Tensor:
    """datasets/ray_utils.py:5-24.  kornia.create_meshgrid(H, W, normalized_coordinates=False) is
    x = 0..W-1 along columns, y = 0..H-1 along rows (kornia 0.2.0 utils/grid.py), restated with arange."""
    i = torch.arange(W, dtype=torch.float32).expand(H, W)
    j = torch.arange(H, dtype=torch.float32).unsqueeze(1).expand(H, W)
    return torch.stack([(i - W / 2) / focal, -(j - H / 2) / focal, -torch.ones_like(i)], -1)


def get_rays(directions: Tensor, c2w: Tensor) -> Tuple[Tensor, Tensor]:
    """datasets/ray_utils.py:27-51"""
    rays_d = directions @ c2w[:, :3].T
    rays_d = rays_d / torch.norm(rays_d, dim=-1, keepdim=True)
    rays_o = c2w[:, 3].expand(rays_d.shape)
    return rays_o.reshape(-1, 3), rays_d.reshape(-1, 3)


def get_ndc_rays(H: int, W: int, focal: float, near, rays_o: Tensor, rays_d: Tensor) -> Tuple[Tensor, Tensor]:
    """datasets/ray_utils.py:52-93"""
    t = -(near + rays_o[..., 2]) / rays_d[..., 2]
    rays_o = rays_o + t[..., None] * rays_d
    ox_oz = rays_o[..., 0] / rays_o[..., 2]
    oy_oz = rays_o[..., 1] / rays_o[..., 2]
    o0 = -1. / (W / (2. * focal)) * ox_oz
    o1 = -1. / (H / (2. * focal)) * oy_oz
    o2 = 1. + 2. * near / rays_o[..., 2]
    d0 = -1. / (W / (2. * focal)) * (rays_d[..., 0] / rays_d[..., 2] - ox_oz)
    d1 = -1. / (H / (2. * focal)) * (rays_d[..., 1] / rays_d[..., 2] - oy_oz)
    d2 = 1 - o2
    return torch.stack([o0, o1, o2], -1), torch.stack([d0, d1, d2], -1)


def dataset_rays(c2w: Tensor, H: int, W: int, focal: float, near: float, far: float, ndc: bool = False) -> Tensor:
    """The (H*W, 8) rows a dataset hands to render_rays: datasets/blender.py:80-84; datasets/llff.py:232-245
    (NDC: get_ndc_rays with near plane `near`, depth bounds [0, 1])."""
    o, d = get_rays(get_ray_directions(H, W, focal), c2w)
    if ndc:
        o, d = get_ndc_rays(H, W, focal, near, o, d)
        near, far = 0.0, 1.0
    one = torch.ones_like(o[:, :1])
    return torch.cat([o, d, near * one, far * one], 1)


def mse_loss(inputs: Dict[str, Tensor], targets: Tensor, prefix: str = "rgb") -> Tensor:
    """losses.py:4-26 (MSELoss for prefix 'rgb', SMMSELoss for 'sm')"""
    loss = torch.mean((inputs[f"{prefix}_coarse"] - targets) ** 2)
    if f"{prefix}_fine" in inputs:
        loss = loss + torch.mean((inputs[f"{prefix}_fine"] - targets) ** 2)
    return loss


def psnr(image_pred: Tensor, image_gt: Tensor, valid_mask: Optional[Tensor] = None) -> Tensor:
    """metrics.py:4-13"""
    value = (image_pred - image_gt) ** 2
    if valid_mask is not None:
        value = value[valid_mask]
    return -10 * torch.log10(torch.mean(value))


def ssim(image_pred: Tensor, image_gt: Tensor, reduction: str = "mean") -> Tensor:
    """metrics.py:15-20: 1 - 2 * kornia.losses.ssim(pred, gt, 3, reduction).  kornia 0.2.0 (requirements.txt:5) is absent
    here: PARITY UNPINNED.  Published algorithm restated with explicit window loops in float64: per channel, 3x3 Gaussian
    window (sigma 1.5, normalised), zero padding, C1 = 0.01^2, C2 = 0.03^2, dssim = clamp(1 - ssim_map, 0, 1) / 2."""
    a = image_pred.detach().double().numpy()
    b = image_gt.detach().double().numpy()
    g = np.exp(-np.array([-1.0, 0.0, 1.0]) ** 2 / (2 * 1.5 ** 2))
    g /= g.sum()
    win = np.outer(g, g)

    def blur(x):
        pad = np.pad(x, ((0, 0), (0, 0), (1, 1), (1, 1)))
        out = np.zeros_like(x)
        for dy in range(3):
            for dx in range(3):
                out += win[dy, dx] * pad[:, :, dy:dy + x.shape[2], dx:dx + x.shape[3]]
        return out

    mu1, mu2 = blur(a), blur(b)
    s1, s2, s12 = blur(a * a) - mu1 ** 2, blur(b * b) - mu2 ** 2, blur(a * b) - mu1 * mu2
    smap = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 ** 2 + mu2 ** 2 + 1e-4) * (s1 + s2 + 9e-4))
    d = np.clip(1.0 - smap, 0.0, 1.0) / 2.0
    if reduction == "mean":
        d = d.mean()
    elif reduction == "sum":
        d = d.sum()
    return torch.as_tensor(1 - 2 * d)


def adam_step(param: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, step: int, lr: float,
              betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
    """One step of the optimizer utils/__init__.py:17-19 constructs: `Adam(parameters, lr, eps=1e-8, weight_decay)`.
    The algorithm lives in the reference's third-party dependency torch (requirements.txt pins torch==1.4.0): Kingma & Ba
    Adam with the weight decay folded into the gradient (L2, not decoupled), bias correction on both moments, no amsgrad.
    Restated here; tests pin it against torch.optim.Adam of the installed torch.  Returns (param, exp_avg, exp_avg_sq)."""
    b1, b2 = betas
    g = grad + weight_decay * param if weight_decay != 0 else grad
    m = b1 * exp_avg + (1 - b1) * g
    v = b2 * exp_avg_sq + (1 - b2) * g * g
    step_size = lr / (1 - b1 ** step)
    denom = v.sqrt() / math.sqrt(1 - b2 ** step) + eps
    return param - step_size * m / denom, m, v


def sgd_step(param: Tensor, grad: Tensor, buf: Optional[Tensor], lr: float, momentum: float = 0.0,
             weight_decay: float = 0.0):
    """One step of `SGD(parameters, lr, momentum, weight_decay)` (utils/__init__.py:14-16; algorithm in the reference's
    dependency torch): L2 term added to the gradient, momentum buffer seeded with the first gradient (buf None),
    dampening 0, no nesterov.  Returns (param, buf)."""
    g = grad + weight_decay * param if weight_decay != 0 else grad
    if momentum != 0:
        buf = g.clone() if buf is None else momentum * buf + g
        g = buf
    return param - lr * g, buf


def _radam_scalars(step: int, beta1: float, beta2: float):
    """(N_sma, N_sma_max, rectified step size, sgd-like step size) in Python floats (utils/optimizers.py:66-78)."""
    beta2_t = beta2 ** step
    n_sma_max = 2 / (1 - beta2) - 1
    n_sma = n_sma_max - 2 * step * beta2_t / (1 - beta2_t)
    rect = None
    if n_sma > 4:
        rect = math.sqrt((1 - beta2_t) * (n_sma - 4) / (n_sma_max - 4) * (n_sma - 2) / n_sma * n_sma_max /
                         (n_sma_max - 2)) / (1 - beta1 ** step)
    return n_sma, rect, 1.0 / (1 - beta1 ** step)


def radam_step(param: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, step: int, lr: float,
               betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0, degenerated_to_sgd: bool = True):
    """utils/optimizers.py:29-95.  Returns (param, exp_avg, exp_avg_sq)."""
    b1, b2 = betas
    v = exp_avg_sq * b2 + (1 - b2) * grad * grad
    m = exp_avg * b1 + (1 - b1) * grad
    n_sma, rect, plain = _radam_scalars(step, b1, b2)
    p = param
    if n_sma >= 5:
        if weight_decay != 0:
            p = p + (-weight_decay * lr) * p
        p = p + (-rect * lr) * (m / (v.sqrt() + eps))
    elif degenerated_to_sgd:
        if weight_decay != 0:
            p = p + (-weight_decay * lr) * p
        p = p + (-plain * lr) * m
    return p, m, v


def ranger_step(param: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, slow: Tensor, step: int, lr: float,
                alpha: float = 0.5, k: int = 6, n_sma_threshold: float = 5, betas=(0.95, 0.999), eps: float = 1e-5,
                weight_decay: float = 0.0):
    """utils/optimizers.py:322-404; `slow` starts as a copy of the initial parameters (:347-348).
    Returns (param, exp_avg, exp_avg_sq, slow)."""
    b1, b2 = betas
    v = exp_avg_sq * b2 + (1 - b2) * grad * grad
    m = exp_avg * b1 + (1 - b1) * grad
    n_sma, rect, plain = _radam_scalars(step, b1, b2)
    p = param
    if weight_decay != 0:
        p = p + (-weight_decay * lr) * p
    if n_sma > n_sma_threshold:
        p = p + (-rect * lr) * (m / (v.sqrt() + eps))
    else:
        p = p + (-plain * lr) * m
    if step % k == 0:
        slow = slow + alpha * (p - slow)
        p = slow
    return p, m, v, slow
