"""Drop-in for datasets/ray_utils.py (reference :5-93) on the GPU, plus the fused generator.

The reference builds an (H,W,3) direction grid on the host (`get_ray_directions`), rotates it per image
(`get_rays`), optionally maps to NDC (`get_ndc_rays`) and the datasets concatenate `[o, d, near, far]` rows
(datasets/blender.py:80-84, datasets/llff.py:232-245).  `gen_rays` does all of that in ONE kernel straight from the
3x4 pose: 32 B written per ray, nothing else touches memory.  The three reference functions keep their signatures;
they are thin views over the same kernel so that a dataset written against the reference runs unchanged."""
import torch

from .. import _lib
from ..ops import _p, _stream


def _c2w_host(c2w):
    m = torch.as_tensor(c2w, dtype=torch.float32).detach().cpu().contiguous()
    if m.shape != (3, 4):
        raise RuntimeError(f"c2w must be a (3, 4) camera-to-world matrix, got {tuple(m.shape)}")
    return m


def gen_rays(c2w, H, W, focal, near, far, pix=None, n=None, ndc=False, device=None):
    """rays (n,8) = [o, d/|d|, near, far] for the pixels `pix` (flat row*W+col, int64 on the device) or, when pix is
    None, the first n (default H*W) pixels in row-major order.  ndc=True: NDC rays with near plane `near`
    (reference get_ndc_rays(H, W, focal, near, ...)), bounds written as [0, 1]."""
    dev = torch.device(device) if device is not None else (pix.device if pix is not None else torch.device("cuda"))
    if dev.type != "cuda":
        raise RuntimeError("gen_rays runs on a CUDA device only (no CPU path)")
    m = _c2w_host(c2w)
    if pix is not None:
        if pix.dtype != torch.int64 or pix.device != dev:
            raise RuntimeError("pix must be an int64 tensor on the target device")
        pix = pix.contiguous()
        n = pix.numel()
    elif n is None:
        n = H * W
    rays = torch.empty(n, 8, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        _lib.call("nrf_gen_rays", m.data_ptr(), int(H), int(W), float(focal), _p(pix), n, float(near), float(far),
                  int(bool(ndc)), _p(rays), _stream())
    return rays


class _Directions(torch.Tensor):
    """(H,W,3) camera-space directions that remember (H, W, focal), so get_rays can use the fused kernel."""


def get_ray_directions(H, W, focal, device=None):
    """directions (H,W,3) in camera coordinates (reference :5-24; no +0.5 pixel centring).  Built once per dataset with
    a few elementwise ops; the result remembers (H, W, focal) so that get_rays can run the fused kernel."""
    dev = torch.device(device) if device is not None else torch.device("cuda")
    i = torch.arange(W, device=dev, dtype=torch.float32).expand(H, W)
    j = torch.arange(H, device=dev, dtype=torch.float32).unsqueeze(1).expand(H, W)
    d = torch.stack([(i - W / 2) / focal, -(j - H / 2) / focal, -torch.ones_like(i)], -1).as_subclass(_Directions)
    d._nrf = (int(H), int(W), float(focal))
    return d


def get_rays(directions, c2w):
    """rays_o, rays_d (H*W,3) in world coordinates, rays_d normalised (reference :27-51)."""
    meta = getattr(directions, "_nrf", None)
    if meta is None or directions.device.type != "cuda":
        raise RuntimeError("get_rays expects the directions returned by nerf_pl_b200.datasets.ray_utils.get_ray_directions")
    H, W, focal = meta
    rays = gen_rays(c2w, H, W, focal, 0.0, 1.0, device=directions.device)
    return rays[:, 0:3], rays[:, 3:6]


def get_ndc_rays(H, W, focal, near, rays_o, rays_d):
    """World rays -> NDC (reference :52-93); `near` may be a float or an (N,) tensor.  Elementwise torch ops on the
    device: the fused path for datasets is gen_rays(..., ndc=True)."""
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
