"""Drop-in for models/shadow_mapping_utils.py (reference :8-174): whole-image differentiable shadow mapping
(used by train_shadow_mapping.py) on the device kernels of csrc/shadow.cu.

Differences from models/efficient_shadow_mapping.py that this module keeps, like the reference: no 1e-5 in the
range -> w division (:65) nor in the min-max normalisation (:8-9), images are (w, h) grids instead of ray lists,
the shadow image is stacked on the last axis (w, h, 3).  Only square images (reference :3-4)."""
import torch

from .. import _lib
from ..ops import _dev, _p, _stream
from ._shadow import LightGather, ProjectToLight, ShadowMap, _f32, _inverse3

_GRID = {}


def _pixel_grid(res, dev):
    """(w*h, 3) rows [i+.5, j+.5, 1] in the row-major order of the reference's (w, h) meshgrid (:46-58): the column
    index is the first coordinate."""
    w, h = int(res[0]), int(res[1])
    key = (w, h, str(dev))
    if key not in _GRID:
        i = (torch.arange(h, dtype=torch.float32) + 0.5).view(1, h).expand(w, h)
        j = (torch.arange(w, dtype=torch.float32) + 0.5).view(w, 1).expand(w, h)
        _GRID[key] = torch.stack([i, j, torch.ones(w, h)], -1).reshape(-1, 3).contiguous().to(dev)
    return _GRID[key]


def _check_square(res):
    if int(res[0]) != int(res[1]):
        raise RuntimeError("only square images are supported, like the reference (shadow_mapping_utils.py:3-4)")


class _WMapping(torch.autograd.Function):
    @staticmethod
    def forward(ctx, range_cam, pixels, M):
        r = _dev(range_cam, "range_cam").reshape(-1)
        if r.numel() != pixels.shape[0]:
            raise RuntimeError(f"range image has {r.numel()} pixels, expected {pixels.shape[0]}")
        out = torch.empty_like(r)
        _lib.call("nrf_w_mapping", _p(pixels), _p(r), _p(M), r.numel(), 0.0, _p(out), _stream())
        ctx.keep = (pixels, M)
        return out

    @staticmethod
    def backward(ctx, g):
        pixels, M = ctx.keep
        g = _dev(g, "grad").reshape(-1)
        gi = torch.empty_like(g)
        _lib.call("nrf_w_mapping", _p(pixels), _p(g), _p(M), g.numel(), 0.0, _p(gi), _stream())
        return gi, None, None


def normalize_min_max(tensor, new_max=1.0, new_min=0.0):
    """(t - min) / (max - min) * (new_max - new_min) + new_min over the whole tensor (reference :8-9)."""
    return ShadowMap.apply(tensor.reshape(-1), None, 'shadow_method_2', 1e-2, 0.0, new_min, new_max, False, False,
                           True).view(tensor.shape)


def differentiable_w_mapping(res, camera, range_cam, device='cpu', use_numpy_meshgrid=False):
    """range (w,h) -> w (w,h) = range / |M [i+.5, j+.5, 1]^T| (reference :39-66)."""
    _check_square(res)
    dev = range_cam.device
    w, h = int(res[0]), int(res[1])
    return _WMapping.apply(range_cam.reshape(-1), _pixel_grid(res, dev), _f32(camera.camera, dev)).view(w, h)


def get_differentiable_projections(w_from_camera, res, from_camera, to_camera, device='cpu', use_numpy_meshgrid=False):
    """K (w,h,3) = [ul, vl, wl] of every pixel of `from_camera` in `to_camera`'s image (reference :68-104)."""
    _check_square(res)
    dev = w_from_camera.device
    w, h = int(res[0]), int(res[1])
    LM_src = to_camera.camera
    K = ProjectToLight.apply(w_from_camera.reshape(-1), _pixel_grid(res, dev),
                             _f32(from_camera.camera, dev).reshape(1, 3, 3), _f32(from_camera.eye_pos, dev).reshape(1, 3),
                             _inverse3(LM_src, dev)[0], _f32(to_camera.eye_pos, dev), 1)
    return K.view(w, h, 3)


def get_differentiable_depths(res, K, w_light, device='cpu', use_numpy_meshgrid=False):
    """(wl, w_light_bounded, ul_, vl_): K's depth, the light map read at K's clamped, truncated pixel and the
    unclamped coordinates (reference :106-147).  `K[j, i]` of the reference is the identity for square images."""
    _check_square(res)
    w, h = int(res[0]), int(res[1])
    wlb = LightGather.apply(K.detach().reshape(-1, 3), w_light.reshape(-1), res).view(w, h)
    return K[..., 2], wlb, K[..., 0], K[..., 1]


def generate_shadow_map(wl, w_light_bounded, delta=1e-2, epsilon=0.0, new_min=0.0, new_max=1.0, sigmoid=False,
                        mode='shadow_method_1', device='cpu'):
    """(w,h,3) shadow image from depth-in-light-space wl and the light map's depth there (reference :149-174);
    new_min / new_max are not forwarded to normalize_min_max by the reference (:167) and have no effect."""
    return ShadowMap.apply(wl.reshape(-1), w_light_bounded.reshape(-1), mode, delta, epsilon, new_min, new_max, sigmoid,
                           True, True).view(tuple(wl.shape) + (3,))


def run_shadow_mapping(res, camera, light_cam, range_cam, range_light, device, mode, delta=1e-2, epsilon=0.0,
                       new_min=0.0, new_max=1.0, sigmoid=False, use_numpy_meshgrid=True):
    """Whole-image stack (reference :11-36): range images (w,h) of the camera and the light -> (w,h,3)."""
    _check_square(res)
    dev = range_cam.device
    w, h = int(res[0]), int(res[1])
    grid = _pixel_grid(res, dev)
    LM_src = light_cam.camera
    w_light = _WMapping.apply(range_light.reshape(-1), grid, _inverse3(LM_src, dev)[1])
    # range -> w (no epsilon) -> projection, one kernel
    K = ProjectToLight.apply(range_cam.reshape(-1), grid, _f32(camera.camera, dev).reshape(1, 3, 3),
                             _f32(camera.eye_pos, dev).reshape(1, 3), _inverse3(LM_src, dev)[0],
                             _f32(light_cam.eye_pos, dev), 2)
    wlb = LightGather.apply(K.detach(), w_light, res)
    return ShadowMap.apply(K[:, 2], wlb, mode, delta, epsilon, new_min, new_max, sigmoid, True, True).view(w, h, 3)
