"""Drop-in for models/efficient_shadow_mapping.py (reference :19-130) on the device kernels.

`run_shadow_mapping` is the composition the trainers use (via efficient_sm); it is one fused
autograd node here.  The steps of that composition which the light-sampler trainer calls one by one
(train_light_sampler.py:15, :141-196: normalize_min_max, get_normed_w, generate_shadow_map; plus
get_projections / get_diff_projections / get_projected_depths) are separate device kernels with
their own backward."""
import torch

from .. import _lib
from ..ops import _dev, _p, _stream
from ._shadow import LightGather, ProjectToLight, ShadowMap, ShadowTest, _f32, _inverse3

EPSILON = 1e-5


class _NormedW(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pixel_depth, M):
        pd = _dev(pixel_depth, "pixel_depth")
        if pd.dim() != 2 or pd.shape[1] != 4:
            raise RuntimeError("pixel_depth must be (num_rays, 4) [i, j, 1, depth]")
        out = torch.empty_like(pd)
        _lib.call("nrf_normed_w", _p(pd), _p(pd), _p(pd), _p(M), pd.shape[0], _p(out), _stream())
        ctx.save_for_backward(pd, M)
        return out

    @staticmethod
    def backward(ctx, g):
        pd, M = ctx.saved_tensors
        g = _dev(g, "grad")
        gi = torch.empty_like(pd)
        _lib.call("nrf_normed_w", _p(pd), _p(g), _p(g), _p(M), pd.shape[0], _p(gi), _stream())
        return gi, None


def get_normed_w(camera, pixel_depth, device='cpu'):
    """[i, j, 1, depth] -> [i, j, 1, depth / (|M [i,j,1]^T| + 1e-5)] (reference :47-62).
    Gradients flow to the depth column (pixel coordinates are treated as constants)."""
    dev = pixel_depth.device
    return _NormedW.apply(pixel_depth, _f32(camera.camera, dev))


def run_shadow_mapping(res, camera, light_cam, batched_mesh_range_cam, meshed_normed_light_cam, device,
                       mode='shadow_method_1', delta=1e-2, epsilon=0.0, new_min=0.0, new_max=1.0, sigmoid=False,
                       use_numpy_meshgrid=True):
    """Shadow value (n,3) for rays of ONE camera against the normalised light map (reference :19-44).
    batched_mesh_range_cam (n,4) [i,j,1,depth]; meshed_normed_light_cam (H*W,4) from get_normed_w."""
    mrc = batched_mesh_range_cam
    dev = mrc.device
    n = mrc.shape[0]
    cam_M = _f32(camera.camera, dev).reshape(1, 3, 3).expand(n, 3, 3).contiguous()
    eye = _f32(camera.eye_pos, dev).reshape(1, 3).expand(n, 3).contiguous()
    LM_src = light_cam.camera
    static = (mrc[:, :3].detach().contiguous(), cam_M, eye, _inverse3(LM_src, dev)[0], None,
              _f32(light_cam.eye_pos, dev), None)
    return ShadowTest.apply(mrc[:, 3], meshed_normed_light_cam[:, 3], static, res, mode, delta, epsilon, new_min,
                            new_max, sigmoid, 0.0)


def normalize_min_max(tensor, new_max=1.0, new_min=0.0):
    """(t - min) / (max - min + 1e-5) * (new_max - new_min) + new_min over the whole tensor (reference :10-11)."""
    return ShadowMap.apply(tensor.reshape(-1), None, 'shadow_method_2', 1e-2, 0.0, new_min, new_max, False,
                           False).view(tensor.shape)


def get_diff_projections(pixels, w_cam, from_camera, to_camera, device='cpu'):
    """K (n,3) = [ul, vl, wl]: pixels (n,3) with normalised depth w_cam (n) of `from_camera` seen by `to_camera`
    (reference :65-83)."""
    dev = w_cam.device
    LM_src = to_camera.camera
    return ProjectToLight.apply(w_cam, _f32(pixels, dev), _f32(from_camera.camera, dev).reshape(1, 3, 3),
                                _f32(from_camera.eye_pos, dev).reshape(1, 3), _inverse3(LM_src, dev)[0],
                                _f32(to_camera.eye_pos, dev), True)


def get_projections(camera, light_cam, batched_mesh_range_cam, device):
    """get_normed_w followed by get_diff_projections (reference :13-17), one kernel."""
    mrc = batched_mesh_range_cam
    dev = mrc.device
    LM_src = light_cam.camera
    return ProjectToLight.apply(mrc[:, 3], mrc[:, :3].detach().contiguous(), _f32(camera.camera, dev).reshape(1, 3, 3),
                                _f32(camera.eye_pos, dev).reshape(1, 3), _inverse3(LM_src, dev)[0],
                                _f32(light_cam.eye_pos, dev), False)


def get_projected_depths(res, K, w_light, device='cpu'):
    """(wl, w_light_bounded): K's depth column and the light map read at K's clamped, truncated pixel
    (reference :85-100)."""
    return K[:, 2], LightGather.apply(K.detach(), w_light, res)


def generate_shadow_map(wl, w_light_bounded, delta=1e-2, epsilon=0.0, new_min=0.0, new_max=1.0, sigmoid=False,
                        mode='shadow_method_1', device='cpu'):
    """(n,3) shadow value from depth-in-light-space wl and the light map's depth there (reference :103-130)."""
    return ShadowMap.apply(wl, w_light_bounded, mode, delta, epsilon, new_min, new_max, sigmoid, True)
