"""Drop-in for models/efficient_shadow_mapping.py (reference :19-130) on the device kernels.

`run_shadow_mapping` is the composition the trainers use (via efficient_sm); it is one fused
autograd node here.  The reference's intermediate helpers that only exist as steps of that
composition (get_diff_projections, get_projected_depths, generate_shadow_map) are fused inside
it and are not exposed separately."""
import torch

from .. import _lib
from ..ops import _dev, _p, _stream
from ._shadow import ShadowTest, _f32

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
    LM = _f32(light_cam.camera, dev)
    static = (mrc[:, :3].detach().contiguous(), cam_M, eye, torch.inverse(LM).contiguous(), None,
              _f32(light_cam.eye_pos, dev), None)
    return ShadowTest.apply(mrc[:, 3], meshed_normed_light_cam[:, 3], static, res, mode, delta, epsilon, new_min,
                            new_max, sigmoid, 0.0)
