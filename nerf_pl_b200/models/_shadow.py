"""Shadow test on the device: autograd node over nrf_shadow_fwd/bwd and the efficient_sm glue
(models/rendering_shadows.py:359-482, models/rendering_rgb_sm.py:356-480)."""
import torch

from .. import _lib
from ..ops import _dev, _p, _stream

EPSILON = 1e-5  # rendering_shadows.py:357

_MODES = {"shadow_method_1": 1, "shadow_method_2": 2}


class ShadowTest(torch.autograd.Function):
    """out (n,3) = clip(shadow(cam_depth, light_depth), 0, 1) + out_bias.

    static = (cam_pixels (n,3), cam_M (n,3,3), cam_eye (n,3), light_Minv (3,3), light_M (3,3) or None,
              light_eye (3), light_pixels (HW,3) or None); light_M/light_pixels None => light_depth is
    already normalised.  Differentiable w.r.t. cam_depth and light_depth."""

    @staticmethod
    def forward(ctx, cam_depth, light_depth, static, res, mode, delta, epsilon, new_min, new_max, sigmoid, out_bias):
        if mode not in _MODES:
            raise ValueError("{} not found".format(mode))
        cam_pixels, cam_M, cam_eye, light_Minv, light_M, light_eye, light_pixels = static
        cam_depth, light_depth = _dev(cam_depth, "cam_depth"), _dev(light_depth, "light_depth")
        n = cam_depth.numel()
        w, h = int(res[0]), int(res[1])
        if light_depth.numel() != w * h:
            raise RuntimeError(f"light depth map has {light_depth.numel()} texels, expected {w}x{h}")
        lib = _lib.load()
        nb = lib.nrf_shadow_workspace_bytes(n, w * h)
        ws = torch.empty(nb, dtype=torch.uint8, device=cam_depth.device)
        out = torch.empty(n, 3, device=cam_depth.device)
        args = (_p(cam_pixels), _p(cam_depth), _p(cam_M), _p(cam_eye), _p(light_Minv), _p(light_M), _p(light_eye),
                _p(light_pixels), _p(light_depth), n, w, h, _MODES[mode], float(delta), float(epsilon),
                float(new_min), float(new_max), int(bool(sigmoid)))
        _lib.call("nrf_shadow_fwd", *args, float(out_bias), _p(out), _p(ws), nb, _stream())
        ctx.args, ctx.static, ctx.nb = args, static, nb
        ctx.save_for_backward(cam_depth, light_depth)
        return out

    @staticmethod
    def backward(ctx, g):
        cam_depth, light_depth = ctx.saved_tensors
        g = _dev(g, "grad")
        need_light = ctx.needs_input_grad[1]
        g_cam = torch.empty_like(cam_depth)
        g_light = torch.zeros_like(light_depth) if need_light else None
        ws = torch.empty(ctx.nb, dtype=torch.uint8, device=cam_depth.device)
        _lib.call("nrf_shadow_bwd", *ctx.args, _p(g), _p(g_cam), _p(g_light), _p(ws), ctx.nb, _stream())
        return (g_cam, g_light) + (None,) * 9


def _f32(t, dev):
    return torch.as_tensor(t, dtype=torch.float32).to(dev).contiguous()


def _light_static(light_camera, dev):
    M = _f32(light_camera.camera, dev)
    return torch.inverse(M).contiguous(), M, _f32(light_camera.eye_pos, dev)


def efficient_sm_impl(prefix, cam_pixels, light_pixels, cam_results, light_results, ppc, light_ppc, image_shape,
                      fine_sampling, Light_N_importance, shadow_method):
    """Writes `{prefix}_coarse` / `{prefix}_fine` into cam_results and returns it.  `ppc` =
    {'eye_pos': (n,3), 'camera': (n,3,3)} per ray; `light_ppc` is a Camera.  Same-camera segments
    (the reference's Python loop, rendering_shadows.py:385-406) are found on the device."""
    dev = cam_results["depth_coarse"].device
    eye = ppc["eye_pos"]
    if isinstance(eye, (list, tuple)):
        eye = torch.stack([torch.as_tensor(e) for e in eye])
    cam_M = ppc["camera"]
    if isinstance(cam_M, (list, tuple)):
        cam_M = torch.stack([torch.as_tensor(c) for c in cam_M])
    n = cam_results["depth_coarse"].numel()
    eye = _f32(eye, dev).reshape(n, 3)
    cam_M = _f32(cam_M, dev).reshape(n, 3, 3)
    inv, LM, Leye = _light_static(light_ppc, dev)
    static = (_f32(cam_pixels, dev), cam_M, eye, inv, LM, Leye, _f32(light_pixels, dev))

    def run(cam_depth, light_depth):
        return ShadowTest.apply(cam_depth.reshape(-1), light_depth.reshape(-1), static, image_shape, shadow_method,
                                1e-2, 0.0, 0.0, 1.0, False, EPSILON)

    cam_results[f"{prefix}_coarse"] = run(cam_results["depth_coarse"], light_results["depth_coarse"])
    if fine_sampling:
        ld = light_results["depth_fine"] if Light_N_importance else light_results["depth_coarse"]
        cam_results[f"{prefix}_fine"] = run(cam_results["depth_fine"], ld)
    return cam_results


def get_K_impl(cam_pixels, cam_results, ppc, light_camera, fine_sampling):
    raise NotImplementedError("get_K (rendering_shadows.py:484-546) belongs to the light-sampler trainer, "
                              "SURVEY §8f rank 2 ('next'); not built yet")
