"""Golden vectors for the whole-image shadow mapping of models/shadow_mapping_utils.py (SURVEY §8f rank 2;
train_shadow_mapping.py via rendering_shadows.shadow_mapping): every function of the module, values and gradients,
produced by the reference's own code.

    python tests/make_golden_shadow_image.py        # needs /root/reference (build container only)"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference  # noqa: E402


def main():
    _, _, rendering_shadows, _, _, camera = import_reference()
    import models.shadow_mapping_utils as smu
    H = Wd = 16
    res = (Wd, H)
    g = torch.Generator().manual_seed(47)

    def lookat_c2w(eye):
        return torch.from_numpy(camera.Camera.c2w_from_lookat(np.array(eye, dtype=np.float32),
                                                              np.zeros(3, dtype=np.float32)))[:3, :4]

    cam = camera.Camera(30, res)
    cam.set_pose_using_blender_matrix(lookat_c2w([0.0, 1.0, 4.0]))
    light = camera.Camera(30, res)
    light.set_pose_using_blender_matrix(lookat_c2w([1.0, 4.0, 2.0]))
    rec = {"res": np.array(res), "cam_eye": cam.eye_pos.numpy(), "cam_M": cam.camera.numpy(),
           "light_eye": light.eye_pos.numpy(), "light_M": light.camera.numpy()}

    def leaf(lo, span):
        return (lo + span * torch.rand(Wd, H, generator=g)).requires_grad_()

    # differentiable_w_mapping
    rc = leaf(3.0, 2.0)
    w_cam = smu.differentiable_w_mapping(res, cam, rc, use_numpy_meshgrid=True)
    pr = torch.rand(Wd, H, generator=g)
    (w_cam * pr).sum().backward()
    rec.update({"wmap/range": rc.detach().numpy(), "wmap/w": w_cam.detach().numpy(), "wmap/probe": pr.numpy(),
                "wmap/g_range": rc.grad.numpy()})
    # get_differentiable_projections
    wc = (w_cam.detach() * (0.9 + 0.2 * torch.rand(Wd, H, generator=g))).requires_grad_()
    K = smu.get_differentiable_projections(wc, res, cam, light, use_numpy_meshgrid=True)
    pr3 = torch.rand(Wd, H, 3, generator=g)
    (K * pr3).sum().backward()
    rec.update({"proj/w_cam": wc.detach().numpy(), "proj/K": K.detach().numpy(), "proj/probe": pr3.numpy(),
                "proj/g_w_cam": wc.grad.numpy()})
    # get_differentiable_depths: pixels inside, outside (clamped) and on the border of the light image
    Kp = torch.stack([torch.rand(Wd, H, generator=g) * (Wd + 6) - 3, torch.rand(Wd, H, generator=g) * (H + 6) - 3,
                      3 + torch.rand(Wd, H, generator=g)], 2)
    Kp[0, 0, :2] = torch.tensor([0.0, 0.0])
    Kp[0, 1, :2] = torch.tensor([Wd - 1.0, H - 1.0])
    Kp[0, 2, :2] = torch.tensor([Wd - 0.001, 7.999])
    Kp = Kp.requires_grad_()
    w_light = leaf(0.01, 0.01)
    wl, wlb, ul_, vl_ = smu.get_differentiable_depths(res, Kp, w_light, use_numpy_meshgrid=True)
    pr = torch.rand(Wd, H, generator=g)
    ((wl * pr).sum() + 3 * (wlb * pr).sum()).backward()
    assert torch.equal(ul_, Kp[..., 0]) and torch.equal(vl_, Kp[..., 1])
    rec.update({"depths/K": Kp.detach().numpy(), "depths/w_light": w_light.detach().numpy(), "depths/probe": pr.numpy(),
                "depths/wl": wl.detach().numpy(), "depths/wlb": wlb.detach().numpy(), "depths/g_K": Kp.grad.numpy(),
                "depths/g_w_light": w_light.grad.numpy()})
    # generate_shadow_map / normalize_min_max
    for tag, kw in (("m1", dict(mode="shadow_method_1")), ("m1_eps", dict(mode="shadow_method_1", delta=0.05, epsilon=0.2)),
                    ("m2", dict(mode="shadow_method_2")), ("m2_sig", dict(mode="shadow_method_2", sigmoid=True)),
                    ("m2_range", dict(mode="shadow_method_2", new_min=-0.5, new_max=1.5))):
        a, b = leaf(3.0, 1.0), leaf(3.0, 1.0)
        sm = smu.generate_shadow_map(a, b, **kw)
        pr3 = torch.rand(Wd, H, 3, generator=g)
        (sm * pr3).sum().backward()
        rec.update({f"map_{tag}/wl": a.detach().numpy(), f"map_{tag}/wlb": b.detach().numpy(), f"map_{tag}/probe": pr3.numpy(),
                    f"map_{tag}/out": sm.detach().numpy(), f"map_{tag}/g_wl": a.grad.numpy(), f"map_{tag}/g_wlb": b.grad.numpy()})
    t = torch.randn(7, 9, generator=g).requires_grad_()
    y = smu.normalize_min_max(t, new_max=2.0, new_min=-1.0)
    pr2 = torch.rand(7, 9, generator=g)
    (y * pr2).sum().backward()
    rec.update({"norm/t": t.detach().numpy(), "norm/probe": pr2.numpy(), "norm/out": y.detach().numpy(),
                "norm/g_t": t.grad.numpy()})
    # run_shadow_mapping, both modes, gradients to both range images
    for tag, mode in (("run_m1", "shadow_method_1"), ("run_m2", "shadow_method_2")):
        rc, rl = leaf(3.0, 2.0), leaf(3.0, 2.0)
        sm = smu.run_shadow_mapping(res, cam, light, rc, rl, "cpu", mode, use_numpy_meshgrid=True)
        pr3 = torch.rand(Wd, H, 3, generator=g)
        (sm * pr3).sum().backward()
        rec.update({f"{tag}/range_cam": rc.detach().numpy(), f"{tag}/range_light": rl.detach().numpy(),
                    f"{tag}/probe": pr3.numpy(), f"{tag}/out": sm.detach().numpy(), f"{tag}/g_range_cam": rc.grad.numpy(),
                    f"{tag}/g_range_light": rl.grad.numpy()})
    # the reference's shadow_mapping() wrapper raises on its own call (documented in _shadow.shadow_mapping_impl)
    try:
        ppc = [{"eye_pos": cam.eye_pos[None], "camera": cam.camera[None]}]
        lppc = {"eye_pos": light.eye_pos[None], "camera": light.camera[None]}
        rendering_shadows.shadow_mapping({"depth_coarse": rc.detach().reshape(-1)}, {"depth_coarse": rl.detach().reshape(-1)},
                                         torch.zeros(H * Wd, 8), ppc, lppc, res, 1, False)
        rec["wrapper_raises"] = np.array(0)
    except RuntimeError as e:
        print("reference shadow_mapping() raises:", e)
        rec["wrapper_raises"] = np.array(1)
    out = os.path.join(HERE, "golden", "shadow_image.npz")
    np.savez_compressed(out, **rec)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
