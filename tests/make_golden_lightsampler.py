"""Golden vectors for the shadow-mapping pieces the light-sampler trainer calls one by one (SURVEY §8f rank 2;
train_light_sampler.py:15, :141-196): get_K, get_projections, get_diff_projections, get_projected_depths,
generate_shadow_map, normalize_min_max - values and gradients, produced by the reference's own functions.

    python tests/make_golden_lightsampler.py        # needs /root/reference (build container only)"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference  # noqa: E402


def main():
    _, _, rendering_shadows, _, esm, camera = import_reference()
    H = Wd = 16
    res = (Wd, H)
    g = torch.Generator().manual_seed(31)

    def lookat_c2w(eye):
        return torch.from_numpy(camera.Camera.c2w_from_lookat(np.array(eye, dtype=np.float32),
                                                              np.zeros(3, dtype=np.float32)))[:3, :4]

    cams = []
    for eye in ([0.0, 1.0, 4.0], [2.0, 1.5, 3.0], [-1.0, 2.0, 3.5]):
        c = camera.Camera(30, res)
        c.set_pose_using_blender_matrix(lookat_c2w(eye))
        cams.append(c)
    light = camera.Camera(30, res)
    light.set_pose_using_blender_matrix(lookat_c2w([1.0, 4.0, 2.0]))
    ii, jj = torch.meshgrid(torch.arange(Wd).float(), torch.arange(H).float(), indexing="xy")
    all_pixels = torch.stack([ii.reshape(-1) + .5, jj.reshape(-1) + .5, torch.ones(H * Wd)], 1)
    rec = {"res": np.array(res), "light_eye": light.eye_pos.numpy(), "light_camera": light.camera.numpy()}

    # ---- get_K: three same-camera runs (the first camera comes back at the end), and a single-camera batch
    for tag, seg in (("split", [0] * 20 + [1] * 13 + [2] * 9 + [0] * 6), ("single", [1] * 32)):
        n = len(seg)
        ppc = {"eye_pos": torch.stack([cams[s].eye_pos for s in seg]),
               "camera": torch.stack([cams[s].camera for s in seg])}
        cam_pixels = all_pixels[torch.randperm(H * Wd, generator=g)[:n]].clone()
        cd = {k: (3.0 + 2.0 * torch.rand(n, generator=g)).requires_grad_() for k in ("depth_coarse", "depth_fine")}
        Kc, Kf = rendering_shadows.get_K(cam_pixels, dict(cd), ppc, light, True)
        Kc_only, none = rendering_shadows.get_K(cam_pixels, dict(cd), ppc, light, False)
        assert none is None and torch.equal(Kc_only, Kc)
        probe = torch.rand(n, 3, generator=g)
        ((Kc * probe).sum() + 2 * (Kf * probe).sum()).backward()
        rec.update({f"getK_{tag}/eye_pos": ppc["eye_pos"].numpy(), f"getK_{tag}/camera": ppc["camera"].numpy(),
                    f"getK_{tag}/cam_pixels": cam_pixels.numpy(), f"getK_{tag}/probe": probe.numpy(),
                    f"getK_{tag}/K_coarse": Kc.detach().numpy(), f"getK_{tag}/K_fine": Kf.detach().numpy()})
        for k in cd:
            rec[f"getK_{tag}/{k}"] = cd[k].detach().numpy()
            rec[f"getK_{tag}/g_{k}"] = cd[k].grad.numpy()

    # ---- the single-camera helpers
    n = 48
    cam = cams[0]
    pix = all_pixels[torch.randperm(H * Wd, generator=g)[:n]].clone()
    depth = (3.0 + 2.0 * torch.rand(n, generator=g)).requires_grad_()
    mrc = torch.cat([pix, depth.view(-1, 1)], 1)
    K = esm.get_projections(cam, light, mrc, "cpu")
    probe = torch.rand(n, 3, generator=g)
    (K * probe).sum().backward()
    rec.update({"proj/cam_M": cam.camera.numpy(), "proj/cam_eye": cam.eye_pos.numpy(), "proj/pix": pix.numpy(),
                "proj/depth": depth.detach().numpy(), "proj/K": K.detach().numpy(), "proj/probe": probe.numpy(),
                "proj/g_depth": depth.grad.numpy()})
    w_cam = (0.5 + torch.rand(n, generator=g)).requires_grad_()
    K2 = esm.get_diff_projections(pix, w_cam, cam, light)
    (K2 * probe).sum().backward()
    rec.update({"diffproj/w_cam": w_cam.detach().numpy(), "diffproj/K": K2.detach().numpy(),
                "diffproj/g_w_cam": w_cam.grad.numpy()})

    # get_projected_depths: K with pixels inside, outside (clamped) and on the border of the light image
    Kp = torch.stack([torch.rand(n, generator=g) * (Wd + 6) - 3, torch.rand(n, generator=g) * (H + 6) - 3,
                      3 + torch.rand(n, generator=g)], 1)
    Kp[0, :2] = torch.tensor([0.0, 0.0])
    Kp[1, :2] = torch.tensor([Wd - 1.0, H - 1.0])
    Kp[2, :2] = torch.tensor([Wd - 0.001, 7.999])
    Kp = Kp.requires_grad_()
    w_light = (3.0 + 2.0 * torch.rand(H * Wd, generator=g)).requires_grad_()
    wl, wlb = esm.get_projected_depths(res, Kp, w_light)
    pr = torch.rand(n, generator=g)
    ((wl * pr).sum() + 3 * (wlb * pr).sum()).backward()
    rec.update({"gather/K": Kp.detach().numpy(), "gather/w_light": w_light.detach().numpy(), "gather/probe": pr.numpy(),
                "gather/wl": wl.detach().numpy(), "gather/wlb": wlb.detach().numpy(), "gather/g_K": Kp.grad.numpy(),
                "gather/g_w_light": w_light.grad.numpy()})

    # generate_shadow_map / normalize_min_max
    for tag, kw in (("m1", dict(mode="shadow_method_1")), ("m1_eps", dict(mode="shadow_method_1", delta=0.05, epsilon=0.2)),
                    ("m2", dict(mode="shadow_method_2")), ("m2_sig", dict(mode="shadow_method_2", sigmoid=True)),
                    ("m2_range", dict(mode="shadow_method_2", new_min=-0.5, new_max=1.5))):
        a = (3.0 + torch.rand(n, generator=g)).requires_grad_()
        b = (3.0 + torch.rand(n, generator=g)).requires_grad_()
        with torch.no_grad():
            b[5] = a[5]                                  # a zero difference (torch.max tie for method 1)
        sm = esm.generate_shadow_map(a, b, **kw)
        pr3 = torch.rand(n, 3, generator=g)
        (sm * pr3).sum().backward()
        rec.update({f"map_{tag}/wl": a.detach().numpy(), f"map_{tag}/wlb": b.detach().numpy(), f"map_{tag}/probe": pr3.numpy(),
                    f"map_{tag}/out": sm.detach().numpy(), f"map_{tag}/g_wl": a.grad.numpy(), f"map_{tag}/g_wlb": b.grad.numpy()})
    t = torch.randn(7, 9, generator=g).requires_grad_()
    y = esm.normalize_min_max(t, new_max=2.0, new_min=-1.0)
    pr2 = torch.rand(7, 9, generator=g)
    (y * pr2).sum().backward()
    rec.update({"norm/t": t.detach().numpy(), "norm/probe": pr2.numpy(), "norm/out": y.detach().numpy(),
                "norm/g_t": t.grad.numpy()})
    out = os.path.join(HERE, "golden", "lightsampler.npz")
    np.savez_compressed(out, **rec)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
