"""GPU parity of the whole-image shadow mapping (models/shadow_mapping_utils.py:8-174 and
rendering_shadows.shadow_mapping :283-353) against vectors produced by the reference's own functions
(tests/make_golden_shadow_image.py) and against the CPU oracle at 128x128.
Tolerances: gathered depths bit-exact; w 1e-6 relative; projections 1e-5; shadow images 1e-5; gradients 1e-4 relative
(1e-3 where the min / max terms sum 49 152 products in another order)."""
import os

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")

MAP_CASES = (("m1", dict(mode="shadow_method_1")), ("m1_eps", dict(mode="shadow_method_1", delta=0.05, epsilon=0.2)),
             ("m2", dict(mode="shadow_method_2")), ("m2_sig", dict(mode="shadow_method_2", sigmoid=True)),
             ("m2_range", dict(mode="shadow_method_2", new_min=-0.5, new_max=1.5)))


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as ge
    ge.build()
    from nerf_pl_b200 import _lib
    _lib.check(_lib.load().nrf_check_device(), "nrf_check_device")
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def z():
    return np.load(os.path.join(G, "shadow_image.npz"))


def _cams(z):
    from nerf_pl_b200.models.camera import Camera
    return (Camera.from_camera_eyepos(torch.from_numpy(z["cam_eye"]), torch.from_numpy(z["cam_M"])),
            Camera.from_camera_eyepos(torch.from_numpy(z["light_eye"]), torch.from_numpy(z["light_M"])))


def _close(a, ref, rtol, atol, msg=""):
    np.testing.assert_allclose(a.detach().cpu().numpy(), ref, rtol=rtol, atol=atol, err_msg=msg)


def test_w_mapping_and_projections_match_reference_golden(dev, z):
    from nerf_pl_b200.models import shadow_mapping_utils as smu
    T = lambda k: torch.from_numpy(z[k]).to(dev)
    res = tuple(int(v) for v in z["res"])
    cam, light = _cams(z)
    rc = T("wmap/range").requires_grad_()
    w = smu.differentiable_w_mapping(res, cam, rc, device=dev, use_numpy_meshgrid=True)
    assert w.shape == rc.shape
    _close(w, z["wmap/w"], 1e-6, 0)
    (w * T("wmap/probe")).sum().backward()
    _close(rc.grad, z["wmap/g_range"], 1e-6, 0)
    wc = T("proj/w_cam").requires_grad_()
    K = smu.get_differentiable_projections(wc, res, cam, light, device=dev, use_numpy_meshgrid=True)
    assert K.shape == wc.shape + (3,)
    _close(K, z["proj/K"], 1e-5, 1e-5)
    (K * T("proj/probe")).sum().backward()
    _close(wc.grad, z["proj/g_w_cam"], 1e-4, 1e-5 * float(np.abs(z["proj/g_w_cam"]).max()))


def test_differentiable_depths_bit_exact(dev, z):
    from nerf_pl_b200.models import shadow_mapping_utils as smu
    T = lambda k: torch.from_numpy(z[k]).to(dev)
    res = tuple(int(v) for v in z["res"])
    K, w_light = T("depths/K").requires_grad_(), T("depths/w_light").requires_grad_()
    wl, wlb, ul_, vl_ = smu.get_differentiable_depths(res, K, w_light, device=dev)
    assert np.array_equal(wl.detach().cpu().numpy(), z["depths/wl"])
    assert np.array_equal(wlb.detach().cpu().numpy(), z["depths/wlb"])
    assert torch.equal(ul_, K[..., 0]) and torch.equal(vl_, K[..., 1])
    pr = T("depths/probe")
    ((wl * pr).sum() + 3 * (wlb * pr).sum()).backward()
    _close(K.grad, z["depths/g_K"], 0, 0)
    _close(w_light.grad, z["depths/g_w_light"], 1e-6, 1e-7)


@pytest.mark.parametrize("tag,kw", MAP_CASES)
def test_image_shadow_map_matches_reference_golden(dev, z, tag, kw):
    from nerf_pl_b200.models import shadow_mapping_utils as smu
    T = lambda k: torch.from_numpy(z[k]).to(dev)
    a, b = T(f"map_{tag}/wl").requires_grad_(), T(f"map_{tag}/wlb").requires_grad_()
    sm = smu.generate_shadow_map(a, b, device=dev, **kw)
    assert sm.shape == a.shape + (3,)
    _close(sm, z[f"map_{tag}/out"], 1e-6, 1e-6)
    (sm * T(f"map_{tag}/probe")).sum().backward()
    _close(a.grad, z[f"map_{tag}/g_wl"], 1e-4, 1e-5, "g_wl")
    _close(b.grad, z[f"map_{tag}/g_wlb"], 1e-4, 1e-5, "g_wlb")


def test_image_normalize_min_max_matches_reference_golden(dev, z):
    from nerf_pl_b200.models import shadow_mapping_utils as smu
    t = torch.from_numpy(z["norm/t"]).to(dev).requires_grad_()
    y = smu.normalize_min_max(t, new_max=2.0, new_min=-1.0)
    assert y.shape == t.shape
    _close(y, z["norm/out"], 1e-6, 1e-6)
    (y * torch.from_numpy(z["norm/probe"]).to(dev)).sum().backward()
    _close(t.grad, z["norm/g_t"], 1e-4, 1e-5)


@pytest.mark.parametrize("tag,mode", [("run_m1", "shadow_method_1"), ("run_m2", "shadow_method_2")])
def test_run_shadow_mapping_matches_reference_golden(dev, z, tag, mode):
    from nerf_pl_b200.models import shadow_mapping_utils as smu
    T = lambda k: torch.from_numpy(z[k]).to(dev)
    res = tuple(int(v) for v in z["res"])
    cam, light = _cams(z)
    rc, rl = T(f"{tag}/range_cam").requires_grad_(), T(f"{tag}/range_light").requires_grad_()
    sm = smu.run_shadow_mapping(res, cam, light, rc, rl, dev, mode, use_numpy_meshgrid=True)
    assert sm.shape == (res[0], res[1], 3)
    _close(sm, z[f"{tag}/out"], 1e-5, 1e-5)
    (sm * T(f"{tag}/probe")).sum().backward()
    for g, k in ((rc.grad, "g_range_cam"), (rl.grad, "g_range_light")):
        _close(g, z[f"{tag}/{k}"], 1e-3, 1e-5 * float(np.abs(z[f"{tag}/{k}"]).max()), k)


def _posed(eye, res):
    from nerf_pl_b200.models.camera import Camera
    c2w = torch.from_numpy(Camera.c2w_from_lookat(np.array(eye, dtype=np.float32), np.zeros(3, dtype=np.float32)))
    e, M = O.pose_camera(O.camera_matrix(30, res), c2w[:3, :4])
    return M, e


@pytest.mark.parametrize("module", ["rendering_shadows", "rendering_rgb_sm"])
def test_shadow_mapping_wrapper_128_vs_oracle(dev, module):
    """train_shadow_mapping.py's call: two full 128x128 camera images against one light image, method 2, gradients to
    camera and light depths; errors for non-square images and CPU tensors."""
    import importlib
    mod = importlib.import_module(f"nerf_pl_b200.models.{module}")
    g = torch.Generator().manual_seed(13)
    res, B = (128, 128), 2
    n = res[0] * res[1]
    cams = [_posed(e, res) for e in ([0., 1., 4.], [2., 1.5, 3.])]
    LM, Leye = _posed([1.0, 4.0, 2.0], res)
    cd = (3.0 + 2.0 * torch.rand(B * n, generator=g)).requires_grad_()
    ld = (3.0 + 2.0 * torch.rand(B * n, generator=g)).requires_grad_()
    ref = torch.cat([O.image_run_shadow_mapping(res, cams[i][0], cams[i][1], LM, Leye, cd.view(B, *res)[i],
                                                ld.view(B, *res)[i], "shadow_method_2") for i in range(B)], 0).view(-1, 3)
    probe = torch.rand(B * n, 3, generator=g)
    (ref * probe).sum().backward()
    ppc = [{"eye_pos": cams[i][1][None], "camera": cams[i][0][None]} for i in range(B)]
    lppc = {"eye_pos": Leye[None], "camera": LM[None]}
    cg, lg = cd.detach().to(dev).requires_grad_(), ld.detach().to(dev).requires_grad_()
    out = mod.shadow_mapping({"depth_coarse": cg, "depth_fine": cg}, {"depth_coarse": lg, "depth_fine": lg},
                             torch.zeros(B * n, 8, device=dev), ppc, lppc, res, B, True)
    assert out["rgb_fine"] is out["rgb_coarse"]            # reference :351
    _close(out["rgb_coarse"], ref.detach().numpy(), 1e-4, 1e-5)
    (out["rgb_coarse"] * probe.to(dev)).sum().backward()
    for got, want, k in ((cg.grad, cd.grad, "cam"), (lg.grad, ld.grad, "light")):
        _close(got, want.numpy(), 1e-3, 1e-4 * float(want.abs().max()), k)
    from nerf_pl_b200.models import shadow_mapping_utils as smu
    from nerf_pl_b200.models.camera import Camera
    cam, light = Camera.from_camera_eyepos(cams[0][1], cams[0][0]), Camera.from_camera_eyepos(Leye, LM)
    with pytest.raises(RuntimeError):
        smu.run_shadow_mapping((128, 64), cam, light, cg[:8192].view(128, 64), lg[:8192].view(128, 64), dev, "shadow_method_2")
    with pytest.raises(RuntimeError):
        smu.run_shadow_mapping(res, cam, light, cd.detach()[:n].view(res), ld.detach()[:n].view(res), "cpu", "shadow_method_2")
    with pytest.raises(ValueError):
        smu.run_shadow_mapping(res, cam, light, cg[:n].view(res), lg[:n].view(res), dev, "shadow_method_3")
