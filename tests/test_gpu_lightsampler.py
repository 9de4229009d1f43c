"""GPU parity of the shadow-mapping pieces the light-sampler trainer calls one by one (SURVEY §8f rank 2;
train_light_sampler.py:15, :141-196) against vectors produced by the reference's own functions
(tests/make_golden_lightsampler.py) and against the CPU oracle at larger sizes.
Tolerances: gathered depths / indices bit-exact; projections 1e-5 relative (fp32 3x3 chains in another order);
shadow maps 1e-6; gradients 1e-4 relative."""
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
    return np.load(os.path.join(G, "lightsampler.npz"))


def _cam(z, M, eye):
    from nerf_pl_b200.models.camera import Camera
    return Camera.from_camera_eyepos(torch.from_numpy(z[eye]), torch.from_numpy(z[M]))


def _close(a, ref, rtol, atol, msg=""):
    np.testing.assert_allclose(a.detach().cpu().numpy(), ref, rtol=rtol, atol=atol, err_msg=msg)


@pytest.mark.parametrize("tag", ["split", "single"])
@pytest.mark.parametrize("module", ["rendering_shadows", "rendering_rgb_sm"])
def test_get_K_matches_reference_golden(dev, z, tag, module):
    import importlib
    mod = importlib.import_module(f"nerf_pl_b200.models.{module}")
    T = lambda k: torch.from_numpy(z[k]).to(dev)
    light = _cam(z, "light_camera", "light_eye")
    cd = {k: T(f"getK_{tag}/{k}").requires_grad_() for k in ("depth_coarse", "depth_fine")}
    ppc = {"eye_pos": torch.from_numpy(z[f"getK_{tag}/eye_pos"]), "camera": torch.from_numpy(z[f"getK_{tag}/camera"])}
    pix = torch.from_numpy(z[f"getK_{tag}/cam_pixels"])
    Kc, Kf = mod.get_K(pix, dict(cd), ppc, light, True)
    _close(Kc, z[f"getK_{tag}/K_coarse"], 1e-5, 1e-5)
    _close(Kf, z[f"getK_{tag}/K_fine"], 1e-5, 1e-5)
    probe = T(f"getK_{tag}/probe")
    ((Kc * probe).sum() + 2 * (Kf * probe).sum()).backward()
    for k in cd:
        _close(cd[k].grad, z[f"getK_{tag}/g_{k}"], 1e-4, 1e-5, k)
    Kc2, none = mod.get_K(pix, {k: v.detach() for k, v in cd.items()}, ppc, light, False)
    assert none is None and torch.equal(Kc2, Kc.detach())


def test_get_K_many_runs_vs_oracle(dev):
    """4096 rays drawn from 5 cameras in ragged same-camera runs (1..400 rays, cameras recurring): the device-side run
    detection must reproduce the reference's host loop, which uses the FIRST ray's camera of every run."""
    from nerf_pl_b200.models.camera import Camera
    from nerf_pl_b200.models import rendering_shadows
    g = torch.Generator().manual_seed(5)
    n, res = 4096, (128, 128)
    def posed(eye):
        c2w = torch.from_numpy(Camera.c2w_from_lookat(np.array(eye, dtype=np.float32), np.zeros(3, dtype=np.float32)))
        e, M = O.pose_camera(O.camera_matrix(30, res), c2w[:3, :4])
        return M, e

    cams = [posed(e) for e in ([0., 1., 4.], [2., 1.5, 3.], [-1., 2., 3.5], [3., 0.5, -2.], [0.5, 3., 3.])]
    seg, lens = [], [1, 1, 2, 400, 33, 1, 257]
    while len(seg) < n:
        c = int(torch.randint(0, 5, (1,), generator=g))
        if seg and seg[-1] == c:
            continue
        seg += [c] * lens[len(seg) % len(lens)]
    seg = seg[:n]
    ppc = {"eye_pos": torch.stack([cams[s][1] for s in seg]), "camera": torch.stack([cams[s][0] for s in seg])}
    LM, Leye = posed([1.0, 4.0, 2.0])
    pix = torch.cat([torch.rand(n, 2, generator=g) * 128, torch.ones(n, 1)], 1)
    cd = {k: (3.0 + 2.0 * torch.rand(n, generator=g)).requires_grad_() for k in ("depth_coarse", "depth_fine")}
    Kc_ref, Kf_ref = O.get_K(pix, cd, ppc, LM, Leye, True)
    probe = torch.rand(n, 3, generator=g)
    ((Kc_ref * probe).sum() + (Kf_ref * probe).sum()).backward()
    light = Camera.from_camera_eyepos(Leye, LM)
    gd = {k: v.detach().to(dev).requires_grad_() for k, v in cd.items()}
    Kc, Kf = rendering_shadows.get_K(pix, dict(gd), ppc, light, True)
    scale = float(Kc_ref.detach().abs().max())
    _close(Kc, Kc_ref.detach().numpy(), 1e-4, 1e-5 * scale)
    _close(Kf, Kf_ref.detach().numpy(), 1e-4, 1e-5 * scale)
    ((Kc * probe.to(dev)).sum() + (Kf * probe.to(dev)).sum()).backward()
    for k in cd:
        ref = cd[k].grad.numpy()
        _close(gd[k].grad, ref, 1e-3, 1e-5 * float(np.abs(ref).max()), k)


def test_projection_helpers_match_reference_golden(dev, z):
    from nerf_pl_b200.models import efficient_shadow_mapping as esm
    T = lambda k: torch.from_numpy(z[k]).to(dev)
    cam, light = _cam(z, "proj/cam_M", "proj/cam_eye"), _cam(z, "light_camera", "light_eye")
    depth = T("proj/depth").requires_grad_()
    K = esm.get_projections(cam, light, torch.cat([T("proj/pix"), depth.view(-1, 1)], 1), dev)
    _close(K, z["proj/K"], 1e-5, 1e-5)
    (K * T("proj/probe")).sum().backward()
    _close(depth.grad, z["proj/g_depth"], 1e-4, 1e-5)
    w_cam = T("diffproj/w_cam").requires_grad_()
    K2 = esm.get_diff_projections(T("proj/pix"), w_cam, cam, light, dev)
    _close(K2, z["diffproj/K"], 1e-5, 1e-5)
    (K2 * T("proj/probe")).sum().backward()
    _close(w_cam.grad, z["diffproj/g_w_cam"], 1e-4, 1e-5)
    # get_normed_w: the (n,4) [pixels, w] of the same chain
    nw = esm.get_normed_w(cam, torch.cat([T("proj/pix"), T("proj/depth").view(-1, 1)], 1), dev)
    ref = O.normed_w(torch.from_numpy(z["proj/cam_M"]),
                     torch.cat([torch.from_numpy(z["proj/pix"]), torch.from_numpy(z["proj/depth"]).view(-1, 1)], 1))
    _close(nw, ref.numpy(), 1e-6, 1e-6)


def test_projected_depths_bit_exact(dev, z):
    from nerf_pl_b200.models import efficient_shadow_mapping as esm
    T = lambda k: torch.from_numpy(z[k]).to(dev)
    res = tuple(int(v) for v in z["res"])
    K, w_light = T("gather/K").requires_grad_(), T("gather/w_light").requires_grad_()
    wl, wlb = esm.get_projected_depths(res, K, w_light, dev)
    assert np.array_equal(wl.detach().cpu().numpy(), z["gather/wl"])
    assert np.array_equal(wlb.detach().cpu().numpy(), z["gather/wlb"])
    pr = T("gather/probe")
    ((wl * pr).sum() + 3 * (wlb * pr).sum()).backward()
    _close(K.grad, z["gather/g_K"], 0, 0)
    _close(w_light.grad, z["gather/g_w_light"], 1e-6, 1e-7)
    # larger, with many collisions on one texel: indices bit-exact against the oracle, scatter within fp32 add order
    g = torch.Generator().manual_seed(9)
    n, res = 100000, (64, 64)
    Kb = torch.stack([torch.rand(n, generator=g) * 80 - 8, torch.rand(n, generator=g) * 80 - 8, torch.rand(n, generator=g)], 1)
    wlt = torch.rand(res[0] * res[1], generator=g).requires_grad_()
    _, ref = O.projected_depths(res, Kb, wlt)
    pr = torch.rand(n, generator=g)
    (ref * pr).sum().backward()
    wg = wlt.detach().to(dev).requires_grad_()
    _, got = esm.get_projected_depths(res, Kb.to(dev), wg, dev)
    assert torch.equal(got.detach().cpu(), ref.detach())
    (got * pr.to(dev)).sum().backward()
    _close(wg.grad, wlt.grad.numpy(), 1e-4, 1e-4)


@pytest.mark.parametrize("tag,kw", MAP_CASES)
def test_generate_shadow_map_matches_reference_golden(dev, z, tag, kw):
    from nerf_pl_b200.models import efficient_shadow_mapping as esm
    T = lambda k: torch.from_numpy(z[k]).to(dev)
    a, b = T(f"map_{tag}/wl").requires_grad_(), T(f"map_{tag}/wlb").requires_grad_()
    sm = esm.generate_shadow_map(a, b, device=dev, **kw)
    assert sm.shape == (a.numel(), 3)
    _close(sm, z[f"map_{tag}/out"], 1e-6, 1e-6)
    (sm * T(f"map_{tag}/probe")).sum().backward()
    _close(a.grad, z[f"map_{tag}/g_wl"], 1e-4, 1e-5, "g_wl")
    _close(b.grad, z[f"map_{tag}/g_wlb"], 1e-4, 1e-5, "g_wlb")


def test_generate_shadow_map_large_vs_oracle_and_errors(dev):
    from nerf_pl_b200.models import efficient_shadow_mapping as esm
    g = torch.Generator().manual_seed(2)
    n = 640000
    a, b = 3 + torch.rand(n, generator=g), 3 + torch.rand(n, generator=g)
    pr = torch.rand(n, 3, generator=g)
    for kw in (dict(mode="shadow_method_1"), dict(mode="shadow_method_2", sigmoid=True)):
        ar, br = a.clone().requires_grad_(), b.clone().requires_grad_()
        ref = O.shadow_map(ar, br, **kw)
        (ref * pr).sum().backward()
        ag, bg = a.to(dev).requires_grad_(), b.to(dev).requires_grad_()
        sm = esm.generate_shadow_map(ag, bg, device=dev, **kw)
        _close(sm, ref.detach().numpy(), 1e-5, 1e-6)
        (sm * pr.to(dev)).sum().backward()
        # method 2's gradient carries the global min / max terms: sums of 1.9 M products, fp32 order differs
        _close(ag.grad, ar.grad.numpy(), 1e-3, 1e-4 * float(ar.grad.abs().max()))
        _close(bg.grad, br.grad.numpy(), 1e-3, 1e-4 * float(br.grad.abs().max()))
    with pytest.raises(ValueError):
        esm.generate_shadow_map(a.to(dev), b.to(dev), mode="shadow_method_3")
    with pytest.raises(RuntimeError):
        esm.generate_shadow_map(a, b)                      # CPU tensors: no fallback


def test_normalize_min_max_matches_reference_golden(dev, z):
    from nerf_pl_b200.models import efficient_shadow_mapping as esm
    t = torch.from_numpy(z["norm/t"]).to(dev).requires_grad_()
    y = esm.normalize_min_max(t, new_max=2.0, new_min=-1.0)
    assert y.shape == t.shape
    _close(y, z["norm/out"], 1e-6, 1e-6)
    (y * torch.from_numpy(z["norm/probe"]).to(dev)).sum().backward()
    _close(t.grad, z["norm/g_t"], 1e-4, 1e-5)
