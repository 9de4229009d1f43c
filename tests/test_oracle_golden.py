"""Pin the CPU oracle (oracle/nerf_oracle.py) against outputs of the reference itself.

The fixtures under tests/golden/ were produced by tests/make_golden.py, which imports
/root/reference/models/*.py unmodified and injects the random tensors.  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O

G = os.path.join(os.path.dirname(__file__), "golden")


def _case(npz, name):
    pre = name + "/"
    return {k[len(pre):]: npz[k] for k in npz.files if k.startswith(pre)}


RENDER_CASES = [("rgb_train", "rgb"), ("rgb_test", "rgb"), ("rgb_disp_fork_bounds", "rgb"),
                ("depth_train", "depth"), ("rgb_sm_train", "rgb_sm")]


@pytest.mark.parametrize("name,variant", RENDER_CASES)
def test_render_rays_matches_reference(name, variant):
    c = _case(np.load(os.path.join(G, "render_rays.npz")), name)
    N, Nc, Nf, W, use_disp, white_back, test_time = [int(v) for v in c["cfg"]]
    perturb, noise_std, near, far = [float(v) for v in c["cfgf"]]
    params = []
    for s in (10, 11):
        p = O.init_mlp_params(W=W, generator=torch.Generator().manual_seed(s))
        params.append({k: v.requires_grad_() for k, v in p.items()})
    noise = {k[len("noise_"):]: torch.from_numpy(v) for k, v in c.items() if k.startswith("noise_")}
    res = O.render_rays(params, torch.from_numpy(c["rays"]), n_samples=Nc, use_disp=bool(use_disp),
                        perturb=perturb, noise_std=noise_std, n_importance=Nf, white_back=bool(white_back),
                        test_time=bool(test_time), variant=variant, noise=noise)
    assert sorted(res.keys()) == list(c["keys"])
    loss = 0
    for k in res:
        np.testing.assert_allclose(res[k].detach().numpy(), c["out_" + k], rtol=1e-6, atol=1e-7, err_msg=k)
        loss = loss + (res[k] * torch.from_numpy(c["probe_" + k])).sum()
    loss.backward()
    for mi, p in enumerate(params):
        for pn, pv in p.items():
            g = pv.grad.numpy() if pv.grad is not None else np.zeros(pv.shape, np.float32)
            ref = c[f"grad{mi}_{pn}"]
            np.testing.assert_allclose(g, ref, rtol=1e-4, atol=1e-6 * max(1.0, np.abs(ref).max()), err_msg=f"{mi}/{pn}")


def test_embedding_and_nerf_match_reference():
    c = np.load(os.path.join(G, "nerf_modules.npz"))
    x3 = torch.from_numpy(c["x3"])
    assert np.array_equal(O.embed(x3, 10).numpy(), c["emb10"])
    assert np.array_equal(O.embed(x3, 4).numpy(), c["emb4"])
    assert np.array_equal(O.embed(x3, 4, logscale=False).numpy(), c["emb4_lin"])
    p = O.init_mlp_params(generator=torch.Generator().manual_seed(20))
    assert abs(float(sum(v.double().sum() for v in p.values())) - float(c["nerf256_wsum"][0])) < 1e-9
    assert O.mlp_param_names() == list(c["nerf_keys"])
    assert [str(tuple(p[k].shape)) for k in O.mlp_param_names()] == list(c["nerf_shapes"])
    x90 = torch.from_numpy(c["x90"])
    np.testing.assert_allclose(O.mlp_forward(p, x90).numpy(), c["nerf256_full"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(O.mlp_forward(p, x90[:, :63], sigma_only=True).numpy(), c["nerf256_sigma"],
                               rtol=1e-6, atol=1e-7)


def test_sample_pdf_matches_reference_bit_exact():
    c = np.load(os.path.join(G, "sample_pdf.npz"))
    z = O.sample_pdf(torch.from_numpy(c["rays"]), torch.from_numpy(c["w"]), c["u"].shape[1],
                     torch.from_numpy(c["u"]), torch.from_numpy(c["jitter"]))
    assert np.array_equal(z.numpy(), c["z"])


def test_searchsorted_known_answers():
    c = np.load(os.path.join(G, "searchsorted.npz"))
    for i in range(int(c["n_cases"])):
        for side in ("left", "right"):
            out = O.searchsorted_rows(c[f"c{i}_a"], c[f"c{i}_v"], side)
            assert out.dtype == np.int64 and np.array_equal(out, c[f"c{i}_{side}"])
            # torch.searchsorted is what the depth-only variant calls (rendering_shadows.py:41)
            if c[f"c{i}_a"].shape[0] == c[f"c{i}_v"].shape[0]:
                t = torch.searchsorted(torch.from_numpy(c[f"c{i}_a"]), torch.from_numpy(c[f"c{i}_v"]),
                                       right=(side == "right"))
                assert np.array_equal(t.numpy(), out)
    assert np.array_equal(O.searchsorted_rows(c["tie_a"], c["tie_v"], "right"), c["tie_right"])
    assert np.array_equal(O.searchsorted_rows(c["tie_a"], c["tie_v"], "left"), c["tie_left"])


def test_searchsorted_reference_extension_if_built():
    """oracle/_ref/ holds the reference's own C++ searchsorted compiled from
    /root/reference/torchsearchsorted/src/cpu (oracle/build_ref.py).  When present it must
    agree with the restatement on the golden vectors."""
    so = os.path.join(os.path.dirname(G), "..", "oracle", "_ref")
    from oracle import ref_loader
    ext = ref_loader.load_reference_searchsorted()
    if ext is None:
        pytest.skip("oracle/_ref not built")
    c = np.load(os.path.join(G, "searchsorted.npz"))
    for i in range(int(c["n_cases"])):
        a, v = torch.from_numpy(c[f"c{i}_a"]), torch.from_numpy(c[f"c{i}_v"])
        for side in ("left", "right"):
            out = torch.empty(max(a.shape[0], v.shape[0]), v.shape[1], dtype=torch.long)
            ext.searchsorted_cpu_wrapper(a, v, out, side == "left")
            assert np.array_equal(out.numpy(), c[f"c{i}_{side}"])
    a, v = torch.from_numpy(c["tie_a"]), torch.from_numpy(c["tie_v"])
    out = torch.empty(1, v.shape[1], dtype=torch.long)
    ext.searchsorted_cpu_wrapper(a, v, out, False)
    assert np.array_equal(out.numpy(), c["tie_right"])


@pytest.mark.parametrize("method", ["shadow_method_1", "shadow_method_2"])
@pytest.mark.parametrize("lni", [0, 8])
def test_efficient_sm_matches_reference(method, lni):
    c = np.load(os.path.join(G, "shadow.npz"))
    tag = f"{method}_lni{lni}"
    t = lambda k: torch.from_numpy(c[k])
    cd = {k: t(f"{tag}/cam_{k}").requires_grad_() for k in ("depth_coarse", "depth_fine")}
    ld = {k: t(f"{tag}/light_{k}").requires_grad_() for k in ("depth_coarse", "depth_fine")}
    out = O.efficient_sm(t("cam_pixels"), t("light_pixels"), dict(cd), dict(ld),
                         {"eye_pos": t("eye_pos"), "camera": t("camera")}, t("light_camera"), t("light_eye"),
                         tuple(int(v) for v in c["res"]), True, lni, method)
    np.testing.assert_allclose(out["rgb_coarse"].detach().numpy(), c[f"{tag}/rgb_coarse"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(out["rgb_fine"].detach().numpy(), c[f"{tag}/rgb_fine"], rtol=1e-6, atol=1e-7)
    probe = t(f"{tag}/probe")
    ((out["rgb_coarse"] * probe).sum() + 2 * (out["rgb_fine"] * probe).sum()).backward()
    for k in cd:
        for grp, d in (("gcam", cd), ("glight", ld)):
            g = d[k].grad.numpy() if d[k].grad is not None else np.zeros(d[k].shape, np.float32)
            ref = c[f"{tag}/{grp}_{k}"]
            np.testing.assert_allclose(g, ref, rtol=1e-4, atol=1e-6 * max(1.0, np.abs(ref).max()), err_msg=f"{grp}_{k}")


def test_camera_algebra_matches_reference():
    c = np.load(os.path.join(G, "shadow.npz"))
    np.testing.assert_allclose(O.camera_matrix(30, (16, 16)).numpy(), c["cam_init_30_16"], rtol=1e-6)
    R, Q = O.transformation_to(torch.from_numpy(c["cam0_M"]), torch.from_numpy(c["cam0_eye"]),
                               torch.from_numpy(c["light_camera"]), torch.from_numpy(c["light_eye"]))
    np.testing.assert_allclose(R.numpy(), c["R01"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(Q.numpy(), c["Q01"], rtol=1e-6, atol=1e-7)


# ---- the rows either side of the path: ray generation, MSE loss, psnr -------------------

def _rays_loss():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rays_loss.npz"))


@pytest.mark.parametrize("name", ["blender_small", "fork_bounds", "llff_ndc"])
def test_oracle_ray_generation_vs_reference(name):
    """oracle get_ray_directions / get_rays / get_ndc_rays / dataset rows == datasets/ray_utils.py run by
    tests/make_golden_rays.py (same torch ops: bit-exact)."""
    z = _rays_loss()
    H, W, focal, near, far, ndc = z[f"{name}/hwf"]
    H, W, ndc = int(H), int(W), bool(ndc)
    c2w = torch.from_numpy(z[f"{name}/c2w"])
    d = O.get_ray_directions(H, W, float(focal))
    np.testing.assert_array_equal(d.numpy(), z[f"{name}/directions"])
    o, dd = O.get_rays(d, c2w)
    np.testing.assert_array_equal(o.numpy(), z[f"{name}/rays_o"])
    np.testing.assert_array_equal(dd.numpy(), z[f"{name}/rays_d"])
    if ndc:
        no, nd = O.get_ndc_rays(H, W, float(focal), float(near), o, dd)
        np.testing.assert_array_equal(no.numpy(), z[f"{name}/ndc_o"])
        np.testing.assert_array_equal(nd.numpy(), z[f"{name}/ndc_d"])
    rays = O.dataset_rays(c2w, H, W, float(focal), float(near), float(far), ndc)
    np.testing.assert_array_equal(rays.numpy(), z[f"{name}/rays"])


@pytest.mark.parametrize("name", ["loss_small", "loss_batch"])
def test_oracle_mse_psnr_vs_reference(name):
    z = _rays_loss()
    a = torch.from_numpy(z[f"{name}/a"]).requires_grad_()
    b = torch.from_numpy(z[f"{name}/b"]).requires_grad_()
    t = torch.from_numpy(z[f"{name}/t"])
    loss = O.mse_loss({"rgb_coarse": a, "rgb_fine": b}, t)
    loss.backward()
    np.testing.assert_allclose(loss.detach().numpy(), z[f"{name}/mse2"], rtol=1e-6)
    np.testing.assert_allclose(a.grad.numpy(), z[f"{name}/ga"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(b.grad.numpy(), z[f"{name}/gb"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(O.mse_loss({"rgb_coarse": a.detach()}, t).numpy(), z[f"{name}/mse1"], rtol=1e-6)
    np.testing.assert_allclose(O.mse_loss({"sm_coarse": a.detach(), "sm_fine": b.detach()}, t, "sm").numpy(), z[f"{name}/sm2"], rtol=1e-6)
    np.testing.assert_allclose(O.psnr(a.detach(), t).numpy(), z[f"{name}/psnr"], rtol=1e-6)
    np.testing.assert_allclose(O.psnr(a.detach(), t, torch.from_numpy(z[f"{name}/mask"])).numpy(), z[f"{name}/psnr_masked"], rtol=1e-6)


def _adam():
    return np.load(os.path.join(G, "adam.npz"))


@pytest.mark.parametrize("name", ["wd0", "wd1e-4"])
def test_oracle_adam_vs_torch_golden(name):
    """oracle.adam_step against five recorded torch.optim.Adam steps (tests/make_golden_adam.py).  fp32 evaluation
    order differs from torch's lerp/addcdiv formulation, so the bar is a few ulp of the update, not bit-exact."""
    z = _adam()
    p = torch.from_numpy(z[f"{name}.p0"])
    m = torch.zeros_like(p)
    v = torch.zeros_like(p)
    lr, wd = float(z[f"{name}.lr"]), float(z[f"{name}.weight_decay"])
    for t in range(z[f"{name}.grads"].shape[0]):
        p, m, v = O.adam_step(p, torch.from_numpy(z[f"{name}.grads"][t]), m, v, t + 1, lr, eps=1e-8, weight_decay=wd)
        want = torch.from_numpy(z[f"{name}.params"][t])
        assert torch.allclose(p, want, rtol=0, atol=4e-7 * lr / 5e-4), (t, (p - want).abs().max())
    # moments are signed sums: the error scales with the largest gradient seen per element, not with the sum
    gmax = torch.from_numpy(np.abs(z[f"{name}.grads"]).max(0)) + wd * 1.0
    assert ((m - torch.from_numpy(z[f"{name}.exp_avg"])).abs() <= 1e-6 * gmax).all()
    assert ((v - torch.from_numpy(z[f"{name}.exp_avg_sq"])).abs() <= 1e-6 * gmax * gmax).all()


def test_oracle_adam_vs_live_torch():
    """Same restatement against the installed torch.optim.Adam on a fresh seed (pins the golden file's generator)."""
    g = torch.Generator().manual_seed(5)
    p = torch.nn.Parameter(torch.randn(1000, generator=g))
    opt = torch.optim.Adam([p], lr=1e-3, eps=1e-8, weight_decay=1e-2)
    q, m, v = p.detach().clone(), torch.zeros(1000), torch.zeros(1000)
    for t in range(3):
        gr = torch.randn(1000, generator=g)
        p.grad = gr.clone()
        opt.step()
        q, m, v = O.adam_step(q, gr, m, v, t + 1, 1e-3, eps=1e-8, weight_decay=1e-2)
        assert torch.allclose(q, p.detach(), rtol=0, atol=1e-6)


# ---- the light-sampler trainer's shadow-mapping pieces -----------------------------------

def _ls():
    z = np.load(os.path.join(G, "lightsampler.npz"))
    return z, (lambda k: torch.from_numpy(z[k]))


@pytest.mark.parametrize("tag", ["split", "single"])
def test_oracle_get_K_vs_reference(tag):
    z, T = _ls()
    cd = {k: T(f"getK_{tag}/{k}").clone().requires_grad_() for k in ("depth_coarse", "depth_fine")}
    ppc = {"eye_pos": T(f"getK_{tag}/eye_pos"), "camera": T(f"getK_{tag}/camera")}
    Kc, Kf = O.get_K(T(f"getK_{tag}/cam_pixels"), cd, ppc, T("light_camera"), T("light_eye"), True)
    assert torch.allclose(Kc, T(f"getK_{tag}/K_coarse"), rtol=1e-5, atol=1e-5)
    assert torch.allclose(Kf, T(f"getK_{tag}/K_fine"), rtol=1e-5, atol=1e-5)
    probe = T(f"getK_{tag}/probe")
    ((Kc * probe).sum() + 2 * (Kf * probe).sum()).backward()
    for k in cd:
        assert torch.allclose(cd[k].grad, T(f"getK_{tag}/g_{k}"), rtol=1e-4, atol=1e-5), k
    assert O.get_K(T(f"getK_{tag}/cam_pixels"), cd, ppc, T("light_camera"), T("light_eye"), False)[1] is None


def test_oracle_shadow_pieces_vs_reference():
    z, T = _ls()
    res = tuple(int(v) for v in z["res"])
    K = O.projections(T("proj/cam_M"), T("proj/cam_eye"), T("light_camera"), T("light_eye"),
                      torch.cat([T("proj/pix"), T("proj/depth").view(-1, 1)], 1))
    assert torch.allclose(K, T("proj/K"), rtol=1e-5, atol=1e-5)
    R, Q = O.transformation_to(T("proj/cam_M"), T("proj/cam_eye"), T("light_camera"), T("light_eye"))
    assert torch.allclose(O.diff_projections(T("proj/pix"), T("diffproj/w_cam"), R, Q), T("diffproj/K"), rtol=1e-5, atol=1e-5)
    wl, wlb = O.projected_depths(res, T("gather/K"), T("gather/w_light"))
    assert torch.equal(wl, T("gather/wl")) and torch.equal(wlb, T("gather/wlb"))
    for tag, kw in (("m1", dict(mode="shadow_method_1")), ("m1_eps", dict(mode="shadow_method_1", delta=0.05, epsilon=0.2)),
                    ("m2", dict(mode="shadow_method_2")), ("m2_sig", dict(mode="shadow_method_2", sigmoid=True)),
                    ("m2_range", dict(mode="shadow_method_2", new_min=-0.5, new_max=1.5))):
        a, b = T(f"map_{tag}/wl").clone().requires_grad_(), T(f"map_{tag}/wlb").clone().requires_grad_()
        sm = O.shadow_map(a, b, **kw)
        assert torch.allclose(sm, T(f"map_{tag}/out"), rtol=1e-6, atol=1e-7), tag
        (sm * T(f"map_{tag}/probe")).sum().backward()
        assert torch.allclose(a.grad, T(f"map_{tag}/g_wl"), rtol=1e-4, atol=1e-5), tag
        assert torch.allclose(b.grad, T(f"map_{tag}/g_wlb"), rtol=1e-4, atol=1e-5), tag
    assert torch.allclose(O.normalize_min_max(T("norm/t"), new_max=2.0, new_min=-1.0), T("norm/out"), rtol=1e-6, atol=1e-7)


# ---- whole-image shadow mapping (models/shadow_mapping_utils.py), tests/make_golden_shadow_image.py ----------------------
def _si():
    z = np.load(os.path.join(G, "shadow_image.npz"))
    return z, (lambda k: torch.from_numpy(z[k]))


def test_oracle_shadow_image_pieces_vs_reference():
    z, T = _si()
    res = tuple(int(v) for v in z["res"])
    assert int(z["wrapper_raises"]) == 1        # rendering_shadows.shadow_mapping() raises in the reference
    assert torch.allclose(O.image_w_mapping(res, T("cam_M"), T("wmap/range")), T("wmap/w"), rtol=1e-6, atol=0)
    K = O.image_projections(res, T("proj/w_cam"), T("cam_M"), T("cam_eye"), T("light_M"), T("light_eye"))
    assert torch.allclose(K, T("proj/K"), rtol=1e-5, atol=1e-5)
    wl, wlb, ul_, vl_ = O.image_depths(res, T("depths/K"), T("depths/w_light"))
    assert torch.equal(wl, T("depths/wl")) and torch.equal(wlb, T("depths/wlb"))
    for tag, kw in (("m1", dict(mode="shadow_method_1")), ("m1_eps", dict(mode="shadow_method_1", delta=0.05, epsilon=0.2)),
                    ("m2", dict(mode="shadow_method_2")), ("m2_sig", dict(mode="shadow_method_2", sigmoid=True)),
                    ("m2_range", dict(mode="shadow_method_2"))):
        a, b = T(f"map_{tag}/wl").clone().requires_grad_(), T(f"map_{tag}/wlb").clone().requires_grad_()
        sm = O.image_shadow_map(a, b, **kw)
        assert torch.allclose(sm, T(f"map_{tag}/out"), rtol=1e-6, atol=1e-7), tag
        (sm * T(f"map_{tag}/probe")).sum().backward()
        assert torch.allclose(a.grad, T(f"map_{tag}/g_wl"), rtol=1e-4, atol=1e-5), tag
        assert torch.allclose(b.grad, T(f"map_{tag}/g_wlb"), rtol=1e-4, atol=1e-5), tag
    assert torch.allclose(O.image_normalize_min_max(T("norm/t"), new_max=2.0, new_min=-1.0), T("norm/out"), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("tag,mode", [("run_m1", "shadow_method_1"), ("run_m2", "shadow_method_2")])
def test_oracle_shadow_image_run_vs_reference(tag, mode):
    z, T = _si()
    res = tuple(int(v) for v in z["res"])
    rc, rl = T(f"{tag}/range_cam").clone().requires_grad_(), T(f"{tag}/range_light").clone().requires_grad_()
    sm = O.image_run_shadow_mapping(res, T("cam_M"), T("cam_eye"), T("light_M"), T("light_eye"), rc, rl, mode)
    assert torch.allclose(sm, T(f"{tag}/out"), rtol=1e-5, atol=1e-6)
    (sm * T(f"{tag}/probe")).sum().backward()
    for g, k in ((rc.grad, "g_range_cam"), (rl.grad, "g_range_light")):
        ref = T(f"{tag}/{k}")
        assert torch.allclose(g, ref, rtol=1e-4, atol=1e-5 * float(ref.abs().max())), k


# ---- sgd / radam / ranger (utils/__init__.py:14-27, utils/optimizers.py), tests/make_golden_optim.py ---------------------
OPT_CASES = {
    "sgd_plain": ("sgd", dict(momentum=0.0, weight_decay=0.0)),
    "sgd_mom_wd": ("sgd", dict(momentum=0.9, weight_decay=1e-4)),
    "radam_wd0": ("radam", dict(weight_decay=0.0)),
    "radam_wd": ("radam", dict(weight_decay=1e-4)),
    "radam_nosgd": ("radam", dict(weight_decay=1e-4, degenerated_to_sgd=False)),
    "ranger_wd0": ("ranger", dict(eps=1e-8, weight_decay=0.0)),
    "ranger_wd": ("ranger", dict(eps=1e-8, weight_decay=1e-4)),
    "ranger_k2": ("ranger", dict(eps=1e-8, alpha=0.3, k=2, weight_decay=0.0)),
}


@pytest.mark.parametrize("name", list(OPT_CASES))
def test_oracle_optimizers_vs_reference_golden(name):
    """oracle.sgd_step / radam_step / ranger_step against eight recorded steps of torch.optim.SGD and of the reference's
    own RAdam / Ranger classes.  Tolerance: 1e-6 of the parameter scale per step (fp32 evaluation order)."""
    z = np.load(os.path.join(G, "optim.npz"))
    kind, kw = OPT_CASES[name]
    lr = float(z["lr"])
    p = torch.from_numpy(z[f"{name}.p0"])
    n = p.numel()
    m, v, slow, buf = torch.zeros(n), torch.zeros(n), p.clone(), None
    for t in range(z[f"{name}.grads"].shape[0]):
        g = torch.from_numpy(z[f"{name}.grads"][t])
        if kind == "sgd":
            p, buf = O.sgd_step(p, g, buf, lr, **kw)
        elif kind == "radam":
            p, m, v = O.radam_step(p, g, m, v, t + 1, lr, **kw)
        else:
            p, m, v, slow = O.ranger_step(p, g, m, v, slow, t + 1, lr, **kw)
        ref = torch.from_numpy(z[f"{name}.params"][t])
        assert torch.allclose(p, ref, rtol=1e-5, atol=1e-7), (name, t, float((p - ref).abs().max()))
    gs = torch.from_numpy(np.abs(z[f"{name}.grads"]).max(0))       # per-element gradient scale (six decades)
    if kind != "sgd":
        assert bool(((m - torch.from_numpy(z[f"{name}.exp_avg"])).abs() <= 1e-6 * gs).all())
        assert bool(((v - torch.from_numpy(z[f"{name}.exp_avg_sq"])).abs() <= 1e-6 * gs * gs).all())
    if kind == "ranger":
        assert torch.allclose(slow, torch.from_numpy(z[f"{name}.slow_buffer"]), rtol=1e-5, atol=1e-7)
    if name == "sgd_mom_wd":
        assert bool(((buf - torch.from_numpy(z[f"{name}.momentum_buffer"])).abs() <= 1e-5 * (gs + 1e-5)).all())
    if name == "radam_nosgd":      # no parameter change while N_sma < 5 (steps 1-5)
        assert np.array_equal(z[f"{name}.params"][4], z[f"{name}.p0"])
