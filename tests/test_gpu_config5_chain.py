"""Config 5 as a PATH (BASELINE.json configs[4]; reference train_efficient_sm.py:139-208): depth-only render_rays of a
4096-ray camera batch (contiguous pixels of one 128 x 128 view, bounds 1..200) + the 16 384-ray light-view depth render +
efficient_sm(shadow_method_2) (models/rendering_shadows.py:359-482) + MSE loss + backward() into the weights of both
NeRF models, against oracle.render_rays x 2 -> oracle.efficient_sm on identical injected random tensors, for
Light_N_importance in {0, 64} x grad_on_light in {False, True} (train_efficient_sm.py:150-172), in fp32 mode and with the
bf16 tensor-core MLP; plus a batch that straddles three cameras (the per-segment loop of rendering_shadows.py:385-406).

The oracle side is evaluated in ray chunks with the two-phase chain rule of tests/_parity.py (the min-max normalisation
of shadow_method_2 couples all rays of a segment).  Shadow mapping is DIScontinuous in the depths (the light texel a
camera ray lands in is `int(ul), int(vl)`, efficient_shadow_mapping.py:85-100): with the bf16 MLP a small fraction of the
rays lands in a neighbouring texel; that fraction is measured, bounded and the values are compared on the others."""
import math

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
import _parity as P

pytestmark = pytest.mark.gpu
RES = (128, 128)


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as ge
    ge.build()
    return torch.device("cuda:0")


def _blender_pose(theta, phi, radius=4.0):
    th, ph = math.radians(theta), math.radians(phi)
    t_r = torch.tensor([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, radius], [0, 0, 0, 1.0]])
    r_ph = torch.tensor([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0], [0, math.sin(ph), math.cos(ph), 0], [0, 0, 0, 1.0]])
    r_th = torch.tensor([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0], [math.sin(th), 0, math.cos(th), 0], [0, 0, 0, 1.0]])
    flip = torch.tensor([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1.0]])
    return (flip @ r_th @ r_ph @ t_r)[:3, :4].contiguous()


def _view(theta, phi):
    """(M (3,3), eye (3)) of a 30-degree 128 x 128 view posed by a Blender matrix - the ORACLE's camera restatement
    (oracle.camera_matrix / pose_camera, models/camera.py:20-31, :69-89)."""
    eye, M = O.pose_camera(O.camera_matrix(30.0, RES), _blender_pose(theta, phi))
    return M, eye


def _rays_through(M, eye, pix_ij):
    """Rays through pixel centres (datasets/blender_efficient_sm.py:98-110): d = normalize(M [i+.5, j+.5, 1]), bounds 1..200."""
    p = torch.cat([pix_ij + 0.5, torch.ones_like(pix_ij[:, :1])], 1)
    d = torch.nn.functional.normalize(p @ M.T, dim=-1)
    o = eye.expand_as(d)
    return torch.cat([o, d, torch.full_like(d[:, :1], 1.0), torch.full_like(d[:, :1], 200.0)], 1).contiguous(), p


def _scene(segments):
    """segments: list of (theta, phi, first_pixel, n_pixels) camera runs; the light is one fixed view."""
    ij = torch.stack(torch.meshgrid(torch.arange(128.0), torch.arange(128.0), indexing="xy"), -1).reshape(-1, 2)
    rays, pix, cams, eyes = [], [], [], []
    for th, ph, first, n in segments:
        M, eye = _view(th, ph)
        r, p = _rays_through(M, eye, ij[first:first + n])
        rays.append(r); pix.append(p)
        cams.append(M.expand(n, 3, 3)); eyes.append(eye.expand(n, 3))
    LM, Leye = _view(75.0, -50.0)
    lrays, lpix = _rays_through(LM, Leye, ij)
    return (torch.cat(rays), torch.cat(pix), {"eye_pos": torch.cat(eyes).contiguous(), "camera": torch.cat(cams).contiguous()},
            lrays, lpix, LM, Leye)


_CACHE = {}


def _oracle_chain(name, segments, lni):
    """Oracle forward of the chain and the two gradient sets (grad_on_light False / True)."""
    if name in _CACHE:
        return _CACHE[name]
    rays_c, pix_c, ppc, rays_l, pix_l, LM, Leye = _scene(segments)
    n_c, n_l = rays_c.shape[0], rays_l.shape[0]
    params = P.leaf_params()
    noise_c = O.make_noise(n_c, 64, 64, seed=31)
    noise_l = O.make_noise(n_l, 64, lni, seed=32)
    kw_c = dict(n_samples=64, use_disp=False, perturb=1.0, noise_std=1.0, n_importance=64, white_back=False, variant="depth")
    kw_l = dict(kw_c, n_importance=lni)
    cam, aux_c = P.oracle_forward(params, rays_c, noise_c, **kw_c)
    light, aux_l = P.oracle_forward(params, rays_l, noise_l, **kw_l)
    tgt = torch.rand(n_c, 3, generator=torch.Generator().manual_seed(33))
    # phase 2: the loss's gradient w.r.t. the four depth maps
    leaves_c = {k: cam[k].clone().requires_grad_() for k in ("depth_coarse", "depth_fine")}
    leaves_l = {k: light[k].clone().requires_grad_() for k in light if k.startswith("depth")}
    out = O.efficient_sm(pix_c, pix_l, dict(leaves_c), dict(leaves_l), ppc, LM, Leye, RES, True, lni, "shadow_method_2")
    loss = ((out["rgb_coarse"] - tgt) ** 2).mean() + ((out["rgb_fine"] - tgt) ** 2).mean()      # losses.py:12-18
    loss.backward()
    g_c = {k: v.grad for k, v in leaves_c.items()}
    g_l = {k: (v.grad if v.grad is not None else None) for k, v in leaves_l.items()}
    # phase 3: camera render backward (= grad_on_light False), then the light render's contribution on top
    P.oracle_backward(params, rays_c, dict(noise_c, z_fine=aux_c["z_fine"]), g_c, **kw_c)
    grads_nolight = [{k: (v.grad.clone() if v.grad is not None else torch.zeros_like(v)) for k, v in p.items()} for p in params]
    nz_l = dict(noise_l, z_fine=aux_l["z_fine"]) if lni else noise_l
    P.oracle_backward(params, rays_l, nz_l, g_l, **kw_l)
    grads_light = [{k: (v.grad.clone() if v.grad is not None else torch.zeros_like(v)) for k, v in p.items()} for p in params]
    _CACHE[name] = dict(params=params, scene=(rays_c, pix_c, ppc, rays_l, pix_l, LM, Leye), noise=(noise_c, noise_l),
                        aux=(aux_c, aux_l), cam=cam, light=light, out={k: out[k].detach() for k in ("rgb_coarse", "rgb_fine")},
                        loss=float(loss.detach()), tgt=tgt, grads={False: grads_nolight, True: grads_light}, kw=(kw_c, kw_l))
    return _CACHE[name]


def _gpu_chain(dev, precision, case, lni, grad_on_light):
    import nerf_pl_b200
    from nerf_pl_b200.losses import loss_dict
    from nerf_pl_b200.models import rendering_shadows
    from nerf_pl_b200.models.camera import Camera
    from nerf_pl_b200.models.nerf import Embedding
    nerf_pl_b200.set_mlp_precision(precision)
    rays_c, pix_c, ppc, rays_l, pix_l, LM, Leye = case["scene"]
    noise_c, noise_l = case["noise"]
    aux_c, aux_l = case["aux"]
    models = P.gpu_models(case["params"], dev)
    emb = [Embedding(3, 10), Embedding(3, 4)]
    inj_c = {k: v.to(dev) for k, v in noise_c.items()}
    inj_c["z_fine"] = aux_c["z_fine"].to(dev)
    inj_l = {k: v.to(dev) for k, v in noise_l.items()}
    if lni:
        inj_l["z_fine"] = aux_l["z_fine"].to(dev)
    # train_efficient_sm.py:146 / :158-168 / :176-183
    cam = rendering_shadows.render_rays(models, emb, rays_c.to(dev), 64, False, 1.0, 1.0, 64, 1024 * 32, False, _inject=inj_c)
    with torch.set_grad_enabled(grad_on_light):
        light = rendering_shadows.render_rays(models, emb, rays_l.to(dev), 64, False, 1.0, 1.0, lni, 1024 * 32, False,
                                              _inject=inj_l)
    light_cam = Camera.from_camera_eyepos(Leye, LM)
    depths = {"cam": {k: cam[k].detach().cpu() for k in cam}, "light": {k: light[k].detach().cpu() for k in light}}
    out = rendering_shadows.efficient_sm(pix_c.to(dev), pix_l.to(dev), cam, light, {k: v.to(dev) for k, v in ppc.items()},
                                         light_cam, RES, True, lni, "shadow_method_2")
    loss = loss_dict["mse"]()(out, case["tgt"].to(dev))
    loss.backward()
    return models, depths, out, float(loss)


def _check(dev, name, case, precision, lni, grad_on_light):
    models, depths, out, loss = _gpu_chain(dev, precision, case, lni, grad_on_light)
    rec = {"loss_ref": case["loss"], "loss_got": loss}
    aux_c, aux_l = case["aux"]
    # (1) the depth maps that feed the shadow test
    d_tol = 2e-5 * 200.0 if precision == "fp32" else 3e-2        # bf16 at bounds 1..200: measured 1.6e-2 (tests/_parity.py)
    for side, ref, aux in (("cam", case["cam"], aux_c), ("light", case["light"], aux_l)):
        stable = P.last_sample_stable(aux)
        rec[f"stable_fraction/{side}"] = float(stable.mean())
        assert stable.mean() >= 0.95
        for k in ref:
            if not k.startswith("depth"):
                continue
            e = np.abs(depths[side][k].numpy() - ref[k].numpy())[stable].max()
            rec[f"max_abs_err/{side}_{k}"] = float(e)
            assert e < d_tol, f"{name} {side} {k}: {e:.3e}"
    # (2) the shadow values: rays that landed in the same light texel agree; the others are counted
    s_tol = 1e-4 if precision == "fp32" else 5e-3
    flip_bound = 1e-3 if precision == "fp32" else 0.01           # measured: fp32 0, bf16 0.2-0.3 % of the camera rays
    for k in ("rgb_coarse", "rgb_fine"):
        e = (out[k].detach().cpu() - case["out"][k]).abs().max(-1)[0].numpy()
        moved = e >= s_tol
        rec[f"texel_flip_fraction/{k}"] = float(moved.mean())
        rec[f"max_abs_err_same_texel/{k}"] = float(e[~moved].max())
        rec[f"median_abs_err/{k}"] = float(np.median(e))
        assert moved.mean() <= flip_bound, f"{name} {k}: {moved.mean():.3f} of the rays changed light texel"
    rec["loss_rel_err"] = abs(loss - case["loss"]) / abs(case["loss"])
    assert rec["loss_rel_err"] < (1e-4 if precision == "fp32" else 3e-2)
    # (3) parameter gradients of both models
    ref_grads = case["grads"][grad_on_light]
    # fp32: measured 2.2e-5 per tensor / 5e-6 per model.  bf16: the rays that changed light texel (above) carry a different
    # shadow gradient, and shadow_method_2's min-max normalisation couples every ray to the two extremal ones: measured up to
    # 0.20 per tensor / 6.8e-2 per model against 5e-2 / 2e-2 for plain render_rays
    tol = {"fp32": {"tensor": 1e-3, "model": 1e-4}, "bf16": {"tensor": 0.35, "model": 0.12}}[precision]
    for mi, m in enumerate(models):
        num = den = 0.0
        worst, worst_name = 0.0, ""
        for pn, pv in m.named_parameters():
            r = ref_grads[mi][pn]
            g = pv.grad.detach().cpu() if pv.grad is not None else torch.zeros_like(r)
            d, n = float((g - r).norm()), float(r.norm())
            num, den = num + d * d, den + n * n
            if n > 0 and d / n > worst:
                worst, worst_name = d / n, pn
            if n == 0:
                assert d == 0.0, f"{name} model {mi} {pn}: gradient where the reference has none"
        whole = (num / max(den, 1e-60)) ** 0.5
        rec[f"grad_rel_l2/model{mi}/whole"] = whole
        rec[f"grad_rel_l2/model{mi}/worst_tensor"] = worst
        rec[f"grad_worst_tensor_name/model{mi}"] = worst_name
        rec[f"grad_norm/model{mi}"] = den ** 0.5
        if den > 0:
            assert whole < tol["model"], f"{name} model {mi}: whole-gradient relative L2 error {whole:.2e}"
            assert worst < tol["tensor"], f"{name} model {mi} {worst_name}: relative L2 error {worst:.2e}"
    P.record(f"config5_chain/{name}/{precision}/lni={lni}/grad_on_light={int(grad_on_light)}", **rec)


ONE_VIEW = [(30.0, -30.0, 6144, 4096)]                                         # 32 contiguous rows of one 128 x 128 view
THREE_VIEWS = [(30.0, -30.0, 15000, 1384), (120.0, -20.0, 0, 1500), (200.0, -40.0, 9000, 1212)]   # 4096 rays, 3 cameras


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("grad_on_light", [False, True])
@pytest.mark.parametrize("lni", [0, 64])
def test_config5_chain_vs_oracle(dev, lni, grad_on_light, precision):
    case = _oracle_chain(f"one_view/lni{lni}", ONE_VIEW, lni)
    try:
        _check(dev, "one_view", case, precision, lni, grad_on_light)
    finally:
        import nerf_pl_b200
        nerf_pl_b200.set_mlp_precision("bf16")


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_config5_chain_three_camera_segments(dev, precision):
    case = _oracle_chain("three_views/lni0", THREE_VIEWS, 0)
    try:
        _check(dev, "three_views", case, precision, 0, True)
    finally:
        import nerf_pl_b200
        nerf_pl_b200.set_mlp_precision("bf16")
