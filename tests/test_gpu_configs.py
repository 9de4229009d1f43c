"""End-to-end parity at the shapes of BASELINE.json configs 4 and 5 that test_gpu_render.py does not reach:
LLFF-style NDC rays (504x378, near 0 / far 1, un-normalised directions) and use_disp with spheric bounds, both with
N_importance = 128 (192 fine evaluations per ray), in fp32 mode (outputs 2e-5; gradients 1e-2 per tensor / 2e-4 per model, see the note in the test) and with the bf16
tensor-core MLP (outputs 2e-3; gradients 5e-2 per tensor / 2e-2 per model), and the
rgb + disparity variant (rendering_rgb_sm) at full width.  Fine depths are injected (bin flips excluded, SURVEY 8c iii)."""
import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as ge
    ge.build()
    return torch.device("cuda:0")


def _llff_rays(n, ndc, near, far, seed):
    H, W, f = 378, 504, 407.6
    c2w = torch.tensor([[0.995, -0.03, 0.09, 0.12], [0.02, 0.99, 0.1, -0.08], [-0.1, -0.1, 0.99, 0.05]])
    rays = O.dataset_rays(c2w, H, W, f, 1.0 if ndc else near, far, ndc=ndc)
    pick = torch.randperm(H * W, generator=torch.Generator().manual_seed(seed))[:n]
    return rays[pick].contiguous()


def _run(dev, precision, variant, rays, Nc, Nf, use_disp, seed):
    import nerf_pl_b200
    from nerf_pl_b200.models import rendering, rendering_rgb_sm
    from nerf_pl_b200.models.nerf import Embedding, NeRF
    nerf_pl_b200.set_mlp_precision(precision)
    params = [O.init_mlp_params(generator=torch.Generator().manual_seed(s)) for s in (0, 1)]
    for p in params:
        for k in p:
            p[k].requires_grad_()
    N = rays.shape[0]
    noise = O.make_noise(N, Nc, Nf, seed=seed)
    ref, aux = O.render_rays(params, rays, n_samples=Nc, use_disp=use_disp, perturb=1.0, noise_std=1.0, n_importance=Nf,
                             white_back=False, variant=variant, noise=noise, return_aux=True)
    models = []
    for p in params:
        m = NeRF().to(dev)
        m.load_state_dict({k: v.detach() for k, v in p.items()})
        models.append(m)
    inj = {k: v.to(dev) for k, v in noise.items()}
    inj["z_fine"] = aux["z_fine"].to(dev)
    mod = {"rgb": rendering, "rgb_sm": rendering_rgb_sm}[variant]
    got = mod.render_rays(models, [Embedding(3, 10), Embedding(3, 4)], rays.to(dev), Nc, use_disp, 1.0, 1.0, Nf,
                          1024 * 32, False, False, _inject=inj)
    return params, models, ref, got, aux


CASES = [("ndc", True, 0.0, 1.0, False), ("disp_spheric", False, 1.33, 10.6, True)]


@pytest.mark.parametrize("name,ndc,near,far,use_disp", CASES)
def test_config4_fp32_vs_oracle(dev, name, ndc, near, far, use_disp):
    rays = _llff_rays(96, ndc, near, far, 7)
    params, models, ref, got, _ = _run(dev, "fp32", "rgb", rays, 64, 128, use_disp, 8)
    assert sorted(got) == sorted(ref)
    loss_r = loss_g = 0
    g = torch.Generator().manual_seed(5)
    for k in sorted(ref):
        r = ref[k].detach().numpy()
        np.testing.assert_allclose(got[k].detach().cpu().numpy(), r, rtol=2e-5, atol=1e-5 * max(1.0, np.abs(r).max()), err_msg=k)
        pr = torch.rand(ref[k].shape, generator=g)
        loss_r = loss_r + (ref[k] * pr).sum()
        loss_g = loss_g + (got[k] * pr.to(dev)).sum()
    loss_r.backward()
    loss_g.backward()
    # Per tensor 1e-2, over a model's whole gradient 2e-4.  The early layers' gradients nearly cancel here (|dW1| = 9e-3
    # against |dW6| = 1.3): the reference algorithm evaluated in fp32 differs from its own fp64 evaluation by 6.4e-3 on
    # xyz_encoding_1.0.weight for these very rays (measured with the oracle on the CPU), so fp32-vs-fp32 agreement of
    # 3.6e-3 on that tensor is the conditioning of the problem, not of the kernels.
    for mi, m in enumerate(models):
        num = den = 0.0
        for pn, pv in m.named_parameters():
            r = params[mi][pn].grad.numpy()
            d = np.linalg.norm(pv.grad.cpu().numpy() - r)
            num, den = num + d * d, den + np.linalg.norm(r) ** 2
            assert d / (np.linalg.norm(r) + 1e-12) < 1e-2, f"{name} grad {mi}/{pn}: relative L2 error {d / np.linalg.norm(r):.2e}"
        assert (num / den) ** 0.5 < 2e-4, f"{name} model {mi}: whole-gradient relative L2 error {(num / den) ** 0.5:.2e}"


@pytest.mark.parametrize("name,ndc,near,far,use_disp", CASES)
def test_config4_bf16_vs_oracle(dev, name, ndc, near, far, use_disp):
    rays = _llff_rays(256, ndc, near, far, 9)
    params, models, ref, got, aux = _run(dev, "bf16", "rgb", rays, 64, 128, use_disp, 10)
    # every ray is checked: 2e-3 absolute away from the last-sample discontinuity, one of its two legal outcomes on it
    import _parity
    rec = _parity.compare_outputs(f"config4_shape/{name}", ref, got, aux, rays, False, "bf16")
    _parity.record(f"config4_shape_256_rays/{name}/bf16", **rec)
    stable = _parity.last_sample_stable(aux)
    loss_r = sum((ref[k][torch.from_numpy(stable)] ** 2).sum() for k in ("rgb_coarse", "rgb_fine"))
    loss_g = sum((got[k][torch.from_numpy(stable).to(dev)] ** 2).sum() for k in ("rgb_coarse", "rgb_fine"))
    loss_r.backward()
    loss_g.backward()
    # bf16 mode against the fp32 oracle: 5e-2 per tensor (as test_full_width_bf16_vs_oracle; the near-cancelling
    # first-layer gradients sit at ~2e-2), 2e-2 over a model's whole gradient (SURVEY 8c iv)
    for mi, m in enumerate(models):
        num = den = 0.0
        for pn, pv in m.named_parameters():
            r = params[mi][pn].grad.numpy()
            d = np.linalg.norm(pv.grad.cpu().numpy() - r)
            num, den = num + d * d, den + np.linalg.norm(r) ** 2
            assert d / (np.linalg.norm(r) + 1e-12) < 5e-2, f"{name} grad {mi}/{pn}: relative L2 error {d / np.linalg.norm(r):.2e}"
        assert (num / den) ** 0.5 < 2e-2, f"{name} model {mi}: whole-gradient relative L2 error {(num / den) ** 0.5:.2e}"


def test_rgb_sm_full_width_fp32_vs_oracle(dev):
    """rendering_rgb_sm.render_rays (rgb + disp_map_*, reference :87-278) at W = 256 with the fork's bounds 1..200."""
    rays = O.blender_rays(96, seed=3, near=1.0, far=200.0)
    _, _, ref, got, _ = _run(dev, "fp32", "rgb_sm", rays, 64, 64, False, 4)
    assert sorted(got) == sorted(ref) and "disp_map_fine" in got
    for k in sorted(ref):
        r = ref[k].detach().numpy()
        np.testing.assert_allclose(got[k].detach().cpu().numpy(), r, rtol=5e-5, atol=1e-5 * max(1.0, np.abs(r).max()), err_msg=k)
