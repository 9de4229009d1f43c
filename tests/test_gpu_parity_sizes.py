"""GPU-vs-oracle parity at the sizes and bounds that are benchmarked:

* one comparison at each BASELINE.json size, so that the steady state of the persistent kernels (tens of tiles per SM,
  next-tile prefetch, stash ring wrap, wgrad granule stealing) is compared with the reference arithmetic:
  4096 x (64+64) forward + backward (config 2), 8192 x (64+128) forward on LLFF NDC rays (config 4), one 32768-ray
  test-time chunk (config 3);
* the bf16 tensor-core path end to end at the fork's own bounds near 1 / far 200 (datasets/blender.py:40-41,
  blender_efficient_sm.py:71-76), all three render_rays variants, with and without use_disp;
* the reference's call sites (train.py:49-71, eval.py:58-86) driven through `install_as_models()`.

Tolerances: TOL below, "north_star" = BASELINE.json's statement, "stated" = what is asserted where the north-star figure
cannot hold for a reason given beside it.  Every achieved number goes to profiles/r02_parity.json (tests/_parity.py)."""
import sys
from collections import defaultdict

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
import _parity as P

pytestmark = pytest.mark.gpu

# Output tolerances: tests/_parity.py BF16_TOL (north_star 2e-3 absolute on rgb / depth / opacity; depth 3e-2 at the fork's
# 1..200 bounds, measured 1.6e-2) and the fp32-mode 2e-5 relative bound.  Parameter gradients (relative L2 against the fp32
# oracle): SURVEY 8c(iv) 2e-2 over a model's gradient; per tensor the near-cancelling first-layer gradients reach 1.5e-2 at
# config 2 and 8.4e-2 at bounds 1..200 with use_disp (measured, profiles/r02_parity.json): stated 5e-2 / 1.5e-1.
GRAD_TOL_BF16 = {"tensor": 5e-2, "model": 2e-2}      # 2e-2 over a model's gradient; 5e-2 per tensor
GRAD_TOL_FP32 = {"tensor": 1e-3, "model": 1e-4}       # measured 8e-5 / 9e-7 at 4096 rays
GRAD_TOL_BF16_FORK = {"tensor": 1.5e-1, "model": 2e-2}


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as ge
    ge.build()
    return torch.device("cuda:0")


_ORACLE_CACHE = {}


def _oracle_case(name, rays, Nc, Nf, variant, use_disp, test_time, white_back, seed, want_grad):
    """Oracle forward (chunked) + MSE-loss backward (chunked, two-phase) of one case, cached per module run."""
    if name in _ORACLE_CACHE:
        return _ORACLE_CACHE[name]
    N = rays.shape[0]
    params = P.leaf_params()
    perturb, noise_std = (0.0, 0.0) if test_time else (1.0, 1.0)
    noise = O.make_noise(N, Nc, Nf, seed=seed, perturb=perturb > 0)
    kw = dict(n_samples=Nc, use_disp=use_disp, perturb=perturb, noise_std=noise_std, n_importance=Nf, white_back=white_back,
              test_time=test_time, variant=variant)
    ref, aux = P.oracle_forward(params, rays, noise, **kw)
    tgt = torch.rand(N, 3, generator=torch.Generator().manual_seed(seed + 100))
    if want_grad:
        # losses.py:12-18: mean((rgb_coarse - t)^2) + mean((rgb_fine - t)^2)
        noise_z = dict(noise, z_fine=aux["z_fine"])
        g = {k: 2.0 * (ref[k] - tgt) / (N * 3) for k in ("rgb_coarse", "rgb_fine")}
        P.oracle_backward(params, rays, noise_z, g, **kw)
    _ORACLE_CACHE[name] = (params, noise, ref, aux, tgt, kw)
    return _ORACLE_CACHE[name]


def _gpu_render(dev, precision, variant, params, rays, noise, aux, kw):
    import nerf_pl_b200
    from nerf_pl_b200.models import rendering, rendering_rgb_sm, rendering_shadows
    from nerf_pl_b200.models.nerf import Embedding
    nerf_pl_b200.set_mlp_precision(precision)
    models = P.gpu_models(params, dev)
    inj = {k: v.to(dev) for k, v in noise.items()}
    if "z_fine" in aux:
        inj["z_fine"] = aux["z_fine"].to(dev)         # bin flips excluded (SURVEY 8c iii); their rate is tested separately
    mod = {"rgb": rendering, "depth": rendering_shadows, "rgb_sm": rendering_rgb_sm}[variant]
    got = mod.render_rays(models, [Embedding(3, 10), Embedding(3, 4)], rays.to(dev), kw["n_samples"], kw["use_disp"],
                          kw["perturb"], kw["noise_std"], kw["n_importance"], 1024 * 32, kw["white_back"], kw["test_time"],
                          _inject=inj)
    return models, got


def _mse_backward(models, got, tgt, dev):
    from nerf_pl_b200.losses import loss_dict
    loss_dict["mse"]()(got, tgt.to(dev)).backward()


# ------------------------------------------------------------------------------------------------------------------
# BASELINE sizes
# ------------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_config2_4096_rays_fwd_bwd_vs_oracle(dev, precision):
    """config 2: 4096 Blender rays, 64+64, perturb 1, noise_std 1, white_back; outputs and MSE-loss parameter gradients.
    41 tiles per SM on the fine pass: the persistent steady state."""
    rays = O.blender_rays(4096, seed=1)
    params, noise, ref, aux, tgt, kw = _oracle_case("c2", rays, 64, 64, "rgb", False, False, True, 2, True)
    try:
        models, got = _gpu_render(dev, precision, "rgb", params, rays, noise, aux, kw)
        rec = P.compare_outputs(f"config2/{precision}", ref, got, aux, rays, True, precision)
        _mse_backward(models, got, tgt, dev)
        tolg = GRAD_TOL_BF16 if precision == "bf16" else GRAD_TOL_FP32
        for mi, (worst, whole, name) in enumerate(P.grad_errors(params, models)):
            rec[f"grad_rel_l2/model{mi}/worst_tensor"] = worst
            rec[f"grad_rel_l2/model{mi}/whole"] = whole
            rec[f"grad_worst_tensor_name/model{mi}"] = name
            assert worst < tolg["tensor"], f"model {mi} {name}: {worst:.2e}"
            assert whole < tolg["model"], f"model {mi}: whole-gradient relative L2 error {whole:.2e}"
        P.record(f"config2_4096x(64+64)_fwd_bwd/{precision}", **rec)
    finally:
        import nerf_pl_b200
        nerf_pl_b200.set_mlp_precision("bf16")


def _llff_ndc_rays(n, seed):
    H, W, f = 378, 504, 407.6
    c2w = torch.tensor([[0.995, -0.03, 0.09, 0.12], [0.02, 0.99, 0.1, -0.08], [-0.1, -0.1, 0.99, 0.05]])
    rays = O.dataset_rays(c2w, H, W, f, 1.0, 1.0, ndc=True)
    pick = torch.randperm(H * W, generator=torch.Generator().manual_seed(seed))[:n]
    return rays[pick].contiguous()


def test_config4_8192_rays_fwd_vs_oracle(dev):
    """config 4: 8192 LLFF NDC rays (near 0, far 1), 64+128 samples, forward, bf16 MLP."""
    rays = _llff_ndc_rays(8192, 4)
    params, noise, ref, aux, tgt, kw = _oracle_case("c4", rays, 64, 128, "rgb", False, False, False, 5, False)
    with torch.no_grad():
        models, got = _gpu_render(dev, "bf16", "rgb", params, rays, noise, aux, kw)
    rec = P.compare_outputs("config4/bf16", ref, got, aux, rays, False, "bf16")
    P.record("config4_8192x(64+128)_fwd/bf16", **rec)


def test_config3_32768_ray_test_chunk_vs_oracle(dev):
    """config 3: one 32768-ray test_time chunk (perturb 0, noise 0, white_back, coarse pass sigma-only), bf16 MLP, through
    the stash-less inference kernels (torch.no_grad, as eval.py:57 decorates batched_inference)."""
    rays = O.blender_rays(32768, seed=7)
    params, noise, ref, aux, tgt, kw = _oracle_case("c3", rays, 64, 64, "rgb", False, True, True, 8, False)
    with torch.no_grad():
        models, got = _gpu_render(dev, "bf16", "rgb", params, rays, noise, aux, kw)
    rec = P.compare_outputs("config3/bf16", ref, got, aux, rays, True, "bf16")
    P.record("config3_32768_test_chunk/bf16", **rec)


def test_no_grad_render_writes_no_stash(dev):
    """a render under torch.no_grad() must run the inference kernels (no 5.4 KB/sample training stash): the
    peak allocation of a 32768 x (64+64) test-time call stays far below the ~30 GB the stash would take."""
    from nerf_pl_b200.models.nerf import Embedding, NeRF
    from nerf_pl_b200.models.rendering import render_rays
    models = [NeRF().to(dev), NeRF().to(dev)]
    rays = O.blender_rays(32768, seed=3).to(dev)
    emb = [Embedding(3, 10), Embedding(3, 4)]
    with torch.no_grad():
        render_rays(models, emb, rays[:256], 64, False, 0, 0, 64, 1024 * 32, True, True)
        torch.cuda.synchronize()
        torch.cuda.reset_peak_memory_stats()
        base = torch.cuda.memory_allocated()
        render_rays(models, emb, rays, 64, False, 0, 0, 64, 1024 * 32, True, True)
        torch.cuda.synchronize()
    peak = torch.cuda.max_memory_allocated() - base
    P.record("no_grad_render_peak_bytes", peak_bytes=peak, stash_would_be=32768 * 192 * 5412.0)
    assert peak < 2e9, f"test-time render allocated {peak / 1e9:.1f} GB: the training stash is being written"


# ------------------------------------------------------------------------------------------------------------------
# bf16 at the fork's bounds
# ------------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("variant,use_disp,test_time", [("rgb", False, False), ("rgb", True, False), ("rgb", False, True),
                                                        ("depth", False, False), ("depth", True, False),
                                                        ("rgb_sm", False, False)])
def test_bf16_fork_bounds_1_200_vs_oracle(dev, variant, use_disp, test_time):
    """near 1 / far 200 (datasets/blender.py:40-41, blender_efficient_sm.py:71-76), 512 rays, 64+64, bf16 MLP, all
    variants; parameter gradients of a probe loss over every output for the training cases."""
    rays = O.blender_rays(512, seed=11, near=1.0, far=200.0)
    name = f"fork/{variant}/{int(use_disp)}/{int(test_time)}"
    params, noise, ref, aux, tgt, kw = _oracle_case(name, rays, 64, 64, variant, use_disp, test_time, True, 12, False)
    grad_mode = torch.enable_grad() if not test_time else torch.no_grad()
    with grad_mode:
        models, got = _gpu_render(dev, "bf16", variant, params, rays, noise, aux, kw)
    rec = P.compare_outputs(name, ref, got, aux, rays, True, "bf16", far200="test_time" if test_time else True)
    if not test_time:
        # gradient of a smooth probe over the stable rays' bounded outputs (rgb, opacity; depth / far)
        stable = torch.from_numpy(P.last_sample_stable(aux))
        g = torch.Generator().manual_seed(5)
        out_grads = {}
        for k in sorted(ref):
            if k.startswith("disp_map"):
                continue
            pr = torch.rand(ref[k].shape, generator=g) * (1.0 / 200.0 if k.startswith("depth") else 1.0)
            pr = pr * (stable if pr.dim() == 1 else stable[:, None])
            out_grads[k] = pr
        for p in params:
            for t in p.values():
                t.grad = None
        P.oracle_backward(params, rays, dict(noise, z_fine=aux["z_fine"]), out_grads, **kw)
        sum((got[k] * out_grads[k].to(dev)).sum() for k in out_grads).backward()
        for mi, (worst, whole, nm) in enumerate(P.grad_errors(params, models)):
            rec[f"grad_rel_l2/model{mi}/worst_tensor"] = worst
            rec[f"grad_rel_l2/model{mi}/whole"] = whole
            rec[f"grad_worst_tensor_name/model{mi}"] = nm
            assert whole < GRAD_TOL_BF16_FORK["model"], f"{name} model {mi}: whole-gradient relative L2 error {whole:.2e}"
            assert worst < GRAD_TOL_BF16_FORK["tensor"], f"{name} model {mi} {nm}: {worst:.2e}"
    P.record(f"bf16_bounds_1_200/{variant}/use_disp={int(use_disp)}/test_time={int(test_time)}", **rec)


# ------------------------------------------------------------------------------------------------------------------
# the reference's call sites through install_as_models()
# ------------------------------------------------------------------------------------------------------------------

def test_reference_call_sites_through_models_alias(dev):
    """`from models.rendering import render_rays` etc. (train.py:11-13, eval.py:9-10) resolve to this implementation after
    install_as_models(); the restated train.py:49-71 forward loop and eval.py:58-86 batched_inference loop (chunked calls,
    per-key torch.cat, .cpu()) run on it and agree with one un-chunked call / the oracle."""
    import nerf_pl_b200
    saved = {k: v for k, v in sys.modules.items() if k in ("models", "torchsearchsorted", "losses", "metrics") or k.startswith("models.")}
    try:
        for k in saved:
            del sys.modules[k]
        nerf_pl_b200.install_as_models()
        from models.nerf import Embedding, NeRF                 # noqa: the reference's import lines
        from models.rendering import render_rays
        from losses import loss_dict
        from metrics import psnr

        embeddings = [Embedding(3, 10), Embedding(3, 4)]
        params = [O.init_mlp_params(generator=torch.Generator().manual_seed(s)) for s in (0, 1)]
        models = []
        for p in params:
            m = NeRF().to(dev)
            m.load_state_dict(p)
            models.append(m)

        # eval.py:57-86
        @torch.no_grad()
        def batched_inference(models, embeddings, rays, N_samples, N_importance, use_disp, chunk, white_back):
            B = rays.shape[0]
            results = defaultdict(list)
            for i in range(0, B, chunk):
                rendered = render_rays(models, embeddings, rays[i:i + chunk], N_samples, use_disp, 0, 0, N_importance, chunk,
                                       white_back, test_time=True)
                for k, v in rendered.items():
                    results[k] += [v]
            return {k: torch.cat(v, 0) for k, v in results.items()}

        rays = O.blender_rays(40000, seed=21)                  # two chunks: 32768 + 7232 (ragged tail)
        res = batched_inference(models, embeddings, rays.to(dev), 64, 64, False, 1024 * 32, True)
        assert sorted(res) == ["depth_fine", "opacity_coarse", "opacity_fine", "rgb_fine"]
        img = res["rgb_fine"].view(200, 200, 3).cpu().numpy()        # eval.py:112-116
        assert img.shape == (200, 200, 3) and np.isfinite(img).all()
        # oracle on a 1024-ray slice that straddles the chunk boundary
        lo = 32768 - 512
        sl = rays[lo:lo + 1024]
        # (the sampler draws its own uniforms here, so the oracle comparison injects nothing and compares the COARSE
        #  opacity, which is deterministic at perturb = 0)
        with torch.no_grad():
            ref_c = O.render_rays(params, sl, n_samples=64, perturb=0.0, noise_std=0.0, n_importance=0, white_back=True,
                                  test_time=True, noise={})
        err = (res["opacity_coarse"][lo:lo + 1024].cpu() - ref_c["opacity_coarse"]).abs()
        stable = err < 0.5                                      # a ray on the last-sample discontinuity flips 0 <-> 1
        assert stable.float().mean() > 0.98 and err[stable].max() < 2e-3
        P.record("call_sites/eval_batched_inference", opacity_coarse_max_abs_err=float(err[stable].max()),
                 stable_fraction=float(stable.float().mean()))

        # train.py:49-71 with hparams.chunk smaller than the batch, then loss / psnr / backward as training_step does
        class H:
            N_samples, use_disp, perturb, noise_std, N_importance, chunk = 64, False, 1.0, 1.0, 64, 1500

        def forward(rays):
            B = rays.shape[0]
            results = defaultdict(list)
            for i in range(0, B, H.chunk):
                rendered = render_rays(models, embeddings, rays[i:i + H.chunk], H.N_samples, H.use_disp, H.perturb,
                                       H.noise_std, H.N_importance, H.chunk, True)
                for k, v in rendered.items():
                    results[k] += [v]
            return {k: torch.cat(v, 0) for k, v in results.items()}

        batch_rays = rays[:4096].to(dev)
        rgbs = torch.rand(4096, 3, device=dev)
        results = forward(batch_rays)
        assert sorted(results) == ["depth_coarse", "depth_fine", "opacity_coarse", "opacity_fine", "rgb_coarse", "rgb_fine"]
        loss = loss_dict["mse"]()(results, rgbs)
        loss.backward()
        with torch.no_grad():
            ps = psnr(results["rgb_fine"], rgbs)
        assert torch.isfinite(loss) and torch.isfinite(ps)
        assert all(p.grad is not None and torch.isfinite(p.grad).all() and p.grad.abs().sum() > 0
                   for m in models for p in m.parameters())
    finally:
        for k in list(sys.modules):
            if k in ("models", "torchsearchsorted", "losses", "metrics") or k.startswith("models."):
                del sys.modules[k]
        sys.modules.update(saved)
