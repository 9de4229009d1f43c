"""Diagnostic: bounds 1..200, no noise, random init - opacity_coarse error of the sigma-only inference kernel (test_time)
against the full kernel (test_time False) and the fp32 mode, each against the CPU oracle."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from oracle import nerf_oracle as O
import _parity as P
import nerf_pl_b200
from nerf_pl_b200.models import rendering
from nerf_pl_b200.models.nerf import Embedding

dev = torch.device("cuda:0")
rays = O.blender_rays(512, seed=11, near=1.0, far=200.0)
params = P.leaf_params()
noise = O.make_noise(512, 64, 64, seed=12, perturb=False)
for tt in (True, False):
    kw = dict(n_samples=64, use_disp=False, perturb=0.0, noise_std=0.0, n_importance=64, white_back=True, test_time=tt, variant="rgb")
    ref, aux = P.oracle_forward(params, rays, noise, **kw)
    raw = aux["raw_coarse"]
    print(f"test_time={tt}: oracle raw_coarse: min {float(raw.min()):.3e} max {float(raw.max()):.3e} mean|.| {float(raw.abs().mean()):.3e}; "
          f"opacity_coarse max {float(ref['opacity_coarse'].max()):.3e}")
    for prec in ("bf16", "fp32"):
        nerf_pl_b200.set_mlp_precision(prec)
        models = P.gpu_models(params, dev)
        inj = {k: v.to(dev) for k, v in noise.items()}
        inj["z_fine"] = aux["z_fine"].to(dev)
        for grad in (False, True):
            with torch.set_grad_enabled(grad):
                got = rendering.render_rays(models, [Embedding(3, 10), Embedding(3, 4)], rays.to(dev), 64, False, 0.0, 0.0, 64,
                                            32768, True, tt, _inject=inj)
            errs = {k: float((got[k].detach().cpu() - ref[k]).abs().max()) for k in sorted(ref)}
            print(f"   {prec} grad={grad}: " + "  ".join(f"{k} {v:.2e}" for k, v in errs.items()))
nerf_pl_b200.set_mlp_precision("bf16")
