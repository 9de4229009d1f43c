"""Shared helpers of the GPU-vs-oracle parity tests (test infrastructure).

* `record(test, **numbers)`: every achieved error / flip rate / excluded-ray fraction the tests measure is written to
  `profiles/r02_parity.json` (and to `gpurun_out/`, the directory that travels back from the GPU box), so the numbers
  behind DESIGN.md's tolerance table are the ones the tests produced.
* `oracle_forward` / `oracle_backward`: the CPU oracle evaluated in ray chunks, so that a 4096..32768-ray comparison
  needs a few GB of host memory instead of tens.  render_rays is independent per ray; a loss that couples rays (a mean,
  the min-max normalisation of shadow_method_2) is handled in two phases: forward all chunks without autograd, take the
  loss's gradient w.r.t. the per-ray outputs, then re-run each chunk with autograd and back-propagate that gradient
  (the chain rule; the forward is deterministic because every random tensor is injected).
"""
import json
import os

import numpy as np
import torch

from oracle import nerf_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_PATHS = [os.path.join(ROOT, "profiles", "r02_parity.json"), os.path.join(ROOT, "gpurun_out", "r02_parity.json")]


def record(test, **vals):
    clean = {}
    for k, v in vals.items():
        if isinstance(v, dict):
            clean[k] = {kk: float(vv) for kk, vv in v.items()}
        elif isinstance(v, (str, bool)):
            clean[k] = v
        else:
            clean[k] = float(v)
    for path in _PATHS:
        if not os.path.isdir(os.path.dirname(path)):
            continue
        try:
            data = json.load(open(path)) if os.path.exists(path) else {}
        except Exception:
            data = {}
        data.setdefault(test, {}).update(clean)
        with open(path, "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)


def _slice_noise(noise, lo, hi):
    return {k: v[lo:hi] for k, v in noise.items()}


def oracle_forward(params, rays, noise, chunk=1024, **kw):
    """(res, aux) of O.render_rays over all rays, evaluated chunk by chunk without autograd."""
    res, aux = {}, {}
    with torch.no_grad():
        for lo in range(0, rays.shape[0], chunk):
            r, a = O.render_rays(params, rays[lo:lo + chunk], noise=_slice_noise(noise, lo, lo + chunk), return_aux=True, **kw)
            for k, v in r.items():
                res.setdefault(k, []).append(v)
            for k, v in a.items():
                aux.setdefault(k, []).append(v)
    return {k: torch.cat(v, 0) for k, v in res.items()}, {k: torch.cat(v, 0) for k, v in aux.items()}


def oracle_backward(params, rays, noise, out_grads, chunk=1024, **kw):
    """Accumulates d(sum_k <out_k, out_grads[k]>)/d(params) into the parameters' .grad, chunk by chunk."""
    for lo in range(0, rays.shape[0], chunk):
        r = O.render_rays(params, rays[lo:lo + chunk], noise=_slice_noise(noise, lo, lo + chunk), **kw)
        loss = 0
        for k, g in out_grads.items():
            if g is not None:
                loss = loss + (r[k] * g[lo:lo + chunk]).sum()
        loss.backward()


def leaf_params(seeds=(0, 1), **kw):
    params = [O.init_mlp_params(generator=torch.Generator().manual_seed(s), **kw) for s in seeds]
    for p in params:
        for k in p:
            p[k].requires_grad_()
    return params


def gpu_models(params, dev):
    from nerf_pl_b200.models.nerf import NeRF
    models = []
    for p in params:
        m = NeRF().to(dev)
        m.load_state_dict({k: v.detach() for k, v in p.items()})
        models.append(m)
    return models


def grad_errors(params, models):
    """Per model: (worst per-tensor relative L2 error, whole-gradient relative L2 error, name of the worst tensor)."""
    out = []
    for p, m in zip(params, models):
        num = den = 0.0
        worst, worst_name = 0.0, ""
        for name, t in m.named_parameters():
            ref = p[name].grad
            if ref is None:
                ref = torch.zeros_like(p[name])
            got = t.grad.detach().cpu() if t.grad is not None else torch.zeros_like(ref)
            d, n = float((got - ref).norm()), float(ref.norm())
            num, den = num + d * d, den + n * n
            if n > 0 and d / n > worst:
                worst, worst_name = d / n, name
            elif n == 0 and d > 0:
                worst, worst_name = float("inf"), name
        out.append((worst, (num / max(den, 1e-60)) ** 0.5, worst_name))
    return out


def last_sample_stable(aux, margin=2e-3):
    """Rays away from the reference's last-sample discontinuity: the final interval is 1e10 long (rendering.py:171), so a
    ray's last alpha is 0 or 1 with the SIGN of the last raw sigma (+ noise).  |raw_last| > margin in every pass."""
    ok = aux["raw_coarse"][:, -1].abs() > margin
    if "raw_fine" in aux:
        ok = ok & (aux["raw_fine"][:, -1].abs() > margin)
    return ok.numpy()


def opacity_candidates(weights):
    """Opacity of a ray with its last alpha forced to 0 and to 1 (for the rays `last_sample_stable` excludes):
    A = sum of the other weights, B = A + transmittance reaching the last sample."""
    a = weights[:, :-1].sum(-1)
    return a, torch.ones_like(a)     # alpha_last = 1 absorbs everything that is left: opacity = 1 (up to 1e-10 terms)
