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


def discontinuity_candidates(aux, pas, rays, white_back, rows):
    """The two outcomes the reference's last-sample discontinuity allows for the rays `rows` of pass `pas` ('coarse' /
    'fine'): the oracle's own compositing (O.composite) of the oracle's per-sample sigma / rgb with the LAST raw sigma
    forced to -1 (last alpha = 0) and to +1 (last alpha = 1).  Returns two dicts {kind: tensor}."""
    z = aux["z_coarse"] if pas == "coarse" else aux["z_all"]
    raw = aux[f"raw_{pas}"][rows]
    rgbs = aux[f"rgbs_{pas}"][rows] if f"rgbs_{pas}" in aux else None
    outs = []
    for forced in (-1.0, 1.0):
        r2 = raw.clone()
        r2[:, -1] = forced
        c = O.composite(r2, rgbs, z[rows], rays[rows, 3:6], None, 0.0, white_back)
        outs.append({"rgb": c.get("rgb"), "depth": c["depth"], "opacity": c["opacity"], "disp_map": c["disp"]})
    return outs


# bf16 tensor-core MLP, absolute errors.  rgb / opacity / depth: BASELINE.json north_star "within 2e-3 absolute".
# depth = sum w_i z_i scales with `far`: at the fork's bounds 1..200 the measured maximum is 1.6e-2 (profiles/
# r02_parity.json), stated bound 3e-2 = 1.5e-4 of the depth range.
# Without the training noise (test_time, noise_std 0) a freshly initialised network's sigma stays ~1e-2 along the whole ray,
# the transmittance stays ~1 out to z = 200 and all 128 samples carry weight there: depth error = sum z_i delta_i dsigma_i
# with z 200, delta 1.5, dsigma 1e-4 (bf16) reaches 5.5e-2 (measured) - stated bound 1e-1 = 5e-4 of the depth range.
# Same case, opacity / rgb: d(alpha_i) = delta_i * d(sigma_i) with delta = 3.16 on the 64-sample coarse grid of these bounds
# and no saturation anywhere (raw sigma within +-0.09): measured 9.7e-3 on opacity_coarse (1.5e-3 on the fine pass, fp32
# mode 1.7e-6 on the same rays) - stated 2e-2.
BF16_TOL = {"rgb": 2e-3, "opacity": 2e-3, "depth": 2e-3, "depth@200": 3e-2, "depth@200/test_time": 1e-1,
            "opacity@200/test_time": 2e-2, "rgb@200/test_time": 2e-2, "disp_map": None}     # 1 / max(1e-10, depth / opacity): unbounded where opacity -> 0; reported only


def compare_outputs(tag, ref, got, aux, rays, white_back, precision, far200=False, margin=2e-3):
    """Every ray is checked (far200: False, True = bounds 1..200 with training noise, "test_time" = 1..200 without).
    Rays whose last raw sigma is farther than `margin` from zero in every pass: |got - ref| <
    tol per key.  The others sit on the reference's last-sample discontinuity (final interval 1e10 long, rendering.py:171:
    the last alpha is 0 or 1 with the SIGN of a number the bf16 MLP reproduces to ~1e-4): each must match, on all keys of
    the pass at once, one of the two outcomes the discontinuity allows.  Returns the record of achieved numbers."""
    stable_all = torch.from_numpy(last_sample_stable(aux, margin))
    rec = {"stable_fraction": float(stable_all.float().mean()), "n_rays": float(stable_all.numel())}
    assert sorted(got) == sorted(ref), (sorted(got), sorted(ref))
    stable_of = {pas: aux[f"raw_{pas}"][:, -1].abs() > margin for pas in ("coarse", "fine") if f"raw_{pas}" in aux}

    def tol_of(kind, rng):
        if precision == "fp32":
            return 2e-5 * max(1.0, rng)              # fp32 mode: 1e-5-class, relative to the key's range
        if far200 == "test_time" and f"{kind}@200/test_time" in BF16_TOL:
            return BF16_TOL[f"{kind}@200/test_time"]
        if kind == "depth" and far200:
            return BF16_TOL["depth@200"]
        return BF16_TOL[kind]

    fails = []
    for k in sorted(ref):
        r, g = ref[k].detach(), got[k].detach().cpu()
        kind, pas = k.rsplit("_", 1)
        stable = stable_of[pas]                      # a key depends on its own pass's last sample only (z_fine is injected)
        ok = stable if r.dim() == 1 else stable[:, None].expand_as(r)
        ok = ok & ~torch.isnan(r)
        err = float((g - r).abs()[ok].max()) if ok.any() else 0.0
        rng = float(r.abs()[ok].max()) if ok.any() else 0.0
        tol = tol_of(kind, rng)
        rec[f"max_abs_err/{k}"], rec[f"range/{k}"], rec[f"tol/{k}"] = err, rng, (-1.0 if tol is None else tol)
        if tol is not None and not err < tol:
            fails.append(f"{k}: max abs err {err:.3e} >= {tol:.1e}")
    assert not fails, f"{tag}: " + "; ".join(fails)       # after the loop, so that one run shows every key's number
    # rays on the discontinuity: per pass, one of the two legal outcomes must explain ALL of the pass's keys
    for pas in ("coarse", "fine"):
        if f"raw_{pas}" not in aux:
            continue
        band = (aux[f"raw_{pas}"][:, -1].abs() <= margin).nonzero().flatten()
        rec[f"discontinuity_rays/{pas}"] = float(band.numel())
        if band.numel() == 0:
            continue
        cands = discontinuity_candidates(aux, pas, rays, white_back, band)
        worst = torch.zeros(band.numel())
        best = torch.full((band.numel(),), float("inf"))
        for cand in cands:
            score = torch.zeros(band.numel())
            for k in ref:
                kind, kp = k.rsplit("_", 1)
                if kp != pas or cand.get(kind) is None:
                    continue
                tol = tol_of(kind, float(ref[k].abs().max()))
                if tol is None:
                    continue
                e = (got[k].detach().cpu()[band] - cand[kind]).abs()
                e = e if e.dim() == 1 else e.max(-1)[0]
                score = torch.maximum(score, e / tol)
            best = torch.minimum(best, score)
        rec[f"discontinuity_rays_worst_err_over_tol/{pas}"] = float(best.max())
        assert float(best.max()) < 1.0, f"{tag} {pas}: a ray on the last-sample discontinuity matches neither legal outcome"
    return rec
