#!/usr/bin/env python
"""Tensor-memory / N-split forward kernel (mlp_tc_ts.cu) against the shared-memory one (mlp_tc.cu) on the same inputs
(debug build): outputs, stash images and sign masks must be IDENTICAL.  usage: ts_check.py [rays] [samples]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib
from nerf_pl_b200.mlp import forward_rays
from nerf_pl_b200.models.nerf import NeRF

IMPL = os.environ.get("TS_IMPL", "ts")   # "ts": debug/mlp_tc_ts.cu, "m": mlp_tc_tmem.cu
B = int(sys.argv[1]) if len(sys.argv) > 1 else 300
S = int(sys.argv[2]) if len(sys.argv) > 2 else 64
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = NeRF().to(dev)
rays = torch.cat([torch.randn(B, 3) * 0.1, torch.nn.functional.normalize(torch.randn(B, 3), dim=-1),
                  torch.full((B, 1), 2.0), torch.full((B, 1), 6.0)], -1).to(dev)
z = torch.sort(torch.rand(B, S, device=dev) * 4 + 2, -1)[0]
params = [p.detach() for p in m.ordered_parameters()]
packed = m.packed_weights()
ok = True
for sigma_only in (False, True):
    for train in (False, True):
        res = {}
        for impl in ("1cta", IMPL):
            os.environ["NRF_FWD_IMPL"] = impl
            sigma, rgb, saved = forward_rays(m.spec, params, rays, z, sigma_only, train, "bf16", packed=packed)
            torch.cuda.synchronize()
            res[impl] = (sigma.clone(), None if rgb is None else rgb.clone(), None if not train else saved[2].clone())
        a, b = res["1cta"], res[IMPL]
        d_sig = (a[0] - b[0]).abs().max().item()
        d_rgb = 0.0 if a[1] is None else (a[1] - b[1]).abs().max().item()
        d_st = -1 if a[2] is None else int((a[2] != b[2]).sum().item())
        print(f"sigma_only={sigma_only} train={train}: max|dsigma|={d_sig:.3e} max|drgb|={d_rgb:.3e} stash bytes differing={d_st}",
              flush=True)
        ok &= d_sig == 0 and d_rgb == 0 and d_st <= 0
# speed
os.environ["NRF_FWD_IMPL"] = "1cta"
Bb, Sb = 4096, 128
raysb = torch.cat([torch.zeros(Bb, 3), torch.nn.functional.normalize(torch.randn(Bb, 3), dim=-1), torch.full((Bb, 1), 2.0),
                   torch.full((Bb, 1), 6.0)], -1).to(dev)
zb = torch.sort(torch.rand(Bb, Sb, device=dev) * 4 + 2, -1)[0]
for impl in ("1cta", IMPL):
    os.environ["NRF_FWD_IMPL"] = impl
    for train in (False, True):
        for _ in range(3):
            forward_rays(m.spec, params, raysb, zb, False, train, "bf16", packed=packed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            forward_rays(m.spec, params, raysb, zb, False, train, "bf16", packed=packed)
        e1.record()
        torch.cuda.synchronize()
        print(f"{impl} train={train}: {e0.elapsed_time(e1) / 10:.3f} ms for {Bb}x{Sb} rows", flush=True)
print("IDENTICAL" if ok else "MISMATCH")
