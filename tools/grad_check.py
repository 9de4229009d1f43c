#!/usr/bin/env python
"""bf16 tensor-core backward vs fp32 kernels: per-parameter relative L2 error (debug aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200.mlp import forward_rays, backward_rays
from nerf_pl_b200.models.nerf import NeRF

torch.manual_seed(0)
dev = torch.device("cuda:0")
m = NeRF().to(dev)
N, S = int(sys.argv[1]) if len(sys.argv) > 1 else 512, 64
rays = torch.randn(N, 8, device=dev)
z = torch.sort(torch.rand(N, S, device=dev) * 4 + 2, -1)[0]
params = [p.detach() for p in m.ordered_parameters()]
ds = torch.randn(N, S, device=dev)
dr = torch.randn(N, S, 3, device=dev)
out = {}
for prec in ("fp32", "bf16"):
    sig, rgb, saved = forward_rays(m.spec, params, rays, z, False, True, prec, packed=m.packed_weights() if prec == "bf16" else None)
    out[prec] = backward_rays(m.spec, params, rays, z, False, saved, ds, dr, rgb)
torch.cuda.synchronize()
for n, a, b in zip(m.spec.names, out["fp32"], out["bf16"]):
    rel = float((a - b).norm() / (a.norm() + 1e-12))
    flag = "  <<<<" if rel > 3e-2 else ""
    print(f"{n:32s} rel {rel:.3e}{flag}")
    if flag and a.dim() == 1:
        d = ((a - b).abs() / (a.abs().mean() + 1e-12))
        print("   worst cols:", torch.topk(d, 8).indices.tolist(), " ratio b/a at those:", (b / a)[torch.topk(d, 8).indices].tolist())
