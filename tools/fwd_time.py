#!/usr/bin/env python
"""Forward kernel alone (4096 rays x 128 samples, full model): inference and training launches, CUDA-event timed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200.mlp import forward_rays, backward_rays
from nerf_pl_b200.models.nerf import NeRF

dev = torch.device("cuda:0")
torch.manual_seed(0)
m = NeRF().to(dev)
B, S = 4096, 128
rays = torch.randn(B, 8, device=dev)
z = torch.sort(torch.rand(B, S, device=dev) * 4 + 2, -1)[0]
params = [p.detach() for p in m.ordered_parameters()]
packed = m.packed_weights()
tag = sys.argv[1] if len(sys.argv) > 1 else ""
out = []
for train in (False, True):
    for _ in range(3):
        forward_rays(m.spec, params, rays, z, False, train, "bf16", packed=packed)
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        forward_rays(m.spec, params, rays, z, False, train, "bf16", packed=packed)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    out.append(f"{'train' if train else 'infer'} {ts[len(ts) // 2]:.3f} ms")
print(f"{tag:24s} fwd 524288 rows: " + ", ".join(out), flush=True)
