#!/usr/bin/env python
"""Short program for the ncu captures (profiles/r02_ncu_*): two warm-up training steps, ONE training step (4096 rays, 64+64,
bf16: 2 forward + 2 dgrad + 2 wgrad launches of the tensor-core MLP kernels), then three stash-less inference forwards of
4096 x 128 rows.  With `-k regex:mlp_ -s 12 -c 9` ncu sees the step's six launches and the three inference launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200.dist import GradBucket
from nerf_pl_b200.mlp import forward_rays
from nerf_pl_b200.models.nerf import Embedding, NeRF
from nerf_pl_b200.models.rendering import render_rays
from nerf_pl_b200.synthetic import blender_rays

dev = torch.device("cuda:0")
torch.manual_seed(0)
models = [NeRF().to(dev), NeRF().to(dev)]
emb = [Embedding(3, 10), Embedding(3, 4)]
bucket = GradBucket(models)
B = 4096
rays = blender_rays(B, seed=1).to(dev)
tgt = torch.rand(B, 3, device=dev)
for _ in range(3):
    bucket.zero()
    out = render_rays(models, emb, rays, 64, False, 1.0, 1.0, 64, 1024 * 32, True)
    (((out["rgb_coarse"] - tgt) ** 2).mean() + ((out["rgb_fine"] - tgt) ** 2).mean()).backward()
torch.cuda.synchronize()
z = torch.sort(torch.rand(B, 128, device=dev) * 4 + 2, -1)[0]
params = [p.detach() for p in models[1].ordered_parameters()]
packed = models[1].packed_weights()
for _ in range(3):
    forward_rays(models[1].spec, params, rays, z, False, False, "bf16", packed=packed)
torch.cuda.synchronize()
print("ok")
