#!/usr/bin/env python
"""Per-kernel time of one training step (4096 rays, 64+64, bf16): forward / dgrad / wgrad launches from the library's
event pairs (nrf_prof_*), the eager step time, and the inference forward alone.  One line; a tag can be given."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib
from nerf_pl_b200.dist import GradBucket
from nerf_pl_b200.mlp import forward_rays
from nerf_pl_b200.models.nerf import Embedding, NeRF
from nerf_pl_b200.models.rendering import render_rays

dev = torch.device("cuda:0")
lib = _lib.load()
torch.manual_seed(0)
models = [NeRF().to(dev), NeRF().to(dev)]
emb = [Embedding(3, 10), Embedding(3, 4)]
bucket = GradBucket(models)
B = 4096
rays = torch.cat([torch.zeros(B, 3), torch.nn.functional.normalize(torch.randn(B, 3), dim=-1), torch.full((B, 1), 2.0),
                  torch.full((B, 1), 6.0)], -1).to(dev)
tgt = torch.rand(B, 3, device=dev)


def step():
    bucket.zero()
    out = render_rays(models, emb, rays, 64, False, 1.0, 1.0, 64, 1024 * 32, True)
    (((out["rgb_coarse"] - tgt) ** 2).mean() + ((out["rgb_fine"] - tgt) ** 2).mean()).backward()


for _ in range(3):
    step()
torch.cuda.synchronize()
K = 8
lib.nrf_prof_enable(1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(K):
    step()
e1.record()
torch.cuda.synchronize()
lib.nrf_prof_enable(0)
res = {}
for kid, name in ((3, "fwd"), (5, "dgrad"), (7, "wgrad")):
    tot, n, rows = ctypes.c_double(), ctypes.c_int64(), ctypes.c_int64()
    lib.nrf_prof_read(kid, ctypes.byref(tot), ctypes.byref(n), ctypes.byref(rows))
    res[name] = tot.value / K
z = torch.sort(torch.rand(B, 128, device=dev) * 4 + 2, -1)[0]
params = [p.detach() for p in models[1].ordered_parameters()]
packed = models[1].packed_weights()
for _ in range(3):
    forward_rays(models[1].spec, params, rays, z, False, False, "bf16", packed=packed)
i0, i1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
i0.record()
for _ in range(10):
    forward_rays(models[1].spec, params, rays, z, False, False, "bf16", packed=packed)
i1.record()
torch.cuda.synchronize()
tag = sys.argv[1] if len(sys.argv) > 1 else ""
print(f"{tag:28s} step(eager) {e0.elapsed_time(e1) / K:.3f} ms | " + " ".join(f"{k}={v:.3f}" for k, v in res.items())
      + f" | sum {sum(res.values()):.3f} | infer fwd 524288 rows {i0.elapsed_time(i1) / 10:.3f} ms", flush=True)
