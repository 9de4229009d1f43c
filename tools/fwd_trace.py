#!/usr/bin/env python
"""Event timeline of CTA 0 of the tensor-core forward kernel (inference, full)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib
from nerf_pl_b200.mlp import forward_rays
from nerf_pl_b200.models.nerf import NeRF

dev = torch.device("cuda:0")
m = NeRF().to(dev)
B, S = 4096, 128
rays = torch.randn(B, 8, device=dev)
z = torch.sort(torch.rand(B, S, device=dev) * 4 + 2, -1)[0]
params = [p.detach() for p in m.ordered_parameters()]
packed = m.packed_weights()
lib = _lib.load()
train = len(sys.argv) > 1 and sys.argv[1] == "train"
for _ in range(3):
    forward_rays(m.spec, params, rays, z, False, train, "bf16", packed=packed)
CAP = 4096
cnt = torch.zeros(148 * 8 + 3 * CAP, dtype=torch.int64, device=dev)
lib.nrf_dbg_set_counters(cnt.data_ptr())
forward_rays(m.spec, params, rays, z, False, train, "bf16", packed=packed)
torch.cuda.synchronize()
lib.nrf_dbg_set_counters(None)
c = cnt.cpu().numpy()
tr = [c[148 * 8 + r * CAP: 148 * 8 + (r + 1) * CAP] for r in range(3)]
t0 = min(int(t[t > 0].min()) for t in tr)
# tile index 2 (steady state): producer events 39/tile; mma events 3 per chunk = 117/tile; epilogue: 10 acc waits + 36 signals = 46/tile
tile = 2
prod = tr[0][tile * 39:(tile + 1) * 39] - t0
mma = (tr[1][tile * 117:(tile + 1) * 117] - t0).reshape(39, 3)
epi = tr[2][tile * 46:(tile + 1) * 46] - t0
base = int(mma[0, 0])
print("chunk: prod_issue | mma: before_wait_full  after_wait_full  after_issue   (cycles rel. to tile start)")
for i in range(39):
    print(f"{i:2d}: {int(prod[i]) - base:7d} | {int(mma[i,0]) - base:7d} {int(mma[i,1]) - base:7d} {int(mma[i,2]) - base:7d}")
print("epilogue events (acc-wait done, then 4 chunk signals per layer; L9 has none):")
print([int(e) - base for e in epi])
