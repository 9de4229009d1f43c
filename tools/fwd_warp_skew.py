#!/usr/bin/env python
"""Per-warp epilogue timeline of CTA 0's third tile in the tensor-core forward kernel (debug build): when each of the 16
epilogue warps saw the accumulator complete and when it finished its two hand-overs, per layer.  Shows the skew between
the warps of the four SM sub-partitions (warp % 4)."""
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
base = 148 * 8 + 4 * CAP + 148 * 32
cnt = torch.zeros(base + 17 * 64, dtype=torch.int64, device=dev)
os.environ["NRF_DBG_FLAGS"] = "16"          # light mode: only these trace points
lib.nrf_dbg_set_counters(cnt.data_ptr())
forward_rays(m.spec, params, rays, z, False, train, "bf16", packed=packed)
torch.cuda.synchronize()
lib.nrf_dbg_set_counters(None)
w = cnt[base:].view(17, 16, 4).cpu().numpy()
t0 = w[:, 1, 0].min()
print("layer: per epilogue warp (sub-partition = (warp+2) % 4): [acc seen, hand-over 1, hand-over 2] relative to the first warp's acc-seen of layer 1")
for L in range(1, 9):
    m0, m1, m2, m3 = (int(w[16, L, e] - t0) for e in range(4))
    print(f"L{L}  UMMA issuer: first A seen {m0:6d}  first group issued {m1:6d}  last A seen {m2:6d}  commit issued {m3:6d}")
    for ew in range(16):
        a0, h1, h2 = (int(w[ew, L, e] - t0) for e in range(3))
        print(f"   warp {ew + 2:2d} smsp {(ew + 2) % 4}: acc {a0:6d}  h1 {h1:6d} (+{h1 - a0:5d})  h2 {h2:6d} (+{h2 - a0:5d})")
