#!/usr/bin/env python
"""Timeline of CTA 0's third tile in the tensor-memory forward kernel (debug build, NRF_FWD_IMPL=ts): per layer, when the
UMMA issuer had operands + weights for the two stages of half 0 and issued the two accumulator commits, and when the
epilogue warps saw each accumulator half and finished its hand-over (min .. max over the 16 warps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["NRF_FWD_IMPL"] = "ts"
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
cnt = torch.zeros(base + 18 * 64, dtype=torch.int64, device=dev)
os.environ["NRF_DBG_FLAGS"] = os.environ.get("TS_FLAGS", "16")
lib.nrf_dbg_set_counters(cnt.data_ptr())
forward_rays(m.spec, params, rays, z, False, train, "bf16", packed=packed)
torch.cuda.synchronize()
lib.nrf_dbg_set_counters(None)
w = cnt[base:].view(18, 16, 4).cpu().numpy()
t0 = w[16, 1, 0]
prev = None
for L in range(1, 9):
    i = [int(w[16, L, e] - t0) for e in range(4)]
    ep = [[int(w[ew, L, e] - t0) for ew in range(16)] for e in range(4)]
    line = (f"L{L} issuer: h0 stage0 {i[0]:6d}  h0 stage1 {i[1]:6d}  commit0 {i[2]:6d}  commit1 {i[3]:6d} | epilogue: acc0 seen {min(ep[0]):6d}..{max(ep[0]):6d}"
            f"  hand-over0 {min(ep[1]):6d}..{max(ep[1]):6d}  acc1 seen {min(ep[2]):6d}..{max(ep[2]):6d}  hand-over1 {min(ep[3]):6d}..{max(ep[3]):6d}")
    x = [int(w[17, L, e] - t0) for e in range(4)]
    line += f" | 2nd stage: h0 issued+A ok {x[0]:6d} weights ok {x[1]:6d}; h1 issued+A ok {x[2]:6d} weights ok {x[3]:6d}"
    if prev is not None:
        line += f"   period {i[3] - prev}"
    prev = i[3]
    print(line)
