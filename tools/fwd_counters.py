#!/usr/bin/env python
"""Cycle breakdown of the tensor-core forward kernel (MMA-thread / epilogue waits) on one B200."""
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
for train in (False, True):
    for _ in range(3):
        forward_rays(m.spec, params, rays, z, False, train, "bf16", packed=packed)
    cnt = torch.zeros(148 * 8, dtype=torch.int64, device=dev)
    lib.nrf_dbg_set_counters(cnt.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    forward_rays(m.spec, params, rays, z, False, train, "bf16", packed=packed)
    e1.record()
    torch.cuda.synchronize()
    lib.nrf_dbg_set_counters(None)
    c = cnt.view(148, 8).double().mean(0).tolist()
    tiles = B * S / 128 / 148
    print(f"train={train} ms={e0.elapsed_time(e1):.3f} per-tile cycles: mma_total={c[0]/tiles:.0f} wait_A={c[1]/tiles:.0f} "
          f"wait_W={c[2]/tiles:.0f} | epi_total={c[4]/tiles:.0f} epi_wait_acc={c[5]/tiles:.0f} epi_wait_stash_read={c[6]/tiles:.0f}", flush=True)
