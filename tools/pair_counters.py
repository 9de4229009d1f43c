#!/usr/bin/env python
"""Cycle breakdown of the CTA-pair forward kernel (debug build): UMMA-issuer waits and epilogue waits per slot-layer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib
from nerf_pl_b200.mlp import forward_rays
from nerf_pl_b200.models.nerf import NeRF

os.environ["NRF_FWD_IMPL"] = "pair"
dev = torch.device("cuda:0")
m = NeRF().to(dev)
B, S = 4096, 128
rays = torch.randn(B, 8, device=dev)
z = torch.sort(torch.rand(B, S, device=dev) * 4 + 2, -1)[0]
params = [p.detach() for p in m.ordered_parameters()]
packed = m.packed_weights()
lib = _lib.load()
for fl in [int(x) for x in os.environ.get('PAIR_FLAGS', '0,4').split(',')]:
  os.environ["NRF_DBG_FLAGS"] = str(fl)
  print("NRF_DBG_FLAGS =", fl)
  for train in (False,):
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
      c = cnt.view(148, 8).double()
      lead, allc = c[0::2].mean(0).tolist(), c.mean(0).tolist()
      n = (B * S / 128 / 296 + 0.999) // 1 * 20          # slot-layers per CTA
      print(f"train={train} ms={e0.elapsed_time(e1):.3f} per slot-layer cycles: mma_total={lead[0]/n:.0f} wait_A={lead[1]/n:.0f} "
            f"wait_W={lead[2]/n:.0f} wait_Wpeer={lead[3]/n:.0f} | epi_total={allc[4]/n:.0f} epi_wait_acc={allc[5]/n:.0f} | "
          f"MMA slot-layer (16 MMAs, weights resident) issue->complete = {lead[6]/max(lead[7],1):.0f} cycles", flush=True)
