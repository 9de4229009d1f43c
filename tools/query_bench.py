#!/usr/bin/env python
"""Query-mode MLP (extract_color_mesh.py:113-133): sigma + rgb on a dense N^3 grid, points/s and TFLOP/s."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nerf_pl_b200.models.nerf import NeRF

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
m = NeRF().to(dev)
lin = torch.linspace(-1.2, 1.2, N, device=dev)
xyz = torch.stack(torch.meshgrid(lin, lin, lin, indexing="xy"), -1).reshape(-1, 3).contiguous()
for sigma_only, flop in ((False, 1186816), (True, 982528)):
    m.query(xyz[: 1 << 20], sigma_only=sigma_only)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    out = m.query(xyz, sigma_only=sigma_only)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"N={N}^3 = {xyz.shape[0]} points, sigma_only={sigma_only}: {ms:.2f} ms, {xyz.shape[0] / ms / 1e3:.1f} M points/s, "
          f"{xyz.shape[0] * flop / ms / 1e9:.0f} TFLOP/s (incl. staging the points as one-sample rays)", flush=True)
