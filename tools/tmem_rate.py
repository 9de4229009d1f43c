#!/usr/bin/env python
"""TMEM read rate (tcgen05.ld) per SM with 4 / 8 / 16 reader warps, alone and under concurrent tcgen05.mma traffic.
Needs a debug build (NRF_DEBUG_BUILD=1 python -m nerf_pl_b200._build --force)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib

dev = torch.device("cuda:0")
reps = 200
for grid in (1, 148):
    for mode in (0, 1, 2, 3):
        for warps in (4, 8, 16):
            out = torch.zeros(grid * 2, dtype=torch.int64, device=dev)
            for _ in range(2):
                _lib.call("nrf_dbg_tmem_rate", warps, reps, mode, grid, out.data_ptr(), None)
            torch.cuda.synchronize()
            o = out.view(grid, 2).double()
            cyc = o[:, 0].mean().item()
            nbytes = reps * 128 * 512 * 4
            mm = o[:, 1].mean().item() if mode & 2 else 0
            print(f"grid={grid:3d} ld={'x32' if mode & 1 else 'x16'} mma={'on ' if mode & 2 else 'off'} warps={warps:2d}: "
                  f"{cyc / reps:8.0f} cycles per 256 KB pass = {nbytes / cyc:6.1f} B/clk/SM"
                  + (f", {mm * 128 / cyc:5.2f} of the MMA rate sustained meanwhile" if mode & 2 else ""), flush=True)

print("shared-memory traffic from the same warps (64 KB per pass and CTA) against the MMA operand fetch:")
for mode, name in ((4, "STS.128 mma=off"), (6, "STS.128 mma=on "), (8, "LDS.128 mma=off"), (10, "LDS.128 mma=on ")):
    for warps in (4, 8, 16):
        out = torch.zeros(148 * 2, dtype=torch.int64, device=dev)
        for _ in range(2):
            _lib.call("nrf_dbg_tmem_rate", warps, 400, mode, 148, out.data_ptr(), None)
        torch.cuda.synchronize()
        o = out.view(148, 2).double()
        cyc = o[:, 0].mean().item()
        mm = o[:, 1].mean().item() if mode & 2 else 0
        print(f"{name} warps={warps:2d}: {400 * 65536 / cyc:6.1f} B/clk/SM"
              + (f", {mm * 128 / cyc:5.2f} of the MMA rate sustained meanwhile" if mode & 2 else ""), flush=True)
