#!/usr/bin/env python
"""Bulk-store rate shared memory -> HBM (cp.async.bulk s2g), one SM alone and all 148 together."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib

dev = torch.device("cuda:0")
steps = 512
buf = torch.empty(148 * steps * 65536, dtype=torch.uint8, device=dev)
for grid in (1, 8, 148):
    for piece in (16384, 65536):
        for depth in (1, 2, 4):
            out = torch.zeros(grid, dtype=torch.int64, device=dev)
            for _ in range(2):
                _lib.call("nrf_dbg_store_rate", buf.data_ptr(), steps * 65536, piece, steps, depth, grid, out.data_ptr(), None)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.call("nrf_dbg_store_rate", buf.data_ptr(), steps * 65536, piece, steps, depth, grid, out.data_ptr(), None)
            e1.record()
            torch.cuda.synchronize()
            cyc = out.double().mean().item()
            ms = e0.elapsed_time(e1)
            print(f"grid={grid:3d} piece={piece:5d} depth={depth}: {steps * piece / cyc:6.1f} B/clk/SM, "
                  f"{grid * steps * piece / ms / 1e6:7.0f} GB/s aggregate", flush=True)
