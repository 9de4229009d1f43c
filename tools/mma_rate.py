#!/usr/bin/env python
"""Cycles per tcgen05.mma (M=128, N=256, K=16, bf16 -> fp32) on this GPU, one CTA per SM and a single CTA."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib

lib = _lib.load()
dev = torch.device("cuda:0")
for grid in (1, 148):
    for reps in (64, 1024, 16384):
        out = torch.zeros(grid, dtype=torch.int64, device=dev)
        for _ in range(2):
            _lib.call("nrf_dbg_mma_rate", reps, grid, out.data_ptr(), None)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.call("nrf_dbg_mma_rate", reps, grid, out.data_ptr(), None)
        e1.record()
        torch.cuda.synchronize()
        cyc = out.double().mean().item()
        ms = e0.elapsed_time(e1)
        flop = 2.0 * 128 * 256 * 16 * reps * grid
        print(f"grid={grid:3d} reps={reps:5d}: {cyc / reps:7.1f} cycles per MMA  ({2 * 128 * 256 * 16 / (cyc / reps):.0f} FLOP/clk/SM), "
              f"kernel {ms:.3f} ms -> {flop / ms / 1e9:.0f} TFLOP/s, implied clock {cyc / ms / 1e3:.0f} MHz", flush=True)

print("N sweep (148 CTAs, 4096 MMAs each): cycles per MMA and per 256 output columns")
for n in (256, 128, 64, 32):
    out = torch.zeros(148, dtype=torch.int64, device=dev)
    for _ in range(2):
        _lib.call("nrf_dbg_mma_rate_n", 4096, n, 148, out.data_ptr(), None)
    torch.cuda.synchronize()
    cyc = out.double().mean().item() / 4096
    print(f"N={n:3d}: {cyc:6.1f} cycles per MMA = {cyc * 256 / n:6.1f} per 256 columns ({2 * 128 * n * 16 / cyc:.0f} FLOP/clk/SM)", flush=True)
