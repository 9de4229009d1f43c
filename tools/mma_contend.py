#!/usr/bin/env python
"""tcgen05.mma rate (M128 N256 K16) with the A operand from shared or tensor memory, alone and while the weight stream
(bulk copies landing) and the epilogue's shared stores run in the same SM (debug build: nrf_dbg_mma_contend)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib

dev = torch.device("cuda:0")
src = torch.zeros((1 << 20) + 148 * 65536, dtype=torch.uint8, device=dev)
out = torch.zeros(148 * 3, dtype=torch.int64, device=dev)
reps = 4096
names = {0: "A smem", 1: "A tmem", 2: "A smem + landing", 3: "A tmem + landing", 4: "A smem + stores", 5: "A tmem + stores",
         6: "A smem + landing + stores", 7: "A tmem + landing + stores"}
for mode in list(range(8)) + [8, 9, 10, 11, 16, 17, 32 + 9, 32 + 8, 32 + 1, 32 + 11, 64 + 9, 64 + 8, 64 + 1, 64 + 0, 128 + 0, 128 + 2, 2, 256, 256 + 128]:
    for _ in range(2):
        out.zero_()
        _lib.call("nrf_dbg_mma_contend", reps, mode, 148, src.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
    o = out.view(148, 3).double()
    cyc = o[:, 0].mean().item() / reps
    land = (o[:, 1] * 32768 / o[:, 0]).mean().item()
    stor = (o[:, 2] * 32768 / o[:, 0]).mean().item()
    nm = ("TMEM ld/st by 16 warps, " if mode & 32 else "") + ("FMA loops in 16 warps, " if mode & 64 else "") + ("bulk stores reading the A tile, " if mode & 128 else "") + ("cp.async landing, " if mode & 256 else "") + names[mode & 7] + (", N=128" if mode & 8 else (", N=64" if mode & 16 else ", N=256"))
    print(f"mode {mode:2d} ({nm:70s}): {cyc:7.1f} cycles per MMA, landing {land:6.1f} B/clk/SM, bulk stores {stor:5.1f} B/clk/SM", flush=True)
