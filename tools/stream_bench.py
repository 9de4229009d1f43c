#!/usr/bin/env python
"""Bulk-copy streaming benchmark (HBM and L2 -> shared memory) on one B200; prints GB/s per pattern."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib

KB, MB = 1024, 1024 * 1024
dev = torch.device("cuda:0")
buf = torch.empty(6 * 1024 * MB, dtype=torch.uint8, device=dev)
buf.zero_()
st = torch.cuda.current_stream().cuda_stream
GRID = 148


def run(name, cta_stride, stream_stride, tile_stride, piece, ns, ppc, wrap, steps, stages, grid=GRID):
    args = (buf.data_ptr(), cta_stride, stream_stride, tile_stride, piece, ns, ppc, wrap, steps, stages, grid, st)
    _lib.call("nrf_dbg_stream", *args)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _lib.call("nrf_dbg_stream", *args)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    gb = grid * steps * ns * piece / 1e9
    print(json.dumps({"pattern": name, "GB": round(gb, 3), "ms": round(ms, 4), "GB/s": round(gb / ms * 1e3, 1)}), flush=True)


T = 40  # tiles per CTA
# HBM patterns -------------------------------------------------------------------------------------
run("hbm tile-major stash: 8 streams x 4KB pieces (16KB chunks), tile stride 640KB, 6 stages",
    T * 640 * KB, 16 * KB, 640 * KB, 4 * KB, 8, 4, 0, T * 4, 6)
run("hbm layer-major: 8 streams x 4KB, tile stride 128KB (contiguous), 6 stages",
    T * 128 * KB, 16 * KB, 128 * KB, 4 * KB, 8, 4, 0, T * 4 * 5, 6)
run("hbm tile-major: 8 streams x 8KB pieces, tile stride 640KB, 3 stages",
    T * 640 * KB, 16 * KB, 640 * KB, 8 * KB, 8, 2, 0, T * 2, 3)
run("hbm layer-major: 8 streams x 8KB, tile stride 128KB, 3 stages",
    T * 128 * KB, 16 * KB, 128 * KB, 8 * KB, 8, 2, 0, T * 2 * 5, 3)
run("hbm layer-major: 4 streams x 16KB, tile stride 64KB, 3 stages",
    T * 5 * 64 * KB, 16 * KB, 64 * KB, 16 * KB, 4, 1, 0, T * 5, 3)
run("hbm sequential: 1 stream x 32KB, 6 stages", 32 * MB, 0, 32 * KB, 32 * KB, 1, 1, 0, 1024, 6)
run("hbm sequential: 1 stream x 64KB, 3 stages", 32 * MB, 0, 64 * KB, 64 * KB, 1, 1, 0, 512, 3)
run("hbm sequential: 8 streams x 4KB adjacent (32KB/stage), 6 stages", 32 * MB, 4 * KB, 32 * KB, 4 * KB, 8, 1, 0, 1024, 6)
run("hbm interleaved CTAs: 1 stream x 32KB, cta stride 32KB, tile stride 148*32KB", 32 * KB, 0, 148 * 32 * KB, 32 * KB, 1, 1, 0, 1024, 6)
# L2 patterns (every CTA reads the same 1.17 MB blob) ----------------------------------------------------
for piece, ns, stages in ((32 * KB, 1, 3), (16 * KB, 1, 3), (16 * KB, 2, 3), (32 * KB, 1, 5), (8 * KB, 4, 3), (16 * KB, 1, 6), (32 * KB, 2, 3), (16 * KB, 4, 3), (8 * KB, 2, 6)):
    wrap = (1184 * KB) // (piece * ns)
    run(f"L2 shared blob 1.16MB: {ns} x {piece // KB}KB per stage, {stages} stages", 0, piece, piece * ns, piece, ns, 1, wrap,
        wrap * 60, stages)
