#!/usr/bin/env python
"""L2 -> shared memory bulk-copy rate when every SM streams the SAME packed-weight blob (1.16 MB), against R replicas of
it at different addresses (CTA b reads replica b % R): does one hot blob limit the weight stream of the MLP kernels?"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_pl_b200 import _lib

KB, MB = 1024, 1024 * 1024
dev = torch.device("cuda:0")
buf = torch.zeros(64 * MB, dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream
BLOB = 1184 * KB
for piece, ns, stages in ((32 * KB, 1, 4), (16 * KB, 2, 4)):
    for rep in (0, 2, 4, 8, 16, 37):
        wrap = BLOB // (piece * ns)
        steps = wrap * 60
        _lib.call("nrf_dbg_stream_replicas", rep)
        args = (buf.data_ptr(), (1280 * KB) if rep else 0, piece, piece * ns, piece, ns, 1, wrap, steps, stages, 148, st)
        _lib.call("nrf_dbg_stream", *args)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.call("nrf_dbg_stream", *args)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        gb = 148 * steps * ns * piece / 1e9
        print(f"{ns} x {piece // KB} KB per stage, {stages} stages, replicas {rep or 1:2d}: {gb / ms * 1e3:8.1f} GB/s = {gb / ms * 1e3 / 148 / 1.965:6.1f} B/clk/SM", flush=True)
_lib.call("nrf_dbg_stream_replicas", 0)
