#!/usr/bin/env python
"""wgrad kernel time of one training step under the NRF_DBG_FLAGS experiments (1 = no side work, 2 = no MMAs,
8 = 2 units per SM instead of 8); results are WRONG under flags 1/2, only the timing is of interest."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nerf_pl_b200
from nerf_pl_b200 import _lib
from nerf_pl_b200.dist import GradBucket
from nerf_pl_b200.models.nerf import Embedding, NeRF
from nerf_pl_b200.models.rendering import render_rays

dev = torch.device("cuda:0")
lib = _lib.load()
models = [NeRF().to(dev), NeRF().to(dev)]
emb = [Embedding(3, 10), Embedding(3, 4)]
bucket = GradBucket(models)
B = 4096
rays = torch.cat([torch.zeros(B, 3), torch.nn.functional.normalize(torch.randn(B, 3), dim=-1), torch.full((B, 1), 2.0),
                  torch.full((B, 1), 6.0)], -1).to(dev)
tgt = torch.rand(B, 3, device=dev)


def step():
    bucket.zero()
    out = render_rays(models, emb, rays, 64, False, 1.0, 1.0, 64, 1024 * 32, True)
    (((out["rgb_coarse"] - tgt) ** 2).mean() + ((out["rgb_fine"] - tgt) ** 2).mean()).backward()


flags = [int(x) for x in sys.argv[1:]] or [0, 1, 2, 3, 8]
ROUNDS = int(os.environ.get("PROBE_ROUNDS", "1"))      # > 1: round-robin over the flags (clock drift hits all alike)
acc = {fl: [] for fl in flags}
for rnd in range(ROUNDS):
    for fl in flags:
        os.environ["NRF_DBG_FLAGS"] = str(fl)
        for _ in range(2):
            step()
        torch.cuda.synchronize()
        lib.nrf_prof_enable(1)
        for _ in range(5):
            step()
        torch.cuda.synchronize()
        lib.nrf_prof_enable(0)
        res = {}
        for kid, name in ((3, "fwd"), (5, "dgrad"), (7, "wgrad")):
            tot, n, rows = ctypes.c_double(), ctypes.c_int64(), ctypes.c_int64()
            lib.nrf_prof_read(kid, ctypes.byref(tot), ctypes.byref(n), ctypes.byref(rows))
            res[name] = tot.value / 5
        acc[fl].append(res)
for fl in flags:
    med = {k: sorted(r[k] for r in acc[fl])[len(acc[fl]) // 2] for k in ("fwd", "dgrad", "wgrad")}
    ratio = sorted(r["wgrad"] / r["fwd"] for r in acc[fl])[len(acc[fl]) // 2]
    print(f"flags={fl}: per step ms (median of {ROUNDS}) " + " ".join(f"{k}={v:.3f}" for k, v in med.items()) + f"  wgrad/fwd={ratio:.3f}", flush=True)
