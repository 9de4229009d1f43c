"""fp32 parity-mode MLP (mlp_fp32.cu) timing: BASELINE.json config 1 (1024 rays, 64+64, fwd+bwd) and the 4096-ray batch.
    python tools/fp32_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import __graft_entry__ as ge
ge.build()
import nerf_pl_b200
from nerf_pl_b200.dist import GradBucket
from nerf_pl_b200.losses import loss_dict
from nerf_pl_b200.models import rendering
from nerf_pl_b200.models.nerf import Embedding, NeRF
from oracle import nerf_oracle as O   # input generator only

dev = torch.device("cuda:0")
torch.manual_seed(0)
models = [NeRF().to(dev), NeRF().to(dev)]
emb = [Embedding(3, 10), Embedding(3, 4)]
bucket = GradBucket(models)
mse = loss_dict["mse"]()
nerf_pl_b200.set_mlp_precision("fp32")
for n in (1024, 4096):
    rays = O.blender_rays(n, seed=1).to(dev)
    tgt = torch.rand(n, 3, device=dev)

    def f():
        bucket.zero()
        out = rendering.render_rays(models, emb, rays, 64, False, 1.0, 1.0, 64, 1024 * 32, True)
        mse(out, tgt).backward()

    for _ in range(2):
        f()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); f(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    print(f"fp32 mode, {n} rays fwd+bwd: {ms:.2f} ms -> {n / ms:.1f} k rays/s, {n * 669.9e6 / ms / 1e9:.1f} TFLOP/s (fp32 FFMA)", flush=True)
