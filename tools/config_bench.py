#!/usr/bin/env python
"""BASELINE.json configs 1, 3, 4, 5 on one B200 (config 2 is bench.py).  Synthetic inputs per SURVEY §8d: random-init
NeRF D=8 W=256, Blender-style / LLFF-style / shadow rays.  CUDA-event timed, 3 warm-ups, median of 7."""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import nerf_pl_b200
from nerf_pl_b200.datasets.ray_utils import gen_rays
from nerf_pl_b200.dist import GradBucket
from nerf_pl_b200.losses import loss_dict
from nerf_pl_b200.models import rendering, rendering_shadows
from nerf_pl_b200.models.camera import Camera
from nerf_pl_b200.models.nerf import Embedding, NeRF

dev = torch.device("cuda:0")
torch.manual_seed(0)
models = [NeRF().to(dev), NeRF().to(dev)]
emb = [Embedding(3, 10), Embedding(3, 4)]
bucket = GradBucket(models)
mse = loss_dict["mse"]()


def timed(fn, reps=7, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def blender_pose(theta=30.0, phi=-30.0, radius=4.0):
    th, ph = math.radians(theta), math.radians(phi)
    t_r = torch.tensor([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, radius], [0, 0, 0, 1.0]])
    r_ph = torch.tensor([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0], [0, math.sin(ph), math.cos(ph), 0], [0, 0, 0, 1.0]])
    r_th = torch.tensor([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0], [math.sin(th), 0, math.cos(th), 0], [0, 0, 0, 1.0]])
    flip = torch.tensor([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1.0]])
    return (flip @ r_th @ r_ph @ t_r)[:3, :4].contiguous()


def train_step(mod, rays, tgt, Nc, Nf, use_disp=False):
    def f():
        bucket.zero()
        out = mod.render_rays(models, emb, rays, Nc, use_disp, 1.0, 1.0, Nf, 1024 * 32, True)
        mse(out, tgt).backward()
    return f


focal800 = 0.5 * 800 / math.tan(0.5 * 0.6911112070083618)
# ---- config 1: 1024 rays fwd+bwd, fp32 MLP -----------------------------------------------------------------
pix = torch.randperm(800 * 800, generator=torch.Generator().manual_seed(1))[:1024].to(dev)
rays1 = gen_rays(blender_pose(), 800, 800, focal800, 2.0, 6.0, pix=pix)
tgt1 = torch.rand(1024, 3, device=dev)
nerf_pl_b200.set_mlp_precision("fp32")
ms = timed(train_step(rendering, rays1, tgt1, 64, 64))
print(f"config 1  1024 rays fwd+bwd fp32 MLP: {ms:.3f} ms  -> {1024 / ms * 1e3 / 1e3:.1f} k rays/s", flush=True)
nerf_pl_b200.set_mlp_precision("bf16")
ms = timed(train_step(rendering, rays1, tgt1, 64, 64))
print(f"config 1b 1024 rays fwd+bwd bf16 MLP: {ms:.3f} ms  -> {1024 / ms * 1e3 / 1e3:.1f} k rays/s", flush=True)

# ---- config 3: 800x800 test-time render, chunks of 32768 rays ----------------------------------------------
frame = gen_rays(blender_pose(), 800, 800, focal800, 2.0, 6.0, device=dev)
for Nf in (64, 128):
    def render():
        with torch.no_grad():
            outs = [rendering.render_rays(models, emb, frame[i:i + 32768], 64, False, 0, 0, Nf, 1024 * 32, True, True)
                    for i in range(0, frame.shape[0], 32768)]
            return torch.cat([o["rgb_fine"] for o in outs], 0)
    ms = timed(render, reps=5, warm=2)
    flop = 640000 * (64 * 982528 + (64 + Nf) * 1186816)
    print(f"config 3  800x800 test render, N_importance={Nf}, 20 calls of 32768 rays: {ms:.1f} ms/frame "
          f"({flop / ms / 1e9:.0f} TFLOP/s, {640000 / ms * 1e3 / 1e6:.2f} M rays/s)", flush=True)

# ---- config 4: LLFF 504x378 NDC, N_importance=128, batch 8192 ----------------------------------------------
H, W, f4 = 378, 504, 407.6
c2w4 = torch.tensor([[0.995, -0.03, 0.09, 0.12], [0.02, 0.99, 0.1, -0.08], [-0.1, -0.1, 0.99, 0.05]])
pix4 = torch.randperm(H * W, generator=torch.Generator().manual_seed(4))[:8192].to(dev)
rays4a = gen_rays(c2w4, H, W, f4, 1.0, 1.0, pix=pix4, ndc=True)
tgt4 = torch.rand(8192, 3, device=dev)
ms = timed(train_step(rendering, rays4a, tgt4, 64, 128))
print(f"config 4a LLFF NDC (near=0, far=1), 8192 rays, 64+128: {ms:.3f} ms -> {8192 / ms * 1e3 / 1e6:.3f} M rays/s", flush=True)
rays4b = gen_rays(c2w4, H, W, f4, 1.33, 10.6, pix=pix4)
ms = timed(train_step(rendering, rays4b, tgt4, 64, 128, use_disp=True))
print(f"config 4b use_disp, spheric bounds 1.33..10.6, 8192 rays, 64+128: {ms:.3f} ms -> {8192 / ms * 1e3 / 1e6:.3f} M rays/s", flush=True)

# ---- config 5: shadow-mapping step at 128x128: 4096 camera rays + 16384 light rays + shadow test ------------
res = (128, 128)
cam, light = Camera(30, res), Camera(30, res)
cam.set_pose_using_blender_matrix(blender_pose(30, -30))
light.set_pose_using_blender_matrix(blender_pose(75, -50))


def cam_rays(c, pix_ij):
    """rays through pixel centres of a models.camera.Camera (datasets/blender_efficient_sm.py:98-110)."""
    M = torch.as_tensor(c.camera, dtype=torch.float32)
    p = torch.cat([pix_ij + 0.5, torch.ones_like(pix_ij[:, :1])], 1)
    d = torch.nn.functional.normalize(p @ M.T, dim=-1)
    o = torch.as_tensor(c.eye_pos, dtype=torch.float32).expand_as(d)
    return torch.cat([o, d, torch.full_like(d[:, :1], 1.0), torch.full_like(d[:, :1], 200.0)], 1).contiguous()


ij = torch.stack(torch.meshgrid(torch.arange(128.0), torch.arange(128.0), indexing="xy"), -1).reshape(-1, 2)
cam_pix = torch.cat([ij[:4096] + 0.5, torch.ones(4096, 1)], 1)
light_pix = torch.cat([ij + 0.5, torch.ones(16384, 1)], 1)
rays_c, rays_l = cam_rays(cam, ij[:4096]).to(dev), cam_rays(light, ij).to(dev)
ppc = {"eye_pos": torch.as_tensor(cam.eye_pos, dtype=torch.float32).expand(4096, 3).contiguous(),
       "camera": torch.as_tensor(cam.camera, dtype=torch.float32).expand(4096, 3, 3).contiguous()}
# the batch reaches the step on the device, as Lightning delivers it (train_efficient_sm.py:186-200)
cam_pix, light_pix = cam_pix.to(dev), light_pix.to(dev)
ppc = {k: v.to(dev) for k, v in ppc.items()}
tgt5 = torch.rand(4096, 3, device=dev)
sm_loss = loss_dict["mse"]()
for lni, grad_light in ((0, False), (64, True)):
    def f():
        bucket.zero()
        cr = rendering_shadows.render_rays(models, emb, rays_c, 64, False, 1.0, 0.0, 64, 1024 * 32, False)
        with torch.set_grad_enabled(grad_light):
            lr = rendering_shadows.render_rays(models, emb, rays_l, 64, False, 0.0, 0.0, lni, 1024 * 32, False)
        out = rendering_shadows.efficient_sm(cam_pix, light_pix, cr, lr, ppc, light, res, True, lni, "shadow_method_2")
        sm_loss(out, tgt5).backward()
    ms = timed(f)
    print(f"config 5  shadow step 128x128: 4096 camera rays (64+64) + 16384 light rays (64+{lni}), grad_on_light={grad_light}: "
          f"{ms:.2f} ms -> {4096 / ms * 1e3 / 1e3:.0f} k camera rays/s", flush=True)
