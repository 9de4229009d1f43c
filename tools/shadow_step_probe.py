"""Where does the config-5 shadow-mapping step (one GPU) spend its time?  Device time (CUDA events) and host wall clock
per phase: camera render, light render, efficient_sm, loss + backward.      python tools/shadow_step_probe.py"""
import math, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import __graft_entry__ as ge
ge.build()
import nerf_pl_b200
from nerf_pl_b200.dist import GradBucket
from nerf_pl_b200.losses import loss_dict
from nerf_pl_b200.models import rendering_shadows as RS
from nerf_pl_b200.models.camera import Camera
from nerf_pl_b200.models.nerf import Embedding, NeRF
from tools.shadow_dist_check import pose

dev = torch.device("cuda:0")
nerf_pl_b200.set_mlp_precision("bf16")
torch.manual_seed(0)
models = [NeRF().to(dev), NeRF().to(dev)]
emb = [Embedding(3, 10), Embedding(3, 4)]
bucket = GradBucket(models)
mse = loss_dict["mse"]()
res = (128, 128)
cam, light = Camera(30, res), Camera(30, res)
cam.set_pose_using_blender_matrix(pose(30, -30))
light.set_pose_using_blender_matrix(pose(75, -50))


def cam_rays(c, pix_ij):
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
tgt = torch.rand(4096, 3, device=dev)
phases = ["camera render", "light render", "efficient_sm", "loss+backward"]
acc = {p: [0.0, 0.0] for p in phases}


def phase(name, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    out = fn()
    e1.record(); t1 = time.perf_counter()
    return out, (name, e0, e1, t1 - t0)


for it in range(8):
    torch.cuda.synchronize()
    T0 = time.perf_counter()
    bucket.zero()
    rec = []
    cr, r = phase(phases[0], lambda: RS.render_rays(models, emb, rays_c, 64, False, 1.0, 0.0, 64, 1024 * 32, False)); rec.append(r)
    with torch.no_grad():
        lr, r = phase(phases[1], lambda: RS.render_rays(models, emb, rays_l, 64, False, 0.0, 0.0, 0, 1024 * 32, False)); rec.append(r)
    out, r = phase(phases[2], lambda: RS.efficient_sm(cam_pix, light_pix, cr, lr, ppc, light, res, True, 0, "shadow_method_2")); rec.append(r)
    _, r = phase(phases[3], lambda: mse(out, tgt).backward()); rec.append(r)
    torch.cuda.synchronize()
    wall = time.perf_counter() - T0
    if it >= 3:
        for name, e0, e1, host in rec:
            acc[name][0] += e0.elapsed_time(e1); acc[name][1] += host * 1e3
        acc.setdefault("wall", [0.0, 0.0])[0] += wall * 1e3
for p in phases:
    print(f"{p:16s} device-span {acc[p][0] / 5:7.3f} ms   host {acc[p][1] / 5:7.3f} ms")
print(f"whole step wall {acc['wall'][0] / 5:.3f} ms")

if os.environ.get("PROFILE_SM"):
    import cProfile, pstats
    cr = RS.render_rays(models, emb, rays_c, 64, False, 1.0, 0.0, 64, 1024 * 32, False)
    with torch.no_grad():
        lr = RS.render_rays(models, emb, rays_l, 64, False, 0.0, 0.0, 0, 1024 * 32, False)
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(20):
        RS.efficient_sm(cam_pix, light_pix, dict(cr), lr, ppc, light, res, True, 0, "shadow_method_2")
    pr.disable()
    torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
