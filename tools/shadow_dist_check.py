#!/usr/bin/env python
"""Shadow-mapping step on N GPUs (BASELINE.json config 5): camera rays sharded over the ranks, the light-view
depth render (a) replicated on every rank or (b) sharded with dist.render_light_sharded (all-gather with gradient).
Both must give the same averaged parameter gradients; prints the relative L2 difference and the step times.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/shadow_dist_check.py
"""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def pose(theta, phi, radius=4.0):
    th, ph = math.radians(theta), math.radians(phi)
    t_r = torch.tensor([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, radius], [0, 0, 0, 1.0]])
    r_ph = torch.tensor([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0], [0, math.sin(ph), math.cos(ph), 0], [0, 0, 0, 1.0]])
    r_th = torch.tensor([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0], [math.sin(th), 0, math.cos(th), 0], [0, 0, 0, 1.0]])
    flip = torch.tensor([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1.0]])
    return (flip @ r_th @ r_ph @ t_r)[:3, :4].contiguous()


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()
    import nerf_pl_b200
    from nerf_pl_b200.dist import GradBucket, render_light_sharded, shard_bounds
    from nerf_pl_b200.losses import loss_dict
    from nerf_pl_b200.models import rendering_shadows as RS
    from nerf_pl_b200.models.camera import Camera
    from nerf_pl_b200.models.nerf import Embedding, NeRF
    precision = os.environ.get("PRECISION", "bf16")
    nerf_pl_b200.set_mlp_precision(precision)
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
    n_cam = 4096
    lo, hi = shard_bounds(n_cam, rank, world)
    cam_pix = torch.cat([ij[lo:hi] + 0.5, torch.ones(hi - lo, 1)], 1)
    light_pix = torch.cat([ij + 0.5, torch.ones(16384, 1)], 1)
    rays_c, rays_l = cam_rays(cam, ij[lo:hi]).to(dev), cam_rays(light, ij).to(dev)
    ppc = {"eye_pos": torch.as_tensor(cam.eye_pos, dtype=torch.float32).expand(hi - lo, 3).contiguous(),
           "camera": torch.as_tensor(cam.camera, dtype=torch.float32).expand(hi - lo, 3, 3).contiguous()}
    tgt = torch.rand(n_cam, 3, generator=torch.Generator().manual_seed(3))[lo:hi].to(dev)

    gi = torch.Generator(device=dev).manual_seed(5 + rank)      # fixed sampler uniforms: the two modes must see the same rays
    inj = {"u_fine": torch.rand(hi - lo, 64, device=dev, generator=gi), "jitter_fine": torch.rand(hi - lo, 64, device=dev, generator=gi)}

    def light_fn(r):
        return RS.render_rays(models, emb, r, 64, False, 0.0, 0.0, 0, 1024 * 32, False)

    def step(sharded):
        bucket.zero()
        cr = RS.render_rays(models, emb, rays_c, 64, False, 0.0, 0.0, 64, 1024 * 32, False, _inject=inj)
        lr = render_light_sharded(light_fn, rays_l) if sharded else light_fn(rays_l)
        out = RS.efficient_sm(cam_pix, light_pix, cr, lr, ppc, light, res, True, 0, "shadow_method_2")
        mse(out, tgt).backward()
        bucket.allreduce_mean()

    def timed(sharded):
        for _ in range(2):
            step(sharded)
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0.record(); step(sharded); e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t))
        return float(np.median(ts)), bucket.flat.clone()

    ms_rep, g_rep = timed(False)
    ms_sh, g_sh = timed(True)
    ms_rep2, g_rep2 = timed(False)          # run-to-run noise of the fp32 atomics, for scale
    rel = float((g_rep - g_sh).norm() / g_rep.norm())
    noise = float((g_rep - g_rep2).norm() / g_rep.norm())
    if rank == 0:
        print(json.dumps({"n_gpus": world, "precision": precision, "camera_rays_per_rank": hi - lo, "light_rays": 16384,
                          "ms_replicated_light": ms_rep, "ms_sharded_light": ms_sh,
                          "grad_rel_l2_sharded_vs_replicated": rel, "grad_rel_l2_replicated_run_to_run": noise,
                          "grad_norm": float(g_rep.norm())}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
