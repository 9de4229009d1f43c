#!/usr/bin/env python
"""BASELINE.json config 3 on N GPUs: one 800x800 test-time frame (perturb=0, noise_std=0, white_back, chunk 32768,
test_time=True), rays sharded contiguously over the ranks, per-ray outputs all-gathered (dist.render_sharded: the only
collective).  Device-timed (CUDA events), max over ranks, 3 warm-ups, median of 7.  With N > 1 the gathered frame is
compared bit for bit with rank 0's own single-GPU render of the whole frame (rays are independent: sharding must not
change a pixel).  The fork's sample_pdf draws fresh uniforms on every call, test time included (rendering.py:36-41), so
the comparison injects one frame-sized set of uniforms that both renders slice by ray index.

    python tools/frame_bench.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/frame_bench.py
"""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


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
    from nerf_pl_b200.datasets.ray_utils import gen_rays
    from nerf_pl_b200.dist import render_sharded
    from nerf_pl_b200.models import rendering
    from nerf_pl_b200.models.nerf import Embedding, NeRF
    nerf_pl_b200.set_mlp_precision("bf16")
    torch.manual_seed(0)                      # same weights on every rank
    models = [NeRF().to(dev), NeRF().to(dev)]
    emb = [Embedding(3, 10), Embedding(3, 4)]
    th, ph, radius = math.radians(30.0), math.radians(-30.0), 4.0
    t_r = torch.tensor([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, radius], [0, 0, 0, 1.0]])
    r_ph = torch.tensor([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0], [0, math.sin(ph), math.cos(ph), 0], [0, 0, 0, 1.0]])
    r_th = torch.tensor([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0], [math.sin(th), 0, math.cos(th), 0], [0, 0, 0, 1.0]])
    flip = torch.tensor([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1.0]])
    c2w = (flip @ r_th @ r_ph @ t_r)[:3, :4].contiguous()
    focal = 0.5 * 800 / math.tan(0.5 * 0.6911112070083618)
    frame = gen_rays(c2w, 800, 800, focal, 2.0, 6.0, device=dev)
    out_lines = []
    for Nf in (64, 128):
        def render_slice(rays):
            with torch.no_grad():
                outs = [rendering.render_rays(models, emb, rays[i:i + 32768], 64, False, 0, 0, Nf, 1024 * 32, True, True)
                        for i in range(0, rays.shape[0], 32768)]
                return {k: torch.cat([o[k] for o in outs], 0) for k in outs[0]}

        def frame_once():
            return render_sharded(render_slice, frame)

        for _ in range(3):
            got = frame_once()
        ts = []
        for _ in range(7):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0.record()
            got = frame_once()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t))
        ms = float(np.median(ts))
        same = None
        if world > 1:
            gu = torch.Generator(device=dev).manual_seed(77)
            u_all = torch.rand(frame.shape[0], Nf, device=dev, generator=gu)
            j_all = torch.rand(frame.shape[0], Nf, device=dev, generator=gu)

            def render_injected(rays):
                first = rays.storage_offset() // 8          # index of the slice's first ray in `frame`
                with torch.no_grad():
                    outs = []
                    for i in range(0, rays.shape[0], 32768):
                        a, b = first + i, first + min(i + 32768, rays.shape[0])
                        outs.append(rendering.render_rays(models, emb, rays[i:i + 32768], 64, False, 0, 0, Nf, 1024 * 32, True,
                                                          True, _inject={"u_fine": u_all[a:b], "jitter_fine": j_all[a:b]}))
                    return {k: torch.cat([o[k] for o in outs], 0) for k in outs[0]}

            gathered = render_sharded(render_injected, frame)
            if rank == 0:
                whole = render_injected(frame)
                assert sorted(whole) == sorted(gathered)
                same = all(torch.equal(whole[k], gathered[k]) for k in whole)
        if rank == 0:
            flop = 640000 * (64 * 982528 + (64 + Nf) * 1186816)
            out_lines.append({"config": "800x800 test-time frame, 64 + %d samples, chunk 32768" % Nf, "n_gpus": world,
                              "ms_per_frame": ms, "rays_per_s": 640000 / ms * 1e3, "tflops": flop / ms / 1e9,
                              "keys": sorted(got), "sharded_equals_single_gpu_bitwise": same})
    if rank == 0:
        for l in out_lines:
            print(json.dumps(l), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
