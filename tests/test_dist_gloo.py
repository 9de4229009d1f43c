"""World-size-2 checks of the multi-process host logic on the CPU (gloo, 127.0.0.1): ray sharding bounds, the single flat
gradient all-reduce of dist.GradBucket, and the sharded full-frame render + gather of dist.render_sharded (ragged frame
sizes included).  The CUDA kernels are not involved: render_fn is a deterministic per-ray stand-in, which is all the
gather logic can see."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nerf_pl_b200.dist import GradBucket, gather_rows, render_light_sharded, render_sharded, shard_bounds


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _stand_in_render(rays):
    return {"rgb_fine": rays[:, :3] * 2.0 + rays[:, 3:6], "depth_fine": rays[:, 6] + 0.5 * rays[:, 7],
            "opacity_coarse": rays[:, 0].abs()}


def _worker(rank, world, port, n_rays, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)                                   # same model on every rank
        from nerf_pl_b200.models.nerf import NeRF
        models = [NeRF(), NeRF()]
        bucket = GradBucket(models)
        assert bucket.flat.numel() == 2 * 595844
        off = 0
        for p in bucket.params:                                # .grad of every parameter is a view of the flat buffer
            assert p.grad.data_ptr() == bucket.flat.data_ptr() + 4 * off
            off += p.numel()
        bucket.zero()
        g = torch.Generator().manual_seed(100 + rank)
        local = torch.randn(bucket.flat.numel(), generator=g)
        bucket.flat.copy_(local)
        bucket.allreduce_mean()
        want = sum(torch.randn(bucket.flat.numel(), generator=torch.Generator().manual_seed(100 + r)) for r in range(world)) / world
        assert torch.allclose(bucket.flat, want, rtol=1e-6, atol=1e-7)
        assert torch.equal(bucket.params[-1].grad.reshape(-1), bucket.flat[-bucket.params[-1].numel():])   # views saw it
        # sharded render: every rank holds the whole frame's rays, renders its slice, rank 0 receives the frame
        rays = torch.randn(n_rays, 8, generator=torch.Generator().manual_seed(7))
        lo, hi = shard_bounds(n_rays, rank, world)
        seen = {}

        def render_fn(r):
            seen["n"] = r.shape[0]
            return _stand_in_render(r)

        got = render_sharded(render_fn, rays)
        assert seen["n"] == hi - lo
        if rank == 0:
            ref = _stand_in_render(rays)
            assert sorted(got) == sorted(ref)
            for k in ref:
                assert torch.equal(got[k], ref[k]), k
        else:
            assert got is None
        # light-map sharding with gradients: every rank's loss uses the WHOLE gathered map; the gradient that reaches a
        # rank's slice is the sum of all ranks' gradients on those texels (= the replicated computation)
        n_light = max(n_rays // 3, 1)
        light = torch.randn(n_light, 8, generator=torch.Generator().manual_seed(11))
        w = torch.nn.Parameter(torch.tensor([0.7, -1.3]))
        llo, lhi = shard_bounds(n_light, rank, world)

        def light_fn(r):
            return {"depth_coarse": w[0] * r[:, 6] + w[1] * r[:, 7] ** 2}

        full = render_light_sharded(light_fn, light)["depth_coarse"]
        assert full.shape == (n_light,)
        assert torch.allclose(full.detach(), light_fn(light)["depth_coarse"].detach())
        probe = torch.randn(n_light, generator=torch.Generator().manual_seed(200 + rank))     # rank-specific use of the map
        (full * probe).sum().backward()
        total_probe = sum(torch.randn(n_light, generator=torch.Generator().manual_seed(200 + r)) for r in range(world))
        wr = torch.tensor([0.7, -1.3], requires_grad=True)
        part = (wr[0] * light[llo:lhi, 6] + wr[1] * light[llo:lhi, 7] ** 2)
        (part * total_probe[llo:lhi]).sum().backward()
        assert torch.allclose(w.grad, wr.grad, rtol=1e-5, atol=1e-6)
        assert gather_rows(part.detach(), n_light).shape == (n_light,)
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_rays", [1000, 1001, 1])
def test_two_ranks_gloo(tmp_path, n_rays):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, n_rays, str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}") for r in range(world))


def test_shard_bounds_cover_and_balance():
    for n in (0, 1, 7, 4096, 640000, 640001):
        for world in (1, 2, 3, 8):
            spans = [shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_single_process_paths_need_no_process_group():
    rays = torch.randn(33, 8)
    out = render_sharded(_stand_in_render, rays)
    assert torch.equal(out["rgb_fine"], _stand_in_render(rays)["rgb_fine"])
    from nerf_pl_b200.models.nerf import NeRF
    b = GradBucket([NeRF()])
    assert b.allreduce_mean() is None
