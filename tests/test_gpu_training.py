"""The whole path learns: 400 graphed training steps (render_rays fwd + fused MSE + bwd into the flat bucket + fused Adam +
operand re-pack, bf16 tensor-core MLP, fresh jitter and noise every step) on a synthetic scene - a solid sphere of
radius 1 whose colour is a smooth function of position, seen from random cameras on a sphere of radius 4 - must cut the
loss several-fold, keep every parameter finite, and leave a model whose test-time render of held-out rays is closer to
the target than the initial model's.  Also a soak of the barrier protocols: ~5000 kernel launches without a hang."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _scene_rays(n, gen):
    """Random cameras on the radius-4 sphere looking near the origin; analytic target: ray/sphere intersection."""
    o = torch.randn(n, 3, generator=gen)
    o = 4.0 * o / o.norm(dim=1, keepdim=True)
    d = torch.nn.functional.normalize(-o + 0.35 * torch.randn(n, 3, generator=gen), dim=1)
    b = (o * d).sum(1)
    disc = b * b - ((o * o).sum(1) - 1.0)
    hit = disc > 0
    t = -b - torch.sqrt(disc.clamp_min(0))
    p = o + d * t[:, None]
    colour = 0.5 + 0.5 * torch.stack([torch.sin(2 * p[:, 0]), torch.cos(2 * p[:, 1]), torch.sin(2 * p[:, 2] + 1)], 1)
    rgb = torch.where(hit[:, None], colour, torch.ones(n, 3))            # white background
    rays = torch.cat([o, d, torch.full((n, 1), 2.0), torch.full((n, 1), 6.0)], 1)
    return rays.contiguous(), rgb.contiguous()


def test_graphed_training_learns_a_scene():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as ge
    ge.build()
    import nerf_pl_b200
    from nerf_pl_b200.graph import GraphedStep
    from nerf_pl_b200.losses import loss_dict
    from nerf_pl_b200.metrics import psnr
    from nerf_pl_b200.models.nerf import Embedding, NeRF
    from nerf_pl_b200.models.rendering import render_rays
    from nerf_pl_b200.optim import FusedAdam
    dev = torch.device("cuda:0")
    nerf_pl_b200.set_mlp_precision("bf16")
    torch.manual_seed(0)
    models = [NeRF().to(dev), NeRF().to(dev)]
    emb = [Embedding(3, 10), Embedding(3, 4)]
    opt = FusedAdam(models, lr=5e-4, capturable=True)
    mse = loss_dict["mse"]()
    gen = torch.Generator().manual_seed(1)
    B, steps = 2048, 400
    pool_rays, pool_rgb = _scene_rays(B * 16, gen)
    test_rays, test_rgb = (t.to(dev) for t in _scene_rays(4096, gen))

    def evaluate():
        with torch.no_grad():
            out = render_rays(models, emb, test_rays, 64, False, 0, 0, 64, 1024 * 32, True, True)
        return float(psnr(out["rgb_fine"], test_rgb))

    psnr0 = evaluate()

    def step(rays, rgbs):
        opt.zero_grad()
        out = render_rays(models, emb, rays, 64, False, 1.0, 1.0, 64, 1024 * 32, True)
        loss = mse(out, rgbs)
        loss.backward()
        opt.step()
        return loss

    hr = [pool_rays[i * B:(i + 1) * B].pin_memory() for i in range(16)]
    hc = [pool_rgb[i * B:(i + 1) * B].pin_memory() for i in range(16)]
    g = GraphedStep(step, (hr[0].to(dev), hc[0].to(dev)), warmup=2, optimizers=[opt])
    losses, pending = [], None
    for i in range(steps):
        nxt = g.submit(hr[i % 16], hc[i % 16])
        if pending is not None:
            losses.append(float(pending.result()))
        pending = nxt
    losses.append(float(pending.result()))
    g.close()
    assert all(math.isfinite(v) for v in losses)
    first, last = sum(losses[:10]) / 10, sum(losses[-10:]) / 10
    print(f"loss {first:.4f} -> {last:.4f}; test psnr {psnr0:.2f} dB", end="")
    assert last < 0.4 * first, (first, last)
    assert all(bool(torch.isfinite(p).all()) for m in models for p in m.parameters())
    psnr1 = evaluate()
    print(f" -> {psnr1:.2f} dB")
    assert psnr1 > psnr0 + 3.0, (psnr0, psnr1)
