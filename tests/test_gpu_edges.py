"""Edge cases and full-size, size-independent properties of render_rays on the GPU (bf16 tensor-core mode unless stated):
empty and ragged batches, coarse-only renders, and - at BASELINE.json's sizes (4096-ray training batch, 32768-ray test
chunk), where the CPU oracle would take minutes - ray-permutation equivariance, chunk/tile-position invariance, output
ranges, determinism under injected randoms, and additivity of the parameter gradients over a split batch."""
import pytest
import torch

from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as ge
    ge.build()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def setup(dev):
    import nerf_pl_b200
    from nerf_pl_b200.models.nerf import Embedding, NeRF
    nerf_pl_b200.set_mlp_precision("bf16")
    models = []
    for s in (0, 1):
        m = NeRF().to(dev)
        m.load_state_dict(O.init_mlp_params(generator=torch.Generator().manual_seed(s)))
        models.append(m)
    return models, [Embedding(3, 10), Embedding(3, 4)]


def _inject(n, nc, nf, dev, seed=2):
    return {k: v.to(dev) for k, v in O.make_noise(n, nc, nf, seed=seed).items()}


def _sel(inj, idx):
    return {k: v[idx].contiguous() for k, v in inj.items()}


@pytest.mark.parametrize("variant", ["rgb", "depth", "rgb_sm"])
@pytest.mark.parametrize("n", [0, 1, 129])
def test_empty_single_and_ragged_batches(dev, setup, variant, n):
    """N = 0 returns empty tensors under the reference's keys (torch ops on empty inputs do the same there); N = 1 and
    N = 129 (one row past a 128-row tile) equal the same rays rendered inside a larger batch, bit for bit."""
    from nerf_pl_b200.models import rendering, rendering_rgb_sm, rendering_shadows
    mod = {"rgb": rendering, "depth": rendering_shadows, "rgb_sm": rendering_rgb_sm}[variant]
    models, emb = setup
    big = O.blender_rays(512, seed=5).to(dev)
    inj = _inject(512, 64, 64, dev)
    with torch.no_grad():
        ref = mod.render_rays(models, emb, big, 64, False, 1.0, 1.0, 64, 1024 * 32, True, _inject=inj)
        idx = torch.arange(200, 200 + n, device=dev)
        got = mod.render_rays(models, emb, big[idx].contiguous(), 64, False, 1.0, 1.0, 64, 1024 * 32, True,
                              _inject=_sel(inj, idx))
    assert sorted(got) == sorted(ref)
    for k in ref:
        assert got[k].shape == (n,) + tuple(ref[k].shape[1:]), k
        assert torch.equal(got[k], ref[k][idx]), k


def test_coarse_only_and_test_time_keys(dev, setup):
    from nerf_pl_b200.models import rendering
    models, emb = setup
    rays = O.blender_rays(300, seed=6).to(dev)
    with torch.no_grad():
        a = rendering.render_rays(models[:1], emb, rays, 64, False, 0, 0, 0, 1024 * 32, True)
        b = rendering.render_rays(models, emb, rays, 64, False, 0, 0, 64, 1024 * 32, True, True)
        c = rendering.render_rays(models, emb, rays, 64, False, 0, 0, 64, 1024 * 32, True)
    assert sorted(a) == ["depth_coarse", "opacity_coarse", "rgb_coarse"]                     # rendering.py:237-247
    assert sorted(b) == ["depth_fine", "opacity_coarse", "opacity_fine", "rgb_fine"]         # test_time: :237-241
    assert torch.equal(a["rgb_coarse"], c["rgb_coarse"]) and torch.equal(a["depth_coarse"], c["depth_coarse"])
    # test_time runs the coarse model sigma-only: same opacity as the full coarse pass
    assert torch.equal(b["opacity_coarse"], c["opacity_coarse"])


@pytest.mark.parametrize("n,test_time", [(4096, False), (32768, True)])
def test_full_size_permutation_chunk_range_determinism(dev, setup, n, test_time):
    from nerf_pl_b200.models import rendering
    models, emb = setup
    rays = O.blender_rays(n, seed=1).to(dev)
    inj = _inject(n, 64, 64, dev)
    perturb, noise_std = (0.0, 0.0) if test_time else (1.0, 1.0)

    def run(r, i, chunk=1024 * 32):
        with torch.no_grad():
            return rendering.render_rays(models, emb, r, 64, False, perturb, noise_std, 64, chunk, True, test_time, _inject=i)

    a = run(rays, inj)
    b = run(rays, inj)
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(3)).to(dev)
    c = run(rays[perm].contiguous(), _sel(inj, perm))
    d = run(rays, inj, chunk=4096)
    near, far = 2.0, 6.0
    for k in a:
        assert torch.isfinite(a[k]).all(), k
        assert torch.equal(a[k], b[k]), f"{k}: two runs differ"
        assert torch.equal(a[k][perm], c[k]), f"{k}: not equivariant under a permutation of the rays"
        assert torch.equal(a[k], d[k]), f"{k}: depends on the chunk argument"
    for s in (("coarse", "fine") if not test_time else ("fine",)):
        op, dp, rgb = a[f"opacity_{s}"], a[f"depth_{s}"], a[f"rgb_{s}"]
        assert float(op.min()) >= 0.0 and float(op.max()) <= 1.0 + 1e-4
        assert bool((dp >= near * op - 1e-3).all()) and bool((dp <= far * op + 1e-3).all())      # depth = sum w z
        assert float(rgb.min()) >= -1e-4 and float(rgb.max()) <= 1.0 + 1e-3                     # sum w c + (1 - opacity)


def test_full_size_gradient_additivity(dev, setup):
    """Parameter gradients of the 4096-ray batch = sum of the gradients of its two halves (the loss is a sum over rays).
    fp32 atomics reorder the sums: relative L2 per tensor <= 2e-4."""
    from nerf_pl_b200.models import rendering
    models, emb = setup
    n = 4096
    rays = O.blender_rays(n, seed=1).to(dev)
    inj = _inject(n, 64, 64, dev)
    tgt = torch.rand(n, 3, generator=torch.Generator().manual_seed(9)).to(dev)
    params = [p for m in models for p in m.parameters()]

    def grads(lo, hi):
        for p in params:
            p.grad = None
        idx = torch.arange(lo, hi, device=dev)
        out = rendering.render_rays(models, emb, rays[lo:hi], 64, False, 1.0, 1.0, 64, 1024 * 32, True, _inject=_sel(inj, idx))
        loss = ((out["rgb_coarse"] - tgt[lo:hi]) ** 2).sum() + ((out["rgb_fine"] - tgt[lo:hi]) ** 2).sum()
        loss.backward()
        return [p.grad.detach().clone() for p in params]

    whole = grads(0, n)
    h1, h2 = grads(0, n // 2), grads(n // 2, n)
    for p in params:
        p.grad = None
    for w, a, b in zip(whole, h1, h2):
        rel = float((w - (a + b)).norm() / (w.norm() + 1e-12))
        assert rel <= 2e-4, rel
