"""nrf_shadow_fwd_static (include/nerf_b200.h): the shadow test with its scratch in a file-scope `static __device__` array,
for callers that may not allocate.  No counterpart in the reference.

Called through the C-ABI (ctypes, raw device pointers) and compared with nrf_shadow_fwd + a caller workspace on the same
inputs, on two calls in a row that share the array: the second call has other inputs, another mode and FEWER rays, so
anything it picked up from the first call's left-over contents would show.  Both entry points launch the same four
kernels in the same order (the per-segment min / max are order-independent atomics), so the bar is bit equality, the
one the shadow kernels' own tests use against each other.  Collected last on purpose
(file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_RAYS, MAX_TEXELS = 4096, 16384          # NRF_SHADOW_STATIC_MAX_RAYS / _TEXELS


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__ as ge
    ge.build()
    return torch.device("cuda:0")


def _case(dev, n, res, n_views, seed):
    """Camera pixels / depths / per-ray camera matrices and eye positions in `n_views` runs, one light view."""
    g = torch.Generator().manual_seed(seed)
    hw = res * res
    ij = torch.stack(torch.meshgrid(torch.arange(float(res)), torch.arange(float(res)), indexing="xy"), -1).reshape(-1, 2)
    pick = torch.randint(0, hw, (n,), generator=g)
    cam_pixels = torch.cat([ij[pick] + 0.5, torch.ones(n, 1)], 1)
    f = 0.5 * res / 0.2679
    K = torch.tensor([[f, 0, res / 2], [0, f, res / 2], [0, 0, 1.0]])
    views = []
    for v in range(n_views + 1):                       # the last one is the light
        R = torch.linalg.qr(torch.randn(3, 3, generator=g))[0]
        views.append((R @ torch.linalg.inv(K), torch.randn(3, generator=g) * 0.3 + torch.tensor([0.0, 0.0, -4.0])))
    per = -(-n // n_views)
    cam_M = torch.cat([views[v][0].expand(per, 3, 3) for v in range(n_views)])[:n]
    cam_eye = torch.cat([views[v][1].expand(per, 3) for v in range(n_views)])[:n]
    LM, Leye = views[-1]
    t = dict(cam_pixels=cam_pixels, cam_depth=1.0 + 5.0 * torch.rand(n, generator=g), cam_M=cam_M, cam_eye=cam_eye,
             light_Minv=torch.linalg.inv(LM), light_M=LM, light_eye=Leye,
             light_pixels=torch.cat([ij + 0.5, torch.ones(hw, 1)], 1), light_depth=1.0 + 5.0 * torch.rand(hw, generator=g))
    return {k: v.contiguous().float().to(dev) for k, v in t.items()}


def _args(t, n, res, mode, sigmoid, out_bias, out):
    from nerf_pl_b200.ops import _p
    return (_p(t["cam_pixels"]), _p(t["cam_depth"]), _p(t["cam_M"]), _p(t["cam_eye"]), _p(t["light_Minv"]), _p(t["light_M"]),
            _p(t["light_eye"]), _p(t["light_pixels"]), _p(t["light_depth"]), n, res, res, mode, 0.05, 1e-5, 0.0, 1.0,
            int(sigmoid), float(out_bias), _p(out))


def test_static_scratch_matches_workspace_entry_point_twice(dev):
    from nerf_pl_b200 import _lib
    from nerf_pl_b200.ops import _p, _stream
    lib = _lib.load()
    # everything either entry point touches is allocated BEFORE the calls: the static one is handed no workspace
    calls = [dict(n=MAX_RAYS, res=128, views=3, mode=2, sigmoid=1, bias=0.0, seed=5),
             dict(n=1000, res=64, views=1, mode=1, sigmoid=0, bias=0.25, seed=6)]
    prepared = []
    for c in calls:
        t = _case(dev, c["n"], c["res"], c["views"], c["seed"])
        nb = lib.nrf_shadow_workspace_bytes(c["n"], c["res"] * c["res"])
        ws = torch.full((nb,), 0xA5, dtype=torch.uint8, device=dev)
        want = torch.full((c["n"], 3), float("nan"), device=dev)
        got = torch.full((c["n"], 3), float("nan"), device=dev)
        prepared.append((c, t, ws, nb, want, got))
    torch.cuda.synchronize()
    for c, t, ws, nb, want, got in prepared:
        _lib.call("nrf_shadow_fwd", *_args(t, c["n"], c["res"], c["mode"], c["sigmoid"], c["bias"], want), _p(ws), nb, _stream())
        _lib.call("nrf_shadow_fwd_static", *_args(t, c["n"], c["res"], c["mode"], c["sigmoid"], c["bias"], got), _stream())
    torch.cuda.synchronize()
    for c, t, ws, nb, want, got in prepared:
        assert torch.isfinite(want).all() and torch.isfinite(got).all()
        if c["mode"] == 2:                              # min-max normalised per segment: the values must spread
            assert float(want.max() - want.min()) > 0.1, "degenerate case: every ray got the same shadow value"
        assert torch.equal(got, want), f"n={c['n']} mode={c['mode']}: max |diff| {float((got - want).abs().max()):.3e}"


def test_static_scratch_rejects_sizes_beyond_the_array(dev):
    from nerf_pl_b200 import _lib
    from nerf_pl_b200.ops import _stream
    for n, res in ((MAX_RAYS + 1, 64), (16, 129)):
        t = _case(dev, n, res, 1, 7)
        out = torch.zeros(n, 3, device=dev)
        with pytest.raises(RuntimeError, match="exceed the static scratch"):
            _lib.call("nrf_shadow_fwd_static", *_args(t, n, res, 2, 1, 0.0, out), _stream())
        torch.cuda.synchronize()
        assert float(out.abs().max()) == 0.0        # nothing was launched
    # n = 0 is accepted like nrf_shadow_fwd accepts it
    t = _case(dev, 4, 16, 1, 8)
    _lib.call("nrf_shadow_fwd_static", *_args(t, 0, 16, 2, 1, 0.0, torch.zeros(1, 3, device=dev)), _stream())
    torch.cuda.synchronize()
