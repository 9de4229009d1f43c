"""GPU parity of the rows either side of render_rays: fused ray generation (datasets/ray_utils.py),
fused MSE loss (losses.py) and psnr (metrics.py), through the reference's own call signatures, against the golden
vectors produced by the reference (tests/make_golden_rays.py) and the oracle at full size."""
import os

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rays_loss.npz")


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from nerf_pl_b200 import _lib
    _lib.check(_lib.load().nrf_check_device(), "nrf_check_device")
    return torch.device("cuda:0")


@pytest.mark.parametrize("name", ["blender_small", "fork_bounds", "llff_ndc"])
def test_gen_rays_vs_reference_golden(dev, name):
    """Tolerance 2e-6 relative: the 3x3 rotation is a torch matmul in the reference (summation order / FMA use of the
    CPU GEMM are not specified); everything else is the same sequence of fp32 roundings."""
    from nerf_pl_b200.datasets import ray_utils as RU
    z = np.load(GOLD)
    H, W, focal, near, far, ndc = z[f"{name}/hwf"]
    H, W, ndc = int(H), int(W), bool(ndc)
    c2w = torch.from_numpy(z[f"{name}/c2w"])
    rays = RU.gen_rays(c2w, H, W, float(focal), float(near), float(far), ndc=ndc, device=dev)
    ref = z[f"{name}/rays"]
    np.testing.assert_allclose(rays.cpu().numpy(), ref, rtol=2e-6, atol=2e-6 * np.abs(ref).max())
    # the reference's own call sequence
    d = RU.get_ray_directions(H, W, float(focal), device=dev)
    np.testing.assert_allclose(d.cpu().numpy(), z[f"{name}/directions"], rtol=2e-7, atol=0)   # torch's CUDA division: 1 ulp
    o, dd = RU.get_rays(d, c2w)
    np.testing.assert_allclose(o.cpu().numpy(), z[f"{name}/rays_o"], rtol=0, atol=0)
    np.testing.assert_allclose(dd.cpu().numpy(), z[f"{name}/rays_d"], rtol=2e-6, atol=2e-7)
    if ndc:
        no, nd = RU.get_ndc_rays(H, W, float(focal), float(near), o, dd)
        np.testing.assert_allclose(no.cpu().numpy(), z[f"{name}/ndc_o"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(nd.cpu().numpy(), z[f"{name}/ndc_d"], rtol=1e-5, atol=1e-6)
    # a pixel subset in arbitrary order (a training batch) selects the same rows
    g = torch.Generator().manual_seed(3)
    pix = torch.randperm(H * W, generator=g)[:57].to(dev)
    sub = RU.gen_rays(c2w, H, W, float(focal), float(near), float(far), pix=pix, ndc=ndc)
    assert torch.equal(sub, rays[pix])


def test_gen_rays_full_frame_vs_oracle(dev):
    """800 x 800 (BASELINE config 3 frame size) against the oracle; empty and single-pixel requests."""
    from nerf_pl_b200.datasets import ray_utils as RU
    H = W = 800
    focal = 0.5 * W / np.tan(0.5 * 0.6911)
    c2w = torch.tensor([[0.6, -0.48, 0.64, 2.56], [0.8, 0.36, -0.48, -1.92], [0.0, 0.8, 0.6, 2.4]])
    ref = O.dataset_rays(c2w, H, W, focal, 2.0, 6.0).numpy()
    got = RU.gen_rays(c2w, H, W, focal, 2.0, 6.0, device=dev).cpu().numpy()
    np.testing.assert_allclose(got, ref, rtol=2e-6, atol=1e-6)
    assert RU.gen_rays(c2w, H, W, focal, 2.0, 6.0, n=0, device=dev).shape == (0, 8)
    one = RU.gen_rays(c2w, H, W, focal, 2.0, 6.0, pix=torch.tensor([H * W - 1], device=dev))
    np.testing.assert_allclose(one.cpu().numpy(), ref[-1:], rtol=2e-6, atol=1e-6)
    with pytest.raises(RuntimeError):
        RU.gen_rays(c2w[:, :3], H, W, focal, 2.0, 6.0, device=dev)


@pytest.mark.parametrize("name", ["loss_small", "loss_batch"])
def test_fused_mse_vs_reference_golden(dev, name):
    """MSELoss / SMMSELoss forward + backward and psnr (masked and not): 1e-6 relative (the reference sums in fp32 with
    torch's blocked order, the kernel in fp64 in a fixed order)."""
    from nerf_pl_b200 import losses, metrics
    z = np.load(GOLD)
    a = torch.from_numpy(z[f"{name}/a"]).to(dev).requires_grad_()
    b = torch.from_numpy(z[f"{name}/b"]).to(dev).requires_grad_()
    t = torch.from_numpy(z[f"{name}/t"]).to(dev)
    loss = losses.loss_dict["mse"]()({"rgb_coarse": a, "rgb_fine": b}, t)
    (loss * 1.0).backward()
    np.testing.assert_allclose(loss.item(), z[f"{name}/mse2"], rtol=1e-6)
    np.testing.assert_allclose(a.grad.cpu().numpy(), z[f"{name}/ga"], rtol=1e-6, atol=1e-10)
    np.testing.assert_allclose(b.grad.cpu().numpy(), z[f"{name}/gb"], rtol=1e-6, atol=1e-10)
    np.testing.assert_allclose(losses.MSELoss()({"rgb_coarse": a.detach()}, t).item(), z[f"{name}/mse1"], rtol=1e-6)
    np.testing.assert_allclose(losses.SMMSELoss()({"sm_coarse": a.detach(), "sm_fine": b.detach()}, t).item(),
                               z[f"{name}/sm2"], rtol=1e-6)
    np.testing.assert_allclose(metrics.psnr(a.detach(), t).item(), z[f"{name}/psnr"], rtol=1e-6)
    mask = torch.from_numpy(z[f"{name}/mask"]).to(dev)
    np.testing.assert_allclose(metrics.psnr(a.detach(), t, mask).item(), z[f"{name}/psnr_masked"], rtol=1e-6)
    np.testing.assert_allclose(metrics.mse(a.detach(), t, mask).item(), z[f"{name}/mse_masked"], rtol=1e-6)
    # an upstream factor reaches both gradients
    a2 = a.detach().clone().requires_grad_()
    (3.0 * losses.fused_mse(a2, None, t)).backward()
    np.testing.assert_allclose(a2.grad.cpu().numpy(), 3.0 * 2.0 * (z[f"{name}/a"] - z[f"{name}/t"]) / a.numel(), rtol=2e-6, atol=1e-10)


def test_fused_mse_full_frame_deterministic(dev):
    """640 000 x 3 elements (a full 800 x 800 frame): many CTAs + last-ticket reduction; two runs are bit-identical and
    match the fp64 value; CPU tensors raise."""
    from nerf_pl_b200 import losses, metrics
    g = torch.Generator().manual_seed(5)
    a = torch.rand(640000, 3, generator=g)
    t = torch.rand(640000, 3, generator=g)
    ref = ((a.double() - t.double()) ** 2).mean().item()
    ad, td = a.to(dev), t.to(dev)
    l1, l2 = losses.mse_pair(ad, None, td).item(), losses.mse_pair(ad, None, td).item()
    assert l1 == l2
    np.testing.assert_allclose(l1, ref, rtol=1e-6)
    np.testing.assert_allclose(metrics.psnr(ad, td).item(), -10 * np.log10(ref), rtol=1e-6)
    with pytest.raises(RuntimeError):
        losses.mse_pair(a, None, t)


def test_ssim_vs_oracle(dev):
    """metrics.py:15-20 (1 - 2 * kornia dssim, 3x3 Gaussian window): the device tensor-op version against the oracle's
    explicit float64 window loops, all three reductions; identical images give exactly 1.  kornia 0.2.0 is absent from
    this image, so both restate its published algorithm: parity unpinned."""
    from nerf_pl_b200 import metrics
    g = torch.Generator().manual_seed(9)
    gt = torch.rand(1, 3, 37, 53, generator=g)
    pred = (gt + 0.1 * torch.randn(1, 3, 37, 53, generator=g)).clamp(0, 1)
    for red in ("mean", "sum", "none"):
        ref = O.ssim(pred, gt, red).numpy()
        got = metrics.ssim(pred.to(dev), gt.to(dev), red).cpu().numpy()
        np.testing.assert_allclose(got, ref, rtol=1e-5 if red == "sum" else 0, atol=1e-4 if red != "sum" else 0)   # fp32 E[x^2] - mu^2
    assert metrics.ssim(gt.to(dev), gt.to(dev)).item() == pytest.approx(1.0, abs=1e-6)
    assert -1.0 <= metrics.ssim(pred.to(dev), gt.to(dev)).item() < 1.0
    with pytest.raises(ValueError):
        metrics.ssim(pred[0].to(dev), gt[0].to(dev))
