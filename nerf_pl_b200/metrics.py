"""Image metrics with the reference's names and arguments (metrics.py:4-20).

`mse` with the default 'mean' reduction and `psnr` run the fused, deterministic reduction kernel (nrf_mse_pair: one
launch, optional boolean pixel mask); any other reduction returns the per-element squared error as plain tensors.
`ssim` is `1 - 2 * kornia.losses.ssim(pred, gt, 3, reduction)` in the reference (metrics.py:15-20); kornia (pinned 0.2.0,
requirements.txt:5) is absent from this image, so its published algorithm is restated here: parity unpinned."""
import torch
import torch.nn.functional as F

from .losses import mse_pair

_DB_PER_DECADE = -10.0


def _squared_error(pred, target, mask):
    err = (pred - target) ** 2
    return err if mask is None else err[mask]


def mse(image_pred, image_gt, valid_mask=None, reduction='mean'):
    if reduction == 'mean':
        return mse_pair(image_pred, None, image_gt, valid_mask)
    return _squared_error(image_pred, image_gt, valid_mask)


def psnr(image_pred, image_gt, valid_mask=None, reduction='mean'):
    """Peak signal-to-noise ratio in dB for images in [0, 1]."""
    return _DB_PER_DECADE * torch.log10(mse(image_pred, image_gt, valid_mask, reduction))


_SSIM_WINDOW, _SSIM_SIGMA = 3, 1.5
_SSIM_C1, _SSIM_C2 = 0.01 ** 2, 0.03 ** 2          # (k * max_val)^2 with max_val = 1


def _gaussian_blur3(t):
    """3x3 normalised Gaussian (sigma 1.5) with zero padding as nine shifted fp32 adds: what kornia's grouped conv2d
    computes, without handing a 3x3 stencil to cuDNN (whose TF32 default would cost three digits)."""
    x = torch.arange(_SSIM_WINDOW, dtype=torch.float64) - _SSIM_WINDOW // 2
    g = torch.exp(-x * x / (2.0 * _SSIM_SIGMA ** 2))
    g = (g / g.sum()).tolist()
    h, w = t.shape[-2:]
    p = F.pad(t, (1, 1, 1, 1))
    out = torch.zeros_like(t)
    for dy in range(_SSIM_WINDOW):
        for dx in range(_SSIM_WINDOW):
            out = out + (g[dy] * g[dx]) * p[..., dy:dy + h, dx:dx + w]
    return out


def ssim(image_pred, image_gt, reduction='mean'):
    """image_pred and image_gt: (1, 3, H, W) in [0, 1].  Structural similarity in [-1, 1] = 1 - 2 * dssim, dssim =
    clamp(1 - ssim_map, 0, 1) / 2 over 3x3 Gaussian windows (sigma 1.5, zero padding), as kornia 0.2.0 computes it."""
    if image_pred.dim() != 4 or image_pred.shape != image_gt.shape:
        raise ValueError("Invalid input shape, we expect BxCxHxW. Got: {} and {}".format(image_pred.shape, image_gt.shape))
    blur = _gaussian_blur3
    mu1, mu2 = blur(image_pred), blur(image_gt)
    var1 = blur(image_pred * image_pred) - mu1 * mu1
    var2 = blur(image_gt * image_gt) - mu2 * mu2
    cov = blur(image_pred * image_gt) - mu1 * mu2
    ssim_map = ((2 * mu1 * mu2 + _SSIM_C1) * (2 * cov + _SSIM_C2)) / ((mu1 * mu1 + mu2 * mu2 + _SSIM_C1) * (var1 + var2 + _SSIM_C2))
    dssim = torch.clamp(1.0 - ssim_map, min=0, max=1) / 2.0
    if reduction == 'mean':
        dssim = dssim.mean()
    elif reduction == 'sum':
        dssim = dssim.sum()
    return 1 - 2 * dssim
