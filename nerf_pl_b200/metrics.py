"""Image metrics with the reference's names and arguments (metrics.py:4-20).

`mse` with the default 'mean' reduction and `psnr` run the fused, deterministic reduction kernel (nrf_mse_pair: one
launch, optional boolean pixel mask); any other reduction returns the per-element squared error as plain tensors.
`ssim` wraps kornia in the reference and is not part of the ray-rendering path."""
import torch

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


def ssim(image_pred, image_gt, reduction='mean'):
    raise NotImplementedError("ssim is kornia's (reference metrics.py:15-20); outside the ray-rendering path")
