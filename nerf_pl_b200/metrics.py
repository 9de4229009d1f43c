"""Drop-in for metrics.py (reference :4-13): mse / psnr through the fused reduction kernel.  ssim (reference :15-20)
wraps kornia, which this tier does not reimplement."""
import torch

from .losses import mse_pair


def mse(image_pred, image_gt, valid_mask=None, reduction='mean'):
    if reduction != 'mean':
        value = (image_pred - image_gt) ** 2
        return value[valid_mask] if valid_mask is not None else value
    return mse_pair(image_pred, None, image_gt, valid_mask)


def psnr(image_pred, image_gt, valid_mask=None, reduction='mean'):
    return -10 * torch.log10(mse(image_pred, image_gt, valid_mask, reduction))


def ssim(image_pred, image_gt, reduction='mean'):
    raise NotImplementedError("ssim is kornia's (reference metrics.py:15-20); outside the ray-rendering path")
