"""Drop-in for models/rendering.py (classic rgb + depth renderer, reference :84-272)."""
from ._render import render_rays_impl, sample_pdf  # noqa: F401

__all__ = ['render_rays']


def render_rays(models, embeddings, rays, N_samples=64, use_disp=False, perturb=0, noise_std=1, N_importance=0,
                chunk=1024 * 32, white_back=False, test_time=False, _inject=None):
    """Returns {'rgb_coarse','depth_coarse','opacity_coarse'[,'rgb_fine','depth_fine','opacity_fine']};
    with test_time=True only 'opacity_coarse' + the fine keys (reference :237-241)."""
    return render_rays_impl("rgb", models, embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance,
                            chunk, white_back, test_time, _inject)
