"""Drop-in for models/rendering_shadows.py: depth-only renderer (sigma-only MLP for both passes,
reference :94-281) and efficient_sm (reference :359-482)."""
from ._render import render_rays_impl, sample_pdf  # noqa: F401
from ._shadow import efficient_sm_impl, get_K_impl, shadow_mapping_impl

__all__ = ['render_rays', 'shadow_mapping', 'efficient_sm', 'get_K']


def render_rays(models, embeddings, rays, N_samples=64, use_disp=False, perturb=0, noise_std=1, N_importance=0,
                chunk=1024 * 32, white_back=False, test_time=False, were_gradients_computed=True, _inject=None):
    """Keys depth_*/opacity_*/disp_map_* (reference :250-255, :277-279); `were_gradients_computed`
    is accepted and ignored like in the reference (:105, :261)."""
    return render_rays_impl("depth", models, embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance,
                            chunk, white_back, test_time, _inject)


def efficient_sm(cam_pixels, light_pixels, cam_results, light_results, ppc, light_ppc, image_shape, fine_sampling,
                 Light_N_importance, shadow_method):
    return efficient_sm_impl("rgb", cam_pixels, light_pixels, cam_results, light_results, ppc, light_ppc, image_shape,
                             fine_sampling, Light_N_importance, shadow_method)


def get_K(cam_pixels, cam_results, ppc, light_camera, fine_sampling):
    return get_K_impl(cam_pixels, cam_results, ppc, light_camera, fine_sampling)


def shadow_mapping(cam_results, light_results, rays, ppc, light_ppc, image_shape, batch_size, fine_sampling):
    return shadow_mapping_impl(cam_results, light_results, rays, ppc, light_ppc, image_shape, batch_size, fine_sampling)
