"""nerf_pl_b200 — B200-native (sm_100a) implementation of nerf_pl's ray-rendering hot path."""
import sys

from .ops import get_mlp_precision, set_mlp_precision  # noqa: F401

__all__ = ["set_mlp_precision", "get_mlp_precision", "install_as_models"]


def install_as_models():
    """Alias `nerf_pl_b200.models` as the top-level package `models`, so that the reference's
    `from models.rendering import render_rays` etc. import this implementation unchanged."""
    import importlib
    pkg = importlib.import_module("nerf_pl_b200.models")
    sys.modules["models"] = pkg
    for sub in ("nerf", "rendering", "rendering_shadows", "rendering_rgb_sm", "efficient_shadow_mapping", "shadow_mapping_utils",
                "camera"):
        sys.modules[f"models.{sub}"] = importlib.import_module(f"nerf_pl_b200.models.{sub}")
    ts = importlib.import_module("nerf_pl_b200.torchsearchsorted")
    sys.modules.setdefault("torchsearchsorted", ts)
    # the steps either side of the path: ray generation, losses, metrics
    sys.modules.setdefault("losses", importlib.import_module("nerf_pl_b200.losses"))
    sys.modules.setdefault("metrics", importlib.import_module("nerf_pl_b200.metrics"))
    # (datasets/ray_utils.py: import nerf_pl_b200.datasets.ray_utils from the dataset classes; the reference's `datasets`
    #  package itself -- file formats, kornia -- stays the caller's)
    return pkg
