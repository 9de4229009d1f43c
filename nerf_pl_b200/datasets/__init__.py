"""Drop-in for the ray-generation half of the reference's `datasets` package (datasets/ray_utils.py)."""
