"""Drop-in mirror of the reference's `models` package for the hot path.

Put the directory that contains this package's parent (`nerf_pl_b200/`) first on sys.path —
or use `nerf_pl_b200.install_as_models()` — and the reference's trainers resolve
`from models.rendering import render_rays`, `from models.nerf import Embedding, NeRF`, ... here."""
