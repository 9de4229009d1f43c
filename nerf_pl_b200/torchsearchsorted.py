"""Drop-in for the `torchsearchsorted` submodule (src/torchsearchsorted/__init__.py):
`from torchsearchsorted import searchsorted`."""
from .ops import searchsorted  # noqa: F401

__all__ = ["searchsorted"]
