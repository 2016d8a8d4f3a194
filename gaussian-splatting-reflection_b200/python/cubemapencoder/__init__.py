"""Drop-in `cubemapencoder` package (reference: C/cubemapencoder/__init__.py)."""
from .cubemap_encoder import CubemapEncoder, MipCubemapEncoder, cubemap_encode  # noqa: F401
