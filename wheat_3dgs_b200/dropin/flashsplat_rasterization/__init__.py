"""Drop-in for the reference's `flashsplat_rasterization` package (label-contribution rasterizer
used by run_3d_seg.py / render_360.py), backed by libwheatsplat.so.
`gaussian_renderer/__init__.py:18-19` imports this unchanged."""
from wheat_3dgs_b200.rasterizer import FlashGaussianRasterizationSettings as GaussianRasterizationSettings  # noqa: F401
from wheat_3dgs_b200.rasterizer import FlashGaussianRasterizer as GaussianRasterizer  # noqa: F401
from wheat_3dgs_b200.rasterizer import flash_rasterize_gaussians as rasterize_gaussians  # noqa: F401
from . import _C  # noqa: F401

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]
