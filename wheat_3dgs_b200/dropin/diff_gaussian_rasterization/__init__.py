"""Drop-in for the reference's `diff_gaussian_rasterization` package (ashawkey fork: color, radii,
depth, alpha), backed by libwheatsplat.so.  Put `wheat_3dgs_b200/dropin` on sys.path (or call
`wheat_3dgs_b200.install()`) and `gaussian_renderer/__init__.py:14` imports this unchanged."""
from wheat_3dgs_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                        rasterize_gaussians)
from . import _C  # noqa: F401

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]
