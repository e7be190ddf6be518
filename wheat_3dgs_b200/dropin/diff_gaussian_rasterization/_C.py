"""`diff_gaussian_rasterization._C`: the three functions the reference's ext.cpp:15-19 binds."""
from wheat_3dgs_b200.ops import (mark_visible, rasterize_gaussians,  # noqa: F401
                                 rasterize_gaussians_backward)
