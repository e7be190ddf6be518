"""`flashsplat_rasterization._C` (reference flashsplat-rasterization/ext.cpp:15-19)."""
from wheat_3dgs_b200.ops import flash_rasterize_gaussians as rasterize_gaussians  # noqa: F401
from wheat_3dgs_b200.ops import flash_rasterize_gaussians_backward as rasterize_gaussians_backward  # noqa: F401
from wheat_3dgs_b200.ops import mark_visible  # noqa: F401
