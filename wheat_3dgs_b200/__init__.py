"""wheat_3dgs_b200 -- B200 (sm_100a) native Gaussian-splatting hot path for Wheat3DGS.

  libwheatsplat.so   hand-written CUDA kernels behind the C ABI of include/wheatsplat.h
  ops                torch-level operators with the reference `_C` signatures
  rasterizer         GaussianRasterizationSettings / GaussianRasterizer (+ flashsplat flavour)
  dropin/            packages named diff_gaussian_rasterization, flashsplat_rasterization, simple_knn
  dist               view-sharded data parallelism (NCCL all-reduce of one gradient bucket)
  scene              seeded synthetic scenes/cameras shaped like the reference workloads

There is no CPU or PyTorch fallback: the ops raise if the CUDA library cannot be loaded.
"""
import os
import sys

DROPIN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dropin")


def install():
    """Makes `import diff_gaussian_rasterization`, `flashsplat_rasterization`, `simple_knn` resolve
    to the drop-in packages (prepends wheat_3dgs_b200/dropin to sys.path)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (DROPIN_DIR, root):
        if p not in sys.path:
            sys.path.insert(0, p)
