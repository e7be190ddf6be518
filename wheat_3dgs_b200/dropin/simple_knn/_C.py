"""`simple_knn._C.distCUDA2` (reference simple-knn/ext.cpp:15-17)."""
from wheat_3dgs_b200.ops import distCUDA2  # noqa: F401
