"""Profiling target for ncu: config-2 (1 M Gaussians, 2048x1536, view 17) forward + backward,
`python tools/profile_step.py [n_steps]` (default 2 warm-up + 1 step; pass the total)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import ops, scene  # noqa: E402

P, W, H = 1000000, 2048, 1536
if len(sys.argv) > 2 and sys.argv[2] == "cfg1":
    P, W, H = 50000, 512, 512
n_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
sc = scene.synthetic_scene(P, seed=0)
cam = scene.overhead_cameras(W, H)[17]
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
g = {k: t(sc[k]) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
view, proj, campos = t(cam["viewmatrix"]), t(cam["projmatrix"]), t(cam["campos"])
bg = torch.zeros(3, device="cuda")
lg = scene.loss_gradients((3, H, W), seed=2)
dc, dd, da = t(lg["dL_dcolor"]), t(lg["dL_ddepth"]), t(lg["dL_dalpha"])
e = torch.Tensor([])
for _ in range(n_steps):
    R, color, depth, alpha, radii, geom, binning, image = ops.rasterize_gaussians(
        bg, g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e, view, proj, cam["tanfovx"],
        cam["tanfovy"], H, W, g["shs"], 3, campos, False, False)
    ops.rasterize_gaussians_backward(bg, g["means3D"], radii, e, g["scales"], g["rotations"], 1.0, e, view, proj,
                                     cam["tanfovx"], cam["tanfovy"], dc, dd, da, g["shs"], 3, campos, geom, R, binning,
                                     image, alpha, False)
torch.cuda.synchronize()
print("ok R=%d launches=%d" % (R, ops.launch_count()))
