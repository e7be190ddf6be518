"""Per-stage CUDA-event times (library-recorded) on config 2, a few views: quick A/B harness.
`python tools/stage_times.py [n_views]`"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import ops, scene  # noqa: E402

P, W, H = 1000000, 2048, 1536
nv = int(sys.argv[1]) if len(sys.argv) > 1 else 12
sc = scene.synthetic_scene(P, seed=0)
cams = scene.overhead_cameras(W, H)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
g = {k: t(sc[k]) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
bg = torch.zeros(3, device="cuda")
lg = scene.loss_gradients((3, H, W), seed=2)
dc, dd, da = t(lg["dL_dcolor"]), t(lg["dL_ddepth"]), t(lg["dL_dalpha"])
e = torch.Tensor([])


def step(cam):
    view, proj, campos = t(cam["viewmatrix"]), t(cam["projmatrix"]), t(cam["campos"])
    R, color, depth, alpha, radii, geom, binning, image = ops.rasterize_gaussians(
        bg, g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e, view, proj, cam["tanfovx"],
        cam["tanfovy"], H, W, g["shs"], 3, campos, False, False)
    ops.rasterize_gaussians_backward(bg, g["means3D"], radii, e, g["scales"], g["rotations"], 1.0, e, view, proj,
                                     cam["tanfovx"], cam["tanfovy"], dc, dd, da, g["shs"], 3, campos, geom, R, binning,
                                     image, alpha, False)


for i in range(3):
    step(cams[i])
torch.cuda.synchronize()
ops.prof_enable(True)
ops.prof_read()
for rep in range(2):
    for i in range(nv):
        step(cams[(3 * i) % 36])
torch.cuda.synchronize()
pr = ops.prof_read()
tot = 0.0
for k, (ms, n) in pr.items():
    print("%-16s %8.4f ms" % (k, ms / (2 * nv)))
    tot += ms / (2 * nv)
print("%-16s %8.4f ms" % ("sum", tot))
