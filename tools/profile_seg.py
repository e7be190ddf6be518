"""Profiling target for the segmentation kernels: one config-3 vote view (500 labels, 2048x1536) with
multi_instance_opt, and one config-4-sized batched eval_obj_labels frame.  `python tools/profile_seg.py`"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import ops, scene, segment  # noqa: E402
from wheat_3dgs_b200.rasterizer import GaussianRasterizationSettings  # noqa: E402

t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
P, nobj = 1000000, 500
sc = scene.synthetic_scene(P, seed=0)
g = {k: t(sc[k]) for k in ("means3D", "opacities", "shs", "scales", "rotations")}


def settings(cam, bg):
    return GaussianRasterizationSettings(
        image_height=cam["image_height"], image_width=cam["image_width"], tanfovx=cam["tanfovx"],
        tanfovy=cam["tanfovy"], bg=bg, scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]),
        projmatrix=t(cam["projmatrix"]), sh_degree=3, campos=t(cam["campos"]), prefiltered=False, debug=False)


cam = scene.overhead_cameras(2048, 1536)[17]
acc = segment.VoteAccumulator(nobj, P)
mask = t(scene.head_masks(sc, cam, n_labels=nobj))
for _ in range(2):
    acc.add_view(mask, settings(cam, torch.zeros(3, device="cuda")), g["means3D"], g["opacities"], shs=g["shs"],
                 scales=g["scales"], rotations=g["rotations"])
labels = acc.labels()
torch.cuda.synchronize()
sets = segment.object_sets(labels)
cam4 = scene.orbit_cameras(1920, 1080)[25]
for _ in range(2):
    pm = segment.eval_obj_labels_tensors(sets, settings(cam4, torch.ones(3, device="cuda")), g["means3D"], g["opacities"],
                                         shs=g["shs"], scales=g["scales"], rotations=g["rotations"])
torch.cuda.synchronize()
print("ok", segment.eval_obj_labels_tensors.last_counts, int((pm > 0).sum()))
