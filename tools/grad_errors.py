"""Measured gradient errors at config 2 (1 M Gaussians, 2048x1536, view 17) against the unmodified reference CUDA
rasterizer run live beside libwheatsplat (oracle/_ref/libref_dgr.so): per gradient tensor the largest element-wise
error relative to the element, relative to the tensor's largest element, the norm-wise relative error, and the same
three figures for the reference against ITSELF (two runs: its float atomics are order-dependent).  The tier the tests
assert (tests/util.py) is a bound; this is the number.  `python tools/grad_errors.py [out.json]`"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpu_util as G  # noqa: E402
import ref as refmod  # noqa: E402
from util import GRAD_NAMES  # noqa: E402
from wheat_3dgs_b200 import scene  # noqa: E402

P, W, H = 1000000, 2048, 1536
sc, cam = scene.synthetic_scene(P, seed=0), scene.overhead_cameras(W, H)[17]
gd, cd = G.dev_inputs(sc, cam)
bg = torch.zeros(3, device="cuda")
lg = scene.loss_gradients((3, H, W), seed=2)
dc, dd, da = G.t(lg["dL_dcolor"]), G.t(lg["dL_ddepth"]), G.t(lg["dL_dalpha"])
rr = refmod.RefRasterizer("dgr")
rr.forward(gd, cd, bg, 3)
r1 = {k: v.double().cpu().numpy().reshape(-1) for k, v in rr.backward(dc, dd, da).items()}
r2 = {k: v.double().cpu().numpy().reshape(-1) for k, v in rr.backward(dc, dd, da).items()}
fw = G.forward(gd, cd, bg, 3)
mine = {k: m.double().cpu().numpy().reshape(-1) for k, m in zip(GRAD_NAMES, G.backward(gd, cd, bg, 3, fw, dc, dd, da))}
mine2 = {k: m.double().cpu().numpy().reshape(-1) for k, m in zip(GRAD_NAMES, G.backward(gd, cd, bg, 3, fw, dc, dd, da))}


def figures(a, b):
    d = np.abs(a - b)
    big = np.abs(b) > 1e-3 * np.abs(b).max() if b.size else np.zeros(0, bool)
    return {"max_abs": float(d.max()), "max_abs_over_max_ref": float(d.max() / max(np.abs(b).max(), 1e-300)),
            "max_rel_on_elements_above_1e-3_of_max": float((d[big] / np.abs(b[big])).max()) if big.any() else 0.0,
            "normwise_rel": float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))}


out = {"config": "cfg2 view 17, P=1000000, 2048x1536, SH degree 3; loss gradients seed 2", "tensors": {}}
for k in GRAD_NAMES:
    if not np.any(r1[k]):
        continue
    out["tensors"][k] = {"ours_vs_reference": figures(mine[k], r1[k]), "reference_vs_itself": figures(r2[k], r1[k]),
                         "ours_vs_itself": figures(mine2[k], mine[k]), "max_abs_ref": float(np.abs(r1[k]).max())}
    print(k, json.dumps(out["tensors"][k]))
if len(sys.argv) > 1:
    json.dump(out, open(sys.argv[1], "w"), indent=1)
