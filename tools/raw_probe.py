"""Diagnostic: raw-parameter entry vs torch getters + standard entry: bitwise equality of the forward, repeatability
of the backward within an arm and across arms."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import test_gpu_raw as T
from wheat_3dgs_b200 import scene
P, W, H = 200000, 1024, 768
sc = scene.synthetic_scene(P, seed=1)
sc["rotations"][:50] = 0.0
sc["opacities"][50:100] = 1.0 - 1e-7
cam = scene.overhead_cameras(W, H)[9]
lg = scene.loss_gradients((3, H, W), seed=2)
runs = [T._both(sc, cam, 3, lg) for _ in range(3)]
a, b = runs[0]["raw"], runs[0]["getters"]
for k in ("color", "depth", "alpha"):
    print(k, "bitwise equal across arms:", np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), "max diff", np.abs(a[k] - b[k]).max())
for k in a["grads"]:
    same_raw = max(np.abs(runs[0]["raw"]["grads"][k] - r["raw"]["grads"][k]).max() for r in runs[1:])
    same_get = max(np.abs(runs[0]["getters"]["grads"][k] - r["getters"]["grads"][k]).max() for r in runs[1:])
    cross = np.abs(a["grads"][k] - b["grads"][k])
    i = int(cross.reshape(-1).argmax())
    print("%-9s run-to-run raw %.3g getters %.3g | across arms max %.3g at flat %d (%.6g vs %.6g), max|b| %.4g"
          % (k, same_raw, same_get, cross.max(), i, a["grads"][k].reshape(-1)[i], b["grads"][k].reshape(-1)[i], np.abs(b["grads"][k]).max()))
d = np.abs(a["grads"]["xyz"] - b["grads"]["xyz"]).max(axis=1)
idx = np.argsort(-d)[:10]
print("worst xyz rows", idx, d[idx])
print("their scales", sc["scales"][idx].max(axis=1), "opac", sc["opacities"][idx, 0], "radii", a["radii"][idx])
