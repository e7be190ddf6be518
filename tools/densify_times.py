"""densify_and_prune at config-2 size (1 M Gaussians, SH degree 3, Adam state present), CUDA-event timing:
ours (one plan + one gather, csrc/ws_densify.cu) against the data movement of the reference's sequence in torch
(cat of the clone rows, cat of the split rows, boolean-mask prune of the split parents, boolean-mask prune --
scene/gaussian_model.py:354-374, :310-352 -- over the 6 parameters and their 12 moment tensors).

    python tools/densify_times.py [P]
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import cases  # noqa: E402
from wheat_3dgs_b200 import densify  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
ATTR = dict(zip(NAMES, ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation")))
dev = torch.device("cuda")
inp = cases.densify_inputs(P=P, seed=4, K=15)


def model():
    m = types.SimpleNamespace()
    groups = []
    for n in NAMES:
        p = torch.nn.Parameter(torch.from_numpy(inp[n]).to(dev).requires_grad_(True))
        setattr(m, ATTR[n], p)
        groups.append({"params": [p], "lr": 1e-3, "name": n})
    m.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    for g in m.optimizer.param_groups:
        p = g["params"][0]
        m.optimizer.state[p] = {"step": torch.tensor(1.0), "exp_avg": torch.randn_like(p) * 1e-3,
                                "exp_avg_sq": torch.rand_like(p) * 1e-6}
    m._which_object = torch.from_numpy(inp["which_object"]).to(dev)
    m.xyz_gradient_accum = torch.from_numpy(inp["xyz_gradient_accum"]).to(dev)
    m.denom = torch.from_numpy(inp["denom"]).to(dev)
    m.max_radii2D = torch.zeros(P, device=dev)
    m.percent_dense = 0.01
    return m


def timed(fn, n=5):
    ms = []
    for _ in range(n + 1):
        m = model()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn(m)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.median(ms[1:])), out


def torch_movement(m):
    flags, _, counts = densify.plan(m, 0.0002, 0.005, 2.0, 20)
    clone = ((flags & 2) != 0) | (((flags & 1) == 0) & ((flags & 8) == 0))   # stand-in masks of the same sizes
    split = (flags & 8) != 0
    arrays = []
    for g in m.optimizer.param_groups:
        p = g["params"][0]
        s = m.optimizer.state[p]
        arrays += [p.detach(), s["exp_avg"], s["exp_avg_sq"]]
    out = []
    for a in arrays:
        a1 = torch.cat((a, a[clone]), 0)
        pad = torch.zeros(a1.shape[0] - split.shape[0], dtype=torch.bool, device=dev)
        s1 = torch.cat((split, pad))
        rep = a1[s1].repeat(2, *([1] * (a.dim() - 1)))
        a2 = torch.cat((a1, rep), 0)
        keep = ~torch.cat((s1, torch.zeros(rep.shape[0], dtype=torch.bool, device=dev)))
        a3 = a2[keep]
        out.append(a3[torch.ones(a3.shape[0], dtype=torch.bool, device=dev)])
    return out[0].shape[0]


if __name__ == "__main__":
    ms, res = timed(lambda m: densify.densify_and_prune(m, 0.0002, 0.005, 2.0, 20, return_rows=True))
    rows = int(res[0].numel())
    words = sum(inp[n][0].size for n in NAMES) * 3 + 1
    gb = (P + rows) * words * 4 / 1e9
    print("ours   densify_and_prune P=%d -> %d rows (counts %s): %.3f ms, %.2f GB moved algorithmically = %.0f GB/s"
          % (P, rows, res[1], ms, gb, gb / ms * 1e3))
    mt, r2 = timed(torch_movement)
    print("torch  cat / mask sequence of the reference over the same 18 tensors: %.3f ms  x%.1f" % (mt, mt / ms))
