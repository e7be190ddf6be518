"""Fused Adam step + densification statistics (ws_optim.cu) vs torch.optim.Adam (CUDA multi-tensor path) and
the reference's masked-index statistics at 1 M Gaussians (59 M parameters): CUDA-event ms per step."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import optim as woptim  # noqa: E402

P = 1000000
SHAPES = [("xyz", (3,), 2e-4), ("f_dc", (1, 3), 0.0025), ("f_rest", (15, 3), 0.000125), ("opacity", (1,), 0.05),
          ("scaling", (3,), 0.005), ("rotation", (4,), 0.001)]


def make(cls):
    ps = [torch.nn.Parameter(torch.randn((P,) + s, device="cuda")) for _, s, _ in SHAPES]
    for p in ps:
        p.grad = torch.randn_like(p) * 1e-3
    return ps, cls([{"params": [p], "lr": lr, "name": n} for p, (n, _, lr) in zip(ps, SHAPES)], lr=0.0, eps=1e-15)


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


_, a = make(woptim.Adam)
_, b = make(torch.optim.Adam)
ta, tb = timeit(a.step), timeit(b.step)
numel = 59 * P
print("fused adam %.4f ms (%.0f GB/s algorithmic, 28 B/element)   torch.optim.Adam %.4f ms   x%.1f"
      % (ta, numel * 28 / ta / 1e6, tb, tb / ta))

radii = (torch.randint(0, 60, (P,), device="cuda") * (torch.rand(P, device="cuda") < 0.8)).int()
g2d = torch.randn(P, 3, device="cuda") * 1e-4
mr, acc, dn = torch.zeros(P, device="cuda"), torch.zeros(P, 1, device="cuda"), torch.zeros(P, 1, device="cuda")


def ref_stats():
    vis = radii > 0                                                  # gaussian_renderer/__init__.py:103
    mr[vis] = torch.max(mr[vis], radii[vis])                         # train_vanilla_3dgs.py:101
    acc[vis] += torch.norm(g2d[vis, :2], dim=-1, keepdim=True)       # scene/gaussian_model.py:463
    dn[vis] += 1


ts = timeit(lambda: woptim.add_densification_stats(radii, g2d, mr, acc, dn))
tr = timeit(ref_stats)
print("fused densify stats %.4f ms   torch masked-index form %.4f ms   x%.1f" % (ts, tr, tr / ts))
