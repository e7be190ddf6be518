import os, sys, time
import numpy as np, torch
sys.path.insert(0, "/root/repo")
from wheat_3dgs_b200 import ops, scene
import bench
sc, cams, (P, W, H, nv) = bench.make_workload("cfg2")
dev = torch.device("cuda", 0)
g, cd = bench.to_dev(sc, cams, dev)
bg = torch.zeros(3, device=dev)
lg = scene.loss_gradients((3, H, W), seed=2)
dc, dd, da = [torch.from_numpy(lg[k]).to(dev) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
e = torch.Tensor([])
def step(i):
    c = cd[i % nv]
    fw = ops.rasterize_gaussians(bg, g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, g["shs"], 3, c["campos"], False, False)
    R, color, depth, alpha, radii, geom, binning, image = fw
    ops.rasterize_gaussians_backward(bg, g["means3D"], radii, e, g["scales"], g["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], dc, dd, da, g["shs"], 3, c["campos"], geom, R, binning, image, alpha, False)
    return fw
for rep in range(3):
    for i in range(5): step(i)
    torch.cuda.synchronize()
    t0 = time.time()
    ts = []
    for i in range(36):
        t = time.time(); step(5 + i); ts.append(time.time() - t)
    torch.cuda.synchronize()
    print("rep", rep, "ms/step", (time.time() - t0) / 36 * 1e3, "max host step ms", max(ts) * 1e3, "median", sorted(ts)[18] * 1e3, flush=True)
    print(torch.cuda.memory_stats()["num_alloc_retries"], torch.cuda.memory_reserved() / 1e9)
