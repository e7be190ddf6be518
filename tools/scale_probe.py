"""N-GPU diagnostic: where a data-parallel step spends its time.  Per rank: fwd+bwd (CUDA events), exchange
(barrier + kernel + barrier), host time per step; with and without the exchange.
torchrun --nproc-per-node N tools/scale_probe.py"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as tdist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from wheat_3dgs_b200 import dist as wdist, ops, scene  # noqa: E402

rank, world, local = wdist.init_from_env("nccl")
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
sc, cams, (P, W, H, nv) = bench.make_workload("cfg2")
M = 16
g, cd = bench.to_dev(sc, cams, dev)
bg = torch.zeros(3, device=dev)
lg = scene.loss_gradients((3, H, W), seed=2)
dc, dd, da = [torch.from_numpy(lg[k]).to(dev) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
e = torch.Tensor([])
bucket, reduce_bucket, how = wdist.make_bucket(wdist.full_bucket_numel(P, M), dev)
n_train = wdist.train_bucket_numel(P, M)
costs = [ops.count_instances(bg, g["means3D"], g["opacities"], g["scales"], g["rotations"], c["viewmatrix"], c["projmatrix"],
                             c["tanfovx"], c["tanfovy"], H, W, g["shs"], 3, c["campos"]) for c in cd]
sched = wdist.balanced_schedule(costs, world, 400)


def step(i, reduce, ev):
    c = cd[sched[i % 400][rank]]
    ev[0].record()
    fw = ops.rasterize_gaussians(bg, g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e, c["viewmatrix"],
                                 c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, g["shs"], 3, c["campos"], False, False)
    ops.rasterize_gaussians_backward(bg, g["means3D"], fw[4], e, g["scales"], g["rotations"], 1.0, e, c["viewmatrix"],
                                     c["projmatrix"], c["tanfovx"], c["tanfovy"], dc, dd, da, g["shs"], 3, c["campos"], fw[5],
                                     fw[0], fw[6], fw[7], fw[3], False, _bucket=bucket)
    ev[1].record()
    if reduce:
        reduce_bucket(n_train)
    ev[2].record()


for reduce in (False, True):
    n = 60
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(n)]
    for i in range(5):
        step(i, reduce, evs[0])
    torch.cuda.synchronize(); tdist.barrier(); torch.cuda.synchronize()
    t0 = time.time()
    for i in range(n):
        step(5 + i, reduce, evs[i])
    torch.cuda.synchronize()
    wall = (time.time() - t0) / n * 1e3
    comp = np.mean([a.elapsed_time(b) for a, b, c in evs])
    red = np.mean([b.elapsed_time(c) for a, b, c in evs])
    gap = np.mean([evs[i][2].elapsed_time(evs[i + 1][0]) for i in range(n - 1)])
    out = [None] * world
    tdist.all_gather_object(out, (rank, round(comp, 3), round(red, 3), round(gap, 3), round(wall, 3)))
    if rank == 0:
        print("exchange=%s (%s): per rank (fwd+bwd ms, exchange ms, gap between steps ms, wall ms/step)" % (reduce, how))
        for o in out:
            print("   ", o)
tdist.destroy_process_group()
