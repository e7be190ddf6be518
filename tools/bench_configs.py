"""BASELINE.json configs 3, 4 and 5 on one B200, this library against the unmodified reference CUDA rasterizers
(oracle/_ref), CUDA-event timing, JSON lines to stdout (committed under profiles/).  bench.py covers config 2.

  cfg3   flashsplat label votes, multi-label form (render_360.py:62-69 / eval_wheatgs.py:106-114): 36 views
         2048x1536 x one mask of ~500 head ids each, votes for 501 rows x 1 M Gaussians accumulated over views
         (reference: fresh 2 GB used_count per view + all_counts += ...; ours: one persistent buffer)
  cfg4   render_360 orbit, 1920x1080, forward only with depth + alpha: MPix/s (every 5th of the 200 frames)
  cfg5   scale stress: 6 M Gaussians, 3840x2160, forward + backward (4 of the 144 views)

    python tools/bench_configs.py cfg3 cfg4 cfg5
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref as refmod  # noqa: E402
from wheat_3dgs_b200 import ops, scene  # noqa: E402

dev = torch.device("cuda")
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
E = torch.Tensor([])


def to_dev(sc, cams):
    g = {k: t(sc[k]) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
    cd = [dict(c, viewmatrix=t(c["viewmatrix"]), projmatrix=t(c["projmatrix"]), campos=t(c["campos"])) for c in cams]
    return g, cd


def timed(fn, n, warm=2):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def ours_fwd(g, c, bg, H, W):
    return ops.rasterize_gaussians(bg, g["means3D"], E, g["opacities"], g["scales"], g["rotations"], 1.0, E,
                                   c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, g["shs"], 3,
                                   c["campos"], False, False)


def cfg4():
    P, W, H = 1000000, 1920, 1080
    sc = scene.synthetic_scene(P, seed=0)
    cams = scene.orbit_cameras(W, H, n_frames=200)[::5]
    g, cd = to_dev(sc, cams)
    bg = torch.ones(3, device=dev)   # render_360.py:146: white
    rr = refmod.RefRasterizer("dgr")
    outs = {}
    R = [0]

    def mine(i):
        R[0] += ours_fwd(g, cd[i % len(cd)], bg, H, W)[0]

    ms = timed(mine, len(cd))
    msr = timed(lambda i: rr.forward(g, cd[i % len(cd)], bg, 3, outputs=outs), len(cd))
    print(json.dumps({"config": "cfg4 render_360 orbit 1920x1080, 1M Gaussians SH3, forward only (color+depth+alpha)",
                      "frames_timed": len(cd), "ours_ms_per_frame": ms, "ours_mpix_s": W * H / ms / 1e3,
                      "reference_cuda_ms_per_frame": msr, "reference_cuda_mpix_s": W * H / msr / 1e3,
                      "speedup": msr / ms, "mean_instances": R[0] / (len(cd) + 2)}))


def cfg3():
    P, W, H, nobj = 1000000, 2048, 1536, 500
    sc = scene.synthetic_scene(P, seed=0)
    cams = scene.overhead_cameras(W, H)
    g, cd = to_dev(sc, cams)
    bg = torch.zeros(3, device=dev)
    views = list(range(0, 36, 3))
    masks = {v: t(scene.head_masks(sc, cams[v], n_labels=nobj)) for v in views}
    counts = torch.zeros((nobj + 1, P), device=dev)

    def mine(i):
        v = views[i % len(views)]
        c = cd[v]
        ops.flash_rasterize_gaussians(masks[v], None, bg, g["means3D"], E, g["opacities"], g["scales"], g["rotations"],
                                      1.0, E, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, g["shs"],
                                      3, c["campos"], False, nobj, False, used_count_out=counts)

    ms = timed(mine, len(views))
    labels = ops.multi_instance_opt(counts)
    torch.cuda.synchronize()
    ml = timed(lambda i: ops.multi_instance_opt(counts), 5)
    rr = refmod.RefRasterizer("fsr")
    all_counts = torch.zeros((nobj + 1, P), device=dev)

    def theirs(i):
        v = views[i % len(views)]
        o = rr.forward(g, cd[v], bg, 3, gt_mask=masks[v], num_obj=nobj, outputs={})   # fresh, zero-filled used_count
        all_counts.add_(o["used_count"])                                             # render_360.py:175 (on the GPU)

    msr = timed(theirs, len(views))

    def ref_labels(i):  # run_3d_seg.py:54-72, torch on the GPU
        s = all_counts.sum(dim=0)
        out = torch.zeros_like(all_counts).bool()
        for k in range(all_counts.shape[0]):
            oc = torch.stack([s - all_counts[k], all_counts[k]], dim=0)
            out[k] = torch.argmax(torch.nn.functional.normalize(oc, dim=0), dim=0)
        return out

    mlr = timed(ref_labels, 2, warm=1)
    print(json.dumps({"config": "cfg3 flashsplat label votes, 36 views 2048x1536 x ~500 head masks (multi-label form), "
                                "1M Gaussians, votes summed over views + multi_instance_opt",
                      "views_timed": len(views), "ours_ms_per_view": ms, "reference_cuda_ms_per_view": msr,
                      "speedup_votes": msr / ms, "ours_multi_instance_opt_ms": ml,
                      "reference_torch_multi_instance_opt_ms": mlr, "speedup_labels": mlr / ml,
                      "labelled_gaussians": int(labels[1:].any(dim=0).sum())}))


def cfg5():
    P, W, H = 6000000, 3840, 2160
    sc = scene.synthetic_scene(P, seed=0)
    cams = scene.overhead_cameras(W, H)
    g, cd = to_dev(sc, cams)
    bg = torch.zeros(3, device=dev)
    lg = scene.loss_gradients((3, H, W), seed=2)
    dc, dd, da = [t(lg[k]) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
    views = [3, 14, 17, 30]
    R = [0]

    def mine(i):
        c = cd[views[i % 4]]
        fw = ours_fwd(g, c, bg, H, W)
        R[0] += fw[0]
        ops.rasterize_gaussians_backward(bg, g["means3D"], fw[4], E, g["scales"], g["rotations"], 1.0, E, c["viewmatrix"],
                                         c["projmatrix"], c["tanfovx"], c["tanfovy"], dc, dd, da, g["shs"], 3, c["campos"],
                                         fw[5], fw[0], fw[6], fw[7], fw[3], False)

    ms = timed(mine, 8)
    rr = refmod.RefRasterizer("dgr")
    outs, grads = {}, {}

    def theirs(i):
        rr.forward(g, cd[views[i % 4]], bg, 3, outputs=outs)
        rr.backward(dc, dd, da, grads=grads)

    msr = timed(theirs, 8)
    print(json.dumps({"config": "cfg5 scale stress: 6M Gaussians SH3, 3840x2160, forward+backward per view",
                      "views_timed": 8, "ours_ms_per_view": ms, "reference_cuda_ms_per_view": msr, "speedup": msr / ms,
                      "mean_instances": R[0] / 10, "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}))


if __name__ == "__main__":
    for name in sys.argv[1:] or ["cfg4", "cfg3", "cfg5"]:
        {"cfg3": cfg3, "cfg4": cfg4, "cfg5": cfg5}[name]()
        torch.cuda.empty_cache()
