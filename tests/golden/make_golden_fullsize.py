"""Full-size goldens from the UNMODIFIED reference CUDA rasterizers (oracle/_ref, built from /root/reference by
oracle/build_ref.py), run on a B200:

    gpurun -- python tests/golden/make_golden_fullsize.py        # writes gpurun_out/fullsize.{json,npz}
    cp gpurun_out/fullsize.json gpurun_out/fullsize.npz tests/golden/

For one named view / frame of BASELINE.json's configs 1, 2, 4, 5: sha256 of the reference's radii, point_list,
ranges and n_contrib (the bit-exact tier), 8192 evenly spaced samples of colour / depth / alpha and of every
gradient tensor (tolerance tiers).  For config 3: the row sums of the label matrix the reference's own
multi_instance_opt (torch, run_3d_seg.py:54-72) gives on the reference flashsplat rasterizer's votes summed over six
views, and how many entries sit within the reference's own run-to-run spread of a tie.  For the training iteration:
the losses of the first three iterations of the reference stack.  bench.py's parity gates and
tests/test_gpu_fullsize.py compare against these; nothing here is needed at bench time.
"""
import hashlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bench  # noqa: E402
import ref as refmod  # noqa: E402
from wheat_3dgs_b200 import scene  # noqa: E402

dev = torch.device("cuda")
VIEWS = {"cfg1": 0, "cfg2": 17, "cfg5": 3, "cfg4": 25}
CFG3_VIEWS = [0, 6, 12, 18, 24, 30]


def sha_np(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def raster_golden(name, samples, with_grads):
    sc, cams, (P, W, H, nv) = bench.make_workload(name)
    g, cd = bench.to_dev(sc, cams, dev)
    bg = torch.ones(3, device=dev) if bench.WORKLOADS[name]["kind"] == "orbit" else torch.from_numpy(bench.BG).to(dev)
    v = VIEWS[name]
    rr = refmod.RefRasterizer("dgr")
    o = rr.forward(g, cd[v], bg, bench.SH_DEGREE)
    torch.cuda.synchronize()
    ent = {"view": v, "num_rendered": int(o["num_rendered"]), "radii": bench.sha(o["radii"]),
           "point_list": sha_np(rr.fetch("point_list")), "ranges": sha_np(rr.fetch("ranges")),
           "n_contrib": sha_np(rr.fetch("n_contrib"))}
    for k in ("color", "depth", "alpha"):
        samples["%s/%s" % (name, k)] = bench.sample(o[k])
    if with_grads:
        lg = scene.loss_gradients((3, H, W), seed=2)
        dc, dd, da = [torch.from_numpy(lg[k]).to(dev) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
        gr = rr.backward(dc, dd, da)
        torch.cuda.synchronize()
        for k in bench.GRAD_KEYS:
            samples["%s/%s" % (name, k)] = bench.sample(gr[k])
    del rr, o, g
    torch.cuda.empty_cache()
    return ent


def votes_golden(name):
    sc, cams, (P, W, H, nv) = bench.make_workload(name)
    nobj = bench.WORKLOADS[name]["labels"]
    g, cd = bench.to_dev(sc, cams, dev)
    bg = torch.from_numpy(bench.BG).to(dev)
    views = [v for v in CFG3_VIEWS if v < nv] or [0]
    masks = {v: torch.from_numpy(m).to(dev) for v, m in bench.make_masks(sc, cams, nobj, views).items()}
    rr = refmod.RefRasterizer("fsr")

    def run():
        allc = torch.zeros((nobj + 1, P), device=dev)
        for v in views:
            o = rr.forward(g, cd[v], bg, bench.SH_DEGREE, gt_mask=masks[v], num_obj=nobj, outputs={})
            allc.add_(o["used_count"])
        return allc

    a1, a2 = run(), run()
    spread = float((a1 - a2).abs().max())
    spread_g = (a1 - a2).abs().sum(dim=0)        # per Gaussian: how far its votes move between two reference runs
    s = a1.sum(dim=0)
    lab = torch.zeros_like(a1).bool()
    for k in range(a1.shape[0]):   # run_3d_seg.py:54-72
        oc = torch.stack([s - a1[k], a1[k]], dim=0)
        lab[k] = torch.argmax(torch.nn.functional.normalize(oc, dim=0), dim=0)
    near = int(((2 * a1 - s[None]).abs() <= 4 * spread_g[None] + 1e-6 * s[None]).logical_and(s[None] > 0).sum())
    return {"views": views, "label_row_sums": lab.sum(dim=1).cpu().tolist(), "near_ties": near,
            "votes_run_to_run_spread": spread, "labelled": int(lab[1:].any(dim=0).sum())}


def train_golden(name, n=3):
    torch_, dev_, iteration, cams, cd, gts_host, gts, (P, W, H, nv) = bench.reference_trainer(0, name)
    losses = [float(iteration(cd[i % nv], gts[i % len(gts)])) for i in range(n)]
    torch.cuda.empty_cache()
    return {"losses": losses}


def main():
    which = sys.argv[1:] or ["cfg1", "cfg2", "cfg4", "cfg5", "cfg3", "train", "tiny3", "tinytrain"]
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    jpath, npath = os.path.join(out_dir, "fullsize.json"), os.path.join(out_dir, "fullsize.npz")
    # start from the committed goldens so that a partial run refreshes only what it names
    try:
        meta = json.load(open(os.path.join(ROOT, "tests", "golden", "fullsize.json")))
        samples = dict(np.load(os.path.join(ROOT, "tests", "golden", "fullsize.npz")))
    except (OSError, ValueError):
        meta, samples = {}, {}
    for name in which:
        kind = bench.WORKLOADS[name]["kind"]
        if kind in ("fwdbwd", "orbit"):
            for k in [k for k in samples if k.startswith(name + "/")]:
                del samples[k]
            meta[name] = raster_golden(name, samples, with_grads=(kind == "fwdbwd"))
        elif kind == "votes":
            meta[name] = votes_golden(name)
        else:
            meta[name] = train_golden(name)
        print(name, json.dumps({k: v for k, v in meta[name].items() if k != "label_row_sums"}), flush=True)
    meta["_made_with"] = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__,
                          "reference": "oracle/_ref/libref_{dgr,fsr}.so built from /root/reference (oracle/build_ref.py)"}
    json.dump(meta, open(jpath, "w"), indent=1)
    np.savez_compressed(npath, **samples)
    print("wrote", jpath, npath)


if __name__ == "__main__":
    main()
