"""Generates tests/golden/*.npz by running the UNMODIFIED reference CUDA rasterizers
(oracle/_ref/libref_{dgr,fsr,knn}.so, built from /root/reference by oracle/build_ref.py) on a B200:

    gpurun -- python tests/golden/make_golden.py        # writes gpurun_out/golden/*.npz
    cp gpurun_out/golden/*.npz tests/golden/

Inputs are the seeded cases of tests/golden/cases.py (a sha256 of the inputs is stored so a test
can tell "inputs regenerated differently" from "outputs differ").  Versions are recorded in
gpurun_out/golden/VERSIONS.txt.
"""
import os
import subprocess
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import cases  # noqa: E402
import ref as refmod  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out", "golden")
DEV = "cuda"
t2n = lambda t: t.detach().cpu().numpy()


def dgr(name):
    g_np, cam_np, lg, deg = cases.dgr_inputs(name)
    g, cam = refmod.to_device(g_np, cam_np, DEV)
    bg = torch.from_numpy(cases.BG).to(DEV)
    dc, dd, da = [torch.from_numpy(lg[k]).to(DEV) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
    rr = refmod.RefRasterizer("dgr")
    ro = rr.forward(g, cam, bg, deg)
    torch.cuda.synchronize()
    runs = []
    for _ in range(3):
        runs.append({k: t2n(v).copy() for k, v in rr.backward(dc, dd, da).items()})
        torch.cuda.synchronize()
    out = dict(num_rendered=np.int64(ro["num_rendered"]), radii=t2n(ro["radii"]), color=t2n(ro["color"]),
               depth=t2n(ro["depth"]), alpha=t2n(ro["alpha"]))
    for k in ("depths", "means2D", "conic_opacity", "rgb", "tiles_touched", "point_list_keys", "point_list", "ranges",
              "n_contrib", "clamped", "cov3D"):
        out[k] = rr.fetch(k)
    for k in runs[0]:
        out["g_" + k] = runs[0][k]
        out["spread_" + k] = np.float64(max(np.abs(runs[0][k] - r[k]).max() if r[k].size else 0.0 for r in runs[1:]))
    out["input_sha256"] = np.array(cases.input_hash([g_np[k] for k in sorted(g_np) if isinstance(g_np[k], np.ndarray)]
                                                    + [cam_np["viewmatrix"], cam_np["projmatrix"], cam_np["campos"]]))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "R=%d V=%d" % (ro["num_rendered"], int((out["radii"] > 0).sum())))


def fsr(name):
    sc, cam_np, mask_np, ul_np, num_obj = cases.fsr_inputs(name)
    g, cam = refmod.to_device(sc, cam_np, DEV)
    bg = torch.zeros(3, device=DEV)
    mask = torch.from_numpy(mask_np).to(DEV)
    ul = None if ul_np is None else torch.from_numpy(ul_np).to(DEV)
    rr = refmod.RefRasterizer("fsr")
    used = []
    for _ in range(3):
        ro = rr.forward(g, cam, bg, 3, gt_mask=mask, unique_label=ul, num_obj=num_obj)
        torch.cuda.synchronize()
        used.append(t2n(ro["used_count"]).copy())
    out = dict(num_rendered=np.int64(ro["num_rendered"]), radii=t2n(ro["radii"]), color=t2n(ro["color"]),
               depth=t2n(ro["depth"]), alpha=t2n(ro["alpha"]), contrib_num=t2n(ro["contrib_num"]),
               used_count=used[0], spread_used_count=np.float64(max(np.abs(used[0] - u).max() for u in used[1:])),
               image_xys=t2n(ro["image_xys"]), gs_depth=t2n(ro["gs_depth"]))
    out["input_sha256"] = np.array(cases.input_hash([sc["means3D"], sc["shs"], mask_np, cam_np["projmatrix"]]))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "R=%d labels=%d" % (ro["num_rendered"], int(mask_np.max())))


def knn(name):
    pts = cases.knn_inputs(name)
    d = t2n(refmod.knn_dist2(torch.from_numpy(pts).to(DEV)))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), dist2=d, input_sha256=np.array(cases.input_hash([pts])))
    print(name, d[:3])


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for n in cases.DGR_CASES:
        dgr(n)
    for n in cases.FSR_CASES:
        fsr(n)
    for n in cases.KNN_CASES:
        knn(n)
    nv = subprocess.run(["nvcc", "--version"], capture_output=True, text=True).stdout.strip().splitlines()[-2:]
    smi = subprocess.run(["nvidia-smi", "--query-gpu=name,driver_version", "--format=csv,noheader"],
                         capture_output=True, text=True).stdout.strip()
    with open(os.path.join(OUT, "VERSIONS.txt"), "w") as f:
        f.write("reference: zdwww/Wheat-3DGS submodules diff-gaussian-rasterization@8829d14f, "
                "flashsplat-rasterization@b7698d27, simple-knn@86710c2d, built -O3 -arch=sm_100 (oracle/build_ref.py)\n")
        f.write("nvcc: %s\ngpu/driver: %s\ntorch: %s numpy: %s\n" % (" | ".join(nv), smi, torch.__version__, np.__version__))
