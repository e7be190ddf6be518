"""Development harness (GPU box): side-by-side run of libwheatsplat against the reference CUDA
rasterizer (oracle/_ref) and the CPU oracle, with verbose mismatch diagnostics and timings.
Not part of the product or of the test suite; `python tools/gpu_check.py [stage ...]`.
"""
import os
import sys
import time
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from wheat_3dgs_b200 import ops, scene  # noqa: E402
import oracle  # noqa: E402
import ref as refmod  # noqa: E402

DEV = "cuda"


def t2n(t):
    return t.detach().cpu().numpy()


def cmp_exact(name, a, b, show=5):
    a, b = np.asarray(a), np.asarray(b)
    if a.shape != b.shape:
        print("  [%s] SHAPE MISMATCH %s vs %s" % (name, a.shape, b.shape))
        return False
    bad = np.flatnonzero(a.reshape(-1) != b.reshape(-1))
    if bad.size == 0:
        print("  [%s] exact OK (%d elems)" % (name, a.size))
        return True
    print("  [%s] MISMATCH %d / %d; first: %s" % (name, bad.size, a.size,
          [(int(i), a.reshape(-1)[i], b.reshape(-1)[i]) for i in bad[:show]]))
    return False


def cmp_tol(name, a, b, atol, rtol=0.0, show=5):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    if a.shape != b.shape:
        print("  [%s] SHAPE MISMATCH %s vs %s" % (name, a.shape, b.shape))
        return False
    err = np.abs(a - b)
    lim = atol + rtol * np.abs(b)
    bad = np.flatnonzero((err > lim).reshape(-1) | ~np.isfinite(a).reshape(-1))
    print("  [%s] max abs err %.3e, mean |ref| %.3e, %s %d/%d beyond atol=%.0e rtol=%.0e" % (
        name, err.max() if err.size else 0, np.abs(b).mean() if b.size else 0, "OK" if bad.size == 0 else "BAD",
        bad.size, a.size, atol, rtol))
    if bad.size:
        print("      first: %s" % [(int(i), a.reshape(-1)[i], b.reshape(-1)[i]) for i in bad[:show]])
    return bad.size == 0


def mine_forward(g, cam, bg, deg):
    e = torch.Tensor([])
    return ops.rasterize_gaussians(bg, g["means3D"], g.get("colors_precomp", e), g["opacities"], g.get("scales", e),
                                   g.get("rotations", e), 1.0, g.get("cov3D_precomp", e), cam["viewmatrix"],
                                   cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], cam["image_height"],
                                   cam["image_width"], g.get("shs", e), deg, cam["campos"], False, False)


def mine_backward(g, cam, bg, deg, fw, dc, dd, da):
    e = torch.Tensor([])
    R, color, depth, alpha, radii, geom, binning, image = fw
    return ops.rasterize_gaussians_backward(bg, g["means3D"], radii, g.get("colors_precomp", e), g.get("scales", e),
                                            g.get("rotations", e), 1.0, g.get("cov3D_precomp", e), cam["viewmatrix"],
                                            cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], dc, dd, da,
                                            g.get("shs", e), deg, cam["campos"], geom, R, binning, image, alpha, False)


def timeit(fn, warm=3, it=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(it):
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(np.min(ts))


def stage_sort():
    print("== sort ==")
    gen = torch.Generator(device="cpu").manual_seed(0)
    for n, bits in [(1, 40), (31, 46), (3072, 46), (3073, 43), (100000, 46), (1 << 20, 47), (5000001, 46), (777, 8), (50000, 33)]:
        keys = torch.randint(0, 1 << 62, (n,), generator=gen, dtype=torch.int64)
        # many duplicates in the sorted bits to exercise stability
        keys = (keys & ((1 << bits) - 1)) >> (bits // 3) << (bits // 3) | (keys & ~((1 << bits) - 1))
        vals = torch.arange(n, dtype=torch.int32)
        ko, vo = ops.sort_pairs(keys.to(DEV), vals.to(DEV), bits)
        torch.cuda.synchronize()
        rk, rv = oracle.sort_pairs(keys.numpy().view(np.uint64), vals.numpy().view(np.uint32), bits)
        ok1 = np.array_equal(t2n(ko).view(np.uint64), rk)
        ok2 = np.array_equal(t2n(vo).view(np.uint32), rv)
        print("  n=%d bits=%d keys %s vals %s" % (n, bits, "OK" if ok1 else "BAD", "OK" if ok2 else "BAD"))
        if not (ok1 and ok2):
            bad = np.flatnonzero(t2n(ko).view(np.uint64) != rk)
            print("     first bad key idx", bad[:5], "of", bad.size)


def run_case(name, P, W, H, cam_idx=17, deg=3, do_oracle=True, do_time=False, save=None, scale_median=0.002):
    print("== case %s: P=%d %dx%d deg=%d ==" % (name, P, W, H, deg))
    sc = scene.synthetic_scene(P, seed=0, scale_median=scale_median)
    cam_np = scene.overhead_cameras(W, H)[cam_idx]
    bg_np = np.array([0.1, 0.2, 0.3], np.float32)
    g, cam = refmod.to_device(sc, cam_np, DEV)
    bg = torch.from_numpy(bg_np).to(DEV)
    lg = scene.loss_gradients((3, H, W))
    dc, dd, da = [torch.from_numpy(lg[k]).to(DEV) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]

    rr = refmod.RefRasterizer("dgr")
    ro = rr.forward(g, cam, bg, deg)
    torch.cuda.synchronize()
    R_ref = ro["num_rendered"]
    print("  ref: R=%d V=%d" % (R_ref, int((ro["radii"] > 0).sum())))
    fw = mine_forward(g, cam, bg, deg)
    torch.cuda.synchronize()
    R, color, depth, alpha, radii, geom, binning, image = fw
    print("  mine: R=%d" % R)
    ok = True
    ok &= cmp_exact("radii", t2n(radii), t2n(ro["radii"]))
    vis = t2n(ro["radii"]) > 0
    md = t2n(ops.debug_fetch("depths", P, W, H, R, geom, binning, image))
    ok &= cmp_exact("depths(bits,visible)", md.view(np.uint32)[vis], rr.fetch("depths").view(np.uint32)[vis])
    mm2 = t2n(ops.debug_fetch("means2D", P, W, H, R, geom, binning, image))
    ok &= cmp_exact("means2D(bits,visible)", mm2.view(np.uint32)[vis], rr.fetch("means2D").view(np.uint32).reshape(P, 2)[vis])
    mco = t2n(ops.debug_fetch("conic_opacity", P, W, H, R, geom, binning, image))
    ok &= cmp_exact("conic_opacity(bits,visible)", mco.view(np.uint32)[vis], rr.fetch("conic_opacity").view(np.uint32).reshape(P, 4)[vis])
    mrgb = t2n(ops.debug_fetch("rgb", P, W, H, R, geom, binning, image))
    cmp_tol("rgb(visible)", mrgb[vis], rr.fetch("rgb").reshape(P, 3)[vis], 1e-5)
    ok &= cmp_exact("tiles_touched", t2n(ops.debug_fetch("tiles_touched", P, W, H, R, geom, binning, image)).view(np.uint32), rr.fetch("tiles_touched"))
    if R == R_ref and R > 0:
        ok &= cmp_exact("point_list_keys", t2n(ops.debug_fetch("point_list_keys", P, W, H, R, geom, binning, image)).view(np.uint64), rr.fetch("point_list_keys"))
        ok &= cmp_exact("point_list", t2n(ops.debug_fetch("point_list", P, W, H, R, geom, binning, image)).view(np.uint32), rr.fetch("point_list"))
    ok &= cmp_exact("ranges", t2n(ops.debug_fetch("ranges", P, W, H, R, geom, binning, image)).view(np.uint32).reshape(-1), rr.fetch("ranges"))
    ok &= cmp_exact("n_contrib", t2n(ops.debug_fetch("n_contrib", P, W, H, R, geom, binning, image)).view(np.uint32), rr.fetch("n_contrib"))
    ok &= cmp_tol("color", t2n(color), t2n(ro["color"]), 1e-4)
    ok &= cmp_tol("depth", t2n(depth), t2n(ro["depth"]), 1e-4)
    ok &= cmp_tol("alpha", t2n(alpha), t2n(ro["alpha"]), 1e-4)
    cmp_exact("color(bits)", t2n(color).view(np.uint32), t2n(ro["color"]).view(np.uint32))

    # backward: reference twice (its own run-to-run spread), then mine
    gr1 = {k: v.clone() for k, v in rr.backward(dc, dd, da).items()}
    torch.cuda.synchronize()
    gr2 = rr.backward(dc, dd, da)
    torch.cuda.synchronize()
    mg = mine_backward(g, cam, bg, deg, fw, dc, dd, da)
    torch.cuda.synchronize()
    names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations"]
    for nm, m in zip(names, mg):
        r1, r2 = t2n(gr1[nm]).reshape(m.shape), t2n(gr2[nm]).reshape(m.shape)
        scale = np.abs(r1).max() + 1e-30
        spread = np.abs(r1 - r2).max()
        print("  (%s) ref run-to-run max spread %.3e (rel to max %.3e)" % (nm, spread, spread / scale))
        ok &= cmp_tol(nm, t2n(m), r1, atol=1e-5 * scale, rtol=1e-3)

    if do_oracle:
        t0 = time.time()
        of = oracle.forward(sc, cam_np, bg_np, sh_degree=deg)
        ob = oracle.backward(sc, cam_np, bg_np, of, lg["dL_dcolor"], lg["dL_ddepth"], lg["dL_dalpha"], sh_degree=deg)
        print("  oracle (CPU, %d threads) fwd+bwd %.2fs" % (oracle.num_threads(), time.time() - t0))
        cmp_exact("oracle radii vs ref", of["radii"], t2n(ro["radii"]))
        if of["num_rendered"] == R_ref:
            cmp_exact("oracle keys vs ref", of["point_list_keys"], rr.fetch("point_list_keys"))
            cmp_exact("oracle point_list vs ref", of["point_list"], rr.fetch("point_list"))
        cmp_exact("oracle ranges vs ref", of["ranges"].reshape(-1), rr.fetch("ranges"))
        cmp_exact("oracle n_contrib vs ref", of["n_contrib"].reshape(-1), rr.fetch("n_contrib"))
        cmp_tol("oracle color vs ref", of["out_color"], t2n(ro["color"]), 1e-4)
        cmp_tol("oracle alpha vs ref", of["out_alpha"], t2n(ro["alpha"]), 1e-4)
        for nm in names:
            r1 = t2n(gr1[nm]).reshape(ob[nm].shape)
            cmp_tol("oracle " + nm, ob[nm], r1, atol=1e-5 * (np.abs(r1).max() + 1e-30), rtol=1e-3)

    if do_time:
        f_ref = timeit(lambda: rr.forward(g, cam, bg, deg, outputs=ro))
        fb_ref = timeit(lambda: (rr.forward(g, cam, bg, deg, outputs=ro), rr.backward(dc, dd, da, grads=gr2)))
        f_me = timeit(lambda: mine_forward(g, cam, bg, deg))

        def fb():
            f = mine_forward(g, cam, bg, deg)
            mine_backward(g, cam, bg, deg, f, dc, dd, da)
        fb_me = timeit(fb)
        print("  TIMING ms (median,min): ref fwd %s  ref fwd+bwd %s | mine fwd %s  mine fwd+bwd %s" % (f_ref, fb_ref, f_me, fb_me))
    print("  CASE %s: %s" % (name, "ALL OK" if ok else "MISMATCHES"))
    if save:
        os.makedirs(os.path.dirname(save), exist_ok=True)
        np.savez_compressed(
            save, radii=t2n(ro["radii"]), point_list_keys=rr.fetch("point_list_keys"), point_list=rr.fetch("point_list"),
            ranges=rr.fetch("ranges"), n_contrib=rr.fetch("n_contrib"), color=t2n(ro["color"]), depth=t2n(ro["depth"]),
            alpha=t2n(ro["alpha"]), depths=rr.fetch("depths"), means2D=rr.fetch("means2D"),
            conic_opacity=rr.fetch("conic_opacity"), rgb=rr.fetch("rgb"),
            **{"g_" + k: t2n(v) for k, v in gr1.items()})
        print("  saved", save)
    return ok


def stage_flash():
    print("== flash ==")
    P, W, H, deg = 50000, 512, 512, 3
    sc = scene.synthetic_scene(P, seed=0)
    cam_np = scene.overhead_cameras(W, H)[17]
    g, cam = refmod.to_device(sc, cam_np, DEV)
    bg = torch.zeros(3, device=DEV)
    e = torch.Tensor([])
    for num_obj, use_ul in [(1, False), (40, False), (1, True)]:
        mask_np = scene.head_masks(sc, cam_np, n_labels=num_obj)
        if num_obj == 1:
            mask_np = (mask_np > 0).astype(np.float32)
        mask = torch.from_numpy(mask_np).to(DEV)
        ul = None
        if use_ul:
            ul_np = np.where(sc["head_id"] == 0, 1, np.where(sc["head_id"] % 3 == 0, -1, 0)).astype(np.int32)
            ul = torch.from_numpy(ul_np).to(DEV)
        rr = refmod.RefRasterizer("fsr")
        ro = rr.forward(g, cam, bg, deg, gt_mask=mask, unique_label=ul, num_obj=num_obj)
        torch.cuda.synchronize()
        out = ops.flash_rasterize_gaussians(mask, ul if ul is not None else torch.tensor([], dtype=torch.int), bg,
                                            g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e,
                                            cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], H, W,
                                            g["shs"], deg, cam["campos"], False, num_obj, False)
        torch.cuda.synchronize()
        R, color, depth, alpha, radii, geom, binning, image, contrib, used, xys, gsd = out
        print(" num_obj=%d unique_label=%s R=%d/%d mask labels=%d" % (num_obj, use_ul, R, ro["num_rendered"], int(mask_np.max())))
        cmp_exact("radii", t2n(radii), t2n(ro["radii"]))
        cmp_tol("color", t2n(color), t2n(ro["color"]), 1e-4)
        cmp_tol("alpha", t2n(alpha), t2n(ro["alpha"]), 1e-4)
        cmp_exact("contrib_num", t2n(contrib), t2n(ro["contrib_num"]))
        ru = t2n(ro["used_count"])
        cmp_tol("used_count", t2n(used), ru, atol=1e-5 * ru.max(), rtol=1e-4)
        cmp_exact("image_xys(bits)", t2n(xys).view(np.uint32), t2n(ro["image_xys"]).view(np.uint32))
        cmp_exact("gs_depth(bits)", t2n(gsd).view(np.uint32), t2n(ro["gs_depth"]).view(np.uint32))
        lab_m = t2n(ops.multi_instance_opt(used))
        lab_r = oracle.multi_instance_opt(ru)
        lab_x = oracle.multi_instance_opt(t2n(used))
        cmp_exact("labels(mine kernel vs oracle on my counts)", lab_m, lab_x)
        d = np.flatnonzero((lab_m != lab_r).reshape(-1))
        print("  labels vs oracle-on-ref-counts: %d differ of %d" % (d.size, lab_r.size))
        uo = oracle.forward(sc, cam_np, np.zeros(3, np.float32), sh_degree=deg, gt_mask=mask_np,
                            unique_label=(ul_np if use_ul else None), num_obj=num_obj)
        cmp_tol("oracle used_count vs ref", uo["used_count"], ru, atol=1e-5 * ru.max(), rtol=1e-4)
        cmp_exact("oracle contrib_num vs ref", uo["contrib_num"], t2n(ro["contrib_num"]))


def stage_knn():
    print("== knn ==")
    for P in (5, 1000, 50000, 300000):
        rng = np.random.default_rng(P)
        pts = (rng.normal(size=(P, 3)) * np.array([1.0, 0.5, 0.2]) + np.array([0.3, -0.2, 1.0])).astype(np.float32)
        tp = torch.from_numpy(pts).to(DEV)
        r = t2n(refmod.knn_dist2(tp))
        m = t2n(ops.distCUDA2(tp))
        torch.cuda.synchronize()
        print(" P=%d" % P)
        cmp_exact("distCUDA2(bits)", m.view(np.uint32), r.view(np.uint32))
        if P <= 50000:
            o = oracle.knn_dist2(pts)
            cmp_exact("oracle distCUDA2(bits) vs ref", o.view(np.uint32), r.view(np.uint32))


def stage_misc():
    print("== misc: mark_visible, colors_precomp / cov3D_precomp path, odd sizes ==")
    P, W, H = 20000, 333, 217
    sc = scene.synthetic_scene(P, seed=3)
    cam_np = scene.overhead_cameras(W, H)[5]
    g, cam = refmod.to_device(sc, cam_np, DEV)
    rr = refmod.RefRasterizer("dgr")
    cmp_exact("mark_visible", t2n(ops.mark_visible(g["means3D"], cam["viewmatrix"], cam["projmatrix"])),
              t2n(rr.mark_visible(g["means3D"], cam["viewmatrix"], cam["projmatrix"])))
    run_case("odd-size sh deg 1", 20000, 333, 217, cam_idx=5, deg=1, do_oracle=False)


def main():
    stages = sys.argv[1:] or ["sort", "small", "cfg1", "flash", "knn", "misc", "cfg2"]
    print("device:", torch.cuda.get_device_name(0), "| stages:", stages)
    gold = os.path.join(ROOT, "gpurun_out", "golden")
    table = {
        "sort": stage_sort,
        "small": lambda: run_case("small", 8000, 256, 256, save=os.path.join(gold, "ref_small.npz")),
        "cfg1": lambda: run_case("cfg1", 50000, 512, 512, do_time=True),
        "flash": stage_flash,
        "knn": stage_knn,
        "misc": stage_misc,
        "cfg2": lambda: run_case("cfg2", 1000000, 2048, 1536, do_oracle=False, do_time=True),
    }
    for s in stages:
        try:
            table[s]()
        except Exception:
            print("STAGE %s FAILED:" % s)
            traceback.print_exc()
        sys.stdout.flush()


if __name__ == "__main__":
    main()
