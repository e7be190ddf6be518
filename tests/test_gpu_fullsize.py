"""GPU (-m gpu): parity at BASELINE.json's full sizes for the workloads round 1 covered at toy size only.

* flashsplat at config-3 size (1 M Gaussians, 2048x1536, ~500 labels) against the unmodified reference FSR run live
  (oracle/_ref/libref_fsr.so): exact tier bit for bit, votes inside the reference's own run-to-run spread, and the
  FINAL LABELS after six views identical except where the reference's own margin is inside that spread.
* the run_3d_seg per-seed loop (segment.HeadSegmenter) against the same loop spelled with the reference rasterizer,
  the reference's multi_instance_opt formula and numpy IoU.
* bench.py's workloads end to end at tiny size (the JSON contract and the parity gates).
* committed full-size goldens (tests/golden/fullsize.json, made by make_golden_fullsize.py from the reference).
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import cases
from util import GRAD_NAMES, assert_bits, assert_exact, assert_grad, assert_image
from wheat_3dgs_b200 import scene

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

if torch.cuda.is_available():
    import gpu_util as G
    from wheat_3dgs_b200 import ops, segment
    import ref as refmod

HAVE_FSR = torch.cuda.is_available() and refmod.available("fsr")


def _ref_labels(all_counts):
    """run_3d_seg.py:54-72 (gamma 0) as the reference runs it: one torch argmax per label row."""
    s = all_counts.sum(dim=0)
    out = torch.zeros_like(all_counts).bool()
    for k in range(all_counts.shape[0]):
        oc = torch.stack([s - all_counts[k], all_counts[k]], dim=0)
        out[k] = torch.argmax(torch.nn.functional.normalize(oc, dim=0), dim=0)
    return out


@pytest.mark.skipif(not HAVE_FSR, reason="oracle/_ref/libref_fsr.so not built")
def test_flash_against_live_reference_cfg3():
    P, W, H, nobj = 1000000, 2048, 1536, 500
    sc = scene.synthetic_scene(P, seed=0)
    cams = scene.overhead_cameras(W, H)
    views = [0, 6, 12, 18, 24, 30]
    bg = G.t(np.zeros(3, np.float32))
    gd, _ = G.dev_inputs(sc, cams[0])
    rr = refmod.RefRasterizer("fsr")
    mine_sum = torch.zeros((nobj + 1, P), device="cuda")
    ref_sum = torch.zeros((nobj + 1, P), device="cuda")
    spread_sum = torch.zeros((P,), device="cuda")
    for v in views:
        _, cd = G.dev_inputs({}, cams[v])
        mask = G.t(scene.head_masks(sc, cams[v], n_labels=nobj))
        r1 = rr.forward(gd, cd, bg, 3, gt_mask=mask, num_obj=nobj, outputs={})
        torch.cuda.synchronize()
        u1 = r1["used_count"].clone()
        r2 = rr.forward(gd, cd, bg, 3, gt_mask=mask, num_obj=nobj, outputs={})   # the reference's own atomic-order spread
        torch.cuda.synchronize()
        spread = (r2["used_count"] - u1).abs()
        out = G.flash_forward(gd, cd, bg, 3, mask, None, nobj)
        R, color, depth, alpha, radii, geom, binning, image, contrib, used, xys, gsd = out
        assert R == r2["num_rendered"]
        assert_exact("radii", G.n(radii), G.n(r2["radii"]))
        assert_exact("contrib_num", G.n(contrib), G.n(r2["contrib_num"]))
        assert_bits("image_xys", G.n(xys), G.n(r2["image_xys"]))
        assert_bits("gs_depth", G.n(gsd), G.n(r2["gs_depth"]))
        for k, t in (("color", color), ("depth", depth), ("alpha", alpha)):
            assert_image(k, G.n(t), G.n(r2[k]))
        # votes: |a - b| <= 1e-3 |b| + max(1e-4 max|b|, 4 x the reference's spread of that entry's Gaussian)
        sp_g = spread.max(dim=0).values
        lim = 1e-3 * u1.abs() + torch.maximum(4.0 * sp_g[None], torch.full_like(sp_g, 1e-4 * float(u1.max()))[None])
        bad = int(((used - u1).abs() > lim).sum())
        assert bad == 0, "view %d: %d used_count entries beyond the tier" % (v, bad)
        mine_sum += used
        ref_sum += u1
        spread_sum += spread.sum(dim=0)
        del r1, r2, out, used, u1, spread, lim
    lab_ref = _ref_labels(ref_sum)
    lab_mine = ops.multi_instance_opt(mine_sum)
    diff = lab_ref != lab_mine
    s = ref_sum.sum(dim=0)
    margin = (2.0 * ref_sum - s[None]).abs()
    # a label the reference itself could flip between two of its runs: margin within its accumulated spread
    near = margin <= 4.0 * spread_sum[None] + 1e-6 * s[None]
    outside = int((diff & ~near).sum())
    print("cfg3 labels: %d of %d entries differ, %d of them outside the reference's own spread; %d near ties; %d "
          "Gaussians labelled" % (int(diff.sum()), diff.numel(), outside, int(near.sum()), int(lab_ref[1:].any(dim=0).sum())))
    assert outside == 0
    assert int(lab_ref[1:].any(dim=0).sum()) > 100000   # the workload does label the heads


@pytest.mark.skipif(not HAVE_FSR, reason="oracle/_ref/libref_fsr.so not built")
def test_flash_unique_label_gate_against_live_reference_cfg3_size():
    """FSR forward.cu:357-362 at full size: Gaussians with a unique label vote only on pixels of that label."""
    P, W, H, nobj = 1000000, 2048, 1536, 500
    sc = scene.synthetic_scene(P, seed=0)
    cam = scene.overhead_cameras(W, H)[17]
    gd, cd = G.dev_inputs(sc, cam)
    bg = G.t(np.zeros(3, np.float32))
    mask = G.t(scene.head_masks(sc, cam, n_labels=nobj))
    hid = sc["head_id"]
    ul_np = np.where(hid < 0, -1, np.where(hid % 3 == 0, hid + 1, np.where(hid % 3 == 1, 0, -1))).astype(np.int32)
    ul = G.t(ul_np)
    rr = refmod.RefRasterizer("fsr")
    r1 = rr.forward(gd, cd, bg, 3, gt_mask=mask, unique_label=ul, num_obj=nobj, outputs={})
    torch.cuda.synchronize()
    u1 = r1["used_count"].clone()
    r2 = rr.forward(gd, cd, bg, 3, gt_mask=mask, unique_label=ul, num_obj=nobj, outputs={})
    torch.cuda.synchronize()
    sp_g = (r2["used_count"] - u1).abs().max(dim=0).values
    out = G.flash_forward(gd, cd, bg, 3, mask, ul, nobj)
    assert_exact("contrib_num", G.n(out[8]), G.n(r2["contrib_num"]))
    for k, i in (("color", 1), ("depth", 2), ("alpha", 3)):
        assert_image(k, G.n(out[i]), G.n(r2[k]))
    lim = 1e-3 * u1.abs() + torch.maximum(4.0 * sp_g[None], torch.full_like(sp_g, 1e-4 * float(u1.max()))[None])
    assert int(((out[9] - u1).abs() > lim).sum()) == 0


@pytest.mark.skipif(not HAVE_FSR, reason="oracle/_ref/libref_fsr.so not built")
def test_head_segmenter_equals_the_reference_loop():
    """segment.HeadSegmenter.segment_seed against run_3d_seg.py:263-315 spelled with the reference rasterizer
    (binary gt_mask votes with obj_num = 1, used_mask subset renders with gt_mask None / obj_num = 2), the reference's
    multi_instance_opt formula and numpy IoU with the bbox-overlap filter (utils/image_helper.py:104-163)."""
    P, W, H, nobj, thr = 30000, 400, 300, 12, 0.3
    sc = scene.synthetic_scene(P, seed=0, n_heads=nobj, scale_median=0.004)
    cams = scene.overhead_cameras(W, H)[12:20]
    nv = len(cams)
    bg = G.t(np.zeros(3, np.float32))
    gd, _ = G.dev_inputs(sc, cams[0])
    cds = [G.dev_inputs({}, c)[1] for c in cams]
    masks_np = [scene.head_masks(sc, c, n_labels=nobj) for c in cams]
    masks = {v: G.t(m) for v, m in enumerate(masks_np)}
    seg = segment.HeadSegmenter(gd, cds, masks, bg, 3, iou_threshold=thr)
    rr = refmod.RefRasterizer("fsr")
    z = gd["means3D"][:, 2]
    pts_filter = z < z.mean()

    def bbox(m):
        ys, xs = np.nonzero(m)
        return None if ys.size == 0 else (xs.min(), ys.min(), xs.max(), ys.max())

    def overlapping(a, b):
        if a is None or b is None:
            return False
        return not (a[2] < b[0] or b[2] < a[0] or a[3] < b[1] or b[3] < a[1])

    def ref_counts(v, k):
        binary = G.t((masks_np[v] == k).astype(np.float32))
        return rr.forward(gd, cds[v], bg, 3, gt_mask=binary, num_obj=1, outputs={})["used_count"].clone()

    def ref_find_match(views, used):
        sub = {k: t[used] for k, t in gd.items()}
        res = []
        for v in views:
            o = rr.forward(sub, cds[v], bg, 3, gt_mask=None, num_obj=2, outputs={})
            pred = G.n(o["alpha"])[0] > 0.5
            pb = bbox(pred)
            best, arg = 0.0, None
            for k in range(1, nobj + 1):
                m = masks_np[v] == k
                if not overlapping(pb, bbox(m)):
                    continue
                u = np.logical_or(m, pred).sum()
                iou = np.logical_and(m, pred).sum() / u if u > 0 else 0.0
                if iou > best:
                    best, arg = iou, k
            if best > thr:
                res.append((v, arg))
        return res

    n_matched = 0
    for seed_view, k in ((0, 1), (0, 5), (3, 2), (5, 9), (2, 12)):
        # reference flow
        counts = ref_counts(seed_view, k)
        counts[1:, pts_filter] = 0
        used = _ref_labels(counts)[1]
        mine = seg.segment_seed(seed_view, k)
        if int(used.sum()) == 0:
            assert mine is None
            continue
        matched = [(seed_view, k)]
        new = ref_find_match([u for u in range(nv) if u != seed_view], used)
        matched += new
        rounds = 0
        while new and rounds < 99:
            rounds += 1
            for u, j in new:
                counts += ref_counts(u, j)
            used = _ref_labels(counts)[1]
            done = {u for u, _ in matched}
            new = ref_find_match([u for u in range(nv) if u not in done], used)
            matched += new
        assert mine is not None
        assert mine["matched"] == matched, (mine["matched"], matched)
        assert mine["rounds"] == rounds
        # the object's Gaussians: identical except entries the reference's atomics could flip (none expected here)
        assert int((mine["used_mask"] != used).sum()) <= 2
        n_matched += len(matched) - 1
        # and the shared multi-label pass gives the same seed counts as the binary render
        shared = seg.view_counts(seed_view, nobj)
        a, b = seg.seed_counts(seed_view, k, shared=shared), seg.seed_counts(seed_view, k)
        assert float((a - b).abs().max()) <= 1e-3 * float(b.abs().max())
    assert n_matched > 0, "the synthetic workload must produce cross-view matches"


def test_subset_alpha_is_the_flash_alpha():
    """find_match reads only alpha of flashsplat_render(used_mask=...): the plain forward gives the same bits."""
    sc, cam, mask, ul, num_obj = cases.fsr_inputs("fsr_multi")
    gd, cd = G.dev_inputs(sc, cam)
    bg = G.t(cases.BG)
    sel = torch.from_numpy(sc["head_id"] == 3).cuda()
    sub = {k: v[sel] for k, v in gd.items()}
    ones = torch.ones((cd["image_height"], cd["image_width"]), device="cuda")
    fo = G.flash_forward(sub, cd, bg, 3, ones, None, 2)
    do = G.forward(sub, cd, bg, 3)
    assert_bits("alpha", G.n(fo[3]), G.n(do[3]))


@pytest.mark.parametrize("workload", ["tiny", "tinytrain", "tiny4", "tiny3"])
def test_bench_workloads_run_and_keep_the_contract(workload):
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", workload, "--steps", "4", "--warmup", "3",
           "--no-cpu-baseline", "--no-extra", "--seeds", "2"]
    env = dict(os.environ, WS_NO_CLOCKS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "dtype", "data", "config", "e2e", "gpu_launches", "parity"):
        assert k in line, k
    assert line["value"] > 0 and line["e2e"]["value"] > 0 and line["gpu_launches"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] > 0 and line["e2e"]["d2h_bytes_per_step"] > 0
    if line["parity"].get("checked"):
        assert line["parity"]["ok"], line["parity"]
    if workload == "tiny3":
        assert line["parity"]["votes_vs_alpha_rel_err"] < 1e-4
        assert line["extra"]["run_3d_seg"]["seeds"] > 0


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "tests", "golden", "fullsize.json")), reason="no full-size goldens yet")
@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg4", "cfg5"])
def test_fullsize_goldens(name):
    """One view / frame of BASELINE.json's configs against hashes and samples of the reference's outputs."""
    sys.path.insert(0, ROOT)
    import bench
    gold = bench.load_golden(name)
    if gold is None:
        pytest.skip("no golden entry for " + name)
    sc, cams, (P, W, H, nv) = bench.make_workload(name)
    g, cd = bench.to_dev(sc, cams, torch.device("cuda"))
    orbit_kind = bench.WORKLOADS[name]["kind"] == "orbit"
    bg = torch.ones(3, device="cuda") if orbit_kind else torch.from_numpy(bench.BG).cuda()
    c = cd[gold["view"]]
    fw = G.forward(g, c, bg, 3)
    sig, samples = bench.fwd_signature(fw, P, W, H)
    grads = None
    if not orbit_kind:
        lg = scene.loss_gradients((3, H, W), seed=2)
        out = G.backward(g, c, bg, 3, fw, G.t(lg["dL_dcolor"]), G.t(lg["dL_ddepth"]), G.t(lg["dL_dalpha"]))
        grads = {k: bench.sample(t) for k, t in zip(bench.GRAD_KEYS, out)}
    par = bench.compare_signature(gold, sig, samples, grads, name=name)
    assert par["ok"], par


def _flash_backward(gd, cd, bg, deg, fw, mask, ul, dc, dd, da):
    """ops.flash_rasterize_gaussians_backward (render_bwd<true>) on the buffers of a flashsplat forward."""
    E = torch.Tensor([])
    R, color, depth, alpha, radii, geom, binning, image = fw[:8]
    return ops.flash_rasterize_gaussians_backward(
        mask, ul if ul is not None else torch.tensor([], dtype=torch.int), bg, gd["means3D"], radii, E, gd["scales"],
        gd["rotations"], 1.0, E, cd["viewmatrix"], cd["projmatrix"], cd["tanfovx"], cd["tanfovy"], dc, dd, da, gd["shs"],
        deg, cd["campos"], geom, R, binning, image, alpha, False)


@pytest.mark.parametrize("name", sorted(cases.FSR_CASES))
def test_flash_backward_matches_oracle(name):
    """The flashsplat backward (FSR backward.cu:455-462,529-536): with unique_label given, a Gaussian whose label is
    >= 0 and differs from the pixel's mask id is skipped by that pixel -- no gradient from it, and it does not take
    part in the pixel's transmittance walk.  Toy size against the CPU oracle (oracle/splat_oracle.c, render backward
    with gt_mask / unique_label), with and without labels."""
    import oracle
    sc, cam, mask, ul, num_obj = cases.fsr_inputs(name)
    if ul is None:   # the multi-label case has no labels of its own: gate a third of the heads, free another third
        hid = sc["head_id"]
        ul = np.where(hid < 0, -1, np.where(hid % 3 == 0, (hid % num_obj) + 1, np.where(hid % 3 == 1, 0, -1))).astype(np.int32)
    H, W = cam["image_height"], cam["image_width"]
    gd, cd = G.dev_inputs(sc, cam)
    bg = torch.zeros(3, device="cuda")
    lg = scene.loss_gradients((3, H, W), seed=4)
    dc, dd, da = G.t(lg["dL_dcolor"]), G.t(lg["dL_ddepth"]), G.t(lg["dL_dalpha"])
    gated_differs = False
    for labels in (ul, None):
        fw = G.flash_forward(gd, cd, bg, 3, G.t(mask), None if labels is None else G.t(labels), num_obj)
        mine = _flash_backward(gd, cd, bg, 3, fw, G.t(mask), None if labels is None else G.t(labels), dc, dd, da)
        of = oracle.forward(sc, cam, np.zeros(3, np.float32), sh_degree=3, gt_mask=mask, unique_label=labels,
                            num_obj=num_obj)
        ob = oracle.backward(sc, cam, np.zeros(3, np.float32), of, lg["dL_dcolor"], lg["dL_ddepth"], lg["dL_dalpha"],
                             sh_degree=3, gt_mask=mask, unique_label=labels)
        for k, m in zip(GRAD_NAMES, mine):
            assert_grad("%s (%s)" % (k, "labels" if labels is not None else "no labels"), G.n(m).reshape(ob[k].shape), ob[k])
        if labels is not None:
            plain = oracle.backward(sc, cam, np.zeros(3, np.float32), of, lg["dL_dcolor"], lg["dL_ddepth"],
                                    lg["dL_dalpha"], sh_degree=3)
            gated_differs = float(np.abs(plain["dL_dopacity"] - ob["dL_dopacity"]).max()) > 1e-3 * float(np.abs(plain["dL_dopacity"]).max())
    assert gated_differs, "the case does not exercise the label gate"


@pytest.mark.skipif(not HAVE_FSR, reason="oracle/_ref/libref_fsr.so not built")
def test_flash_backward_against_live_reference_cfg3_size():
    """render_bwd<true> at P = 1 M, 2048x1536, 500 labels with the unique_label gate, against the unmodified flashsplat
    rasterizer's backward run live.

    Bar: every gradient tensor within 1e-3 norm-wise of the reference's (the gradient tier's relative tolerance).  The
    element-wise form of the tier used for the DGR backward is NOT applied here, for a measured reason: on identical
    inputs WITHOUT labels the two unmodified references themselves -- libref_fsr.so and libref_dgr.so, whose
    backward.cu differ only by the gate -- disagree by 2.6e-4 to 2.8e-4 norm-wise in every gradient tensor
    (deterministically: three FSR runs agree to 1e-6), while this library agrees with the DGR one to 7e-7 and its
    FLASH and plain kernels with each other to 2e-7 (round 3, gpurun diagnostic).  The cause inside the flashsplat build
    was not found; until it is, the reference's own two rasterizers bound how closely the FSR backward can be matched.
    The gate itself is pinned exactly at toy size by test_flash_backward_matches_oracle above."""
    P, W, H, nobj = 1000000, 2048, 1536, 500
    sc = scene.synthetic_scene(P, seed=0)
    cam = scene.overhead_cameras(W, H)[17]
    gd, cd = G.dev_inputs(sc, cam)
    bg = G.t(np.zeros(3, np.float32))
    mask = G.t(scene.head_masks(sc, cam, n_labels=nobj))
    hid = sc["head_id"]
    ul = G.t(np.where(hid < 0, -1, np.where(hid % 3 == 0, hid + 1, np.where(hid % 3 == 1, 0, -1))).astype(np.int32))
    lg = scene.loss_gradients((3, H, W), seed=2)
    dc, dd, da = G.t(lg["dL_dcolor"]), G.t(lg["dL_ddepth"]), G.t(lg["dL_dalpha"])
    nrm = lambda a, b: float((a.reshape(-1).double() - b.reshape(-1).double()).norm() / b.reshape(-1).double().norm())
    for labels in (ul, None):
        rr = refmod.RefRasterizer("fsr")
        rr.forward(gd, cd, bg, 3, gt_mask=mask, unique_label=labels, num_obj=nobj, outputs={})
        torch.cuda.synchronize()
        ref = {k: v.clone() for k, v in rr.backward(dc, dd, da).items()}
        torch.cuda.synchronize()
        fw = G.flash_forward(gd, cd, bg, 3, mask, labels, nobj)
        mine = _flash_backward(gd, cd, bg, 3, fw, mask, labels, dc, dd, da)
        for k, m in zip(GRAD_NAMES, mine):
            assert torch.isfinite(m).all(), k
            assert nrm(m, ref[k]) < 1e-3, "%s (%s): norm-wise %g" % (k, "labels" if labels is not None else "no labels", nrm(m, ref[k]))
        plain = G.backward(gd, cd, bg, 3, fw[:8], dc, dd, da)
        if labels is None:
            # without labels the FLASH kernel is the plain kernel
            for k, m, q in zip(GRAD_NAMES, mine, plain):
                assert nrm(m, q) < 1e-4, k
        else:
            # the gate is live in this case: the ungated walk over the gated forward's buffers (whose alpha no longer
            # matches it) is far outside the bar, or not even finite
            assert not (nrm(plain[2], ref["dL_dopacity"]) < 1e-2)
        del rr, ref, fw, mine, plain
