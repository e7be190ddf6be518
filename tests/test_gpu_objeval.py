"""GPU: batched eval_obj_labels (csrc/ws_objeval.cu through wheat_3dgs_b200.segment, C ABI ws_objeval_*) against the
reference (golden = unmodified reference flashsplat rasterizer on a B200 + the reference's own Python merge,
tests/golden/make_golden_objeval.py), against the CPU oracle, and against the reference's per-object loop run on
this library's own flashsplat drop-in at a larger size."""
import os

import numpy as np
import pytest
import torch

import cases
import segment_oracle

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from wheat_3dgs_b200 import ops, scene, segment
    from wheat_3dgs_b200.rasterizer import GaussianRasterizationSettings

t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _settings(cam, bg):
    return GaussianRasterizationSettings(
        image_height=cam["image_height"], image_width=cam["image_width"], tanfovx=cam["tanfovx"],
        tanfovy=cam["tanfovy"], bg=t(bg), scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]),
        projmatrix=t(cam["projmatrix"]), sh_degree=3, campos=t(cam["campos"]), prefiltered=False, debug=False)


def _run(labels, sc, cam, bg=cases.BG):
    return segment.eval_obj_labels_tensors(t(labels), _settings(cam, bg), t(sc["means3D"]), t(sc["opacities"]),
                                           shs=t(sc["shs"]), scales=t(sc["scales"]), rotations=t(sc["rotations"]))


def test_pred_mask_matches_reference_golden_and_oracle(golden_dir):
    gold = np.load(os.path.join(golden_dir, "objeval.npz"))
    sc, cam, labels = cases.objeval_inputs()
    pred = _run(labels, sc, cam)
    assert pred.shape == (1, cam["image_height"], cam["image_width"]) and pred.dtype == torch.float32
    mine = pred.cpu().numpy()
    assert np.array_equal(mine, gold["pred_mask"]), "%d pixels differ from the reference" % (mine != gold["pred_mask"]).sum()
    assert np.array_equal(mine, segment_oracle.eval_obj_labels(labels, sc, cam, cases.BG, sh_degree=3))
    # both branches of the overlap rule and the multi-membership rows are exercised by the case
    ids = set(np.unique(mine).tolist())
    assert {8.0, 9.0} <= ids and 11.0 not in ids and 10.0 not in ids


def _loop_on_dropin(labels, sc, cam, bg):
    """The reference's algorithm (wheatgs_helper.py:245-271) with this library's flashsplat drop-in as the renderer."""
    H, W = cam["image_height"], cam["image_width"]
    g = {k: t(sc[k]) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
    e = torch.Tensor([])
    pred = torch.zeros((1, H, W), device="cuda")
    min_depth = torch.ones((1, H, W), device="cuda")
    bgt = t(bg)
    for i in range(1, labels.shape[0]):
        m = t(labels[i])
        if int(m.sum()) == 0:
            continue
        out = ops.flash_rasterize_gaussians(
            torch.ones((H, W), device="cuda"), None, bgt, g["means3D"][m], e, g["opacities"][m], g["scales"][m],
            g["rotations"][m], 1.0, e, t(cam["viewmatrix"]), t(cam["projmatrix"]), cam["tanfovx"], cam["tanfovy"], H, W,
            g["shs"][m], 3, t(cam["campos"]), False, 2, False)
        depth, alpha = out[2], out[3]
        _pix = alpha > 0.5
        pix = _pix.clone()
        ov = _pix & (pred > 0)
        if int(ov.sum()) > 0 and float(min_depth[ov].double().mean().float()) < float(depth[ov].double().mean().float()):
            pix[_pix] = ~(pred[_pix] > 0)
        pred[pix] = i
        min_depth[pix] = depth[pix]
    return pred


def test_batched_equals_per_object_loop_on_a_field_scene():
    """60 K Gaussians, 40 heads, a 640x360 orbit frame (heads overlap in an oblique view): the batched pass and the
    reference's per-object loop paint the same image; ragged image size (360 = 22.5 tiles)."""
    P, W, H, nh = 60000, 640, 360, 40
    sc = scene.synthetic_scene(P, seed=4, n_heads=nh, scale_median=0.004, extent=(0.6, 0.4, 0.25))
    cam = scene.orbit_cameras(W, H, n_frames=8, center=(0.0, 0.0, 0.12), extent=(0.6, 0.4, 0.25))[3]
    hid = sc["head_id"]
    labels = np.zeros((nh + 1, P), bool)
    labels[0] = hid < 0
    labels[hid[hid >= 0] + 1, np.nonzero(hid >= 0)[0]] = True
    bg = np.ones(3, np.float32)
    mine = _run(labels, sc, cam, bg)
    ref = _loop_on_dropin(labels, sc, cam, bg)
    assert torch.equal(mine, ref), "%d pixels differ" % int((mine != ref).sum())
    assert len(torch.unique(mine)) > nh // 2
    again = _run(labels, sc, cam, bg)
    assert torch.equal(mine, again), "not deterministic"


def test_objeval_edges():
    sc, cam, labels = cases.objeval_inputs()
    none = np.zeros_like(labels)
    none[0] = True
    assert _run(none, sc, cam) is None                      # wheatgs_helper.py:248,270: no object -> None
    only_bg_row = labels[:1]
    assert _run(only_bg_row, sc, cam) is None
    # objects entirely behind the camera / culled: an all-zero mask, not None
    far = dict(sc, means3D=sc["means3D"] + np.array([0, 0, 100.0], np.float32))
    pm = _run(labels, far, cam)
    assert pm is not None and float(pm.abs().sum()) == 0.0
    with pytest.raises(RuntimeError):
        segment.object_sets(torch.from_numpy(labels))       # CPU tensor: no CPU path
    sets = segment.object_sets(t(labels))
    assert sets.num_rows == labels.shape[0] and int(sets.member_start[-1]) == int(labels[1:].sum())
    with pytest.raises(RuntimeError):
        segment.eval_obj_labels_tensors(sets, _settings(cam, cases.BG), t(sc["means3D"][:100]), t(sc["opacities"][:100]),
                                        shs=t(sc["shs"][:100]), scales=t(sc["scales"][:100]), rotations=t(sc["rotations"][:100]))


def test_vote_accumulator_sums_views_on_the_device():
    """run_3d_seg.py:86-97: all_counts += used_count over views, here without leaving the device."""
    sc, cam_np, mask_np, ul, num_obj = cases.fsr_inputs("fsr_multi")
    cams = scene.overhead_cameras(cam_np["image_width"], cam_np["image_height"])
    acc = segment.VoteAccumulator(num_obj, sc["means3D"].shape[0])
    total = torch.zeros_like(acc.all_counts)
    g = {k: t(sc[k]) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
    for ci in (17, 5):
        cam = cams[ci]
        mask = t(scene.head_masks(sc, cam, n_labels=num_obj))
        rs = _settings(cam, np.zeros(3, np.float32))
        acc.add_view(mask, rs, g["means3D"], g["opacities"], shs=g["shs"], scales=g["scales"], rotations=g["rotations"])
        e = torch.Tensor([])
        single = ops.flash_rasterize_gaussians(mask, None, rs.bg, g["means3D"], e, g["opacities"], g["scales"],
                                               g["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                                               rs.tanfovy, rs.image_height, rs.image_width, g["shs"], 3, rs.campos,
                                               False, num_obj, False)[9]
        total += single
    acc.join()   # the views ran on the accumulator's own streams
    assert torch.allclose(acc.all_counts, total, rtol=1e-5, atol=1e-6)
    assert torch.equal(acc.labels(), segment.multi_instance_opt(acc.all_counts))


def test_vote_accumulator_keeps_dropped_inputs_alive_across_its_streams():
    """The usual loop `mask = load(v).cuda(); acc.add_view(mask, ...)` drops the previous mask while the view's kernels
    (on the accumulator's own stream) may not have read it yet.  Here the freed block is handed out again at once and
    filled with another label on the caller's stream; the votes must still be those of the original masks."""
    sc, cam_np, mask_np, ul, num_obj = cases.fsr_inputs("fsr_multi")
    cams = scene.overhead_cameras(cam_np["image_width"], cam_np["image_height"])
    g = {k: t(sc[k]) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
    P = sc["means3D"].shape[0]
    views = (17, 5, 9, 22, 30, 2)
    masks = [scene.head_masks(sc, cams[ci], n_labels=num_obj) for ci in views]

    def run(streams, clobber):
        acc = segment.VoteAccumulator(num_obj, P, streams=streams)
        for ci, m in zip(views, masks):
            mask = t(m)
            rs = _settings(cams[ci], np.zeros(3, np.float32))
            acc.add_view(mask, rs, g["means3D"], g["opacities"], shs=g["shs"], scales=g["scales"],
                         rotations=g["rotations"])
            if clobber:
                shape = mask.shape
                del mask
                junk = torch.full(shape, float(num_obj), dtype=torch.float32, device="cuda")  # same size: same block if free
                del junk
        acc.join()
        torch.cuda.synchronize()
        return acc.all_counts.clone()

    want = run(1, False)
    got = run(3, True)
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-6)
