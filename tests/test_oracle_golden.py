"""CPU: the oracle (oracle/splat_oracle.c) against the golden outputs of the unmodified reference
CUDA rasterizers (tests/golden/*.npz, made on a B200 by tests/golden/make_golden.py) and against
the reference's own Python multi_instance_opt (tests/golden/labels.npz).  This is what pins the
oracle; the GPU tests then compare libwheatsplat with the oracle AND with the same goldens."""
import os

import numpy as np
import pytest

import cases
import oracle
from util import GRAD_NAMES, assert_bits, assert_exact, assert_grad, assert_image


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


def _check_inputs(g, arrays):
    assert str(g["input_sha256"]) == cases.input_hash(arrays), "golden inputs regenerate differently on this machine"


@pytest.mark.parametrize("name", sorted(cases.DGR_CASES))
def test_forward_and_backward_match_reference(golden_dir, name):
    gold = _load(golden_dir, name)
    g, cam, lg, deg = cases.dgr_inputs(name)
    _check_inputs(gold, [g[k] for k in sorted(g) if isinstance(g[k], np.ndarray)]
                  + [cam["viewmatrix"], cam["projmatrix"], cam["campos"]])
    of = oracle.forward(g, cam, cases.BG, sh_degree=deg)
    assert of["num_rendered"] == int(gold["num_rendered"])
    assert_exact("radii", of["radii"], gold["radii"])
    vis = gold["radii"] > 0
    assert_bits("depths", of["depths"][vis], gold["depths"][vis])
    assert_bits("means2D", of["means2D"][vis], gold["means2D"].reshape(-1, 2)[vis])
    assert_bits("conic_opacity", of["conic_opacity"][vis], gold["conic_opacity"].reshape(-1, 4)[vis])
    assert_exact("tiles_touched", of["tiles_touched"], gold["tiles_touched"])
    assert_exact("point_list_keys", of["point_list_keys"], gold["point_list_keys"])
    assert_exact("point_list", of["point_list"], gold["point_list"])
    assert_exact("ranges", of["ranges"].reshape(-1), gold["ranges"])
    assert_exact("n_contrib", of["n_contrib"].reshape(-1), gold["n_contrib"])
    assert_image("rgb", of["rgb"][vis], gold["rgb"].reshape(-1, 3)[vis], 1e-5)
    assert_image("color", of["out_color"], gold["color"])
    assert_image("depth", of["out_depth"], gold["depth"])
    assert_image("alpha", of["out_alpha"], gold["alpha"])
    ob = oracle.backward(g, cam, cases.BG, of, lg["dL_dcolor"], lg["dL_ddepth"], lg["dL_dalpha"], sh_degree=deg)
    for k in GRAD_NAMES:
        ref = gold["g_" + k]
        assert_grad(k, ob[k].reshape(ref.shape) if ref.size else ref, ref, spread=float(gold["spread_" + k]))


@pytest.mark.parametrize("name", sorted(cases.FSR_CASES))
def test_flashsplat_votes_match_reference(golden_dir, name):
    gold = _load(golden_dir, name)
    sc, cam, mask, ul, num_obj = cases.fsr_inputs(name)
    _check_inputs(gold, [sc["means3D"], sc["shs"], mask, cam["projmatrix"]])
    of = oracle.forward(sc, cam, np.zeros(3, np.float32), sh_degree=3, gt_mask=mask, unique_label=ul, num_obj=num_obj)
    assert_exact("radii", of["radii"], gold["radii"])
    assert_exact("contrib_num", of["contrib_num"], gold["contrib_num"])
    assert_bits("image_xys", of["image_xys"], gold["image_xys"])
    assert_bits("gs_depth", of["gs_depth"], gold["gs_depth"])
    assert_image("color", of["out_color"], gold["color"])
    assert_image("alpha", of["out_alpha"], gold["alpha"])
    ru = gold["used_count"]
    assert_grad("used_count", of["used_count"], ru, spread=float(gold["spread_used_count"]), rtol=1e-4)
    # final label assignment: identical wherever the reference's own vote margin is above its noise
    lab_o, lab_r = oracle.multi_instance_opt(of["used_count"]), oracle.multi_instance_opt(ru)
    rest = ru.sum(axis=0, keepdims=True) - ru
    safe = (np.abs(ru - rest) > 1e-4 * (1.0 + np.abs(ru) + np.abs(rest))) | ((ru == 0) & (rest == 0))
    assert_exact("labels", lab_o[safe], lab_r[safe])
    assert safe.mean() > 0.99


def test_multi_instance_opt_matches_reference_python(golden_dir):
    gold = _load(golden_dir, "labels")
    a = cases.votes_inputs()
    _check_inputs(gold, [a])
    for key in gold.files:
        if key.startswith("labels_g"):
            assert_exact(key, oracle.multi_instance_opt(a, float(key[len("labels_g"):])), gold[key])


@pytest.mark.parametrize("name", sorted(cases.KNN_CASES))
def test_knn_matches_reference(golden_dir, name):
    gold = _load(golden_dir, name)
    pts = cases.knn_inputs(name)
    _check_inputs(gold, [pts])
    assert_bits("distCUDA2", oracle.knn_dist2(pts), gold["dist2"])


def test_sort_is_stable_and_matches_numpy():
    rng = np.random.default_rng(0)
    for n, bits in ((0, 40), (1, 40), (1000, 46), (70000, 43), (5000, 8)):
        keys = rng.integers(0, 1 << 62, size=n, dtype=np.uint64)
        keys = (keys >> np.uint64(20)) << np.uint64(20)  # many duplicates in the low sorted bits
        vals = np.arange(n, dtype=np.uint32)
        ko, vo = oracle.sort_pairs(keys, vals, bits)
        order = np.argsort(keys & np.uint64((1 << bits) - 1), kind="stable")
        assert_exact("keys", ko, keys[order])
        assert_exact("vals", vo, vals[order])


def test_empty_and_fully_culled_scene():
    g, cam, lg, deg = cases.dgr_inputs("dgr_deg1")
    g = dict(g)
    g["means3D"] = g["means3D"].copy()
    g["means3D"][:, 2] += 100.0  # behind every overhead camera
    of = oracle.forward(g, cam, cases.BG, sh_degree=deg)
    assert of["num_rendered"] == 0 and not of["radii"].any()
    H, W = cam["image_height"], cam["image_width"]
    assert_image("color", of["out_color"], np.broadcast_to(cases.BG[:, None, None], (3, H, W)))
    assert not of["out_alpha"].any()


@pytest.mark.parametrize("name", sorted(cases.LOSS_CASES))
def test_loss_oracle_matches_reference_python(golden_dir, name):
    """oracle/loss_oracle.py against the reference's own utils/loss_utils.py + autograd on CPU
    (tests/golden/make_golden_loss.py)."""
    import loss_oracle
    gold = _load(golden_dir, "loss_" + name)
    img, gt, lam = cases.loss_inputs(name)
    _check_inputs(gold, [img, gt])
    o = loss_oracle.loss_and_grad(img, gt, lam)
    assert abs(o["loss"] - float(gold["loss"])) <= 2e-6
    assert abs(o["l1"] - float(gold["l1"])) <= 2e-6
    assert abs(o["ssim"] - float(gold["ssim"])) <= 1e-5
    assert_grad("dloss/dimage", o["grad"], gold["grad"])
    if "ssim_per_image" in gold.files:
        per = o["ssim_per_plane"].reshape(img.shape[0], -1).mean(axis=1)
        assert np.abs(per - gold["ssim_per_image"]).max() <= 1e-5


def _optim_close(name, a, b):
    """Optimiser tier: float32, ~1 ulp per operation; exp_avg cancels (g - m), hence the absolute floor."""
    a, b = np.asarray(a, np.float32).reshape(-1), np.asarray(b, np.float32).reshape(-1)
    lim = 2e-6 * np.abs(b) + 2e-7 * np.abs(b).max()
    bad = np.abs(a - b) > lim
    assert not bad.any(), "%s: %d elements off, max err %g" % (name, bad.sum(), np.abs(a - b).max())


def test_optim_oracle_matches_torch_adam_and_reference_stats(golden_dir):
    """oracle/optim_oracle.py against torch.optim.Adam as GaussianModel.training_setup configures it and the
    reference's own add_densification_stats (tests/golden/make_golden_optim.py)."""
    import optim_oracle
    gold = _load(golden_dir, "optim")
    params, grads, radii, g2d = cases.optim_inputs()
    assert str(gold["input_sha256"]) == cases.input_hash([params[n] for n, _, _ in cases.OPTIM_GROUPS])
    P = cases.OPTIM_P
    st = {n: (params[n], np.zeros_like(params[n]), np.zeros_like(params[n])) for n, _, _ in cases.OPTIM_GROUPS}
    mr, acc, dn = (np.zeros(P, np.float32) for _ in range(3))
    for it in range(cases.OPTIM_STEPS):
        mr, acc, dn = optim_oracle.densify_stats(radii[it], g2d[it], mr, acc, dn)
        assert np.array_equal(mr, gold["step%d_max_radii2D" % it])
        assert np.array_equal(dn, gold["step%d_denom" % it].reshape(-1))
        _optim_close("xyz_gradient_accum", acc, gold["step%d_xyz_gradient_accum" % it])
        for n, _, lr in cases.OPTIM_GROUPS:
            lr = cases.OPTIM_XYZ_LRS[it] if n == "xyz" else lr
            st[n] = optim_oracle.adam_step(st[n][0], grads[it][n], st[n][1], st[n][2], it + 1, lr)
            for suffix, a in zip(("", "_exp_avg", "_exp_avg_sq"), st[n]):
                _optim_close("step %d %s%s" % (it, n, suffix), a, gold["step%d_%s%s" % (it, n, suffix)])


def test_segment_oracle_matches_reference_eval_obj_labels(golden_dir):
    """oracle/segment_oracle.py against the reference: per-object alpha/depth from the unmodified reference
    flashsplat rasterizer (B200) and the pred_mask its own eval_obj_labels paints from them."""
    import segment_oracle
    gold = _load(golden_dir, "objeval")
    sc, cam, labels = cases.objeval_inputs()
    keys = ("means3D", "opacities", "scales", "rotations", "shs")
    assert str(gold["input_sha256"]) == cases.input_hash([sc[k] for k in keys] + [labels.astype(np.uint8)])
    pred, renders = segment_oracle.eval_obj_labels(labels, sc, cam, cases.BG, sh_degree=3, return_renders=True)
    assert sorted(renders) == gold["rendered_rows"].tolist()
    for i, (alpha, depth) in renders.items():
        assert np.abs(alpha - gold["alpha_%d" % i]).max() <= 1e-4
        assert np.abs(depth - gold["depth_%d" % i]).max() <= 1e-4
    assert np.array_equal(pred, gold["pred_mask"])
    # the merge alone, fed the reference's own renders
    again = segment_oracle.merge({i: (gold["alpha_%d" % i], gold["depth_%d" % i]) for i in renders}, pred.shape)
    assert np.array_equal(again, gold["pred_mask"])


DENSIFY_EXACT = ("f_dc", "f_rest", "opacity", "rotation", "which_object")


def _densify_before(gold):
    return {k[len("before/"):]: gold[k] for k in gold.files if k.startswith("before/") and "/step_" not in k}


@pytest.mark.parametrize("tag", list(cases.DENSIFY_CASES))
def test_densify_oracle_matches_reference_class(golden_dir, tag):
    """oracle/densify_oracle.py against the UNMODIFIED reference GaussianModel.densify_and_prune
    (tests/golden/make_golden_densify.py): same rows in the same order, moments of new rows zero, statistics reset.
    Copied rows bit-exact; xyz / scaling of split children pass through exp/log/matmul: 4 ulp."""
    import densify_oracle
    gold = _load(golden_dir, "densify")
    kw = cases.DENSIFY_CASES[tag]
    inp = cases.densify_inputs()
    # "before" = the seeded inputs after the one Adam step that fills the moments
    assert np.array_equal(gold["before/denom"], inp["denom"]) and np.array_equal(gold["before/which_object"], inp["which_object"])
    st, src, S = densify_oracle.densify_and_prune(_densify_before(gold), gold[tag + "/samples"], kw["max_grad"],
                                                  kw["min_opacity"], kw["extent"], kw["max_screen_size"],
                                                  cases.DENSIFY_TRAIN_ARGS["percent_dense"])
    assert 2 * S == gold[tag + "/samples"].shape[0] and S > 0
    rows = gold[tag + "/xyz"].shape[0]
    assert src.shape == (rows,) and rows != inp["xyz"].shape[0]
    for n in densify_oracle.NAMES:
        for pre in ("", "m_", "v_"):
            a, b = st[pre + n], gold["%s/%s%s" % (tag, pre, n)]
            assert a.shape == b.shape, (pre + n, a.shape, b.shape)
            if n in DENSIFY_EXACT or pre:
                assert np.array_equal(a, b), pre + n
            else:
                assert np.allclose(a, b, rtol=5e-7, atol=1e-7), (pre + n, np.abs(a - b).max())
        assert float(gold["%s/step_%s" % (tag, n)]) == float(gold["before/step_" + n]) == 1.0
    assert np.array_equal(st["which_object"], gold[tag + "/which_object"])
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        assert st[k].shape == gold[tag + "/" + k].shape and not st[k].any() and not gold[tag + "/" + k].any()
    # the structure the fused plan relies on: survivors in order, then clones, then the two child blocks
    P = inp["xyz"].shape[0]
    first_new = int(np.argmax(np.diff(src) < 0)) + 1 if (np.diff(src) < 0).any() else rows
    assert (np.diff(src[:first_new]) > 0).all() and src.max() < P
    new_rows = ~st["m_xyz"].any(axis=1)
    assert new_rows[first_new:].all()


def test_reset_opacity_oracle_matches_reference_class(golden_dir):
    import densify_oracle
    gold = _load(golden_dir, "densify")
    got = densify_oracle.reset_opacity(gold["before/opacity"])
    assert np.allclose(got, gold["reset/opacity"], rtol=2e-6, atol=1e-6)
    assert gold["before/m_opacity"].any() and not gold["reset/m_opacity"].any() and not gold["reset/v_opacity"].any()
    assert (1 / (1 + np.exp(-got)) <= 0.0100001).all()
