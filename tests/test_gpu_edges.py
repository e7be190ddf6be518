"""GPU (-m gpu): the radix sort on its own, edge cases (empty, fully culled, single Gaussian, ragged
image sizes, one Gaussian covering every tile), size-independent properties at BASELINE.json's
full config-2 size, and the reference run live beside libwheatsplat when oracle/_ref travelled."""
import numpy as np
import pytest
import torch

import cases
import oracle
from util import GRAD_NAMES, assert_bits, assert_exact, assert_grad, assert_image
from wheat_3dgs_b200 import scene

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import gpu_util as G
    from wheat_3dgs_b200 import ops
    import ref as refmod


@pytest.mark.parametrize("n,bits", [(0, 40), (1, 40), (31, 46), (3072, 46), (3073, 43), (100000, 46), (777, 8),
                                    (50000, 33), (1 << 20, 47), (5000001, 46), (12345, 64)])
def test_radix_sort_bit_exact_and_stable(n, bits):
    gen = torch.Generator().manual_seed(n + bits)
    keys = torch.randint(0, 1 << 62, (n,), generator=gen, dtype=torch.int64)
    if bits < 64:
        low = (keys & ((1 << bits) - 1)) >> (bits // 3) << (bits // 3)  # many equal keys -> exercises stability
        keys = low | (keys & ~((1 << bits) - 1))                          # garbage above end_bit must be ignored
    vals = torch.arange(n, dtype=torch.int32)
    ko, vo = ops.sort_pairs(keys.cuda(), vals.cuda(), bits)
    torch.cuda.synchronize()
    rk, rv = oracle.sort_pairs(keys.numpy().view(np.uint64), vals.numpy().view(np.uint32), bits)
    assert_exact("keys", G.n(ko).view(np.uint64), rk)
    assert_exact("vals", G.n(vo).view(np.uint32), rv)


@pytest.mark.parametrize("n,bits", [(0, 14), (1, 32), (31, 14), (3072, 32), (3073, 15), (4096, 14), (4097, 21), (100000, 14), (777, 1),
                                    (1 << 20, 32), (5000001, 14), (12345, 6)])
def test_radix_sort_u32_bit_exact_and_stable(n, bits):
    # the form depth-first binning uses: Gaussians by 32 depth bits, instances by tile id
    gen = torch.Generator().manual_seed(7 * n + bits)
    keys = torch.randint(0, 1 << 31, (n,), generator=gen, dtype=torch.int64)
    if bits < 32:
        low = (keys & ((1 << bits) - 1)) >> (bits // 3) << (bits // 3)
        keys = low | (keys & ~((1 << bits) - 1))
    keys = keys.to(torch.int32)
    vals = torch.arange(n, dtype=torch.int32)
    ko, vo = ops.sort_pairs(keys.cuda(), vals.cuda(), bits)
    torch.cuda.synchronize()
    k = keys.numpy().view(np.uint32)
    order = np.argsort(k & np.uint32((1 << bits) - 1 if bits < 32 else 0xffffffff), kind="stable")
    assert_exact("keys", G.n(ko).view(np.uint32), k[order])
    assert_exact("vals", G.n(vo).view(np.uint32), vals.numpy().view(np.uint32)[order])


def test_sort_all_equal_and_already_sorted():
    n = 200000
    for keys in (torch.full((n,), 0x1234_0000_0001, dtype=torch.int64), torch.arange(n, dtype=torch.int64) << 20,
                 torch.arange(n, 0, -1, dtype=torch.int64) << 7):
        vals = torch.arange(n, dtype=torch.int32)
        ko, vo = ops.sort_pairs(keys.cuda(), vals.cuda(), 46)
        order = np.argsort(keys.numpy().view(np.uint64) & np.uint64((1 << 46) - 1), kind="stable")
        assert_exact("vals", G.n(vo), vals.numpy()[order])
        assert_exact("keys", G.n(ko), keys.numpy()[order])


def _small(P=300, W=64, H=48, seed=5):
    sc = scene.synthetic_scene(P, seed=seed, scale_median=0.01)
    cam = scene.overhead_cameras(W, H)[17]
    return sc, cam


def test_empty_input_launches_nothing():
    sc, cam = _small()
    gd, cd = G.dev_inputs({k: v[:0] for k, v in sc.items() if isinstance(v, np.ndarray) and v.ndim > 0 and k != "head_centers"}, cam)
    n0 = ops.launch_count()
    R, color, depth, alpha, radii = G.forward(gd, cd, G.t(cases.BG), 3)[:5]
    assert R == 0 and radii.numel() == 0 and ops.launch_count() == n0
    assert float(color.abs().max()) == 0.0  # DGR rasterize_points.cu:82: P == 0 leaves the zero-filled outputs


def test_fully_culled_scene_renders_background():
    sc, cam = _small()
    sc = dict(sc)
    sc["means3D"] = sc["means3D"] + np.array([0, 0, 100.0], np.float32)
    gd, cd = G.dev_inputs(sc, cam)
    fw = G.forward(gd, cd, G.t(cases.BG), 3)
    R, color, depth, alpha, radii = fw[:5]
    assert R == 0 and int(radii.abs().sum()) == 0
    H, W = cam["image_height"], cam["image_width"]
    assert_exact("color", G.n(color), np.broadcast_to(cases.BG[:, None, None], (3, H, W)))
    assert float(alpha.abs().max()) == 0.0 and float(depth.abs().max()) == 0.0
    z = torch.ones_like(color)
    grads = G.backward(gd, cd, G.t(cases.BG), 3, fw, z, z[:1], z[:1])
    for k, t_ in zip(GRAD_NAMES, grads):
        assert float(t_.abs().max()) == 0.0, k


@pytest.mark.parametrize("P,W,H", [(1, 16, 16), (1, 17, 33), (5, 1, 1), (257, 15, 250), (300, 333, 7)])
def test_ragged_sizes_match_oracle(P, W, H):
    sc, cam = _small(P, W, H)
    gd, cd = G.dev_inputs(sc, cam)
    lg = scene.loss_gradients((3, H, W), seed=9)
    fw = G.forward(gd, cd, G.t(cases.BG), 3)
    grads = G.backward(gd, cd, G.t(cases.BG), 3, fw, G.t(lg["dL_dcolor"]), G.t(lg["dL_ddepth"]), G.t(lg["dL_dalpha"]))
    of = oracle.forward(sc, cam, cases.BG, sh_degree=3)
    ob = oracle.backward(sc, cam, cases.BG, of, lg["dL_dcolor"], lg["dL_ddepth"], lg["dL_dalpha"], sh_degree=3)
    assert fw[0] == of["num_rendered"]
    assert_exact("radii", G.n(fw[4]), of["radii"])
    assert_exact("point_list", G.fetch("point_list", fw, P, W, H).view(np.uint32), of["point_list"])
    assert_exact("ranges", G.fetch("ranges", fw, P, W, H).view(np.uint32), of["ranges"])
    assert_exact("n_contrib", G.fetch("n_contrib", fw, P, W, H).view(np.uint32).reshape(H, W), of["n_contrib"])
    assert_image("color", G.n(fw[1]), of["out_color"])
    assert_image("alpha", G.n(fw[3]), of["out_alpha"])
    for k, mine in zip(GRAD_NAMES, grads):
        assert_grad(k, G.n(mine).reshape(ob[k].shape), ob[k])


def test_giant_gaussian_touches_every_tile_and_debug_mode():
    """One Gaussian larger than the image: maximum tile rectangle, one instance per tile; plus many
    coincident Gaussians (equal depth keys -> the sort's stability decides point_list)."""
    W, H, P = 200, 120, 40
    sc, cam = _small(P, W, H)
    sc = dict(sc)
    sc["means3D"] = sc["means3D"].copy()
    sc["scales"] = sc["scales"].copy()
    sc["means3D"][:] = np.array([0.0, 0.0, 0.3], np.float32)  # all coincident: identical depth bits
    sc["scales"][0] = 5.0
    sc["scales"][1:] = 0.02
    gd, cd = G.dev_inputs(sc, cam)
    fw = G.forward(gd, cd, G.t(cases.BG), 3, debug=True)
    of = oracle.forward(sc, cam, cases.BG, sh_degree=3)
    T = ((W + 15) // 16) * ((H + 15) // 16)
    assert int(G.fetch("tiles_touched", fw, P, W, H).view(np.uint32)[0]) == T
    assert fw[0] == of["num_rendered"]
    assert_exact("point_list", G.fetch("point_list", fw, P, W, H).view(np.uint32), of["point_list"])
    assert_exact("n_contrib", G.fetch("n_contrib", fw, P, W, H).view(np.uint32).reshape(H, W), of["n_contrib"])
    assert_image("color", G.n(fw[1]), of["out_color"])


def test_sh_degree_zero_with_single_coefficient_and_scale_modifier():
    sc, cam = _small(500, 96, 80)
    sc = dict(sc)
    sc["shs"] = np.ascontiguousarray(sc["shs"][:, :1])  # M = 1: the scalar SH staging path (3M % 4 != 0)
    gd, cd = G.dev_inputs(sc, cam)
    lg = scene.loss_gradients((3, 80, 96), seed=4)
    fw = G.forward(gd, cd, G.t(cases.BG), 0, scale_modifier=0.7)
    grads = G.backward(gd, cd, G.t(cases.BG), 0, fw, G.t(lg["dL_dcolor"]), G.t(lg["dL_ddepth"]), G.t(lg["dL_dalpha"]),
                       scale_modifier=0.7)
    of = oracle.forward(sc, cam, cases.BG, sh_degree=0, scale_modifier=0.7)
    ob = oracle.backward(sc, cam, cases.BG, of, lg["dL_dcolor"], lg["dL_ddepth"], lg["dL_dalpha"], sh_degree=0,
                         scale_modifier=0.7)
    assert_exact("radii", G.n(fw[4]), of["radii"])
    assert_image("color", G.n(fw[1]), of["out_color"])
    for k, mine in zip(GRAD_NAMES, grads):
        assert_grad(k, G.n(mine).reshape(ob[k].shape), ob[k])


def _cfg2():
    P, W, H = 1000000, 2048, 1536
    sc = scene.synthetic_scene(P, seed=0)
    cam = scene.overhead_cameras(W, H)[17]
    return sc, cam, P, W, H


def test_full_size_properties_config2():
    """BASELINE.json config 2 (1 M Gaussians, 2048x1536): properties that need no oracle run."""
    sc, cam, P, W, H = _cfg2()
    gd, cd = G.dev_inputs(sc, cam)
    bg = G.t(cases.BG)
    fw = G.forward(gd, cd, bg, 3)
    R, color, depth, alpha, radii, geom, binning, image = fw
    T = ((W + 15) // 16) * ((H + 15) // 16)
    tt = G.fetch("tiles_touched", fw, P, W, H).view(np.uint32)
    assert int(tt.astype(np.int64).sum()) == R
    assert_exact("tiles>0 <=> radii>0", tt > 0, G.n(radii) > 0)
    keys = ops.debug_fetch("point_list_keys", P, W, H, R, geom, binning, image)
    plist = ops.debug_fetch("point_list", P, W, H, R, geom, binning, image).long()
    assert bool((keys[1:] >= keys[:-1]).all()), "keys not sorted"
    same = keys[1:] == keys[:-1]
    assert bool((plist[1:][same] > plist[:-1][same]).all()), "equal keys out of emission order (unstable sort)"
    # the sorted list is a permutation of the emitted (tile, depth, id) instances
    depths = ops.debug_fetch("depths", P, W, H, R, geom, binning, image)
    assert bool(((keys & 0xFFFFFFFF).int() == depths.view(torch.int32)[plist]).all())
    assert int(torch.bincount(plist, minlength=P).max()) <= T
    assert_exact("instances per Gaussian", G.n(torch.bincount(plist, minlength=P)), tt.astype(np.int64))
    # ranges partition the list by tile id
    rng_ = ops.debug_fetch("ranges", P, W, H, R, geom, binning, image).long()
    tiles = (keys >> 32)
    cnt = torch.bincount(tiles, minlength=T)
    assert bool(((rng_[:, 1] - rng_[:, 0]) == cnt).all())
    nz = cnt > 0
    assert bool((tiles[rng_[nz, 0]] == torch.nonzero(nz).squeeze(1)).all())
    a = G.n(alpha)
    assert a.min() >= 0.0 and a.max() <= 1.0 and np.isfinite(G.n(color)).all()
    ncon = ops.debug_fetch("n_contrib", P, W, H, R, geom, binning, image).long().view(H, W)
    tile_len = (rng_[:, 1] - rng_[:, 0]).view((H + 15) // 16, (W + 15) // 16)
    per_pix_len = tile_len.repeat_interleave(16, 0).repeat_interleave(16, 1)[:H, :W]
    assert bool((ncon <= per_pix_len).all())
    # backward is linear in the upstream gradients
    lg = scene.loss_gradients((3, H, W), seed=2)
    dc, dd, da = G.t(lg["dL_dcolor"]), G.t(lg["dL_ddepth"]), G.t(lg["dL_dalpha"])
    g1 = [x.clone() for x in G.backward(gd, cd, bg, 3, fw, dc, dd, da)]
    g1b = [x.clone() for x in G.backward(gd, cd, bg, 3, fw, dc, dd, da)]  # own run-to-run spread (float atomics)
    g2 = G.backward(gd, cd, bg, 3, fw, 2 * dc, 2 * dd, 2 * da)
    for k, x1, x1b, x2 in zip(GRAD_NAMES, g1, g1b, g2):
        assert_grad(k + " linearity", G.n(x2), 2.0 * G.n(x1), spread=2.0 * float((x1 - x1b).abs().max()))
    inv = G.n(radii) == 0
    for k, x1 in zip(GRAD_NAMES, g1):
        assert float(x1.view(P, -1)[torch.from_numpy(inv).cuda()].abs().max()) == 0.0, k + ": culled Gaussians must get zero"
    # determinism of the bit-exact tier: a second forward gives the same list
    fw2 = G.forward(gd, cd, bg, 3)
    assert bool((ops.debug_fetch("point_list", P, W, H, fw2[0], fw2[5], fw2[6], fw2[7]).long() == plist).all())
    assert bool((fw2[1] == color).all())


@pytest.mark.skipif(not (torch.cuda.is_available() and refmod.available("dgr")), reason="oracle/_ref not built")
@pytest.mark.parametrize("cfg", ["cfg1", "cfg2"])
def test_against_live_reference(cfg):
    """The unmodified reference CUDA rasterizer run beside libwheatsplat on config 1 / config 2."""
    if cfg == "cfg1":
        P, W, H = 50000, 512, 512
        sc, cam = scene.synthetic_scene(P, seed=0), scene.overhead_cameras(W, H)[17]
    else:
        sc, cam, P, W, H = _cfg2()
    gd, cd = G.dev_inputs(sc, cam)
    bg = G.t(cases.BG)
    lg = scene.loss_gradients((3, H, W), seed=2)
    dc, dd, da = G.t(lg["dL_dcolor"]), G.t(lg["dL_ddepth"]), G.t(lg["dL_dalpha"])
    rr = refmod.RefRasterizer("dgr")
    ro = rr.forward(gd, cd, bg, 3)
    torch.cuda.synchronize()
    fw = G.forward(gd, cd, bg, 3)
    assert fw[0] == ro["num_rendered"]
    assert_exact("radii", G.n(fw[4]), G.n(ro["radii"]))
    assert_exact("point_list_keys", G.fetch("point_list_keys", fw, P, W, H).view(np.uint64), rr.fetch("point_list_keys"))
    assert_exact("point_list", G.fetch("point_list", fw, P, W, H).view(np.uint32), rr.fetch("point_list"))
    assert_exact("ranges", G.fetch("ranges", fw, P, W, H).view(np.uint32).reshape(-1), rr.fetch("ranges"))
    assert_exact("n_contrib", G.fetch("n_contrib", fw, P, W, H).view(np.uint32), rr.fetch("n_contrib"))
    for k, i in (("color", 1), ("depth", 2), ("alpha", 3)):
        assert_image(k, G.n(fw[i]), G.n(ro[k]))
    r1 = {k: v.clone() for k, v in rr.backward(dc, dd, da).items()}
    torch.cuda.synchronize()
    r2 = {k: v.clone() for k, v in rr.backward(dc, dd, da).items()}
    torch.cuda.synchronize()
    r3 = {k: v.clone() for k, v in rr.backward(dc, dd, da).items()}
    torch.cuda.synchronize()
    mine = G.backward(gd, cd, bg, 3, fw, dc, dd, da)
    for k, m in zip(GRAD_NAMES, mine):
        # the reference's own run-to-run spread (its float atomics), the largest of three runs' pairwise differences
        spread = float(max((r1[k] - r2[k]).abs().max(), (r1[k] - r3[k]).abs().max(), (r2[k] - r3[k]).abs().max()))
        assert_grad(k, G.n(m).reshape(-1), G.n(r1[k]).reshape(-1), spread=spread)


@pytest.mark.skipif(not (torch.cuda.is_available() and refmod.available("knn")), reason="oracle/_ref not built")
def test_knn_against_live_reference_300k():
    rng = np.random.default_rng(3)
    pts = (rng.normal(size=(300000, 3)) * np.array([1.0, 0.5, 0.2])).astype(np.float32)
    assert_bits("distCUDA2", G.n(ops.distCUDA2(G.t(pts))), G.n(refmod.knn_dist2(G.t(pts))))
