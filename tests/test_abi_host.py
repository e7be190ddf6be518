"""CPU: the C-ABI library loads and exports every symbol include/wheatsplat.h declares; argument
validation that needs no GPU; the host-side mirror of the reference's Python interface."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import wheat_3dgs_b200
from wheat_3dgs_b200 import dist as wdist
from wheat_3dgs_b200 import lib as wlib
from wheat_3dgs_b200 import rasterizer, scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "wheatsplat.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(ws_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 18
    L = wlib.load()
    for name in declared:
        assert hasattr(L, name), "libwheatsplat.so does not export %s" % name
    assert sorted(wlib.SYMBOLS) == declared, "lib.py SYMBOLS out of sync with the header"
    assert L.ws_abi_version() == wlib.WS_ABI_VERSION


def test_scratch_sizes():
    L = wlib.load()
    assert L.ws_geom_bytes(0) > 0
    assert L.ws_geom_bytes(1000000) >= 1000000 * (64 + 4 + 1 + 48)   # 64-byte rows, tiles_touched, clamped, acc
    assert L.ws_geom_bytes(2000) > L.ws_geom_bytes(1000)
    assert L.ws_image_bytes(2048, 1536) >= 2048 * 1536 * 4 + 12288 * 8
    # keys + values ping-pong only: no per-instance record stream (the blend kernels gather the rows themselves)
    assert (1 << 20) * 24 <= L.ws_binning_bytes(1 << 20, 64, 64) < (1 << 20) * 40
    assert L.ws_binning_bytes(0, 64, 64) > 0
    assert L.ws_sort_bytes(1 << 20) > 0 and L.ws_knn_bytes(1000) > 0


def test_invalid_arguments_are_rejected_without_touching_the_gpu():
    L = wlib.load()
    v, g = wlib.WsView(), wlib.WsGaussians()
    assert L.ws_forward_prepare(None, None, None, None, 0, None, None) == -1
    assert b"NULL" in L.ws_last_error()
    g.P = 0
    assert L.ws_forward_prepare(C.byref(v), C.byref(g), None, None, 0, None, None) == -1
    g.P, v.image_width, v.image_height = 10, 64, 64
    assert L.ws_forward_prepare(C.byref(v), C.byref(g), None, None, 0, None, None) == -1  # NULL view tensors
    assert L.ws_mark_visible(5, None, None, None, None, None) == -1
    assert L.ws_mark_visible(0, None, None, None, None, None) == 0
    assert L.ws_knn_dist2(0, None, None, None, 0, None) == 0
    assert L.ws_sort_pairs_u64(-1, None, None, None, None, 40, None, 0, None) == -1
    assert L.ws_sort_pairs_u32(-1, None, None, None, None, 14, None, 0, None) == -1
    assert L.ws_multi_instance_opt(0, 0, None, 0.0, None, None) == 0
    assert L.ws_debug_fetch(b"nope", 1, 16, 16, 0, None, None, None, None, 0, None) == -1


def test_densify_and_sharded_step_reject_bad_arguments_on_the_host():
    """Argument checks of the densification and fused exchange+Adam entry points run before any CUDA call."""
    L = wlib.load()
    counts = (C.c_uint32 * 4)()
    assert L.ws_densify_plan_bytes(1000) > 2 * 3 * 1000 * 4
    assert L.ws_densify_plan(0, None, None, None, None, 0.0002, 0.005, 0.02, 0.2, 0, 0.0, None, None, 0, counts, None) == 0
    assert L.ws_densify_plan(-1, None, None, None, None, 0.0002, 0.005, 0.02, 0.2, 0, 0.0, None, None, 0, counts, None) == -1
    assert L.ws_densify_plan(10, None, None, None, None, 0.0002, 0.005, 0.02, 0.2, 0, 0.0, None, None, 0, counts, None) == -1
    assert L.ws_densify_plan(10, None, None, None, None, 0.0002, 0.005, 0.02, 0.2, 0, 0.0, None, None, 0, None, None) == -1
    assert L.ws_densify_gather(0, None, None, 0, None, None, None, 0, None, None, None) == 0
    assert L.ws_reset_opacity(0, None, None, None, None) == 0
    assert L.ws_reset_opacity(5, None, None, None, None) == -1
    seg = (wlib.WsShardSegment * 2)(wlib.WsShardSegment(0, 8, 0, 1e-3), wlib.WsShardSegment(12, 16, 8, 1e-3))   # gap at 8..12
    peers = (C.c_void_p * 2)(16, 32)
    m = C.c_void_p(64)
    assert L.ws_sharded_adam_step(0, 2, 16, 2, seg, None, peers, None, peers, m, m, 1, 0.9, 0.999, 1e-15, 0, None) == -1
    assert b"tile" in L.ws_last_error()
    seg[1].begin = 8
    assert L.ws_sharded_adam_step(0, 2, 16, 2, seg, None, peers, None, peers, m, m, 0, 0.9, 0.999, 1e-15, 0, None) == -1  # step 0
    assert L.ws_sharded_adam_step(2, 2, 16, 2, seg, None, peers, None, peers, m, m, 1, 0.9, 0.999, 1e-15, 0, None) == -1  # rank
    assert L.ws_sharded_adam_step(0, 2, 18, 2, seg, None, peers, None, peers, m, m, 1, 0.9, 0.999, 1e-15, 0, None) == -1  # % 4
    assert L.ws_sharded_adam_step(0, 2, 0, 2, seg, None, peers, None, peers, m, m, 1, 0.9, 0.999, 1e-15, 0, None) == 0


def test_both_or_neither_of_an_either_or_pair_is_an_error():
    L = wlib.load()
    v, g = wlib.WsView(), wlib.WsGaussians()
    dummy = 0x1000  # never dereferenced: validation fails first
    v.image_width = v.image_height = 64
    v.bg = v.viewmatrix = v.projmatrix = v.campos = dummy
    g.P, g.M, g.means3D, g.opacities = 10, 16, dummy, dummy
    g.scales = g.rotations = dummy
    assert L.ws_forward_prepare(C.byref(v), C.byref(g), dummy, dummy, 1 << 30, None, None) == -1  # no colour source
    assert b"exactly one of SHs or precomputed colors" in L.ws_last_error()
    g.shs = g.colors_precomp = dummy
    assert L.ws_forward_prepare(C.byref(v), C.byref(g), dummy, dummy, 1 << 30, None, None) == -1
    g.colors_precomp, g.cov3D_precomp = None, dummy
    assert L.ws_forward_prepare(C.byref(v), C.byref(g), dummy, dummy, 1 << 30, None, None) == -1
    assert b"scale/rotation pair or precomputed 3D covariance" in L.ws_last_error()
    g.cov3D_precomp = None
    v.sh_degree = 4
    assert L.ws_forward_prepare(C.byref(v), C.byref(g), dummy, dummy, 1 << 30, None, None) == -1
    v.sh_degree = 3
    g.M = 4  # degree 3 needs 16 coefficients
    assert L.ws_forward_prepare(C.byref(v), C.byref(g), dummy, dummy, 1 << 30, None, None) == -1
    g.M = 16
    assert L.ws_forward_prepare(C.byref(v), C.byref(g), dummy, dummy, 16, None, None) == -2  # scratch too small


def test_dropin_packages_expose_the_reference_names():
    wheat_3dgs_b200.install()
    import diff_gaussian_rasterization as dgr
    import flashsplat_rasterization as fsr
    from simple_knn._C import distCUDA2  # noqa: F401  (scene/gaussian_model.py:20)
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")
    assert fsr.GaussianRasterizationSettings._fields == dgr.GaussianRasterizationSettings._fields + ("mask_grad", "num_obj")
    for mod in (dgr, fsr):
        assert issubclass(mod.GaussianRasterizer, torch.nn.Module)
        assert callable(mod.rasterize_gaussians)
        for f in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
            assert callable(getattr(mod._C, f))


def test_rasterizer_keeps_the_reference_exception_texts():
    rs = rasterizer.GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                                  torch.zeros(3), False, False)
    r = rasterizer.GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(m, m, torch.zeros(4, 1), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))


def test_product_has_no_cpu_path():
    from wheat_3dgs_b200 import ops
    m = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.rasterize_gaussians(torch.zeros(3), m, torch.Tensor([]), torch.zeros(4, 1), m, torch.zeros(4, 4), 1.0,
                                torch.Tensor([]), torch.eye(4), torch.eye(4), 1.0, 1.0, 16, 16, torch.zeros(4, 1, 3), 0,
                                torch.zeros(3), False, False)
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        ops.rasterize_gaussians(torch.zeros(3), torch.zeros(4), torch.Tensor([]), torch.zeros(4, 1), m, torch.zeros(4, 4),
                                1.0, torch.Tensor([]), torch.eye(4), torch.eye(4), 1.0, 1.0, 16, 16, torch.zeros(4, 1, 3),
                                0, torch.zeros(3), False, False)


def test_product_sources_never_import_the_oracle():
    pkg = os.path.join(ROOT, "wheat_3dgs_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", txt, flags=re.M), f
                assert "splat_oracle" not in txt or f.endswith((".cu", ".cuh")) and "see oracle/splat_oracle.c" in txt, f


def test_scene_generator_is_deterministic_and_reference_shaped():
    a, b = scene.synthetic_scene(3000, seed=0), scene.synthetic_scene(3000, seed=0)
    for k in ("means3D", "scales", "rotations", "opacities", "shs"):
        assert a[k].dtype == np.float32 and np.array_equal(a[k], b[k])
    assert a["shs"].shape == (3000, 16, 3) and a["opacities"].shape == (3000, 1)
    assert np.allclose(np.linalg.norm(a["rotations"], axis=1), 1.0, atol=1e-6)
    cams = scene.overhead_cameras(2048, 1536)
    assert len(cams) == 36
    c = cams[0]
    # full projection = view (transposed) @ projection (transposed), scene/cameras.py:57-58
    proj_t = scene.projection_matrix(0.01, 100.0, c["FoVx"], c["FoVy"]).T
    assert np.allclose(c["projmatrix"], c["viewmatrix"] @ proj_t, atol=1e-6)
    assert len(scene.orbit_cameras(1920, 1080)) == 200


def test_view_sharding_and_bucket_layout():
    assert wdist.shard(range(36), 1, 8) == [1, 9, 17, 25, 33]
    assert sum(wdist.shard_counts(36, 8)) == 36 and sum(wdist.shard_counts(200, 8)) == 200
    units = [(v, m) for v in range(3) for m in range(5)]
    got = sorted(sum((wdist.shard(units, r, 4) for r in range(4)), []))
    assert got == units
    P, M = 7, 16
    assert wdist.train_bucket_numel(P, M) == 62 * P  # 62 floats per Gaussian at SH degree 3
    b = torch.arange(wdist.full_bucket_numel(P, M), dtype=torch.float32)
    s = wdist.split_bucket(b, P, M)
    assert s["dL_dmeans3D"].shape == (P, 3) and s["dL_dsh"].shape == (P, M, 3) and s["dL_dcov3D"].shape == (P, 6)
    assert s["dL_dmeans3D"][0, 0] == 0 and s["dL_dmeans2D"][0, 0] == 3 * P and s["dL_dcolors"][0, 0] == 62 * P
    assert wdist.world_size() == 1 and wdist.allreduce_train_grads(b, P, M) is None


def test_balanced_schedule_groups_views_of_similar_cost():
    from wheat_3dgs_b200 import dist as wdist
    costs = [5.0, 9.0, 1.0, 7.0, 3.0, 8.0, 2.0, 6.0]
    sched = wdist.balanced_schedule(costs, 4, 6)
    assert sched[0] == [1, 5, 3, 7] and sched[1] == [0, 4, 6, 2] and sched[2] == sched[0]
    # every view is visited equally often over a whole number of cycles, one view per rank and step
    flat = [v for step in wdist.balanced_schedule(costs, 3, 8) for v in step]
    assert sorted(flat) == sorted(list(range(8)) * 3)
    # the spread of costs inside a step is no larger than that of round robin
    spread = lambda s: max(max(costs[v] for v in st) - min(costs[v] for v in st) for st in s)
    rr = [[(r + i * 4) % 8 for r in range(4)] for i in range(2)]
    assert spread(sched[:2]) <= spread(rr)
