"""numpy front-end of the CPU oracle (oracle/splat_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this module.  It restates, stage by stage, what the reference's
CudaRasterizer::Rasterizer::forward / ::backward do
(submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu:197-446) and the
flashsplat / simple-knn / multi_instance_opt additions, on plain numpy arrays.
"""
import ctypes as C
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
if _HERE not in sys.path:
    sys.path.insert(0, _HERE)
import build as _build  # noqa: E402

_lib = None


class _Scene(C.Structure):
    _fields_ = [("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("W", C.c_int), ("H", C.c_int),
                ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
                ("bg", C.c_void_p), ("view", C.c_void_p), ("proj", C.c_void_p), ("campos", C.c_void_p),
                ("means3D", C.c_void_p), ("shs", C.c_void_p), ("colors_precomp", C.c_void_p),
                ("opacities", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
                ("cov3D_precomp", C.c_void_p)]


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_build.build())
        _lib.so_inclusive_scan.restype = C.c_uint32
        _lib.so_get_higher_msb.restype = C.c_uint32
        _lib.so_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def num_threads():
    return lib().so_num_threads()


def set_num_threads(n):
    lib().so_set_num_threads(C.c_int(n))


def _scene(g, cam, bg, scale_modifier=1.0, sh_degree=None):
    """g: dict(means3D, opacities, and shs|colors_precomp, scales+rotations|cov3D_precomp)."""
    keep = {}
    for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp"):
        keep[k] = _f32(g.get(k))
    keep["bg"] = _f32(bg)
    keep["view"] = _f32(cam["viewmatrix"])
    keep["proj"] = _f32(cam["projmatrix"])
    keep["campos"] = _f32(cam["campos"])
    s = _Scene()
    s.P = keep["means3D"].shape[0]
    s.M = 0 if keep["shs"] is None else keep["shs"].shape[1]
    s.D = int(g.get("sh_degree", 0) if sh_degree is None else sh_degree)
    s.W, s.H = cam["image_width"], cam["image_height"]
    s.tanfovx, s.tanfovy, s.scale_modifier = cam["tanfovx"], cam["tanfovy"], scale_modifier
    for k, v in keep.items():
        setattr(s, k, None if v is None else v.ctypes.data)
    return s, keep


def forward(g, cam, bg, scale_modifier=1.0, sh_degree=None, gt_mask=None, unique_label=None,
            num_obj=None, stop_after=None):
    """Full forward.  Returns a dict with every stage's arrays (names follow the reference's
    GeometryState / BinningState / ImageState members).  If num_obj is given this is the
    flashsplat forward (FSR/rasterize_points.cu:36-131) and used_count / contrib_num /
    image_xys / gs_depth are produced too."""
    L = lib()
    s, keep = _scene(g, cam, bg, scale_modifier, sh_degree)
    P, W, H = s.P, s.W, s.H
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy
    flash = num_obj is not None
    o = {"P": P, "W": W, "H": H, "T": T}
    o["radii"] = np.zeros(P, np.int32)
    o["means2D"] = np.zeros((P, 2), np.float32)
    o["depths"] = np.zeros(P, np.float32)
    o["cov3D"] = np.zeros((P, 6), np.float32)
    o["rgb"] = np.zeros((P, 3), np.float32)
    o["conic_opacity"] = np.zeros((P, 4), np.float32)
    o["clamped"] = np.zeros((P, 3), np.uint8)
    o["tiles_touched"] = np.zeros(P, np.uint32)
    if flash:
        o["image_xys"] = np.full((2, P), -1.0, np.float32)
        o["gs_depth"] = np.full(P, -0.1, np.float32)
    L.so_preprocess(C.byref(s), _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]), _p(o["cov3D"]),
                    _p(o["rgb"]), _p(o["conic_opacity"]), _p(o["clamped"]), _p(o["tiles_touched"]),
                    _p(o.get("image_xys")), _p(o.get("gs_depth")))
    o["point_offsets"] = np.zeros(P, np.uint32)
    R = int(L.so_inclusive_scan(C.c_int(P), _p(o["tiles_touched"]), _p(o["point_offsets"]))) if P else 0
    o["num_rendered"] = R
    if stop_after == "preprocess":
        return o
    o["keys_unsorted"] = np.zeros(R, np.uint64)
    o["values_unsorted"] = np.zeros(R, np.uint32)
    L.so_duplicate_with_keys(C.c_int(P), C.c_int(W), C.c_int(H), _p(o["means2D"]), _p(o["depths"]),
                             _p(o["point_offsets"]), _p(o["radii"]), _p(o["keys_unsorted"]),
                             _p(o["values_unsorted"]))
    bit = int(L.so_get_higher_msb(C.c_uint32(T)))
    o["sort_bits"] = 32 + bit
    o["point_list_keys"] = np.zeros(R, np.uint64)
    o["point_list"] = np.zeros(R, np.uint32)
    L.so_sort_pairs(C.c_uint64(R), _p(o["keys_unsorted"]), _p(o["point_list_keys"]),
                    _p(o["values_unsorted"]), _p(o["point_list"]), C.c_int(32 + bit))
    o["ranges"] = np.zeros((T, 2), np.uint32)
    L.so_identify_tile_ranges(C.c_uint64(R), _p(o["point_list_keys"]), C.c_int(T), _p(o["ranges"]))
    if stop_after == "binning":
        return o
    feats = keep["colors_precomp"] if keep["colors_precomp"] is not None else o["rgb"]
    o["out_color"] = np.zeros((3, H, W), np.float32)
    o["out_depth"] = np.zeros((1, H, W), np.float32)
    o["out_alpha"] = np.zeros((1, H, W), np.float32)
    o["n_contrib"] = np.zeros((H, W), np.uint32)
    gm = ul = uc = cn = None
    if flash:
        if gt_mask is None:  # FSR/flashsplat_rasterization/__init__.py:64-65 (all ones)
            gt_mask = np.ones((H, W), np.float32)
        gm = _f32(gt_mask).reshape(-1)
        if unique_label is not None and len(unique_label):
            ul = np.ascontiguousarray(unique_label, dtype=np.int32)
        uc = np.zeros((num_obj + 1, P), np.float64)
        cn = np.zeros((1, H, W), np.int32)
    L.so_render_forward(C.c_int(W), C.c_int(H), _p(o["ranges"]), _p(o["point_list"]), _p(o["means2D"]),
                        _p(feats), _p(o["depths"]), _p(o["conic_opacity"]), _p(keep["bg"]),
                        _p(o["out_color"]), _p(o["out_depth"]), _p(o["out_alpha"]), _p(o["n_contrib"]),
                        _p(gm), _p(ul), _p(uc), C.c_int(P), _p(cn))
    if flash:
        o["used_count64"] = uc
        o["used_count"] = uc.astype(np.float32)
        o["contrib_num"] = cn
    o["_features"] = feats
    return o


def backward(g, cam, bg, fwd, dL_dcolor, dL_ddepth, dL_dalpha, scale_modifier=1.0, sh_degree=None,
             gt_mask=None, unique_label=None):
    """Backward (rasterizer_impl.cu:343-446).  Returns the 8 tensors the reference returns
    (DGR/rasterize_points.cu:207) plus the two internal ones, as float32; the render-backward
    sums are accumulated in double first (see splat_oracle.c)."""
    L = lib()
    s, keep = _scene(g, cam, bg, scale_modifier, sh_degree)
    P, W, H = s.P, s.W, s.H
    d_mean2D = np.zeros((P, 2), np.float64)
    d_conic = np.zeros((P, 3), np.float64)
    d_opac = np.zeros(P, np.float64)
    d_colors = np.zeros((P, 3), np.float64)
    d_depths = np.zeros(P, np.float64)
    gm = None if gt_mask is None else _f32(gt_mask).reshape(-1)
    ul = None
    if unique_label is not None and len(unique_label):
        ul = np.ascontiguousarray(unique_label, dtype=np.int32)
    dc, dd, da = _f32(dL_dcolor), _f32(dL_ddepth), _f32(dL_dalpha)
    L.so_render_backward(C.c_int(W), C.c_int(H), _p(fwd["ranges"]), _p(fwd["point_list"]), _p(keep["bg"]),
                         _p(fwd["means2D"]), _p(fwd["conic_opacity"]), _p(fwd["_features"]),
                         _p(fwd["depths"]), _p(fwd["out_alpha"]), _p(fwd["n_contrib"]), _p(dc), _p(dd),
                         _p(da), _p(d_mean2D), _p(d_conic), _p(d_opac), _p(d_colors), _p(d_depths),
                         _p(gm), _p(ul))
    out = {}
    out["dL_dmeans2D"] = np.zeros((P, 3), np.float32)
    out["dL_dmeans2D"][:, :2] = d_mean2D
    out["dL_dconic"] = d_conic.astype(np.float32)
    out["dL_dopacity"] = d_opac.astype(np.float32).reshape(P, 1)
    out["dL_dcolors"] = d_colors.astype(np.float32)
    out["dL_ddepths"] = d_depths.astype(np.float32)
    out["dL_dmeans3D"] = np.zeros((P, 3), np.float32)
    out["dL_dcov3D"] = np.zeros((P, 6), np.float32)
    out["dL_dsh"] = np.zeros((P, s.M, 3), np.float32)
    out["dL_dscales"] = np.zeros((P, 3), np.float32)
    out["dL_drotations"] = np.zeros((P, 4), np.float32)
    L.so_preprocess_backward(C.byref(s), _p(fwd["radii"]), _p(fwd["cov3D"]), _p(fwd["clamped"]),
                             _p(out["dL_dmeans2D"]), _p(out["dL_dconic"]), _p(out["dL_dmeans3D"]),
                             _p(out["dL_dcolors"]), _p(out["dL_ddepths"]), _p(out["dL_dcov3D"]),
                             _p(out["dL_dsh"]), _p(out["dL_dscales"]), _p(out["dL_drotations"]))
    return out


def mark_visible(means3D, viewmatrix):
    m, v = _f32(means3D), _f32(viewmatrix)
    out = np.zeros(m.shape[0], np.uint8)
    lib().so_mark_visible(C.c_int(m.shape[0]), _p(m), _p(v), _p(out))
    return out.astype(bool)


def sort_pairs(keys, vals, end_bit):
    k = np.ascontiguousarray(keys, np.uint64)
    v = np.ascontiguousarray(vals, np.uint32)
    ko, vo = np.zeros_like(k), np.zeros_like(v)
    lib().so_sort_pairs(C.c_uint64(k.size), _p(k), _p(ko), _p(v), _p(vo), C.c_int(end_bit))
    return ko, vo


def multi_instance_opt(all_contrib, gamma=0.0):
    a = _f32(all_contrib)
    out = np.zeros(a.shape, np.uint8)
    lib().so_multi_instance_opt(C.c_int(a.shape[0]), C.c_int(a.shape[1]), _p(a), C.c_float(gamma), _p(out))
    return out.astype(bool)


def knn_dist2(points, return_debug=False):
    p = _f32(points)
    P = p.shape[0]
    out = np.zeros(P, np.float32)
    mort = np.zeros(P, np.uint32)
    idx = np.zeros(P, np.uint32)
    lib().so_knn_dist2(C.c_int(P), _p(p), _p(out), _p(mort), _p(idx))
    return (out, mort, idx) if return_debug else out
