"""ctypes binding of libwheatsplat.so (the C ABI in include/wheatsplat.h).

The library is the product: there is no CPU or PyTorch fallback.  If the shared object is
missing or a symbol is absent, import fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# WHEATSPLAT_LIB: another build of the same library (A/B timing of kernel variants, tools/ab.sh)
LIB_PATH = os.environ.get("WHEATSPLAT_LIB") or os.path.join(_HERE, "libwheatsplat.so")

WS_ABI_VERSION = 3

# every symbol include/wheatsplat.h declares (tests check the library exports all of them)
SYMBOLS = [
    "ws_abi_version", "ws_last_error", "ws_geom_bytes", "ws_image_bytes", "ws_binning_bytes",
    "ws_forward_prepare", "ws_forward_render", "ws_backward", "ws_mark_visible", "ws_knn_bytes",
    "ws_knn_dist2", "ws_knn_dist2_static", "ws_multi_instance_opt", "ws_sort_bytes", "ws_sort_pairs_u64", "ws_sort_pairs_u32", "ws_set_binning_mode",
    "ws_loss_bytes", "ws_l1_ssim_forward", "ws_l1_ssim_backward", "ws_debug_fetch",
    "ws_adam_step", "ws_densify_stats", "ws_allreduce_sum_f32", "ws_allreduce_tune", "ws_allreduce_split",
    "ws_sharded_adam_step", "ws_densify_plan_bytes", "ws_densify_plan", "ws_densify_gather", "ws_reset_opacity",
    "ws_objeval_bytes", "ws_objeval_prepare", "ws_objeval_bin", "ws_objeval_render",
    "ws_launch_count", "ws_prof_enable", "ws_prof_read", "ws_prof_stage_name", "ws_debug_count_pairs",
    "ws_debug_fma_peak",
]
WS_PROF_NSTAGES = 8
WS_KNN_STATIC_MAX_POINTS = 300000  # include/wheatsplat.h


class WsView(C.Structure):
    _fields_ = [("image_height", C.c_int32), ("image_width", C.c_int32), ("tanfovx", C.c_float),
                ("tanfovy", C.c_float), ("scale_modifier", C.c_float), ("sh_degree", C.c_int32),
                ("prefiltered", C.c_int32), ("debug", C.c_int32), ("bg", C.c_void_p),
                ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("campos", C.c_void_p)]


class WsGaussians(C.Structure):
    _fields_ = [("P", C.c_int32), ("M", C.c_int32), ("means3D", C.c_void_p), ("shs", C.c_void_p),
                ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
                ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p), ("shs_rest", C.c_void_p),
                ("raw_params", C.c_int32)]


class WsFlash(C.Structure):
    _fields_ = [("num_obj", C.c_int32), ("gt_mask", C.c_void_p), ("unique_label", C.c_void_p),
                ("out_contrib_num", C.c_void_p), ("used_count", C.c_void_p), ("image_xys", C.c_void_p),
                ("gs_depth", C.c_void_p)]


class WsGrads(C.Structure):
    _fields_ = [("dL_dmeans2D", C.c_void_p), ("dL_dcolors", C.c_void_p), ("dL_dopacity", C.c_void_p),
                ("dL_dmeans3D", C.c_void_p), ("dL_dcov3D", C.c_void_p), ("dL_dsh", C.c_void_p),
                ("dL_dscales", C.c_void_p), ("dL_drotations", C.c_void_p), ("dL_dsh_rest", C.c_void_p)]


class WsObjSets(C.Structure):
    _fields_ = [("num_rows", C.c_int32), ("member_start", C.c_void_p), ("member_obj", C.c_void_p)]


class WsShardSegment(C.Structure):
    _fields_ = [("begin", C.c_int64), ("end", C.c_int64), ("grad_offset", C.c_int64), ("lr", C.c_double)]


class WsRowArray(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("width", C.c_int32), ("role", C.c_int32)]


WS_ROW_COPY, WS_ROW_MOMENT, WS_ROW_XYZ, WS_ROW_SCALING = 0, 1, 2, 3


class WsAdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("numel", C.c_int64), ("lr", C.c_double), ("step", C.c_int64)]


_lib = None


def load():
    """Loads libwheatsplat.so (building it first if nvcc and the sources are there and the
    library is stale or missing)."""
    global _lib
    if _lib is not None:
        return _lib
    try:
        from . import build as _build
        _build.build()
    except Exception as e:  # noqa: BLE001 - surfaced below if the library is unusable
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "libwheatsplat.so is missing and could not be built (%s). This package has no "
                "CPU/PyTorch fallback: build it with `python -m wheat_3dgs_b200.build`." % e)
    lib = C.CDLL(LIB_PATH)
    for s in SYMBOLS:
        if not hasattr(lib, s):
            raise ImportError("libwheatsplat.so does not export %s (stale build?)" % s)
    vp, i32, i64, sz, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t, C.c_float
    lib.ws_abi_version.restype = C.c_int
    lib.ws_last_error.restype = C.c_char_p
    lib.ws_launch_count.restype = i64
    lib.ws_geom_bytes.restype = sz
    lib.ws_geom_bytes.argtypes = [i32]
    lib.ws_image_bytes.restype = sz
    lib.ws_image_bytes.argtypes = [i32, i32]
    lib.ws_binning_bytes.restype = sz
    lib.ws_binning_bytes.argtypes = [i64, i32, i32]
    lib.ws_forward_prepare.restype = i64
    lib.ws_forward_prepare.argtypes = [C.POINTER(WsView), C.POINTER(WsGaussians), vp, vp, sz,
                                       C.POINTER(WsFlash), vp]
    lib.ws_forward_render.restype = C.c_int
    lib.ws_forward_render.argtypes = [C.POINTER(WsView), C.POINTER(WsGaussians), vp, vp, vp, sz, vp, sz, i64,
                                      vp, vp, vp, C.POINTER(WsFlash), vp]
    lib.ws_backward.restype = C.c_int
    lib.ws_backward.argtypes = [C.POINTER(WsView), C.POINTER(WsGaussians), vp, vp, vp, vp, i64, vp, vp, vp, vp,
                                C.POINTER(WsGrads), C.POINTER(WsFlash), vp]
    lib.ws_mark_visible.restype = C.c_int
    lib.ws_mark_visible.argtypes = [i32, vp, vp, vp, vp, vp]
    lib.ws_knn_bytes.restype = sz
    lib.ws_knn_bytes.argtypes = [i32]
    lib.ws_knn_dist2.restype = C.c_int
    lib.ws_knn_dist2.argtypes = [i32, vp, vp, vp, sz, vp]
    lib.ws_knn_dist2_static.restype = C.c_int
    lib.ws_knn_dist2_static.argtypes = [i32, vp, vp, vp]
    lib.ws_multi_instance_opt.restype = C.c_int
    lib.ws_multi_instance_opt.argtypes = [i32, i32, vp, f32, vp, vp]
    lib.ws_sort_bytes.restype = sz
    lib.ws_sort_bytes.argtypes = [i64]
    lib.ws_sort_pairs_u64.restype = C.c_int
    lib.ws_sort_pairs_u64.argtypes = [i64, vp, vp, vp, vp, i32, vp, sz, vp]
    lib.ws_loss_bytes.restype = sz
    lib.ws_loss_bytes.argtypes = [i32, i32, i32]
    lib.ws_l1_ssim_forward.restype = C.c_int
    lib.ws_l1_ssim_forward.argtypes = [i32, i32, i32, vp, vp, vp, vp, sz, vp]
    lib.ws_l1_ssim_backward.restype = C.c_int
    lib.ws_l1_ssim_backward.argtypes = [i32, i32, i32, vp, vp, vp, i32, vp, sz, vp, vp]
    lib.ws_objeval_bytes.restype = sz
    lib.ws_objeval_bytes.argtypes = [i32, i32, i32, i64]
    lib.ws_objeval_prepare.restype = i64
    lib.ws_objeval_prepare.argtypes = [C.POINTER(WsView), C.POINTER(WsGaussians), C.POINTER(WsObjSets), vp, vp, sz, vp]
    lib.ws_objeval_bin.restype = i64
    lib.ws_objeval_bin.argtypes = [C.POINTER(WsView), C.POINTER(WsGaussians), C.POINTER(WsObjSets), vp, vp, vp, sz,
                                   vp, sz, i64, vp]
    lib.ws_objeval_render.restype = C.c_int
    lib.ws_objeval_render.argtypes = [C.POINTER(WsView), C.POINTER(WsObjSets), i32, vp, vp, vp, sz, i64, i64, vp, sz, vp, vp]
    lib.ws_allreduce_sum_f32.restype = C.c_int
    lib.ws_allreduce_sum_f32.argtypes = [i32, i32, i64, vp, C.POINTER(vp), i32, vp]
    lib.ws_allreduce_tune.restype = C.c_int
    lib.ws_allreduce_tune.argtypes = [i32, i32]
    lib.ws_allreduce_split.restype = C.c_int
    lib.ws_allreduce_split.argtypes = [i32, i32]
    lib.ws_adam_step.restype = C.c_int
    lib.ws_adam_step.argtypes = [i32, C.POINTER(WsAdamGroup), C.c_double, C.c_double, C.c_double, vp]
    lib.ws_densify_stats.restype = C.c_int
    lib.ws_densify_stats.argtypes = [i32, vp, vp, vp, vp, vp, vp]
    lib.ws_sharded_adam_step.restype = C.c_int
    lib.ws_sharded_adam_step.argtypes = [i32, i32, i64, i32, C.POINTER(WsShardSegment), vp, C.POINTER(vp), vp,
                                         C.POINTER(vp), vp, vp, i64, C.c_double, C.c_double, C.c_double, i32, vp]
    lib.ws_densify_plan_bytes.restype = sz
    lib.ws_densify_plan_bytes.argtypes = [i32]
    lib.ws_densify_plan.restype = C.c_int
    lib.ws_densify_plan.argtypes = [i32, vp, vp, vp, vp, f32, f32, f32, f32, i32, f32, vp, vp, sz,
                                    C.POINTER(C.c_uint32), vp]
    lib.ws_densify_gather.restype = C.c_int
    lib.ws_densify_gather.argtypes = [i32, vp, vp, sz, C.POINTER(C.c_uint32), vp, vp, i32, C.POINTER(WsRowArray), vp, vp]
    lib.ws_reset_opacity.restype = C.c_int
    lib.ws_reset_opacity.argtypes = [i32, vp, vp, vp, vp]
    lib.ws_set_binning_mode.restype = C.c_int
    lib.ws_set_binning_mode.argtypes = [i32]
    lib.ws_sort_pairs_u32.restype = C.c_int
    lib.ws_sort_pairs_u32.argtypes = [i64, vp, vp, vp, vp, i32, vp, sz, vp]
    lib.ws_debug_fetch.restype = i64
    lib.ws_debug_fetch.argtypes = [C.c_char_p, i32, i32, i32, i64, vp, vp, vp, vp, sz, vp]
    lib.ws_prof_enable.restype = None
    lib.ws_prof_enable.argtypes = [i32]
    lib.ws_prof_read.restype = C.c_int
    lib.ws_prof_read.argtypes = [C.POINTER(C.c_double), C.POINTER(i64)]
    lib.ws_prof_stage_name.restype = C.c_char_p
    lib.ws_prof_stage_name.argtypes = [i32]
    lib.ws_debug_count_pairs.restype = C.c_int
    lib.ws_debug_count_pairs.argtypes = [i32, i32, i32, vp, vp, vp, i64, vp, vp]
    lib.ws_debug_fma_peak.restype = i64
    lib.ws_debug_fma_peak.argtypes = [i32, i32, i32, vp, vp]
    if lib.ws_abi_version() != WS_ABI_VERSION:
        raise ImportError("libwheatsplat.so ABI %d != expected %d" % (lib.ws_abi_version(), WS_ABI_VERSION))
    _lib = lib
    return lib


def last_error():
    return load().ws_last_error().decode("utf-8", "replace")


def check(rc, what):
    if rc < 0:
        raise RuntimeError("%s failed (%d): %s" % (what, rc, last_error()))
    return rc
