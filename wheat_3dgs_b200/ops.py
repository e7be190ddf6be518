"""Torch-level operators over the C ABI, with the names, argument order and return tuples of the
pybind modules they replace:

  rasterize_gaussians / rasterize_gaussians_backward / mark_visible
      = diff_gaussian_rasterization._C (reference submodules/diff-gaussian-rasterization/
        rasterize_points.h:18-70, ext.cpp:15-19)
  flash_rasterize_gaussians / flash_rasterize_gaussians_backward
      = flashsplat_rasterization._C (reference submodules/flashsplat-rasterization/
        rasterize_points.h:18-46)
  distCUDA2 = simple_knn._C.distCUDA2 (reference submodules/simple-knn/spatial.cu:15-26)
  multi_instance_opt = run_3d_seg.py:54-72 on the device

PyTorch is plumbing here (allocation, streams); all compute is in libwheatsplat.so.  Optional
inputs follow the reference convention: an empty tensor means "not provided"
(diff_gaussian_rasterization/__init__.py:200-210).
"""
import ctypes as C

import torch

from . import lib as _lib
from .lib import WsFlash, WsGaussians, WsGrads, WsView


def _ptr(t):
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t, name):
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (wheat_3dgs_b200 has no CPU path)" % name)
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class _NoCtx:
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


_NOCTX = _NoCtx()


def _on(dev):
    """`with torch.cuda.device(dev)` only when dev is not already current (the context manager costs ~10 us per use,
    three uses per forward + backward: visible when the GPU work of a view is 0.2 ms)."""
    idx = dev.index
    if idx is None or idx == torch.cuda.current_device():
        return _NOCTX
    return torch.cuda.device(dev)


class _Call:
    """Marshals one rasterizer call: keeps the contiguous tensors alive and builds the structs."""

    def __init__(self, bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                 viewmatrix, projmatrix, tanfovx, tanfovy, image_height, image_width, sh, degree, campos,
                 prefiltered, debug, sh_rest=None, raw=False):
        if means3D.dim() != 2 or means3D.size(1) != 3:  # DGR rasterize_points.cu:57-59
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        self.P = means3D.size(0)
        self.H, self.W = int(image_height), int(image_width)
        self.keep = dict(
            bg=_f32c(bg, "bg"), means3D=_f32c(means3D, "means3D"), colors=_f32c(colors, "colors_precomp"),
            opacity=_f32c(opacity, "opacities"), scales=_f32c(scales, "scales"),
            rotations=_f32c(rotations, "rotations"), cov3D=_f32c(cov3D_precomp, "cov3D_precomp"),
            view=_f32c(viewmatrix, "viewmatrix"), proj=_f32c(projmatrix, "projmatrix"), sh=_f32c(sh, "sh"),
            campos=_f32c(campos, "campos"), sh_rest=_f32c(sh_rest, "features_rest"))
        k = self.keep
        self.M = 0 if k["sh"] is None else k["sh"].size(1)  # DGR rasterize_points.cu:85-89
        if k["sh_rest"] is not None:
            if k["sh"] is None or k["sh"].size(1) != 1:
                raise RuntimeError("features_rest needs features_dc of shape [P,1,3]")
            self.M = 1 + k["sh_rest"].size(1)
        self.raw = bool(raw)
        self.device = means3D.device
        self.view = WsView(self.H, self.W, float(tanfovx), float(tanfovy), float(scale_modifier), int(degree),
                           int(bool(prefiltered)), int(bool(debug)), _ptr(k["bg"]), _ptr(k["view"]),
                           _ptr(k["proj"]), _ptr(k["campos"]))
        self.gauss = WsGaussians(self.P, self.M, _ptr(k["means3D"]), _ptr(k["sh"]), _ptr(k["colors"]),
                                 _ptr(k["opacity"]), _ptr(k["scales"]), _ptr(k["rotations"]), _ptr(k["cov3D"]),
                                 _ptr(k["sh_rest"]), int(self.raw))


_BIN_GRANULE = 64 << 20
_BIN_POOL = {}   # (device, stream) -> [uint8 base tensors]
_USE_COUNT = getattr(torch._C, "_storage_Use_Count", None)


def _calibrate_use_count():
    """The pool below decides reuse from torch's private storage use count.  Check, once at import, that the
    count behaves as `_in_use` assumes (2 for a tensor nobody else references: the tensor + the temporary
    untyped_storage() wrapper; 3 with one view alive); if a torch version counts differently the pool is
    switched off and every call allocates (slower, never wrong)."""
    global _USE_COUNT
    if _USE_COUNT is None:
        return
    try:
        base = torch.empty((64,), dtype=torch.uint8)
        alone = _USE_COUNT(base.untyped_storage()._cdata)
        view = base.narrow(0, 0, 8)
        shared = _USE_COUNT(base.untyped_storage()._cdata)
        del view
        again = _USE_COUNT(base.untyped_storage()._cdata)
        if not (alone == 2 and shared == 3 and again == 2):
            _USE_COUNT = None
    except Exception:  # noqa: BLE001 - private API gone or changed
        _USE_COUNT = None


_calibrate_use_count()


def _in_use(base):
    # references to the storage besides the pool's tensor and the temporary made here: views handed out earlier that
    # are still alive in Python or saved in an autograd graph
    return _USE_COUNT(base.untyped_storage()._cdata) > 2


def _binning_alloc(nbytes, dev):
    """The instance-list scratch changes size with every view (R is data dependent) and is the largest allocation
    of a step (~0.5 GB at config 2).  Asking torch's caching allocator for a different size every view fragments
    its pool and ends in cudaMalloc/cudaFree calls (device synchronisations) inside the training loop, so the
    scratch comes from a small grow-only pool per (device, stream): a buffer is handed out again (as a view) once
    nothing references its storage any more -- neither Python nor a saved-for-backward slot of an autograd graph."""
    if nbytes <= 0:
        return torch.empty((0,), dtype=torch.uint8, device=dev)
    if _USE_COUNT is None:
        return torch.empty(((nbytes + _BIN_GRANULE - 1) // _BIN_GRANULE * _BIN_GRANULE,), dtype=torch.uint8, device=dev)
    key = (str(dev), torch.cuda.current_stream(dev).cuda_stream)
    pool = _BIN_POOL.setdefault(key, [])
    for base in pool:
        if base.numel() >= nbytes and not _in_use(base):
            return base.narrow(0, 0, nbytes)
    pool[:] = [b for b in pool if _in_use(b)]   # unused but too small: give them back to the allocator
    want = (int(nbytes * 1.25) + _BIN_GRANULE - 1) // _BIN_GRANULE * _BIN_GRANULE
    base = torch.empty((want,), dtype=torch.uint8, device=dev)
    pool.append(base)
    return base.narrow(0, 0, nbytes)


def _output_planes(H, W, dev):
    """color [3,H,W], depth [1,H,W], alpha [1,H,W] in ONE allocation, as three tensors of their own on the shared
    storage (Tensor.set_), not views of a common base: autograd marks several views of one tensor returned from a
    Function as outputs of a multi-output node and then refuses in-place edits of them (`image.clamp_()`,
    `depth[m] = 0`), which the reference's independently allocated outputs allow.  Each part starts on a 256-byte
    boundary whatever H*W is."""
    hw = H * W
    pad = lambda n: (n + 63) // 64 * 64   # floats
    o_depth = pad(3 * hw)
    o_alpha = o_depth + pad(hw)
    storage = torch.empty((o_alpha + hw,), dtype=torch.float32, device=dev).untyped_storage()
    part = lambda off, c: torch.empty((0,), dtype=torch.float32, device=dev).set_(storage, off, (c, H, W), (hw, W, 1))
    return part(0, 3), part(o_depth, 1), part(o_alpha, 1)


def _forward(call, flash):
    """Shared forward: returns (R, color, depth, alpha, radii, geom, binning, image)."""
    L = _lib.load()
    dev, P, H, W = call.device, call.P, call.H, call.W
    # three allocations per call: the five output planes (color 3, depth, alpha share one allocation), radii, and one
    # byte buffer carved into the geometry and image scratch (512-byte aligned parts)
    color, depth, alpha = _output_planes(H, W, dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    gb, ib = (L.ws_geom_bytes(P) + 511) // 512 * 512, L.ws_image_bytes(W, H)
    scratch = torch.empty((gb + ib,), dtype=torch.uint8, device=dev)
    geom, image = scratch[:gb], scratch[gb:]
    fl = C.byref(flash) if flash is not None else None
    with _on(dev):
        st = _stream()
        R = _lib.check(L.ws_forward_prepare(C.byref(call.view), C.byref(call.gauss), radii.data_ptr(),
                                            geom.data_ptr(), geom.numel(), fl, st), "ws_forward_prepare")
        binning = _binning_alloc(L.ws_binning_bytes(R, W, H) if R > 0 else 0, dev)
        _lib.check(L.ws_forward_render(C.byref(call.view), C.byref(call.gauss), radii.data_ptr(), geom.data_ptr(),
                                       _ptr(binning), binning.numel(), image.data_ptr(), image.numel(), R,
                                       color.data_ptr(), depth.data_ptr(), alpha.data_ptr(), fl, st),
                   "ws_forward_render")
    return R, color, depth, alpha, radii, geom, binning, image


def rasterize_gaussians(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tanfovx, tanfovy, image_height, image_width, sh, degree,
                        campos, prefiltered, debug):
    """= _C.rasterize_gaussians (DGR rasterize_points.cu:35-119).
    Returns (num_rendered, color[3,H,W], depth[1,H,W], alpha[1,H,W], radii[P] int32, geomBuffer,
    binningBuffer, imgBuffer)."""
    call = _Call(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                 projmatrix, tanfovx, tanfovy, image_height, image_width, sh, degree, campos, prefiltered, debug)
    if call.P == 0:  # DGR rasterize_points.cu:82: nothing is launched, outputs stay zero
        dev = means3D.device
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        e = torch.empty((0,), dtype=torch.uint8, device=dev)
        return 0, z(3, call.H, call.W), z(1, call.H, call.W), z(1, call.H, call.W), \
            torch.zeros((0,), dtype=torch.int32, device=dev), e, e.clone(), e.clone()
    return _forward(call, None)


def _backward(call, radii, geom, binning, image, R, alpha, dL_dcolor, dL_ddepth, dL_dalpha, flash, bucket=None):
    L = _lib.load()
    dev, P, M = call.device, call.P, call.M
    has_sr = call.keep["scales"] is not None
    # one flat allocation: [means3D 3 | means2D 3 | sh 3M | opacity 1 | scales 3 | rot 4 | colors 3 | cov3D 6]
    # The first 14+3M floats per Gaussian are exactly the training-time gradient set, contiguous,
    # so a data-parallel trainer can all-reduce them as ONE bucket (wheat_3dgs_b200/dist.py).
    sizes = [3, 3, 3 * M, 1, 3, 4, 3, 6]
    total = sum(sizes) * P
    if bucket is None:
        bucket = torch.empty((total,), dtype=torch.float32, device=dev)
    elif bucket.numel() < total or bucket.dtype != torch.float32 or not bucket.is_contiguous():
        raise RuntimeError("gradient bucket must be a contiguous float32 tensor of >= %d elements" % total)
    g_means3D, g_means2D, g_sh, g_opac, g_scales, g_rot, g_colors, g_cov = bucket[:total].split([s * P for s in sizes])
    split = call.keep.get("sh_rest") is not None   # dL_dsh = [features_dc P*3 | features_rest P*3(M-1)]
    grads = WsGrads(g_means2D.data_ptr(), g_colors.data_ptr(), g_opac.data_ptr(), g_means3D.data_ptr(),
                    g_cov.data_ptr(), g_sh.data_ptr() if M > 0 else None,
                    g_scales.data_ptr() if has_sr else None, g_rot.data_ptr() if has_sr else None,
                    g_sh[3 * P:].data_ptr() if split else None)
    dc, dd, da = _f32c(dL_dcolor, "dL_dout_color"), _f32c(dL_ddepth, "dL_dout_depth"), _f32c(dL_dalpha, "dL_dout_alpha")
    al = _f32c(alpha, "alphas")
    fl = C.byref(flash) if flash is not None else None
    with _on(dev):
        _lib.check(L.ws_backward(C.byref(call.view), C.byref(call.gauss), radii.data_ptr(), geom.data_ptr(),
                                 _ptr(binning), image.data_ptr(), int(R), al.data_ptr(), dc.data_ptr(),
                                 dd.data_ptr(), da.data_ptr(), C.byref(grads), fl, _stream()), "ws_backward")
    if not has_sr:
        g_scales.zero_()
        g_rot.zero_()
    if split:
        return (g_means2D.view(P, 3), g_colors.view(P, 3), g_opac.view(P, 1), g_means3D.view(P, 3), g_cov.view(P, 6),
                (g_sh[:3 * P].view(P, 1, 3), g_sh[3 * P:].view(P, M - 1, 3)), g_scales.view(P, 3), g_rot.view(P, 4))
    return (g_means2D.view(P, 3), g_colors.view(P, 3), g_opac.view(P, 1), g_means3D.view(P, 3), g_cov.view(P, 6),
            g_sh.view(P, M, 3), g_scales.view(P, 3), g_rot.view(P, 4))


def rasterize_gaussians_backward(bg, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tanfovx, tanfovy, dL_dout_color, dL_dout_depth,
                                 dL_dout_alpha, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer,
                                 alphas, debug, _bucket=None):
    """= _C.rasterize_gaussians_backward (DGR rasterize_points.cu:121-208).  Returns
    (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)."""
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    P = means3D.size(0)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    if P == 0:
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=means3D.device)
        return z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 6), z(0, M, 3), z(0, 3), z(0, 4)
    # opacities are not needed by the backward (they live in the forward's geometry state)
    call = _Call(bg, means3D, colors, None, scales, rotations, scale_modifier, cov3D_precomp,
                 viewmatrix, projmatrix, tanfovx, tanfovy, H, W, sh, degree, campos, False, debug)
    return _backward(call, radii, geomBuffer, binningBuffer, imageBuffer, R, alphas, dL_dout_color,
                     dL_dout_depth, dL_dout_alpha, None, _bucket)


def rasterize_gaussians_raw(bg, xyz, features_dc, features_rest, opacity, scaling, rotation, scale_modifier, viewmatrix,
                            projmatrix, tanfovx, tanfovy, image_height, image_width, degree, campos, debug):
    """rasterize_gaussians on GaussianModel's RAW parameters: the getters of
    scene/gaussian_model.py:101-121 (exp, normalize, sigmoid, torch.cat of the two feature tensors) run inside
    preprocess.  Same 8-tuple as rasterize_gaussians."""
    call = _Call(bg, xyz, None, opacity, scaling, rotation, scale_modifier, None, viewmatrix, projmatrix, tanfovx,
                 tanfovy, image_height, image_width, features_dc, degree, campos, False, debug, sh_rest=features_rest,
                 raw=True)
    if call.P == 0:
        dev = xyz.device
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        e = torch.empty((0,), dtype=torch.uint8, device=dev)
        return 0, z(3, call.H, call.W), z(1, call.H, call.W), z(1, call.H, call.W), \
            torch.zeros((0,), dtype=torch.int32, device=dev), e, e.clone(), e.clone()
    return _forward(call, None)


def rasterize_gaussians_raw_backward(bg, xyz, radii, features_dc, features_rest, scaling, rotation, scale_modifier,
                                     viewmatrix, projmatrix, tanfovx, tanfovy, dL_dout_color, dL_dout_depth,
                                     dL_dout_alpha, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, alphas,
                                     debug, _bucket=None):
    """Gradients w.r.t. the raw parameters: (dL_dmeans2D, dL_dopacity, dL_dxyz, dL_dfeatures_dc, dL_dfeatures_rest,
    dL_dscaling, dL_drotation)."""
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    P = xyz.size(0)
    if P == 0:
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=xyz.device)
        return z(0, 3), z(0, 1), z(0, 3), z(0, 1, 3), z(0, features_rest.size(1), 3), z(0, 3), z(0, 4)
    call = _Call(bg, xyz, None, None, scaling, rotation, scale_modifier, None, viewmatrix, projmatrix, tanfovx, tanfovy,
                 H, W, features_dc, degree, campos, False, debug, sh_rest=features_rest, raw=True)
    g2d, _, gop, gxyz, _, (gdc, grest), gsc, grot = _backward(call, radii, geomBuffer, binningBuffer, imageBuffer, R,
                                                             alphas, dL_dout_color, dL_dout_depth, dL_dout_alpha, None,
                                                             _bucket)
    return g2d, gop, gxyz, gdc, grest, gsc, grot


def count_instances(bg, means3D, opacity, scales, rotations, viewmatrix, projmatrix, tanfovx, tanfovy, image_height,
                    image_width, sh, degree, campos):
    """num_rendered (= sum of tiles_touched, the length of the instance list) of a view without rendering it:
    preprocess + scan only.  The cost estimate dist.balanced_schedule groups views by."""
    call = _Call(bg, means3D, None, opacity, scales, rotations, 1.0, None, viewmatrix, projmatrix, tanfovx, tanfovy,
                 image_height, image_width, sh, degree, campos, False, False)
    if call.P == 0:
        return 0
    L = _lib.load()
    radii = torch.empty((call.P,), dtype=torch.int32, device=call.device)
    geom = torch.empty((L.ws_geom_bytes(call.P),), dtype=torch.uint8, device=call.device)
    with torch.cuda.device(call.device):
        return _lib.check(L.ws_forward_prepare(C.byref(call.view), C.byref(call.gauss), radii.data_ptr(), geom.data_ptr(),
                                               geom.numel(), None, _stream()), "ws_forward_prepare")


def mark_visible(means3D, viewmatrix, projmatrix):
    """= _C.mark_visible (DGR rasterize_points.cu:210-229) -> bool[P]."""
    L = _lib.load()
    m, v, pm = _f32c(means3D, "means3D"), _f32c(viewmatrix, "viewmatrix"), _f32c(projmatrix, "projmatrix")
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.uint8, device=means3D.device)
    if P:
        with torch.cuda.device(means3D.device):
            _lib.check(L.ws_mark_visible(P, m.data_ptr(), v.data_ptr(), _ptr(pm), present.data_ptr(), _stream()),
                       "ws_mark_visible")
    return present.view(torch.bool)


def flash_rasterize_gaussians(gt_mask, unique_label, bg, means3D, colors, opacity, scales, rotations,
                              scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tanfovx, tanfovy,
                              image_height, image_width, sh, degree, campos, prefiltered, num_obj, debug,
                              used_count_out=None):
    """= flashsplat _C.rasterize_gaussians (FSR rasterize_points.cu:36-131).  Returns the 12-tuple
    (num_rendered, color, depth, alpha, radii, geomBuffer, binningBuffer, imgBuffer, out_contrib_num,
    used_count, image_xys, gs_depth).  `used_count_out`: optional persistent [num_obj+1,P] buffer the
    votes are ACCUMULATED into (one buffer across views instead of a 2 GB zero-fill per view)."""
    call = _Call(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                 projmatrix, tanfovx, tanfovy, image_height, image_width, sh, degree, campos, prefiltered, debug)
    dev, P, H, W = means3D.device, call.P, call.H, call.W
    contrib = torch.empty((1, H, W), dtype=torch.int32, device=dev) if P else \
        torch.zeros((1, H, W), dtype=torch.int32, device=dev)
    if used_count_out is None:
        used = torch.zeros((num_obj + 1, P), dtype=torch.float32, device=dev)
    else:
        used = used_count_out
        if used.shape != (num_obj + 1, P) or used.dtype != torch.float32 or not used.is_contiguous():
            raise RuntimeError("used_count_out must be contiguous float32 [num_obj+1, P]")
    xys = torch.full((2, P), -1.0, dtype=torch.float32, device=dev)   # FSR rasterize_points.cu:77
    gsd = torch.full((P,), -0.1, dtype=torch.float32, device=dev)     # FSR rasterize_points.cu:78
    if P == 0:
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        e = torch.empty((0,), dtype=torch.uint8, device=dev)
        return (0, z(3, H, W), z(1, H, W), z(1, H, W), torch.zeros((0,), dtype=torch.int32, device=dev), e,
                e.clone(), e.clone(), contrib, used, xys, gsd)
    gm = _f32c(gt_mask, "gt_mask")
    if gm.numel() != H * W:
        raise RuntimeError("gt_mask must have image_height*image_width elements")
    ul = None
    if unique_label is not None and unique_label.numel() != 0:
        ul = unique_label.to(device=dev, dtype=torch.int32).contiguous()
        if ul.numel() != P:
            raise RuntimeError("unique_label must have P elements")
    flash = WsFlash(int(num_obj), gm.data_ptr(), _ptr(ul), contrib.data_ptr(), used.data_ptr(), xys.data_ptr(),
                    gsd.data_ptr())
    call.keep["flash"] = (gm, ul)
    R, color, depth, alpha, radii, geom, binning, image = _forward(call, flash)
    return R, color, depth, alpha, radii, geom, binning, image, contrib, used, xys, gsd


def flash_rasterize_gaussians_backward(gt_mask, unique_label, bg, means3D, radii, colors, scales, rotations,
                                       scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tanfovx, tanfovy,
                                       dL_dout_color, dL_dout_depth, dL_dout_alpha, sh, degree, campos,
                                       geomBuffer, R, binningBuffer, imageBuffer, alphas, debug):
    """= flashsplat _C.rasterize_gaussians_backward (FSR rasterize_points.cu:134-226)."""
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    P = means3D.size(0)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    if P == 0:
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=means3D.device)
        return z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 6), z(0, M, 3), z(0, 3), z(0, 4)
    call = _Call(bg, means3D, colors, None, scales, rotations, scale_modifier, cov3D_precomp,
                 viewmatrix, projmatrix, tanfovx, tanfovy, H, W, sh, degree, campos, False, debug)
    gm = _f32c(gt_mask, "gt_mask")
    ul = None
    if unique_label is not None and unique_label.numel() != 0:
        ul = unique_label.to(device=means3D.device, dtype=torch.int32).contiguous()
    flash = WsFlash(0, gm.data_ptr(), _ptr(ul), None, None, None, None)
    return _backward(call, radii, geomBuffer, binningBuffer, imageBuffer, R, alphas, dL_dout_color,
                     dL_dout_depth, dL_dout_alpha, flash)


def distCUDA2(points):
    """= simple_knn._C.distCUDA2: mean squared distance to the 3 nearest neighbours, float[P]."""
    L = _lib.load()
    p = _f32c(points, "points")
    P = points.size(0)
    means = torch.zeros((P,), dtype=torch.float32, device=points.device)
    if P:
        scratch = torch.empty((L.ws_knn_bytes(P),), dtype=torch.uint8, device=points.device)
        with torch.cuda.device(points.device):
            _lib.check(L.ws_knn_dist2(P, p.data_ptr(), means.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                      _stream()), "ws_knn_dist2")
    return means


def multi_instance_opt(all_contrib, gamma=0.0):
    """Device version of run_3d_seg.py:54-72: float[num_obj+1, P] votes -> bool[num_obj+1, P]."""
    L = _lib.load()
    a = _f32c(all_contrib, "all_contrib")
    rows, P = all_contrib.shape
    out = torch.zeros((rows, P), dtype=torch.uint8, device=all_contrib.device)
    if rows * P:
        with torch.cuda.device(all_contrib.device):
            _lib.check(L.ws_multi_instance_opt(rows, P, a.data_ptr(), float(gamma), out.data_ptr(), _stream()),
                       "ws_multi_instance_opt")
    return out.view(torch.bool)


def sort_pairs(keys, vals, end_bit):
    """Stable radix sort of (int64-viewed-as-u64 or int32-viewed-as-u32 keys, int32-viewed-as-u32 vals)
    on key bits [0,end_bit)."""
    L = _lib.load()
    fn, fname = (L.ws_sort_pairs_u64, "ws_sort_pairs_u64") if keys.element_size() == 8 else \
        (L.ws_sort_pairs_u32, "ws_sort_pairs_u32")
    n = keys.numel()
    k_in, v_in = keys.clone().contiguous(), vals.clone().contiguous()
    k_out, v_out = torch.empty_like(k_in), torch.empty_like(v_in)
    if n:
        scratch = torch.empty((L.ws_sort_bytes(n),), dtype=torch.uint8, device=keys.device)
        with torch.cuda.device(keys.device):
            _lib.check(fn(n, k_in.data_ptr(), k_out.data_ptr(), v_in.data_ptr(), v_out.data_ptr(),
                          int(end_bit), scratch.data_ptr(), scratch.numel(), _stream()), fname)
    return k_out, v_out


_FETCH = {"depths": (torch.float32, lambda P, T, HW, R: (P,)), "means2D": (torch.float32, lambda P, T, HW, R: (P, 2)),
          "conic_opacity": (torch.float32, lambda P, T, HW, R: (P, 4)), "rgb": (torch.float32, lambda P, T, HW, R: (P, 3)),
          "tiles_touched": (torch.int32, lambda P, T, HW, R: (P,)), "point_list_keys": (torch.int64, lambda P, T, HW, R: (R,)),
          "point_list": (torch.int32, lambda P, T, HW, R: (R,)), "ranges": (torch.int32, lambda P, T, HW, R: (T, 2)),
          "n_contrib": (torch.int32, lambda P, T, HW, R: (HW,))}


def debug_fetch(name, P, W, H, R, geom, binning, image):
    """Internal arrays of a finished forward, for parity tests (see ws_debug_fetch)."""
    L = _lib.load()
    dt, shp = _FETCH[name]
    T = ((W + 15) // 16) * ((H + 15) // 16)
    out = torch.zeros(shp(P, T, H * W, R), dtype=dt, device=geom.device)
    if out.numel():
        with torch.cuda.device(geom.device):
            _lib.check(L.ws_debug_fetch(name.encode(), P, W, H, R, geom.data_ptr(), _ptr(binning), image.data_ptr(),
                                        out.data_ptr(), out.numel() * out.element_size(), _stream()), "ws_debug_fetch")
    return out


def launch_count():
    return int(_lib.load().ws_launch_count())


def prof_enable(on):
    """Stage timing with CUDA events inside the library (bench.py's roofline leg)."""
    _lib.load().ws_prof_enable(1 if on else 0)


def prof_read():
    """{stage name: (total ms, brackets)} accumulated since the last read."""
    L = _lib.load()
    n = _lib.WS_PROF_NSTAGES
    ms = (C.c_double * n)()
    calls = (C.c_int64 * n)()
    _lib.check(L.ws_prof_read(ms, calls), "ws_prof_read")
    return {L.ws_prof_stage_name(i).decode(): (ms[i], int(calls[i])) for i in range(n)}


def count_pairs(P, W, H, R, geom, binning, image):
    """(N_eval, N_contrib, quadrant-warp pairs after the cull, instances with an empty mask) of a finished forward
    (see ws_debug_count_pairs)."""
    L = _lib.load()
    counts = torch.zeros((4,), dtype=torch.int64, device=image.device)
    with torch.cuda.device(image.device):
        _lib.check(L.ws_debug_count_pairs(int(P), W, H, geom.data_ptr(), _ptr(binning), image.data_ptr(), int(R),
                                          counts.data_ptr(), _stream()), "ws_debug_count_pairs")
    return tuple(int(x) for x in counts.cpu().tolist())


def fma_peak(packed, device=None, iters=4096, ctas=148 * 8, repeats=5):
    """Measured FP32 FMA throughput in TFLOP/s (best of `repeats`, CUDA events on the current stream): the blend
    kernels' roofline denominator.  packed=True issues FFMA2 (fma.rn.f32x2), False scalar FFMA."""
    L = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    out = torch.empty((ctas * 256,), dtype=torch.float32, device=dev)
    best = 0.0
    with torch.cuda.device(dev):
        for _ in range(repeats + 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            flops = _lib.check(L.ws_debug_fma_peak(1 if packed else 0, iters, ctas, out.data_ptr(), _stream()),
                               "ws_debug_fma_peak")
            e1.record()
            e1.synchronize()
            best = max(best, flops / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best
