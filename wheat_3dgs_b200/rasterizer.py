"""Host-side mirror of the reference's Python operator API for the rasterizers.

Drop-in surface (same names, argument meaning, return order and exception texts):

  diff_gaussian_rasterization   GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians
      reference: submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:21-223
  flashsplat_rasterization      GaussianRasterizationSettings (+ mask_grad, num_obj), GaussianRasterizer
      (+ gt_mask, unique_label), rasterize_gaussians
      reference: submodules/flashsplat-rasterization/flashsplat_rasterization/__init__.py:21-271

One autograd Function serves both; the flashsplat flavour only adds the label-vote outputs.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import ops


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class FlashGaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    mask_grad: bool
    num_obj: int


_GRAD_BUCKET = None
_ZERO_PLANES = {}


def _zeros(c, H, W, dev):
    """Read-only zero gradient of an output the loss did not use (autograd hands None with
    set_materialize_grads(False)); one cached tensor per shape instead of a fill per backward."""
    key = (c, H, W, dev)
    z = _ZERO_PLANES.get(key)
    if z is None:
        z = _ZERO_PLANES[key] = torch.zeros((c, H, W), dtype=torch.float32, device=dev)
    return z


def set_grad_bucket(bucket):
    """Optional flat float32 buffer the next backward passes write their gradients into (layout:
    wheat_3dgs_b200.dist.split_bucket).  A data-parallel trainer all-reduces its first
    (14+3M)*P floats as one collective; None restores per-call allocation."""
    global _GRAD_BUCKET
    _GRAD_BUCKET = bucket


def _check_bucket_free(leaf):
    """With a gradient bucket set, the gradients autograd hands to the leaves ARE slices of the bucket.  A second
    backward before the first one's .grad tensors are dropped would overwrite them in place and then be ADDED to
    itself by AccumulateGrad (2 x new instead of old + new): one backward per bucket per step."""
    g = getattr(leaf, "grad", None)
    b = _GRAD_BUCKET
    if g is None or b is None or g.device != b.device:
        return
    lo, hi = b.data_ptr(), b.data_ptr() + b.numel() * b.element_size()
    if lo <= g.data_ptr() < hi:
        raise RuntimeError("set_grad_bucket: the previous backward's gradients still alias the bucket; drop them "
                           "(optimizer.zero_grad(set_to_none=True) / p.grad = None) before the next backward -- "
                           "gradient accumulation over several backwards needs set_grad_bucket(None)")


def _snapshot(args, path):
    """debug=True behaviour of the reference: CPU copy of the inputs dumped when the call throws
    (DGR __init__.py:83-90,135-142)."""
    torch.save(tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args), path)


def _guarded(fn, args, debug, path, what):
    if not debug:
        return fn(*args)
    try:
        return fn(*args)
    except Exception:
        _snapshot(args, path)
        print("\nAn error occured in %s. Writing %s for debugging.\n" % (what, path))
        raise


class _Rasterize(torch.autograd.Function):
    """forward(ctx, flash, gt_mask, unique_label, means3D, means2D, sh, colors_precomp, opacities, scales,
    rotations, cov3Ds_precomp, raster_settings)"""

    @staticmethod
    def forward(ctx, flash, gt_mask, unique_label, means3D, means2D, sh, colors_precomp, opacities, scales,
                rotations, cov3Ds_precomp, rs):
        common = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                  rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                  rs.sh_degree, rs.campos, rs.prefiltered)
        ctx.flash = flash
        ctx.rs = rs
        ctx.set_materialize_grads(False)
        if not flash:
            out = _guarded(ops.rasterize_gaussians, common + (rs.debug,), rs.debug, "snapshot_fw.dump", "forward")
            R, color, depth, alpha, radii, geom, binning, image = out
            ctx.num_rendered = R
            ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom,
                                  binning, image, alpha)
            ctx.mark_non_differentiable(radii)
            return color, radii, depth, alpha
        if gt_mask is None:  # FSR __init__.py:64-65
            gt_mask = torch.ones(1, rs.image_width, rs.image_height, device=rs.bg.device)
        gt_mask = gt_mask.float().contiguous()
        args = (gt_mask, unique_label) + common + (rs.num_obj, rs.debug)
        out = _guarded(ops.flash_rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
        R, color, depth, alpha, radii, geom, binning, image, contrib, used, xys, gsd = out
        ctx.num_rendered = R
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning,
                              image, alpha, used, gt_mask, unique_label)
        ctx.mark_non_differentiable(radii, contrib, used, xys, gsd)
        return color, radii, depth, alpha, contrib, used, xys, gsd

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha, *unused):
        rs = ctx.rs
        saved = ctx.saved_tensors
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, image, alpha = saved[:11]
        H, W = rs.image_height, rs.image_width
        dev = means3D.device
        if grad_color is None:
            grad_color = _zeros(3, H, W, dev)
        if grad_depth is None:
            grad_depth = _zeros(1, H, W, dev)
        if grad_alpha is None:
            grad_alpha = _zeros(1, H, W, dev)
        tail = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha, sh,
                rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, image, alpha, rs.debug)
        if not ctx.flash:
            fn = ops.rasterize_gaussians_backward
            if _GRAD_BUCKET is not None:
                _check_bucket_free(means3D)
                fn = lambda *a: ops.rasterize_gaussians_backward(*a, _bucket=_GRAD_BUCKET)  # noqa: E731
            g = _guarded(fn, tail, rs.debug, "snapshot_bw.dump", "backward")
        else:
            used, gt_mask, unique_label = saved[11:14]
            g = _guarded(ops.flash_rasterize_gaussians_backward, (gt_mask, unique_label) + tail, rs.debug,
                         "snapshot_bw.dump", "backward")
        g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_sh, g_scales, g_rot = g
        if ctx.flash and rs.mask_grad:  # FSR __init__.py:168-182
            m = (used != 0)[:, None].float()
            g_sh, g_colors, g_opac = g_sh * m[:, :, None], g_colors * m, g_opac * m
            g_scales, g_rot, g_cov3D = g_scales * m, g_rot * m, g_cov3D * m

        def opt(t, ref):  # no gradient for inputs that were "not provided" (empty tensors)
            return t if ref is not None and ref.numel() != 0 else None

        return (None, None, None, g_means3D, g_means2D, opt(g_sh, sh), opt(g_colors, colors_precomp), g_opac,
                opt(g_scales, scales), opt(g_rot, rotations), opt(g_cov3D, cov3Ds_precomp), None)


class _RasterizeRaw(torch.autograd.Function):
    """forward(ctx, xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, raster_settings) on
    GaussianModel's raw parameters; the activations and their backward run inside the kernels."""

    @staticmethod
    def forward(ctx, xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, rs):
        args = (rs.bg, xyz, features_dc, features_rest, opacity, scaling, rotation, rs.scale_modifier, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, rs.sh_degree, rs.campos, rs.debug)
        R, color, depth, alpha, radii, geom, binning, image = _guarded(ops.rasterize_gaussians_raw, args, rs.debug,
                                                                       "snapshot_fw.dump", "forward")
        ctx.rs, ctx.num_rendered = rs, R
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(xyz, features_dc, features_rest, scaling, rotation, radii, geom, binning, image, alpha)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha):
        rs = ctx.rs
        xyz, features_dc, features_rest, scaling, rotation, radii, geom, binning, image, alpha = ctx.saved_tensors
        H, W, dev = rs.image_height, rs.image_width, xyz.device
        z = lambda c: _zeros(c, H, W, dev)
        grad_color = z(3) if grad_color is None else grad_color
        grad_depth = z(1) if grad_depth is None else grad_depth
        grad_alpha = z(1) if grad_alpha is None else grad_alpha
        args = (rs.bg, xyz, radii, features_dc, features_rest, scaling, rotation, rs.scale_modifier, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha, rs.sh_degree, rs.campos, geom,
                ctx.num_rendered, binning, image, alpha, rs.debug)
        fn = ops.rasterize_gaussians_raw_backward
        if _GRAD_BUCKET is not None:
            _check_bucket_free(xyz)
            fn = lambda *a: ops.rasterize_gaussians_raw_backward(*a, _bucket=_GRAD_BUCKET)  # noqa: E731
        g2d, gop, gxyz, gdc, grest, gsc, grot = _guarded(fn, args, rs.debug, "snapshot_bw.dump", "backward")
        return gxyz, g2d, gdc, grest, gop, gsc, grot, None


def rasterize_gaussians_raw(xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, raster_settings):
    return _RasterizeRaw.apply(xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, raster_settings)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    """diff_gaussian_rasterization.rasterize_gaussians (DGR __init__.py:21-42)."""
    return _Rasterize.apply(False, None, None, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                            cov3Ds_precomp, raster_settings)


def flash_rasterize_gaussians(gt_mask, unique_label, means3D, means2D, sh, colors_precomp, opacities, scales,
                              rotations, cov3Ds_precomp, raster_settings):
    """flashsplat_rasterization.rasterize_gaussians (FSR __init__.py:21-46)."""
    return _Rasterize.apply(True, gt_mask, unique_label, means3D, means2D, sh, colors_precomp, opacities, scales,
                            rotations, cov3Ds_precomp, raster_settings)


def _check_inputs(shs, colors_precomp, scales, rotations, cov3D_precomp):
    # exception texts as in the reference (DGR __init__.py:194-198)
    if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    if ((scales is None or rotations is None) and cov3D_precomp is None) or \
            ((scales is not None or rotations is not None) and cov3D_precomp is not None):
        raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')


def _or_empty(t):
    return torch.Tensor([]) if t is None else t


class GaussianRasterizer(nn.Module):
    """diff_gaussian_rasterization.GaussianRasterizer (DGR __init__.py:174-223)."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return ops.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        _check_inputs(shs, colors_precomp, scales, rotations, cov3D_precomp)
        return rasterize_gaussians(means3D, means2D, _or_empty(shs), _or_empty(colors_precomp), opacities,
                                   _or_empty(scales), _or_empty(rotations), _or_empty(cov3D_precomp),
                                   self.raster_settings)


def _forward_raw(self, xyz, means2D, features_dc, features_rest, opacity, scaling, rotation):
    """Extension (not in the reference API): render GaussianModel's RAW parameters -- `_xyz`, `_features_dc`,
    `_features_rest`, `_opacity`, `_scaling`, `_rotation` (scene/gaussian_model.py:147-161) -- with the getters'
    activations (:101-121) and their backward folded into the kernels.  Returns color, radii, depth, alpha."""
    return rasterize_gaussians_raw(xyz, means2D, features_dc, features_rest, opacity, scaling, rotation,
                                   self.raster_settings)


GaussianRasterizer.forward_raw = _forward_raw


class FlashGaussianRasterizer(nn.Module):
    """flashsplat_rasterization.GaussianRasterizer (FSR __init__.py:217-271): 8 outputs
    (color, radii, depth, alpha, contrib_num, used_count, proj_xy, gs_depth)."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return ops.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, gt_mask=None, unique_label=None):
        _check_inputs(shs, colors_precomp, scales, rotations, cov3D_precomp)
        if unique_label is None:
            unique_label = torch.tensor([], dtype=torch.int)
        return flash_rasterize_gaussians(gt_mask, unique_label, means3D, means2D, _or_empty(shs),
                                         _or_empty(colors_precomp), opacities, _or_empty(scales),
                                         _or_empty(rotations), _or_empty(cov3D_precomp), self.raster_settings)
