"""Host mirror of the segmentation helpers either side of the flashsplat rasterizer:

multi_instance_opt(all_contrib, gamma)     run_3d_seg.py:54-72 = utils/wheatgs_helper.py:68-86, one device pass
VoteAccumulator                            the `all_counts += used_count.cpu()` loops of run_3d_seg.py:86-97,
                                           render_360.py:168-175, eval_wheatgs.py:106-114 kept on the device: the
                                           rasterizer accumulates into ONE persistent [num_obj+1, P] buffer
object_sets(all_obj_labels)                bool [rows, P] -> CSR membership lists for eval_obj_labels
eval_obj_labels(...)                       utils/wheatgs_helper.py:245-271 = eval_wheatgs.py:29-64: the
                                           segmentation image of one view, one batched pass instead of one
                                           flashsplat_render per object (csrc/ws_objeval.cu)

PyTorch is plumbing (allocation, the CSR build); no CPU path.
"""
import ctypes as C
import math

import torch

from . import lib as _lib
from . import ops as _ops
from .lib import WsObjSets


def multi_instance_opt(all_contrib, gamma=0.0):
    return _ops.multi_instance_opt(all_contrib, gamma)


class VoteAccumulator:
    """all_counts of run_3d_seg.py:86-97 on the device.  `add_view` runs one flashsplat forward whose votes are
    accumulated straight into `all_counts` (no per-view zero fill, no .cpu() round trip)."""

    def __init__(self, num_obj, P, device="cuda"):
        self.num_obj = int(num_obj)
        self.all_counts = torch.zeros((self.num_obj + 1, int(P)), dtype=torch.float32, device=device)

    def add_view(self, gt_mask, raster_settings, means3D, opacities, shs=None, colors_precomp=None, scales=None,
                 rotations=None, cov3D_precomp=None, unique_label=None):
        rs = raster_settings
        e = torch.Tensor([])
        o = lambda t: e if t is None else t
        return _ops.flash_rasterize_gaussians(
            gt_mask, unique_label, rs.bg, means3D, o(colors_precomp), opacities, o(scales), o(rotations),
            rs.scale_modifier, o(cov3D_precomp), rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
            rs.image_height, rs.image_width, o(shs), rs.sh_degree, rs.campos, rs.prefiltered, self.num_obj, rs.debug,
            used_count_out=self.all_counts)

    def labels(self, gamma=0.0):
        return multi_instance_opt(self.all_counts, gamma)


class ObjectSets:
    """CSR membership lists of all_obj_labels: Gaussian g belongs to member_obj[member_start[g]:member_start[g+1]]
    (ascending object rows, row 0 = background left out as eval_obj_labels skips it)."""

    def __init__(self, num_rows, member_start, member_obj):
        self.num_rows, self.member_start, self.member_obj = int(num_rows), member_start, member_obj
        self.any = member_obj.numel() > 0


def object_sets(all_obj_labels):
    if all_obj_labels.dim() != 2:
        raise RuntimeError("all_obj_labels must be [num_rows, P]")
    if not all_obj_labels.is_cuda:
        raise RuntimeError("all_obj_labels must be a CUDA tensor (wheat_3dgs_b200 has no CPU path)")
    rows, P = all_obj_labels.shape
    lab = all_obj_labels.bool()
    if rows > 1:
        nz = lab[1:].t().nonzero()          # (g, row-1) sorted by g, then row
        member_obj = (nz[:, 1] + 1).to(torch.int32).contiguous()
        counts = torch.bincount(nz[:, 0], minlength=P)
    else:
        member_obj = torch.zeros((0,), dtype=torch.int32, device=lab.device)
        counts = torch.zeros((P,), dtype=torch.int64, device=lab.device)
    member_start = torch.zeros((P + 1,), dtype=torch.int32, device=lab.device)
    member_start[1:] = torch.cumsum(counts, 0).to(torch.int32)
    if member_obj.numel() == 0:   # keep a valid pointer for the C ABI
        member_obj = torch.zeros((1,), dtype=torch.int32, device=lab.device)[:0]
    return ObjectSets(rows, member_start, member_obj)


def eval_obj_labels_tensors(sets, raster_settings, means3D, opacities, shs=None, colors_precomp=None, scales=None,
                            rotations=None, cov3D_precomp=None):
    """pred_mask float [1,H,W] on the device (object id per pixel, 0 = none), or None when no object has a
    Gaussian (the reference returns None then, wheatgs_helper.py:248,270)."""
    if not isinstance(sets, ObjectSets):
        sets = object_sets(sets)
    if not sets.any:
        return None
    rs = raster_settings
    e = torch.Tensor([])
    o = lambda t: e if t is None else t
    call = _ops._Call(rs.bg, means3D, o(colors_precomp), opacities, o(scales), o(rotations), rs.scale_modifier,
                      o(cov3D_precomp), rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                      rs.image_width, o(shs), rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
    dev, P, H, W = call.device, call.P, call.H, call.W
    if sets.member_start.numel() != P + 1:
        raise RuntimeError("object sets were built for %d Gaussians, the scene has %d" % (sets.member_start.numel() - 1, P))
    pred = torch.zeros((1, H, W), dtype=torch.float32, device=dev)
    if P == 0:
        return pred
    L = _lib.load()
    u8 = dict(dtype=torch.uint8, device=dev)
    mo = sets.member_obj if sets.member_obj.numel() else torch.zeros((1,), dtype=torch.int32, device=dev)
    ws = WsObjSets(sets.num_rows, sets.member_start.data_ptr(), mo.data_ptr())
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    geom = torch.empty((L.ws_geom_bytes(P),), **u8)
    with torch.cuda.device(dev):
        st = _ops._stream()
        R = _lib.check(L.ws_objeval_prepare(C.byref(call.view), C.byref(call.gauss), C.byref(ws), radii.data_ptr(),
                                            geom.data_ptr(), geom.numel(), st), "ws_objeval_prepare")
        binning = _ops._binning_alloc(L.ws_binning_bytes(R, W, H) if R > 0 else 0, dev)
        scratch = torch.empty((L.ws_objeval_bytes(sets.num_rows, W, H, R),), **u8)
        J = _lib.check(L.ws_objeval_bin(C.byref(call.view), C.byref(call.gauss), C.byref(ws), radii.data_ptr(),
                                        geom.data_ptr(), _ops._ptr(binning), binning.numel(), scratch.data_ptr(),
                                        scratch.numel(), R, st), "ws_objeval_bin")
        job_out = torch.empty((max(J, 1) * 512,), dtype=torch.float32, device=dev)
        _lib.check(L.ws_objeval_render(C.byref(call.view), C.byref(ws), _ops._ptr(binning), scratch.data_ptr(),
                                       scratch.numel(), R, J, job_out.data_ptr(), job_out.numel() * 4,
                                       pred.data_ptr(), st), "ws_objeval_render")
    eval_obj_labels_tensors.last_counts = (R, J)
    return pred


def eval_obj_labels(all_obj_labels, viewpoint_cam, gaussians, pipe, background):
    """Drop-in for utils/wheatgs_helper.py:245-271 (same arguments; `all_obj_labels` may also be an ObjectSets
    built once with object_sets()).  Reads from the camera and model what flashsplat_render reads
    (gaussian_renderer/__init__.py:109-200).  The result stays on the device."""
    from .rasterizer import GaussianRasterizationSettings
    if getattr(pipe, "compute_cov3D_python", False) or getattr(pipe, "convert_SHs_python", False):
        raise RuntimeError("eval_obj_labels: compute_cov3D_python / convert_SHs_python are not supported")
    rs = GaussianRasterizationSettings(
        image_height=int(viewpoint_cam.image_height), image_width=int(viewpoint_cam.image_width),
        tanfovx=math.tan(viewpoint_cam.FoVx * 0.5), tanfovy=math.tan(viewpoint_cam.FoVy * 0.5), bg=background,
        scale_modifier=1.0, viewmatrix=viewpoint_cam.world_view_transform,
        projmatrix=viewpoint_cam.full_proj_transform, sh_degree=gaussians.active_sh_degree,
        campos=viewpoint_cam.camera_center, prefiltered=False, debug=False)
    return eval_obj_labels_tensors(all_obj_labels, rs, gaussians.get_xyz, gaussians.get_opacity,
                                   shs=gaussians.get_features, scales=gaussians.get_scaling,
                                   rotations=gaussians.get_rotation)
