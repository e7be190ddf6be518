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

    def __init__(self, num_obj, P, device="cuda", streams=3):
        """`streams` > 1: consecutive add_view calls go on that many CUDA streams in turn (orbit.StreamRing): views are
        independent and the votes are float atomics into one accumulator, so one view's short binning kernels run
        under another's blend kernel (config 3: 729 -> 800 views/s).  `labels()` joins the streams first; the tuple
        add_view returns belongs to the view's stream (order your reads after it: `join()` or a stream wait).
        Lifetimes are handled here: every CUDA tensor the caller passes is recorded on the view's stream
        (Tensor.record_stream), so the usual loop `mask = load(v).cuda(); acc.add_view(mask, ...)` may drop or
        overwrite-by-reallocation the previous mask at once -- the caching allocator does not hand the block out
        again until the view's kernels have read it -- and the returned tensors are recorded on the caller's stream."""
        from .orbit import StreamRing
        self.num_obj = int(num_obj)
        self.all_counts = torch.zeros((self.num_obj + 1, int(P)), dtype=torch.float32, device=device)
        self._ring, self._n = StreamRing(streams, device), 0
        self._ring.fork()

    def add_view(self, gt_mask, raster_settings, means3D, opacities, shs=None, colors_precomp=None, scales=None,
                 rotations=None, cov3D_precomp=None, unique_label=None):
        rs = raster_settings
        e = torch.Tensor([])
        o = lambda t: e if t is None else t
        view_stream, caller_stream = self._ring.stream_of(self._n), torch.cuda.current_stream()
        with self._ring.stream(self._n):
            self._n += 1
            out = _ops.flash_rasterize_gaussians(
                gt_mask, unique_label, rs.bg, means3D, o(colors_precomp), opacities, o(scales), o(rotations),
                rs.scale_modifier, o(cov3D_precomp), rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, o(shs), rs.sh_degree, rs.campos, rs.prefiltered, self.num_obj, rs.debug,
                used_count_out=self.all_counts)
        if view_stream is not None:
            # the inputs were allocated on the caller's stream and are read asynchronously on the view's stream
            for x in (gt_mask, unique_label, rs.bg, means3D, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                      rs.viewmatrix, rs.projmatrix, rs.campos, shs):
                if torch.is_tensor(x) and x.is_cuda:
                    x.record_stream(view_stream)
            # the outputs were allocated on the view's stream and are used on the caller's
            for x in out:
                if torch.is_tensor(x) and x.is_cuda and x is not self.all_counts:
                    x.record_stream(caller_stream)
        return out

    def join(self):
        """The current stream waits for every view added so far (call before reading all_counts)."""
        self._ring.join()

    def labels(self, gamma=0.0):
        self.join()
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
        _lib.check(L.ws_objeval_render(C.byref(call.view), C.byref(ws), P, geom.data_ptr(), _ops._ptr(binning), scratch.data_ptr(),
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


# ---------------------------------------------------------------------------------------------------------------
# run_3d_seg.py on the device: per-mask label votes, find_match over the other views, the growing loop
# ---------------------------------------------------------------------------------------------------------------
_E = torch.Tensor([])


def best_mask_iou(pred, ids, area):
    """find_match's inner search (run_3d_seg.py:134-157) for one view: `pred` bool [H,W] = alpha > 0.5, `ids` int64
    [H,W] instance-id image, `area` = bincount(ids).  Returns float64 [2] = (max IoU over the masks k >= 1, lowest
    k attaining it); (0, 0) when nothing overlaps."""
    inter = torch.bincount(ids[pred], minlength=area.numel()).to(torch.float64)
    union = area.to(torch.float64) + pred.sum().to(torch.float64) - inter
    iou = torch.where(union > 0, inter / union.clamp(min=1.0), torch.zeros_like(union))
    iou[0] = 0.0                                   # id 0 is "no mask"
    m = iou.max()
    first = torch.nonzero(iou == m)[0, 0]          # lowest index among equals (the reference scans with a strict >)
    first = torch.where(m > 0, first, torch.zeros_like(first))
    return torch.stack([m, first.to(torch.float64)])


class HeadSegmenter:
    """The 3D-segmentation loop of run_3d_seg.py:243-315 for ONE seed mask at a time, everything on the device.

    Inputs stand in for what the script reads from disk: `g` = the post-activation Gaussian tensors
    (means3D, opacities, shs, scales, rotations), `cams` = camera dicts with device matrices, `id_masks[v]` =
    float32 [H,W] instance-id image of view v (0 = no mask, k = the k-th YOLO/SAM mask of that view: the
    reference's per-mask PNGs of one view do not overlap in the synthetic workload, so one id image holds them
    all).  `pts_filter` = run_3d_seg.py:208 (Gaussians below the mean height never join an object).

      seed_counts(v, k)        opt_label_w_seg([view], [mask], pts_filter)      :263 -> :74-105
      labels_of(counts)        counts_to_obj_labels -> row 1                    :264-268 -> :107-114
      find_match(views, used)  find_match over the given views                  :275 -> :116-173
      segment_seed(v, k)       the whole loop for one seed, up to 99 rounds     :263-315

    find_match's host side (alpha > 0.5, bbox, box overlap, per-mask IoU in numpy, :130-157) runs as one
    bincount per view on the device: IoU_k = |pred & mask_k| / (|pred| + |mask_k| - |pred & mask_k|) for every
    mask of the view at once.  The YOLO boxes are the masks' bounding boxes, so a mask whose box misses the
    prediction's has IoU 0 and can never win `iou > max_iou` (max_iou starts at 0.0): the box filter only saves
    work and is dropped.  Ties keep the lowest mask index, as the reference's strict `>` scan does."""

    def __init__(self, g, cams, id_masks, bg, sh_degree, iou_threshold=0.6, pts_filter=None):
        self.g, self.cams, self.bg, self.deg = g, cams, bg, int(sh_degree)
        self.id_masks = id_masks
        self.iou_threshold = float(iou_threshold)
        self.P = int(g["means3D"].shape[0])
        self.dev = g["means3D"].device
        if pts_filter is None:   # run_3d_seg.py:206-208
            z = g["means3D"][:, 2]
            pts_filter = z < z.mean()
        self.pts_filter = pts_filter
        self._ids = {}     # view -> (int64 id image, per-id pixel counts)
        self.renders = 0   # rasterizer forwards issued (the unit the reference's cost is counted in)

    # -- votes
    def _flash(self, v, mask, num_obj, used_out=None):
        c = self.cams[v]
        g = self.g
        self.renders += 1
        return _ops.flash_rasterize_gaussians(mask, None, self.bg, g["means3D"], _E, g["opacities"], g["scales"],
                                              g["rotations"], 1.0, _E, c["viewmatrix"], c["projmatrix"], c["tanfovx"],
                                              c["tanfovy"], c["image_height"], c["image_width"], g["shs"], self.deg,
                                              c["campos"], False, int(num_obj), False, used_count_out=used_out)

    def view_counts(self, v, num_obj=None):
        """ONE multi-label vote pass over all masks of view v: used_count [num_obj+1, P]."""
        m = self.id_masks[v]
        if num_obj is None:
            num_obj = int(m.max().item())
        return self._flash(v, m, num_obj)[9]

    def seed_counts(self, v, k, shared=None, filtered=True):
        """all_counts [2,P] of mask k of view v (run_3d_seg.py:86-104).  `shared` = view_counts(v): the masks of a
        view do not overlap, so the binary render's row 1 is the multi-label row k and its row 0 the sum of all
        other rows; without it the reference's binary render (gt_mask in {0,1}, obj_num=1) is issued."""
        if shared is not None:
            counts = torch.stack([shared.sum(dim=0) - shared[k], shared[k]], dim=0)
        else:
            binary = (self.id_masks[v] == float(k)).to(torch.float32)
            counts = self._flash(v, binary, 1)[9].clone()
        if filtered:
            counts[1:, self.pts_filter] = 0
        return counts

    @staticmethod
    def labels_of(counts):
        """Row 1 of counts_to_obj_labels: the Gaussians of the object (bool [P])."""
        return multi_instance_opt(counts, 0.0)[1]

    # -- matching
    def _view_ids(self, v):
        if v not in self._ids:
            ids = self.id_masks[v].to(torch.int64)
            self._ids[v] = (ids, torch.bincount(ids.reshape(-1)))
        return self._ids[v]

    def render_subset_alpha(self, v, used_mask):
        """flashsplat_render(used_mask=...)['alpha'] (gaussian_renderer/__init__.py:154-186): the selected Gaussians
        alone.  gt_mask is None there (all-ones labels, obj_num 2) and only alpha is read, so the plain forward of
        the same blend is issued (alpha is the same arithmetic in both flavours)."""
        g, c = self.g, self.cams[v]
        self.renders += 1
        out = _ops.rasterize_gaussians(self.bg, g["means3D"][used_mask], _E, g["opacities"][used_mask],
                                       g["scales"][used_mask], g["rotations"][used_mask], 1.0, _E, c["viewmatrix"],
                                       c["projmatrix"], c["tanfovx"], c["tanfovy"], c["image_height"], c["image_width"],
                                       g["shs"][used_mask], self.deg, c["campos"], False, False)
        return out[3]

    def find_match(self, views, used_mask):
        """[(view, mask index, iou)] for the views whose best mask has IoU > threshold (run_3d_seg.py:116-173)."""
        if not views:
            return []
        best = torch.zeros((len(views), 2), dtype=torch.float64, device=self.dev)
        for i, v in enumerate(views):
            pred = self.render_subset_alpha(v, used_mask)[0] > 0.5
            ids, area = self._view_ids(v)
            best[i] = best_mask_iou(pred, ids, area)
        res = best.cpu().tolist()                          # one read-back for all views
        return [(v, int(k), iou) for v, (iou, k) in zip(views, res) if iou > self.iou_threshold and iou > 0.0]

    # -- the loop
    def segment_seed(self, v, k, shared=None, max_rounds=99):
        """run_3d_seg.py:263-315 for the seed mask k of view v.  Returns None when no Gaussian is identified (:265),
        else {"used_mask", "all_counts", "matched": [(view, mask)], "rounds"}."""
        all_views = list(range(len(self.cams)))
        counts = self.seed_counts(v, k, shared=shared)
        used = self.labels_of(counts)
        if int(used.sum().item()) == 0:
            return None
        matched = [(v, k)]
        new = self.find_match([u for u in all_views if u != v], used)
        matched += [(u, j) for u, j, _ in new]
        rounds = 0
        while new and rounds < max_rounds:
            rounds += 1
            for u, j, _ in new:                            # opt_label_w_seg(new views, new masks), no pts_filter (:301)
                counts += self.seed_counts(u, j, filtered=False)
            used = self.labels_of(counts)
            done = {u for u, _ in matched}
            new = self.find_match([u for u in all_views if u not in done], used)
            matched += [(u, j) for u, j, _ in new]
        return {"used_mask": used, "all_counts": counts, "matched": matched, "rounds": rounds}
