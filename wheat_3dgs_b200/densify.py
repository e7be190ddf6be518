"""Host mirror of GaussianModel's densification on csrc/ws_densify.cu.

densify_and_prune(model, max_grad, min_opacity, extent, max_screen_size)
        drop-in for GaussianModel.densify_and_prune (scene/gaussian_model.py:440-458); `model` is anything with the
        reference's attributes (_xyz, _features_dc, _features_rest, _opacity, _scaling, _rotation, _which_object,
        optimizer, xyz_gradient_accum, denom, max_radii2D, percent_dense), e.g. the reference's own GaussianModel:
            GaussianModel.densify_and_prune = wheat_3dgs_b200.densify.densify_and_prune
        Clone (:420-438), split (:392-418) and prune (:327-352) with their optimizer surgery
        (cat_tensors_to_optimizer :354-374, _prune_optimizer :310-325) run as one plan + one gather: same rows, same
        order, same Adam state (moments kept for survivors, zero for new rows, "step" untouched), statistics
        re-created as zeros (:387-389).  The split samples are drawn with the reference's own torch.normal call, so
        the same seed gives the same children.
reset_opacity(model)
        drop-in for GaussianModel.reset_opacity (:234-237).

Differences a caller can see: none in values; `torch.cuda.empty_cache()` (:456) is not called.  No CPU path.
"""
import ctypes as C

import torch
from torch import nn

from . import lib as _lib
from .lib import WS_ROW_COPY, WS_ROW_MOMENT, WS_ROW_SCALING, WS_ROW_XYZ, WsRowArray

_GROUP_ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity",
               "scaling": "_scaling", "rotation": "_rotation"}
_ROLE = {"xyz": WS_ROW_XYZ, "scaling": WS_ROW_SCALING}


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32(t, what):
    if not t.is_cuda:
        raise RuntimeError("densify: %s must be a CUDA tensor (wheat_3dgs_b200 has no CPU path)" % what)
    if t.dtype != torch.float32:
        raise RuntimeError("densify: %s must be float32" % what)
    return t.detach().contiguous()


def plan(model, max_grad, min_opacity, extent, max_screen_size):
    """The verdicts only: (flags uint8[P], scratch, counts) with counts = (kept originals, kept clones, kept children
    per copy, split-selected)."""
    L = _lib.load()
    P = int(model._xyz.shape[0])
    dev = model._xyz.device
    flags = torch.empty((P,), dtype=torch.uint8, device=dev)
    scratch = torch.empty((L.ws_densify_plan_bytes(P),), dtype=torch.uint8, device=dev)
    counts = (C.c_uint32 * 4)()
    accum, denom = _f32(model.xyz_gradient_accum, "xyz_gradient_accum"), _f32(model.denom, "denom")
    if accum.numel() != P or denom.numel() != P:
        raise RuntimeError("densify: xyz_gradient_accum / denom must have one entry per Gaussian")
    opacity, scaling = _f32(model._opacity, "_opacity"), _f32(model._scaling, "_scaling")
    use_screen = 1 if max_screen_size else 0
    with torch.cuda.device(dev):
        _lib.check(L.ws_densify_plan(P, accum.data_ptr(), denom.data_ptr(), opacity.data_ptr(), scaling.data_ptr(),
                                     float(max_grad), float(min_opacity), float(model.percent_dense * extent),
                                     float(0.1 * extent), use_screen, float(max_screen_size or 0.0), flags.data_ptr(),
                                     scratch.data_ptr(), scratch.numel(), counts, _stream()), "ws_densify_plan")
    return flags, scratch, counts


@torch.no_grad()
def densify_and_prune(model, max_grad, min_opacity, extent, max_screen_size, return_rows=False, samples=None):
    """`samples` (float32 [2S,3] on the device) replays a recorded draw instead of calling torch.normal;
    `return_rows=True` returns (source Gaussian of every output row, counts, samples)."""
    L = _lib.load()
    P = int(model._xyz.shape[0])
    dev = model._xyz.device
    flags, scratch, counts = plan(model, max_grad, min_opacity, extent, max_screen_size)
    nA, nB, nC, S = [int(c) for c in counts]
    rows = nA + nB + 2 * nC
    if samples is not None:
        samples = _f32(samples, "samples")
        if samples.numel() != 6 * S:
            raise RuntimeError("densify: samples must be [2*%d, 3]" % S)
    elif S > 0:
        # scene/gaussian_model.py:398-400, the reference's own draw (same generator, same shapes)
        sel = (flags & 8).bool()
        stds = torch.exp(model._scaling.detach())[sel].repeat(2, 1)
        means = torch.zeros((stds.size(0), 3), device=dev)
        samples = torch.normal(mean=means, std=stds).contiguous()
    table, keep, new = [], [], {}

    def add(src, role):
        src = src.detach().contiguous()
        w = src.numel() // max(P, 1)
        if src.element_size() not in (4, 8):
            raise RuntimeError("densify: rows must be made of 4- or 8-byte elements")
        dst = torch.empty((rows,) + tuple(src.shape[1:]), dtype=src.dtype, device=dev)
        if w == 0:   # zero-width rows (features_rest [P,0,3] at max_sh_degree 0): nothing to move
            return dst
        table.append(WsRowArray(src.data_ptr(), dst.data_ptr(), w * src.element_size() // 4, role))
        keep.append(src)
        return dst

    opt = model.optimizer
    for group in opt.param_groups:
        if len(group["params"]) != 1:
            raise RuntimeError("densify: one parameter per group (scene/gaussian_model.py:357)")
        name = group["name"]
        p = group["params"][0]
        _f32(p, name)
        st = opt.state.get(p, None)
        ent = {"param": add(p, _ROLE.get(name, WS_ROW_COPY))}
        if st is not None:
            ent["exp_avg"] = add(st["exp_avg"], WS_ROW_MOMENT)
            ent["exp_avg_sq"] = add(st["exp_avg_sq"], WS_ROW_MOMENT)
        new[name] = ent
    which = getattr(model, "_which_object", None)
    new_which = None
    if which is not None and which.numel() > 0:
        if not which.is_cuda:
            raise RuntimeError("densify: _which_object must be a CUDA tensor")
        new_which = add(which, WS_ROW_COPY)
    src_rows = torch.empty((rows,), dtype=torch.int32, device=dev) if return_rows else None
    arr = (WsRowArray * len(table))(*table)
    rot = model._rotation.detach().contiguous()
    with torch.cuda.device(dev):
        _lib.check(L.ws_densify_gather(P, flags.data_ptr(), scratch.data_ptr(), scratch.numel(), counts, rot.data_ptr(),
                                       samples.data_ptr() if samples is not None else None, len(table), arr,
                                       src_rows.data_ptr() if src_rows is not None else None, _stream()),
                   "ws_densify_gather")
    # the optimizer surgery of :310-325 / :354-374: new Parameter objects, the state dict moves with them
    for group in opt.param_groups:
        name = group["name"]
        old = group["params"][0]
        ent = new[name]
        st = opt.state.get(old, None)
        param = nn.Parameter(ent["param"].requires_grad_(True))
        if st is not None:
            st["exp_avg"], st["exp_avg_sq"] = ent["exp_avg"], ent["exp_avg_sq"]
            del opt.state[old]
            opt.state[param] = st
        group["params"][0] = param
        if name in _GROUP_ATTR:
            setattr(model, _GROUP_ATTR[name], param)
    if new_which is not None:
        model._which_object = new_which
    model.xyz_gradient_accum = torch.zeros((rows, 1), device=dev)
    model.denom = torch.zeros((rows, 1), device=dev)
    model.max_radii2D = torch.zeros((rows,), device=dev)
    if return_rows:
        return src_rows, (nA, nB, nC, S), samples
    return None


@torch.no_grad()
def reset_opacity(model):
    L = _lib.load()
    opt = model.optimizer
    for group in opt.param_groups:
        if group["name"] != "opacity":
            continue
        old = group["params"][0]
        new = _f32(old, "_opacity").clone()
        st = opt.state.get(old, None)
        m = v = None
        if st is not None:
            # replace_tensor_to_optimizer :295-308: fresh zero moments
            m, v = torch.empty_like(new), torch.empty_like(new)
        with torch.cuda.device(new.device):
            _lib.check(L.ws_reset_opacity(new.numel(), new.data_ptr(), m.data_ptr() if m is not None else None,
                                          v.data_ptr() if v is not None else None, _stream()), "ws_reset_opacity")
        param = nn.Parameter(new.requires_grad_(True))
        if st is not None:
            st["exp_avg"], st["exp_avg_sq"] = m, v
            del opt.state[old]
            opt.state[param] = st
        group["params"][0] = param
        model._opacity = param
