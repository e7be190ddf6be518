"""Host mirror of the optimiser step and densification bookkeeping the reference runs right after the
rasterizer backward, on the fused kernels of csrc/ws_optim.cu.

Adam                       drop-in for the `torch.optim.Adam(l, lr=0.0, eps=1e-15)` of
                           GaussianModel.training_setup (scene/gaussian_model.py:167-186): same constructor
                           arguments, same `param_groups` (the reference rewrites group["lr"] and swaps
                           group["params"][0] on densification, :188-194, :296-345), same per-parameter state
                           keys ("step", "exp_avg", "exp_avg_sq") so `cat_tensors_to_optimizer`,
                           `_prune_optimizer`, `replace_tensor_to_optimizer`, `capture()`/`restore()` and
                           checkpoints written by torch.optim.Adam keep working.  `step()` updates every
                           group in ONE kernel launch (torch: 7 foreach kernels).
add_densification_stats    train_vanilla_3dgs.py:100-103 + GaussianModel.add_densification_stats
                           (scene/gaussian_model.py:462-464) in one pass without the boolean-mask host sync.

No CPU fallback: CPU parameters raise.
"""
import ctypes as C

import torch

from . import lib as _lib
from .lib import WsAdamGroup

_MAX_GROUPS = 8


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class Adam(torch.optim.Optimizer):
    """torch.optim.Adam semantics (no weight decay, no amsgrad, not maximize) on one fused kernel."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise ValueError("wheat_3dgs_b200.optim.Adam: weight_decay and amsgrad are not implemented "
                             "(the reference uses neither, scene/gaussian_model.py:181)")
        if not 0.0 <= lr:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if not 0.0 <= eps:
            raise ValueError("Invalid epsilon value: {}".format(eps))
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError("Invalid beta parameter at index 0: {}".format(betas[0]))
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError("Invalid beta parameter at index 1: {}".format(betas[1]))
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        L = _lib.load()
        batches = {}   # (beta1, beta2, eps, device) -> [WsAdamGroup]
        keep, bump = [], []
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise RuntimeError("wheat_3dgs_b200.optim.Adam: CPU parameters are not supported (no CPU fallback)")
                if p.grad.is_sparse:
                    raise RuntimeError("Adam does not support sparse gradients")
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("wheat_3dgs_b200.optim.Adam: parameters must be contiguous float32")
                st = self.state[p]
                if len(st) == 0 or "exp_avg" not in st:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                if "step" not in st:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                g = p.grad
                if g.dtype != torch.float32 or not g.is_contiguous():
                    g = g.to(torch.float32).contiguous()
                m, v = st["exp_avg"], st["exp_avg_sq"]
                if not (m.is_contiguous() and v.is_contiguous()):
                    m = st["exp_avg"] = m.contiguous()
                    v = st["exp_avg_sq"] = v.contiguous()
                if (p.data_ptr() | m.data_ptr() | v.data_ptr()) & 15:
                    raise RuntimeError("wheat_3dgs_b200.optim.Adam: parameters and moments must be 16-byte aligned")
                # gradients may be 4-byte aligned only: they are slices of the flat bucket ops._backward() fills, at
                # offsets 3P, 6P, ... (the kernel then loads them with scalar instructions)
                keep.append(g)
                bump.append(st)
                batches.setdefault((float(beta1), float(beta2), float(group["eps"]), p.device), []).append(
                    WsAdamGroup(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(),
                                float(group["lr"]), int(st["step"].item()) + 1))
        for (beta1, beta2, eps, dev), groups in batches.items():
            with torch.cuda.device(dev):
                for i in range(0, len(groups), _MAX_GROUPS):
                    chunk = groups[i:i + _MAX_GROUPS]
                    arr = (WsAdamGroup * len(chunk))(*chunk)
                    _lib.check(L.ws_adam_step(len(chunk), arr, beta1, beta2, eps, _stream()), "ws_adam_step")
        for st in bump:   # only once every launch went through: a raised error leaves the step counts untouched
            st["step"] += 1
        return loss


def add_densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom):
    """In place: for Gaussians with radii > 0, max_radii2D = max(max_radii2D, radii),
    xyz_gradient_accum += |viewspace_grad[:, :2]|, denom += 1.  radii int32 [P]; viewspace_grad float [P,3]
    (= viewspace_point_tensor.grad); max_radii2D [P]; xyz_gradient_accum, denom [P,1]."""
    L = _lib.load()
    P = radii.numel()
    for name, t in (("radii", radii), ("viewspace_grad", viewspace_grad), ("max_radii2D", max_radii2D),
                    ("xyz_gradient_accum", xyz_gradient_accum), ("denom", denom)):
        if not t.is_cuda:
            raise RuntimeError("add_densification_stats: %s must be a CUDA tensor (no CPU fallback)" % name)
        if not t.is_contiguous():
            raise RuntimeError("add_densification_stats: %s must be contiguous" % name)
    if radii.dtype != torch.int32:
        raise RuntimeError("add_densification_stats: radii must be int32")
    if viewspace_grad.shape != (P, 3) or viewspace_grad.dtype != torch.float32:
        raise RuntimeError("add_densification_stats: viewspace_grad must be float32 [P,3]")
    for name, t in (("max_radii2D", max_radii2D), ("xyz_gradient_accum", xyz_gradient_accum), ("denom", denom)):
        if t.numel() != P or t.dtype != torch.float32:
            raise RuntimeError("add_densification_stats: %s must be float32 with P elements" % name)
    with torch.cuda.device(radii.device):
        _lib.check(L.ws_densify_stats(P, radii.data_ptr(), viewspace_grad.data_ptr(), max_radii2D.data_ptr(),
                                      xyz_gradient_accum.data_ptr(), denom.data_ptr(), _stream()),
                   "ws_densify_stats")
