"""CPU restatement (numpy float32) of the optimiser step and densification bookkeeping the reference runs
either side of the rasterizer backward.  TEST INFRASTRUCTURE ONLY: imported by tests/,
never by the product.  Pinned against torch.optim.Adam and the reference's own
GaussianModel.add_densification_stats run in this container (tests/golden/make_golden_optim.py ->
tests/golden/optim.npz, checked in tests/test_oracle_golden.py).

adam_step():      torch/optim/adam.py `_single_tensor_adam` / `_multi_tensor_adam` (torch 2.11; the reference
                  calls it through scene/gaussian_model.py:181 and train_vanilla_3dgs.py:111), float32 tensors,
                  Python-double scalars
densify_stats():  train_vanilla_3dgs.py:101 and scene/gaussian_model.py:462-464
"""
import math

import numpy as np

f32 = np.float32


def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-15):
    """One step on one tensor; returns (param, exp_avg, exp_avg_sq).  `step` counts from 1."""
    p, g, m, v = (np.asarray(a, f32) for a in (param, grad, exp_avg, exp_avg_sq))
    m = m + f32(1.0 - beta1) * (g - m)                      # exp_avg.lerp_(grad, 1 - beta1)
    v = v * f32(beta2) + f32(1.0 - beta2) * (g * g)         # mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    step_size = lr / bc1
    denom = np.sqrt(v) / f32(math.sqrt(bc2)) + f32(eps)     # (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
    p = p + f32(-step_size) * (m / denom)                   # param.addcdiv_(exp_avg, denom, value=-step_size)
    return p.astype(f32), m.astype(f32), v.astype(f32)


def densify_stats(radii, dL_dmeans2D, max_radii2D, xyz_gradient_accum, denom):
    """Returns the three updated accumulators (flat float32 [P])."""
    r = np.asarray(radii)
    vis = r > 0
    g = np.asarray(dL_dmeans2D, f32)
    mr = np.asarray(max_radii2D, f32).reshape(-1).copy()
    acc = np.asarray(xyz_gradient_accum, f32).reshape(-1).copy()
    dn = np.asarray(denom, f32).reshape(-1).copy()
    mr[vis] = np.maximum(mr[vis], r[vis].astype(f32))
    acc[vis] += np.sqrt(g[vis, 0] * g[vis, 0] + g[vis, 1] * g[vis, 1])
    dn[vis] += f32(1.0)
    return mr, acc, dn
