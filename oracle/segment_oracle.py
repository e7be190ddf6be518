"""CPU restatement of eval_obj_labels (utils/wheatgs_helper.py:245-271 = eval_wheatgs.py:29-64) on top of the
CPU oracle of the flashsplat forward (oracle.forward).  TEST INFRASTRUCTURE ONLY: imported by tests/, never by
the product.  Pinned against the reference: per-object alpha/depth from the unmodified reference CUDA rasterizer
(libref_fsr.so on a B200) fed through the reference's own Python eval_obj_labels (tests/golden/
make_golden_objeval.py -> tests/golden/objeval.npz, checked in tests/test_oracle_golden.py).
"""
import numpy as np

import oracle

_KEYS = ("means3D", "opacities", "scales", "rotations", "shs")


def subset_render(g, cam, bg, mask, sh_degree=None):
    """flashsplat_render(..., used_mask=mask) (gaussian_renderer/__init__.py:154-187): alpha, depth [1,H,W]."""
    sub = {k: np.ascontiguousarray(g[k][mask]) for k in _KEYS if g.get(k) is not None}
    if "sh_degree" in g:
        sub["sh_degree"] = g["sh_degree"]
    o = oracle.forward(sub, cam, bg, sh_degree=sh_degree, gt_mask=None, num_obj=2)
    return o["out_alpha"], o["out_depth"]


def merge(renders, shape):
    """wheatgs_helper.py:251-270 given {obj_idx: (alpha, depth)} in ascending obj_idx.  float32 means, as torch."""
    pred = np.zeros(shape, np.float32)
    min_depth = np.ones(shape, np.float32)
    for obj_idx in sorted(renders):
        alpha, depth = renders[obj_idx]
        _pix = alpha > 0.5
        pix = _pix.copy()
        overlap = _pix & (pred > 0)
        if overlap.sum() > 0:
            if np.float32(min_depth[overlap].mean(dtype=np.float64)) < np.float32(depth[overlap].mean(dtype=np.float64)):
                pix[_pix] = ~(pred[_pix] > 0)
        pred[pix] = obj_idx
        min_depth[pix] = depth[pix]
    return pred


def eval_obj_labels(all_obj_labels, g, cam, bg, sh_degree=None, return_renders=False):
    labels = np.asarray(all_obj_labels).astype(bool)
    H, W = cam["image_height"], cam["image_width"]
    renders = {}
    for obj_idx in range(labels.shape[0]):
        if obj_idx == 0 or labels[obj_idx].sum() == 0:   # wheatgs_helper.py:253-255
            continue
        renders[obj_idx] = subset_render(g, cam, bg, labels[obj_idx], sh_degree)
    if not renders:
        return (None, renders) if return_renders else None
    pred = merge(renders, (1, H, W))
    return (pred, renders) if return_renders else pred
