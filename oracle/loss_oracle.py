"""CPU restatement (numpy, float64 accumulation of float32 data) of the reference's training loss
(utils/loss_utils.py:17-62 as combined by train_vanilla_3dgs.py:77-80) and of its gradient w.r.t.
the rendered image.  TEST INFRASTRUCTURE ONLY: imported by tests/ and bench.py's checker legs, never
by the product.  Pinned against the reference's own Python (tests/golden/make_golden_loss.py ->
tests/golden/loss_*.npz, checked in tests/test_oracle_golden.py).

    window(): loss_utils.py:23-31      ssim_map(): :47-57      l1: :17-18
"""
import math

import numpy as np

C1, C2 = 0.01 ** 2, 0.03 ** 2


def window():
    g = np.array([math.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)], np.float32)
    return (g / g.sum(dtype=np.float32)).astype(np.float32)


def _conv(img, w):
    """Zero-padded separable 11x11 convolution of [planes,H,W] (F.conv2d padding=5, groups=channels)."""
    p, H, W = img.shape
    pad = np.zeros((p, H + 10, W + 10), np.float64)
    pad[:, 5:-5, 5:-5] = img
    tmp = np.zeros((p, H + 10, W), np.float64)
    for k in range(11):
        tmp += w[k] * pad[:, :, k:k + W]
    out = np.zeros((p, H, W), np.float64)
    for k in range(11):
        out += w[k] * tmp[:, k:k + H, :]
    return out


def ssim_parts(x, y):
    w = window().astype(np.float64)
    x = x.astype(np.float64)
    y = y.astype(np.float64)
    mu1, mu2 = _conv(x, w), _conv(y, w)
    exx, eyy, exy = _conv(x * x, w), _conv(y * y, w), _conv(x * y, w)
    s1, s2, s12 = exx - mu1 * mu1, eyy - mu2 * mu2, exy - mu1 * mu2
    A1, A2 = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    B1, B2 = mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    m = A1 * A2 / (B1 * B2)
    return w, m, mu1, mu2, A1, A2, B1, B2


def loss_and_grad(img, gt, lambda_dssim):
    """img, gt: float32 [...,H,W].  Returns dict(loss, l1, ssim, grad) with grad = dloss/dimg."""
    shape = img.shape
    x = img.reshape(-1, shape[-2], shape[-1])
    y = gt.reshape(-1, shape[-2], shape[-1])
    n = x.size
    w, m, mu1, mu2, A1, A2, B1, B2 = ssim_parts(x, y)
    ssim = m.mean()
    l1 = np.abs(x.astype(np.float64) - y).mean()
    loss = (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - ssim)
    inv = 1.0 / (B1 * B2)
    dmu1 = 2.0 * inv * (mu2 * (A2 - A1) - mu1 * m * (B2 - B1))
    dexx = -m / B2
    dexy = 2.0 * A1 * inv
    dssim = _conv(dmu1, w) + 2.0 * x * _conv(dexx, w) + y * _conv(dexy, w)  # symmetric window: conv^T = conv
    grad = (-lambda_dssim / n) * dssim + ((1.0 - lambda_dssim) / n) * np.sign(x.astype(np.float64) - y)
    return {"loss": loss, "l1": l1, "ssim": ssim, "grad": grad.reshape(shape).astype(np.float32),
            "ssim_per_plane": m.reshape(m.shape[0], -1).mean(axis=1)}
