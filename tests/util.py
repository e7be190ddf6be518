"""Shared comparison helpers of the parity tests.

Tiers (BASELINE.json north_star / SURVEY.md 8c):
  exact      tile keys, sorted order, tile ranges, radii, n_contrib, label assignments: bit for bit
  image      color / depth / alpha: <= 1e-4 absolute
  gradient   1e-3 relative.  The reference's gradients are float atomics in arbitrary order, so an
             element that is a sum of large cancelling terms moves by far more than 1e-3 of ITSELF
             between two runs of the reference; the tolerance is therefore
                 |a - b| <= 1e-3 |b| + max(1e-4 max|b|, 4 x the reference's own run-to-run spread)
             (the spread is stored with every golden and measured live on the GPU box).
"""
import numpy as np

IMAGE_ATOL = 1e-4
GRAD_RTOL = 1e-3
GRAD_NAMES = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
              "dL_drotations"]


def assert_exact(name, a, b):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, "%s: shape %s vs %s" % (name, a.shape, b.shape)
    bad = np.flatnonzero(a.reshape(-1) != b.reshape(-1))
    assert bad.size == 0, "%s: %d of %d differ, first at %d: %r vs %r" % (
        name, bad.size, a.size, bad[0], a.reshape(-1)[bad[0]], b.reshape(-1)[bad[0]])


def assert_bits(name, a, b):
    """float arrays compared as bit patterns"""
    assert_exact(name, np.ascontiguousarray(a, np.float32).view(np.uint32), np.ascontiguousarray(b, np.float32).view(np.uint32))


def assert_image(name, a, b, atol=IMAGE_ATOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, "%s: shape %s vs %s" % (name, a.shape, b.shape)
    assert np.all(np.isfinite(a)), "%s: non-finite values" % name
    err = np.abs(a - b).max() if a.size else 0.0
    assert err <= atol, "%s: max abs err %g > %g" % (name, err, atol)


def assert_grad(name, a, b, spread=0.0, rtol=GRAD_RTOL):
    a, b = np.asarray(a, np.float64).reshape(-1), np.asarray(b, np.float64).reshape(-1)
    assert a.shape == b.shape, "%s: shape %s vs %s" % (name, a.shape, b.shape)
    if a.size == 0:
        return
    assert np.all(np.isfinite(a)), "%s: non-finite values" % name
    floor = max(1e-4 * np.abs(b).max(), 4.0 * float(spread))
    lim = rtol * np.abs(b) + floor
    bad = np.flatnonzero(np.abs(a - b) > lim)
    assert bad.size == 0, "%s: %d of %d beyond rtol %g + floor %g; worst %g vs %g" % (
        name, bad.size, a.size, rtol, floor, a[bad[0]], b[bad[0]])
    # and the tensor as a whole is far inside the tier
    nrm = np.linalg.norm(b)
    if nrm > 0:
        assert np.linalg.norm(a - b) / nrm < rtol, "%s: norm-wise relative error %g" % (name, np.linalg.norm(a - b) / nrm)
