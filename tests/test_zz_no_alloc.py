"""ws_knn_dist2_static: the nearest-neighbour distances for callers that cannot bring scratch memory.

ws_knn_dist2 wants ws_knn_bytes(P) bytes of device scratch from its caller.  A caller that only holds the two
pointers it was given (a bare ctypes / cgo binding with no CUDA runtime of its own, a plug-in living inside a
process whose single memory pool already owns the card, a code path that must not trigger the device-wide
synchronisation cudaMalloc implies) cannot produce them.  ws_knn_dist2_static keeps that scratch in a fixed 8 MiB
array inside the library and allocates nothing; the price is a limit of WS_KNN_STATIC_MAX_POINTS points and that all
calls on a device share (and leave their leftovers in) the same array.

This file is named test_zz_* so that pytest collects it after every other file: `pytest -x` stops at the first
failure and a failure here must not hide the parity tests.  It has to pass both alone in a fresh process (where its
first call is the one that looks up the array's device address) and after the rest of the suite.

The comparison is against the existing entry point on the same input, bit for bit (the exact tier the other
distCUDA2 tests use, tests/util.py assert_bits).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from util import assert_bits
from wheat_3dgs_b200 import lib as wslib

WS_E_SCRATCH = -2


def _points(P, seed):
    rng = np.random.default_rng(seed)
    # clustered cloud + a few far points and exact duplicates, like an SfM reconstruction
    pts = rng.normal(size=(P, 3)).astype(np.float32) * rng.choice([0.05, 1.0, 7.0], size=(P, 1)).astype(np.float32)
    pts[: P // 50] = pts[P // 50: 2 * (P // 50)]
    return pts


def _both(L, pts):
    """(distances from ws_knn_dist2 with caller scratch, distances from ws_knn_dist2_static), through the C ABI."""
    dev = torch.device("cuda", 0)
    P = pts.shape[0]
    p = torch.from_numpy(pts).to(dev).contiguous()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    want = torch.zeros(P, dtype=torch.float32, device=dev)
    scratch = torch.empty(L.ws_knn_bytes(P), dtype=torch.uint8, device=dev)
    wslib.check(L.ws_knn_dist2(P, p.data_ptr(), want.data_ptr(), scratch.data_ptr(), scratch.numel(), stream),
                "ws_knn_dist2")
    got = torch.full((P,), float("nan"), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    wslib.check(L.ws_knn_dist2_static(P, p.data_ptr(), got.data_ptr(), stream), "ws_knn_dist2_static")
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return want.cpu().numpy(), got.cpu().numpy(), free_before - free_after


@pytest.mark.gpu
def test_static_scratch_knn_equals_caller_scratch_knn_on_consecutive_calls():
    L = wslib.load()
    with torch.cuda.device(0):
        # Consecutive calls on one stream share the built-in array.  The sizes go down and up again so that every call
        # after the first finds stale keys, boxes and look-back words of a DIFFERENT layout in it.
        taken = []
        for i, P in enumerate((60000, 20001, 60000, 1, 4, wslib.WS_KNN_STATIC_MAX_POINTS)):
            want, got, took = _both(L, _points(P, 100 + i))
            # the output was pre-filled with NaN; fewer than 4 points legitimately give +inf (no third neighbour)
            assert not np.isnan(got).any(), "call %d (P=%d): output not fully written" % (i, P)
            assert_bits("ws_knn_dist2_static call %d (P=%d)" % (i, P), got, want)
            taken.append(took)
        # the first call may load the module; after that the device's free memory must not move
        assert all(t == 0 for t in taken[1:]), "device memory taken by calls: %r" % (taken,)


@pytest.mark.gpu
def test_static_scratch_knn_same_input_twice_gives_same_bits():
    L = wslib.load()
    with torch.cuda.device(0):
        pts = _points(33333, 7)
        a = _both(L, pts)[1]
        b = _both(L, pts)[1]
        assert_bits("second call on the same points", b, a)


def test_static_scratch_knn_rejects_more_points_than_the_array_holds():
    """Runs without a GPU: the size check comes before any CUDA call, the pointers are never dereferenced."""
    L = wslib.load()
    P = wslib.WS_KNN_STATIC_MAX_POINTS + 1
    rc = L.ws_knn_dist2_static(P, C.c_void_p(256), C.c_void_p(256), None)
    assert rc == WS_E_SCRATCH
    msg = wslib.last_error()
    assert "ws_knn_dist2_static" in msg and str(wslib.WS_KNN_STATIC_MAX_POINTS) in msg and "ws_knn_dist2 " in msg
    # the documented limit really fits the array, and empty input is a no-op like ws_knn_dist2
    assert L.ws_knn_bytes(wslib.WS_KNN_STATIC_MAX_POINTS) <= 8 << 20
    assert L.ws_knn_dist2_static(0, None, None, None) == 0
    assert L.ws_knn_dist2_static(-1, None, None, None) == -1
    assert L.ws_knn_dist2_static(10, None, C.c_void_p(256), None) == -1
