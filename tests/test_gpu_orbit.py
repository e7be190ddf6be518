"""orbit.render_360 / StreamRing: frames rendered on several CUDA streams in turn are the frames rendered on one stream,
bit for bit (config-4 shape at small size), and VoteAccumulator's streamed views sum to the single-stream votes."""
import numpy as np
import pytest
import torch

from wheat_3dgs_b200 import orbit, scene

pytestmark = pytest.mark.gpu


def _scene(P=30000, W=320, H=240, n_frames=7):
    sc = scene.synthetic_scene(P, seed=5)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    g = {k: t(sc[k]) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    og = scene.overhead_cameras(2 * W, 2 * H)[0]
    cams = orbit.orbit_views(og, 1.0, n_frames)
    for c in cams:
        for k in ("viewmatrix", "projmatrix", "campos"):
            c[k] = t(c[k])
    return g, cams


def test_frames_on_three_streams_equal_frames_on_one():
    g, cams = _scene()
    bg = torch.ones(3, device="cuda")
    one = orbit.render_360(g, cams, bg, 3, streams=1)
    three = orbit.render_360(g, cams, bg, 3, streams=3)
    torch.cuda.synchronize()
    assert [f[0] for f in one] == [f[0] for f in three] == list(range(len(cams)))
    for a, b in zip(one, three):
        for x, y in zip(a[1:4], b[1:4]):
            assert torch.equal(x, y)


def test_frames_sharded_over_ranks_cover_the_orbit():
    g, cams = _scene(n_frames=5)
    bg = torch.ones(3, device="cuda")
    full = {f[0]: f[1] for f in orbit.render_360(g, cams, bg, 3, streams=2)}
    seen = {}
    for r in range(2):
        for f in orbit.render_360(g, cams, bg, 3, rank=r, world=2, streams=2):
            seen[f[0]] = f[1]
    torch.cuda.synchronize()
    assert sorted(seen) == sorted(full)
    for k in full:
        assert torch.equal(seen[k], full[k])


def test_on_frame_callback_runs_under_the_frames_stream():
    g, cams = _scene(n_frames=4)
    bg = torch.ones(3, device="cuda")
    sums = []
    orbit.render_360(g, cams, bg, 3, streams=3, on_frame=lambda i, c, d, a, p: sums.append((i, c.sum())))
    ref = orbit.render_360(g, cams, bg, 3, streams=1)
    torch.cuda.synchronize()
    for (i, s_), f in zip(sums, ref):
        assert i == f[0] and float(s_) == float(f[1].sum())
