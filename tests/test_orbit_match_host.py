"""Host logic either side of the rasterizer for configs 3 and 4, against goldens made by the reference's own
helper functions (tests/golden/make_golden_orbit.py)."""
import os

import numpy as np
import torch

from wheat_3dgs_b200 import orbit, segment

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "orbit.npz"))


def test_camera_path_equals_reference():
    for i in range(3):
        n, dist, cx, cy, cz, el = G["path%d_args" % i]
        mine = orbit.camera_path_fixed_elevation(int(n), dist, [cx, cy, cz], el)
        assert mine.dtype == G["path%d" % i].dtype
        assert np.array_equal(mine, G["path%d" % i])


def test_orbit_views_follow_the_reference_conventions():
    og = {"image_width": 3840, "image_height": 2161, "FoVx": 0.9, "FoVy": 0.6}
    cams = orbit.orbit_views(og, 0.75, 5, elevation=45.0, centroid=(0.0, 0.1, 0.3))
    assert len(cams) == 5 and cams[0]["image_width"] == 1920 and cams[0]["image_height"] == 1080  # floor(/2)
    c = cams[0]
    # camera centre = the orbit position at distance 2 x radius around the WORLD origin (not the centroid)
    assert np.allclose(np.linalg.norm(c["campos"]), 1.5, atol=1e-5)
    # the centroid projects to the image centre
    p = np.array([0.0, 0.1, 0.3, 1.0], np.float32) @ c["projmatrix"]
    assert abs(p[0] / p[3]) < 1e-4 and abs(p[1] / p[3]) < 1e-4
    # first and last frame coincide (linspace includes 2 pi)
    assert np.allclose(cams[0]["viewmatrix"], cams[-1]["viewmatrix"], atol=1e-5)


def test_best_mask_iou_equals_reference_search():
    ids = torch.from_numpy(G["iou_ids"])
    area = torch.bincount(ids.reshape(-1))
    for pred, (iou, k) in zip(G["iou_preds"], G["iou_best"]):
        got = segment.best_mask_iou(torch.from_numpy(pred), ids, area)
        assert float(got[0]) == iou and int(got[1]) == int(k), (got, iou, k)
