"""Generates tests/golden/orbit.npz in THIS container (CPU) by executing the reference's own camera-path and
matching helpers, cut out of their modules with `ast` (the modules import packages that are not installed here)
and run unmodified:

  get_c2w_from_up_and_look_at, get_camera_path_fixed_elevation      utils/wheatgs_helper.py:208-241
  get_bbox_from_mask, is_overlapping, calculate_seg_iou              utils/image_helper.py:104-163

    python tests/golden/make_golden_orbit.py
"""
import ast
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def cut(path, names):
    tree = ast.parse(open(path).read())
    fns = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    ns = {"np": np}
    exec(compile(ast.Module(body=fns, type_ignores=[]), path, "exec"), ns)
    return [ns[n] for n in names]


def iou_cases(seed=5, H=60, W=80, n=24):
    """Prediction / candidate mask pairs: overlapping blobs, disjoint ones, empty masks, exact ties."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W]
    ids = np.zeros((H, W), np.int64)
    for k in range(1, n + 1):
        cx, cy, r = rng.uniform(0, W), rng.uniform(0, H), rng.uniform(3, 9)
        ids[((xx - cx) ** 2 + (yy - cy) ** 2 < r * r) & (ids == 0)] = k
    preds = []
    for _ in range(12):
        cx, cy, r = rng.uniform(0, W), rng.uniform(0, H), rng.uniform(2, 12)
        preds.append(((xx - cx) ** 2 / 1.5 + (yy - cy) ** 2 < r * r))
    preds.append(np.zeros((H, W), bool))          # nothing rendered
    preds.append(ids == 3)                        # a perfect match
    preds.append((ids == 5) | (ids == 6))         # two candidates
    return ids, np.stack(preds)


if __name__ == "__main__":
    c2w, path = cut("/root/reference/utils/wheatgs_helper.py",
                    ["get_c2w_from_up_and_look_at", "get_camera_path_fixed_elevation"])
    path.__globals__["get_c2w_from_up_and_look_at"] = c2w
    out = {}
    for i, (n, dist, center, el) in enumerate([(200, 3.0, [0.1, -0.2, 0.3], 45), (7, 1.5, [0, 0, 0], 0),
                                               (36, 2.2, [0.5, 0.5, 0.1], 70)]):
        out["path%d_args" % i] = np.array([n, dist, center[0], center[1], center[2], el], np.float64)
        out["path%d" % i] = path(n_frames=n, camera_distance=dist, cam_center=center, elevation=el)
    bbox, overlap, iou = cut("/root/reference/utils/image_helper.py",
                             ["get_bbox_from_mask", "is_overlapping", "calculate_seg_iou"])
    ids, preds = iou_cases()
    K = int(ids.max())
    best = []
    for pred in preds:                            # run_3d_seg.py:136-157 with box = bbox of each mask
        pb = bbox(pred.astype(np.uint8))
        max_iou, arg = 0.0, 0
        for k in range(1, K + 1):
            m = ids == k
            if not overlap(pb, bbox(m.astype(np.uint8))):
                continue
            v = iou(m, pred)
            if v > max_iou:
                max_iou, arg = v, k
        best.append((max_iou, arg))
    out["iou_ids"], out["iou_preds"] = ids, preds
    out["iou_best"] = np.array(best, np.float64)
    np.savez_compressed(os.path.join(HERE, "orbit.npz"), **out)
    print({k: np.asarray(v).shape for k, v in out.items()})
