"""Seeded inputs of the golden cases (shared by make_golden.py and the tests).  numpy only."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from wheat_3dgs_b200 import scene  # noqa: E402

BG = np.array([0.1, 0.2, 0.3], np.float32)

# name: (P, W, H, camera index, active SH degree, kind)
DGR_CASES = {
    "dgr_sh3": (4000, 200, 150, 17, 3, "sh"),
    "dgr_deg1": (3000, 97, 61, 5, 1, "sh"),          # M = 16 coefficients stored, degree 1 active
    "dgr_precomp": (2500, 128, 96, 30, 0, "precomp"),  # colors_precomp + cov3D_precomp
}
FSR_CASES = {
    # name: (P, W, H, camera index, num_obj, use unique_label)
    "fsr_multi": (4000, 200, 150, 17, 12, False),
    "fsr_binary_ul": (4000, 200, 150, 17, 1, True),
}
KNN_CASES = {"knn_5000": 5000, "knn_7": 7}


def cov3d_from_scale_rot(scales, rots):
    """float32 numpy version of scene/gaussian_model.py:27-31 (build_covariance_from_scaling_rotation)."""
    r, x, y, z = [rots[:, i].astype(np.float64) for i in range(4)]
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                  2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                  2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], axis=1).reshape(-1, 3, 3)
    L = R * scales.astype(np.float64)[:, None, :]
    S = L @ L.transpose(0, 2, 1)
    return np.ascontiguousarray(np.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]],
                                         axis=1), dtype=np.float32)


def dgr_inputs(name):
    P, W, H, ci, deg, kind = DGR_CASES[name]
    sc = scene.synthetic_scene(P, seed=0, scale_median=0.004)
    cam = scene.overhead_cameras(W, H)[ci]
    lg = scene.loss_gradients((3, H, W), seed=2)
    if kind == "precomp":
        rng = np.random.default_rng(7)
        g = {"means3D": sc["means3D"], "opacities": sc["opacities"],
             "colors_precomp": rng.uniform(0, 1, size=(P, 3)).astype(np.float32),
             "cov3D_precomp": cov3d_from_scale_rot(sc["scales"], sc["rotations"]), "head_id": sc["head_id"]}
    else:
        g = sc
    return g, cam, lg, deg


def fsr_inputs(name):
    P, W, H, ci, num_obj, use_ul = FSR_CASES[name]
    sc = scene.synthetic_scene(P, seed=0, scale_median=0.004)
    cam = scene.overhead_cameras(W, H)[ci]
    mask = scene.head_masks(sc, cam, n_labels=num_obj)
    if num_obj == 1:
        mask = (mask > 0).astype(np.float32)
    ul = None
    if use_ul:
        ul = np.where(sc["head_id"] == 0, 1, np.where(sc["head_id"] % 3 == 0, -1, 0)).astype(np.int32)
    return sc, cam, mask, ul, num_obj


def knn_inputs(name):
    P = KNN_CASES[name]
    rng = np.random.default_rng(P)
    return (rng.normal(size=(P, 3)) * np.array([1.0, 0.5, 0.2]) + np.array([0.3, -0.2, 1.0])).astype(np.float32)


def votes_inputs():
    """Vote matrix for the multi_instance_opt golden: sparse non-negative votes with exact ties."""
    rng = np.random.default_rng(11)
    rows, P = 9, 6000
    a = rng.gamma(0.5, 2.0, size=(rows, P)).astype(np.float32)
    a[rng.uniform(size=a.shape) < 0.6] = 0.0
    a[:, :50] = 0.0                     # never seen
    a[3, 50:100] = a[[0, 1, 2, 4, 5, 6, 7, 8], 50:100].sum(axis=0)  # near-ties
    a[1:, 100:150] = 0.0                # only row 0 voted
    return a


# name: (shape, lambda_dssim).  Images are smooth blobs + noise in [0,1]; "ragged" is not a multiple of the
# loss kernels' 32x32 tile and smaller than the 11x11 window's halo on one side; "batch" is 4-D.
LOSS_CASES = {
    "rgb": ((3, 96, 128), 0.2),
    "ragged": ((3, 37, 75), 0.2),
    "tiny": ((1, 7, 5), 0.5),
    "batch": ((2, 3, 40, 48), 0.2),
}


def loss_inputs(name):
    shape, lam = LOSS_CASES[name]
    rng = np.random.default_rng(len(name) * 1000 + shape[-1])
    H, W = shape[-2], shape[-1]
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float64)
    planes = int(np.prod(shape[:-2]))
    gt = np.zeros((planes, H, W))
    for p in range(planes):
        for _ in range(6):
            cx, cy, r = rng.uniform(0, W), rng.uniform(0, H), rng.uniform(2, max(3, W / 4))
            gt[p] += rng.uniform(0.1, 0.6) * np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2 * r * r))
    gt = np.clip(gt + rng.normal(0, 0.02, gt.shape), 0, 1)
    img = np.clip(gt + rng.normal(0, 0.08, gt.shape) + 0.05, 0, 1)
    # some exactly equal pixels: torch.abs has gradient sign(0) = 0 there
    img[:, ::5, ::7] = gt[:, ::5, ::7]
    return img.reshape(shape).astype(np.float32), gt.reshape(shape).astype(np.float32), lam


def input_hash(arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


# ---- optimiser step + densification statistics
# group name: (shape per Gaussian, learning rate) -- scene/gaussian_model.py:172-179 with arguments/__init__.py defaults
OPTIM_P = 257
OPTIM_GROUPS = [
    ("xyz", (3,), 0.00016 * 1.3),       # position_lr_init * spatial_lr_scale (rescheduled every step)
    ("f_dc", (1, 3), 0.0025),
    ("f_rest", (15, 3), 0.0025 / 20.0),
    ("opacity", (1,), 0.05),
    ("scaling", (3,), 0.005),
    ("rotation", (4,), 0.001),
]
OPTIM_STEPS = 3
OPTIM_XYZ_LRS = [0.00016 * 1.3, 0.00016 * 1.3 * 0.97, 0.00016 * 1.3 * 0.94]  # update_learning_rate() per iteration


def optim_inputs(P=OPTIM_P, seed=11):
    """Parameters, per-step gradients (rows of invisible Gaussians are zero, as the rasterizer backward leaves
    them), radii and screen-space gradients for the densification statistics."""
    rng = np.random.default_rng(seed)
    params = {n: rng.normal(0, 1, (P,) + s).astype(np.float32) for n, s, _ in OPTIM_GROUPS}
    grads, radii, g2d = [], [], []
    for _ in range(OPTIM_STEPS):
        vis = rng.random(P) < 0.7
        grads.append({n: (rng.normal(0, 1, (P,) + s) * 10.0 ** rng.uniform(-6, 0, (P,) + (1,) * len(s))
                          * vis.reshape((P,) + (1,) * len(s))).astype(np.float32) for n, s, _ in OPTIM_GROUPS})
        radii.append((rng.integers(1, 60, P) * vis).astype(np.int32))
        g2d.append((rng.normal(0, 1e-4, (P, 3)) * vis[:, None]).astype(np.float32))
    return params, grads, radii, g2d


# ---- batched eval_obj_labels
OBJEVAL = dict(P=4000, W=200, H=150, cam=17, n_heads=8, extent=(0.3, 0.2, 0.12))


def objeval_inputs():
    """A small dense scene (8 heads in a 30 x 20 cm plot seen from above) and an all_obj_labels matrix that
    exercises every branch of eval_obj_labels: plain heads, an object IN FRONT of an earlier one (the upper half of
    head 0: it repaints), an object BEHIND earlier ones (head 1, the lower half of head 0 and the so far
    unpainted head 7: it only takes free pixels), a Gaussian in several objects, an empty row, and row 0 (background, never rendered)."""
    c = OBJEVAL
    sc = scene.synthetic_scene(c["P"], seed=3, n_heads=c["n_heads"], scale_median=0.004, extent=c["extent"])
    cam = scene.overhead_cameras(c["W"], c["H"], extent=c["extent"])[c["cam"]]
    hid, z = sc["head_id"], sc["means3D"][:, 2]
    zc = sc["head_centers"][:, 2]
    rows = [hid < 0]                                        # 0: background
    rows += [hid == k for k in range(c["n_heads"] - 1)]     # 1..7: heads 0..6
    rows.append((hid == 0) & (z > zc[0]))                   # 8: upper half of head 0 (in front of object 1)
    rows.append((hid == 1) | ((hid == 0) & (z <= zc[0])) | (hid == 7))  # 9: behind objects 2 and 8 where it overlaps
    #                                                         them; head 7 is free, so it is painted there only
    rows.append(np.zeros_like(hid, dtype=bool))             # 10: empty
    rows.append((hid == 2) | (hid == 3))                    # 11: two heads at once
    return sc, cam, np.stack(rows).astype(bool)


# ---- densification: scene/gaussian_model.py:440-458 on a seeded model
DENSIFY_P = 700
DENSIFY_SEED = 5
DENSIFY_TRAIN_ARGS = dict(percent_dense=0.01, position_lr_init=0.00016, position_lr_final=0.0000016,
                          position_lr_delay_mult=0.01, position_lr_max_steps=30000, feature_lr=0.0025,
                          opacity_lr=0.05, scaling_lr=0.005, rotation_lr=0.001)   # arguments/__init__.py:74-82
DENSIFY_CASES = {   # train_vanilla_3dgs.py:106-107: size_threshold = 20 after the first opacity reset, else None
    "screen": dict(max_grad=0.0002, min_opacity=0.005, extent=2.0, max_screen_size=20),
    "noscreen": dict(max_grad=0.0002, min_opacity=0.005, extent=2.0, max_screen_size=None),
}


def densify_inputs(P=DENSIFY_P, seed=DENSIFY_SEED, K=3):
    """Raw parameters, one set of gradients (for the Adam step that fills the moments) and the statistics.
    Scales straddle percent_dense*extent = 0.02 with a tail beyond 0.1*extent, opacities straddle 0.005,
    grads straddle 0.0002, some denominators are 0 (0/0 -> NaN -> 0 as :441-442)."""
    r = np.random.RandomState(seed)
    f = np.float32
    inp = {
        "xyz": r.uniform(-1, 1, (P, 3)).astype(f),
        "f_dc": r.normal(0, 0.5, (P, 1, 3)).astype(f),
        "f_rest": r.normal(0, 0.05, (P, K, 3)).astype(f),
        "opacity": r.normal(-3.0, 2.5, (P, 1)).astype(f),
        "scaling": (np.log(0.02) + r.normal(0, 1.2, (P, 3))).astype(f),
        "rotation": r.normal(0, 1, (P, 4)).astype(f),
        "which_object": r.randint(0, 9, (P, 1)).astype(np.int32),
    }
    for n in ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"):
        inp["g_" + n] = r.normal(0, 1e-3, inp[n].shape).astype(f)
    denom = r.randint(0, 6, (P, 1)).astype(f)
    inp["denom"] = denom
    inp["xyz_gradient_accum"] = (denom * r.uniform(0, 0.0006, (P, 1))).astype(f)
    inp["max_radii2D"] = r.uniform(0, 40, (P,)).astype(f)
    return inp
