"""Seeded synthetic scenes and cameras shaped like the Wheat3DGS workloads (numpy only).

There is no dataset or checkpoint offline, so the bench/tests use synthetic inputs of the
shapes BASELINE.json names: a plot_461-like slab of Gaussians with
ellipsoidal "wheat head" clusters, 36 overhead cameras on a 12x3 grid, and a fixed-elevation
orbit.  Camera matrices follow the reference conventions so the drop-in ops see exactly what
`gaussian_renderer.render()` would hand them:

  * `viewmatrix`  = world-to-view, stored TRANSPOSED (row-vector convention),
    reference scene/cameras.py:56  (getWorld2View2(...).transpose(0, 1))
  * `projmatrix`  = viewmatrix @ projection^T  (full projection, also transposed), :57-58
  * `campos`      = inverse(viewmatrix)[3, :3], :59
  * projection    = reference utils/graphics_utils.py:51-71 (znear 0.01, zfar 100)

Everything is float32 numpy generated with np.random.default_rng(seed) so the same arrays
come out on the build container and on the GPU box.
"""
import math

import numpy as np

C0 = 0.28209479177387814  # reference utils/sh_utils.py:25


def projection_matrix(znear, zfar, fovx, fovy):
    thy, thx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = thy * znear, thx * znear
    P = np.zeros((4, 4), dtype=np.float32)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def look_at_w2c(eye, target, up=(0.0, 1.0, 0.0)):
    """World-to-camera 4x4 (float64) for a COLMAP-style camera: +z forward, +x right, +y down."""
    eye = np.asarray(eye, dtype=np.float64)
    f = np.asarray(target, dtype=np.float64) - eye
    f /= np.linalg.norm(f)
    upv = np.asarray(up, dtype=np.float64)
    r = np.cross(f, upv)
    if np.linalg.norm(r) < 1e-8:
        r = np.cross(f, np.array([1.0, 0.0, 0.0]))
    r /= np.linalg.norm(r)
    d = np.cross(f, r)  # "down"
    R = np.stack([r, d, f], axis=0)  # rows = camera axes in world coords
    w2c = np.eye(4)
    w2c[:3, :3] = R
    w2c[:3, 3] = -R @ eye
    return w2c


def make_camera(w2c, fovx, fovy, width, height, znear=0.01, zfar=100.0):
    """Camera dict in the reference's tensor conventions (all float32 numpy)."""
    view_t = np.ascontiguousarray(np.asarray(w2c, dtype=np.float64).astype(np.float32).T)
    proj_t = projection_matrix(znear, zfar, fovx, fovy).T
    full = (view_t @ proj_t).astype(np.float32)
    campos = np.linalg.inv(view_t.astype(np.float64))[3, :3].astype(np.float32)
    return {
        "image_width": int(width), "image_height": int(height),
        "tanfovx": math.tan(fovx * 0.5), "tanfovy": math.tan(fovy * 0.5),
        "FoVx": fovx, "FoVy": fovy,
        "viewmatrix": view_t, "projmatrix": np.ascontiguousarray(full),
        "campos": np.ascontiguousarray(campos),
    }


def synthetic_scene(P, seed=0, sh_degree=3, n_heads=500, scale_median=0.002, scale_sigma=0.6,
                    extent=(1.5, 1.0, 0.6)):
    """plot_461-shaped scene: 30% of Gaussians uniform in the slab (soil/leaves), 70% in
    `n_heads` ellipsoidal clusters ("wheat heads", sigma ~1.5 cm) in the upper half.

    Returns post-activation tensors exactly as GaussianModel's getters would produce them:
    opacities after sigmoid, scales after exp, rotations normalised (w,x,y,z),
    shs [P,(deg+1)^2,3] (coefficient-major, RGB innermost)."""
    rng = np.random.default_rng(seed)
    ex, ey, ez = extent
    n_bg = int(round(P * 0.3))
    n_hd = P - n_bg
    bg = rng.uniform([-ex / 2, -ey / 2, 0.0], [ex / 2, ey / 2, ez], size=(n_bg, 3))
    centers = rng.uniform([-ex / 2 * 0.95, -ey / 2 * 0.95, ez * 0.5], [ex / 2 * 0.95, ey / 2 * 0.95, ez],
                          size=(n_heads, 3))
    head_sigma = rng.uniform(0.008, 0.015, size=(n_heads, 3)) * np.array([1.0, 1.0, 2.0])
    hid = rng.integers(0, n_heads, size=n_hd)
    hd = centers[hid] + rng.normal(size=(n_hd, 3)) * head_sigma[hid]
    means = np.concatenate([bg, hd], axis=0)
    head_id = np.concatenate([np.full(n_bg, -1, dtype=np.int32), hid.astype(np.int32)])
    perm = rng.permutation(P)  # mix so memory order carries no spatial meaning
    means, head_id = means[perm], head_id[perm]
    scales = np.exp(rng.normal(math.log(scale_median), scale_sigma, size=(P, 3)))
    q = rng.normal(size=(P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = 1.0 / (1.0 + np.exp(-rng.normal(0.0, 2.0, size=(P, 1))))
    M = (sh_degree + 1) ** 2
    shs = rng.normal(0.0, 0.05, size=(P, M, 3))
    shs[:, 0, :] = (rng.uniform(0.0, 1.0, size=(P, 3)) - 0.5) / C0  # RGB2SH
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    return {
        "means3D": f32(means), "scales": f32(scales), "rotations": f32(q), "opacities": f32(opac),
        "shs": f32(shs), "head_id": head_id, "sh_degree": sh_degree,
        "head_centers": f32(centers), "extent": extent,
    }


def overhead_cameras(width, height, n_x=12, n_y=3, seed=1, extent=(1.5, 1.0, 0.6)):
    """12x3 overhead grid 1.0-1.5 m above the canopy, looking down with up to ~15 deg tilt
    (reference README.md:50: 12 cameras x 3 batches).  FoV such that the plot fills the frame
    from the mean height."""
    rng = np.random.default_rng(seed)
    ex, ey, ez = extent
    mean_dist = ez + 1.25 - 0.3
    fovx = 2 * math.atan((ex / 2) / mean_dist)
    fovy = 2 * math.atan(math.tan(fovx / 2) * height / width)
    cams = []
    for j in range(n_y):
        for i in range(n_x):
            cx = (i / (n_x - 1) - 0.5) * ex * 0.8
            cy = (j / (n_y - 1) - 0.5) * ey * 0.6
            h = ez + rng.uniform(1.0, 1.5)
            eye = np.array([cx, cy, h])
            target = np.array([cx * 0.5 + rng.normal(0, 0.02), cy * 0.5 + rng.normal(0, 0.02), 0.3])
            cams.append(make_camera(look_at_w2c(eye, target, up=(0.0, 1.0, 0.0)), fovx, fovy, width, height))
    return cams


def orbit_cameras(width, height, n_frames=200, elevation_deg=45.0, center=(0.0, 0.0, 0.3),
                  radius=None, extent=(1.5, 1.0, 0.6), fovx=None):
    """Fixed-elevation orbit (reference utils/wheatgs_helper.py:273-290: camera_distance =
    2 x scene radius, elevation 45, frames at half resolution are the caller's choice)."""
    if radius is None:
        radius = 2.0 * 0.5 * math.sqrt(extent[0] ** 2 + extent[1] ** 2)
    if fovx is None:
        fovx = 2 * math.atan((extent[0] / 2) / (extent[2] + 1.25 - 0.3))
    fovy = 2 * math.atan(math.tan(fovx / 2) * height / width)
    el = math.radians(elevation_deg)
    c = np.asarray(center, dtype=np.float64)
    cams = []
    for k in range(n_frames):
        az = 2 * math.pi * k / n_frames
        eye = c + radius * np.array([math.cos(el) * math.cos(az), math.cos(el) * math.sin(az), math.sin(el)])
        cams.append(make_camera(look_at_w2c(eye, c, up=(0.0, 0.0, -1.0)), fovx, fovy, width, height))
    return cams


def loss_gradients(color_shape, seed=2):
    """Seeded upstream gradients for a fwd+bwd step: dL/dcolor ~ N(0,1)/numel-scaled,
    dL/ddepth, dL/dalpha ~ small N(0,1e-3) so the fork's depth/alpha paths are exercised."""
    rng = np.random.default_rng(seed)
    _, H, W = color_shape
    return {
        "dL_dcolor": rng.normal(0, 1.0, size=(3, H, W)).astype(np.float32),
        "dL_ddepth": rng.normal(0, 1e-3, size=(1, H, W)).astype(np.float32),
        "dL_dalpha": rng.normal(0, 1e-3, size=(1, H, W)).astype(np.float32),
    }


def head_masks(scene, cam, n_labels=None, dilate=6, solid=False):
    """Synthetic per-view instance-id mask f32[H,W] (0 = background, k = head k) made by
    splatting each head's projected Gaussian centres and dilating -- stands in for the
    YOLO/SAM masks of run_3d_seg.py / the seg2d ids of render_360.py:62-69.  The dilation
    stamps every second pixel; `solid` closes the lattice with a 3x3 fill, which is what a SAM
    mask looks like to find_match's IoU (run_3d_seg.py:134-157)."""
    H, W = cam["image_height"], cam["image_width"]
    hid = scene["head_id"]
    m = scene["means3D"].astype(np.float64)
    hom = np.concatenate([m, np.ones((m.shape[0], 1))], axis=1) @ cam["projmatrix"].astype(np.float64)
    w = hom[:, 3] + 1e-7
    px = ((hom[:, 0] / w + 1.0) * W - 1.0) * 0.5
    py = ((hom[:, 1] / w + 1.0) * H - 1.0) * 0.5
    ok = (hid >= 0) & (w > 0.2) & (px >= 0) & (px < W) & (py >= 0) & (py < H)
    if n_labels is not None:
        ok &= hid < n_labels
    mask = np.zeros((H, W), dtype=np.float32)
    xi, yi, lab = px[ok].astype(np.int64), py[ok].astype(np.int64), hid[ok] + 1
    for dy in range(-dilate, dilate + 1, 2):
        for dx in range(-dilate, dilate + 1, 2):
            x = np.clip(xi + dx, 0, W - 1)
            y = np.clip(yi + dy, 0, H - 1)
            mask[y, x] = lab
    if solid:
        filled = mask.copy()
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                if dx == 0 and dy == 0:
                    continue
                sh = np.zeros_like(mask)
                ys, yd = (slice(max(dy, 0), H + min(dy, 0)), slice(max(-dy, 0), H + min(-dy, 0)))
                xs, xd = (slice(max(dx, 0), W + min(dx, 0)), slice(max(-dx, 0), W + min(-dx, 0)))
                sh[yd, xd] = mask[ys, xs]
                filled = np.where(filled == 0, sh, filled)
        mask = filled
    return mask
