"""Host mirror of the orbit renderer around the rasterizer (BASELINE config 4):

orbit_views(og_view, scene_radius, n_frames, elevation, centroid)
        the camera path of utils/wheatgs_helper.py:273-289 (fixed elevation, distance 2 x scene radius, the
        first training view's intrinsics at HALF resolution, `get_camera_path_fixed_elevation` :224-241 with
        `get_c2w_from_up_and_look_at` :208-222: positions orbit the WORLD origin, look at the centroid, up = +z)
render_360(...)
        the frame loop of :282-300: render(view) per frame, optionally the per-object segmentation of
        eval_obj_labels (one batched pass, segment.eval_obj_labels_tensors) -- frames r::N of an N-GPU job,
        gathered on rank 0 (dist.gather_frames).  PNG / ffmpeg encoding (:300-313) is the caller's.

Frames stay on the device as float32 [3,H,W] (+ depth, alpha); nothing here has a CPU path.
"""
import math

import numpy as np
import torch

from . import ops
from . import scene as _scene
from . import segment as _segment
from .rasterizer import GaussianRasterizationSettings


def _c2w(up, look_at, pos):
    """Camera-to-world [3,4] with +z towards `look_at`, y = down (utils/wheatgs_helper.py:208-222)."""
    z = look_at - pos
    z = z / np.linalg.norm(z)
    x = np.cross(-up / np.linalg.norm(up), z)
    x = x / np.linalg.norm(x)
    y = np.cross(z, x)
    m = np.zeros((3, 4), dtype=np.float32)
    m[:, 0], m[:, 1], m[:, 2], m[:, 3] = x, y, z, pos
    return m


def camera_path_fixed_elevation(n_frames, camera_distance, cam_center, elevation_deg, n_circles=1):
    """[n_frames,3,4] float32 c2w matrices (utils/wheatgs_helper.py:224-241).  The azimuth runs over
    linspace(0, 2 pi) INCLUSIVE (first and last frame coincide) and the positions are NOT offset by the centre."""
    az = np.linspace(0, 2 * np.pi * n_circles, n_frames)
    el = np.deg2rad(elevation_deg)
    pos = np.stack([camera_distance * np.cos(az) * np.cos(el), camera_distance * np.sin(az) * np.cos(el),
                    camera_distance * np.sin(el) * np.ones_like(az)], axis=1)
    up = np.array([0, 0, 1], dtype=np.float32)
    look = np.array(cam_center, dtype=np.float32)
    return np.stack([_c2w(up, look, pos[i]) for i in range(n_frames)], axis=0)


def orbit_views(og_view, scene_radius, n_frames, elevation=45.0, centroid=(0.0, 0.0, 0.0), znear=0.01, zfar=100.0):
    """Camera dicts (scene.make_camera conventions) of the orbit: `og_view` supplies image size and FoV
    (a dict with image_width, image_height, FoVx, FoVy); frames are half resolution (wheatgs_helper.py:277)."""
    width, height = math.floor(og_view["image_width"] / 2), math.floor(og_view["image_height"] / 2)
    c2ws = camera_path_fixed_elevation(n_frames, scene_radius * 2, centroid, elevation)
    cams = []
    for c2w in c2ws:
        full = np.vstack([c2w, [0.0, 0.0, 0.0, 1.0]])
        w2c = np.linalg.inv(np.float64(full))                       # :285
        cams.append(_scene.make_camera(w2c, og_view["FoVx"], og_view["FoVy"], width, height, znear, zfar))
    return cams


def settings_of(cam, bg, sh_degree, scale_modifier=1.0):
    """GaussianRasterizationSettings of a camera dict whose matrices already live on the device."""
    return GaussianRasterizationSettings(
        image_height=int(cam["image_height"]), image_width=int(cam["image_width"]), tanfovx=cam["tanfovx"],
        tanfovy=cam["tanfovy"], bg=bg, scale_modifier=scale_modifier, viewmatrix=cam["viewmatrix"],
        projmatrix=cam["projmatrix"], sh_degree=sh_degree, campos=cam["campos"], prefiltered=False, debug=False)


_E = torch.Tensor([])


class StreamRing:
    """Independent frames on a few CUDA streams in turn.  A forward-only frame is a chain of short, latency-bound
    kernels (prefix sum, radix passes, tile ranges) in front of one long blend kernel; frames are independent, so the
    chain of frame i+1 runs under the blend of frame i when they sit on different streams (config 4: 0.78 -> 0.69
    ms/frame with three streams).  One host thread drives all of them (ws_forward_prepare blocks until a frame's
    instance count is known).  Use: `with ring.stream(i): render_frame(...)`; `ring.fork()` before the first frame
    (the ring waits for the current stream), `ring.join()` after the last (the current stream waits for the ring)."""

    def __init__(self, n=3, device=None):
        self.n = max(1, int(n))
        self.streams = [torch.cuda.Stream(device=device) for _ in range(self.n)] if self.n > 1 else []

    def stream(self, i):
        """Context of frame i's stream.  The stream first waits for what the current stream has queued so far (the
        frame's inputs -- camera, mask -- are usually produced there)."""
        if not self.streams:
            return torch.cuda.stream(torch.cuda.current_stream())
        st = self.streams[i % self.n]
        st.wait_stream(torch.cuda.current_stream())
        return torch.cuda.stream(st)

    def stream_of(self, i):
        """Frame i's stream object, or None when the ring is the current stream only."""
        return self.streams[i % self.n] if self.streams else None

    def fork(self):
        cur = torch.cuda.current_stream()
        for st in self.streams:
            st.wait_stream(cur)

    def join(self):
        cur = torch.cuda.current_stream()
        for st in self.streams:
            cur.wait_stream(st)


@torch.no_grad()
def render_frame(g, cam, bg, sh_degree):
    """gaussian_renderer.render() on device tensors, forward only: (color [3,H,W], depth, alpha, radii, R)."""
    R, color, depth, alpha, radii, _, _, _ = ops.rasterize_gaussians(
        bg, g["means3D"], _E, g["opacities"], g["scales"], g["rotations"], 1.0, _E, cam["viewmatrix"],
        cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], cam["image_height"], cam["image_width"], g["shs"],
        sh_degree, cam["campos"], False, False)
    return color, depth, alpha, radii, R


@torch.no_grad()
def render_360(g, cams, bg, sh_degree, rank=0, world=1, object_sets=None, on_frame=None, streams=3):
    """Frames rank::world of the orbit, on `streams` CUDA streams in turn (StreamRing; 1 = the current stream only).
    Returns [(frame index, color, depth, alpha, pred_mask or None)] unless `on_frame(idx, color, depth, alpha, pred)`
    consumes them (nothing is kept then; it is called under the frame's stream).  The current stream has waited for
    every frame when this returns."""
    out = []
    ring = StreamRing(streams, g["means3D"].device)
    ring.fork()
    for k, idx in enumerate(range(rank, len(cams), world)):
        cam = cams[idx]
        with ring.stream(k):
            color, depth, alpha, _, _ = render_frame(g, cam, bg, sh_degree)
            pred = None
            if object_sets is not None:
                pred = _segment.eval_obj_labels_tensors(object_sets, settings_of(cam, bg, sh_degree), g["means3D"],
                                                        g["opacities"], shs=g["shs"], scales=g["scales"],
                                                        rotations=g["rotations"])
            if on_frame is not None:
                on_frame(idx, color, depth, alpha, pred)
            else:
                out.append((idx, color, depth, alpha, pred))
    ring.join()
    if ring.streams:   # the frames were allocated on the ring's streams and will be read on the current one
        cur = torch.cuda.current_stream()
        for fr in out:
            for t_ in fr[1:]:
                if t_ is not None:
                    t_.record_stream(cur)
    return out
