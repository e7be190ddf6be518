"""Loader for the UNMODIFIED reference CUDA rasterizers built by oracle/build_ref.py
(oracle/_ref/libref_{dgr,fsr,knn}.so).  TEST / BASELINE INFRASTRUCTURE ONLY: used by the GPU
parity tests, by tests/golden/make_golden.py and by `bench.py --impl reference`.

The reference runs on the legacy default stream, which is also torch's default stream, so
ordinary torch ordering applies.
"""
import ctypes as C
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF = os.path.join(_HERE, "_ref")


def available(kind="dgr"):
    return os.path.exists(os.path.join(_REF, "libref_%s.so" % kind))


def _p(t):
    return None if t is None or t.numel() == 0 else C.c_void_p(t.data_ptr())


class RefRasterizer:
    """kind = 'dgr' (diff-gaussian-rasterization) or 'fsr' (flashsplat-rasterization)."""

    def __init__(self, kind="dgr"):
        self.kind = kind
        self.lib = C.CDLL(os.path.join(_REF, "libref_%s.so" % kind))
        self.lib.ref_create.restype = C.c_void_p
        self.lib.ref_forward.restype = C.c_int
        self.lib.ref_copy_out.restype = C.c_longlong
        self.h = C.c_void_p(self.lib.ref_create())
        self.last = None

    def forward(self, g, cam, bg, sh_degree, scale_modifier=1.0, zero_fill=True, gt_mask=None, unique_label=None,
                num_obj=None, outputs=None):
        """g: dict of CUDA float32 tensors (means3D, opacities, shs|colors_precomp,
        scales+rotations|cov3D_precomp); cam: dict with CUDA viewmatrix/projmatrix/campos tensors."""
        dev = g["means3D"].device
        P = g["means3D"].shape[0]
        H, W = cam["image_height"], cam["image_width"]
        shs = g.get("shs")
        M = 0 if shs is None else shs.shape[1]
        o = outputs if outputs is not None else {}
        f32 = dict(dtype=torch.float32, device=dev)
        o.setdefault("color", torch.empty((3, H, W), **f32))
        o.setdefault("depth", torch.empty((1, H, W), **f32))
        o.setdefault("alpha", torch.empty((1, H, W), **f32))
        o.setdefault("radii", torch.empty((P,), dtype=torch.int32, device=dev))
        args = [self.h, C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(bg), C.c_int(W), C.c_int(H),
                _p(g["means3D"]), _p(shs), _p(g.get("colors_precomp")), _p(g["opacities"]), _p(g.get("scales")),
                C.c_float(scale_modifier), _p(g.get("rotations")), _p(g.get("cov3D_precomp")),
                _p(cam["viewmatrix"]), _p(cam["projmatrix"]), _p(cam["campos"]), C.c_float(cam["tanfovx"]),
                C.c_float(cam["tanfovy"]), C.c_int(0), _p(o["color"]), _p(o["depth"]), _p(o["alpha"]),
                _p(o["radii"]), C.c_int(1 if zero_fill else 0)]
        if self.kind == "fsr":
            if gt_mask is None:
                gt_mask = torch.ones((H, W), **f32)
            o.setdefault("contrib_num", torch.empty((1, H, W), dtype=torch.int32, device=dev))
            o.setdefault("used_count", torch.empty((num_obj + 1, P), **f32))
            o["image_xys"] = torch.full((2, P), -1.0, **f32)
            o["gs_depth"] = torch.full((P,), -0.1, **f32)
            self._flash = (gt_mask.float().contiguous(), unique_label)
            args += [_p(self._flash[0]), _p(unique_label), C.c_int(num_obj), _p(o["contrib_num"]),
                     _p(o["used_count"]), _p(o["image_xys"]), _p(o["gs_depth"])]
        R = self.lib.ref_forward(*args)
        o["num_rendered"] = R
        self.last = dict(g=g, cam=cam, bg=bg, sh_degree=sh_degree, scale_modifier=scale_modifier, P=P, M=M, W=W,
                         H=H, R=R, out=o)
        return o

    def backward(self, dL_dcolor, dL_ddepth, dL_dalpha, zero_fill=True, grads=None):
        L = self.last
        g, cam, P, M = L["g"], L["cam"], L["P"], L["M"]
        dev = g["means3D"].device
        f32 = dict(dtype=torch.float32, device=dev)
        gr = grads if grads is not None else {}
        for name, shape in (("dL_dmeans2D", (P, 3)), ("dL_dconic", (P, 2, 2)), ("dL_dopacity", (P, 1)),
                            ("dL_dcolors", (P, 3)), ("dL_ddepths", (P, 1)), ("dL_dmeans3D", (P, 3)),
                            ("dL_dcov3D", (P, 6)), ("dL_dsh", (P, M, 3)), ("dL_dscales", (P, 3)),
                            ("dL_drotations", (P, 4))):
            gr.setdefault(name, torch.empty(shape, **f32))
        args = [self.h, C.c_int(P), C.c_int(L["sh_degree"]), C.c_int(M), _p(L["bg"]), C.c_int(L["W"]), C.c_int(L["H"]),
                _p(g["means3D"]), _p(g.get("shs")), _p(g.get("colors_precomp")), _p(L["out"]["alpha"]),
                _p(g.get("scales")), C.c_float(L["scale_modifier"]), _p(g.get("rotations")),
                _p(g.get("cov3D_precomp")), _p(cam["viewmatrix"]), _p(cam["projmatrix"]), _p(cam["campos"]),
                C.c_float(cam["tanfovx"]), C.c_float(cam["tanfovy"]), _p(L["out"]["radii"]), _p(dL_dcolor),
                _p(dL_ddepth), _p(dL_dalpha), _p(gr["dL_dmeans2D"]), _p(gr["dL_dconic"]), _p(gr["dL_dopacity"]),
                _p(gr["dL_dcolors"]), _p(gr["dL_ddepths"]), _p(gr["dL_dmeans3D"]), _p(gr["dL_dcov3D"]),
                _p(gr["dL_dsh"]) if M > 0 else None, _p(gr["dL_dscales"]), _p(gr["dL_drotations"]),
                C.c_int(1 if zero_fill else 0)]
        if self.kind == "fsr":
            args += [_p(self._flash[0]), _p(self._flash[1])]
        self.lib.ref_backward(*args)
        return gr

    _DT = {"depths": np.float32, "clamped": np.uint8, "means2D": np.float32, "cov3D": np.float32,
           "conic_opacity": np.float32, "rgb": np.float32, "tiles_touched": np.uint32, "point_offsets": np.uint32,
           "n_contrib": np.uint32, "ranges": np.uint32, "point_list": np.uint32, "point_list_unsorted": np.uint32,
           "point_list_keys": np.uint64, "point_list_keys_unsorted": np.uint64}

    def fetch(self, name):
        """Copies one of the reference's internal arrays to a numpy array."""
        torch.cuda.synchronize()
        L = self.last
        per = {"depths": L["P"], "clamped": 3 * L["P"], "means2D": 2 * L["P"], "cov3D": 6 * L["P"],
               "conic_opacity": 4 * L["P"], "rgb": 3 * L["P"], "tiles_touched": L["P"], "point_offsets": L["P"],
               "n_contrib": L["W"] * L["H"], "ranges": 2 * ((L["W"] + 15) // 16) * ((L["H"] + 15) // 16),
               "point_list": L["R"], "point_list_unsorted": L["R"], "point_list_keys": L["R"],
               "point_list_keys_unsorted": L["R"]}[name]
        out = np.zeros(per, self._DT[name])
        if per:
            n = self.lib.ref_copy_out(self.h, name.encode(), C.c_void_p(out.ctypes.data), C.c_longlong(out.nbytes))
            if n != out.nbytes:
                raise RuntimeError("ref_copy_out(%s) returned %d, expected %d" % (name, n, out.nbytes))
        return out

    def mark_visible(self, means3D, viewmatrix, projmatrix):
        out = torch.zeros((means3D.shape[0],), dtype=torch.uint8, device=means3D.device)
        self.lib.ref_mark_visible(C.c_int(means3D.shape[0]), _p(means3D), _p(viewmatrix), _p(projmatrix), _p(out))
        return out.view(torch.bool)


def knn_dist2(points):
    lib = C.CDLL(os.path.join(_REF, "libref_knn.so"))
    p = points.float().contiguous()
    out = torch.zeros((p.shape[0],), dtype=torch.float32, device=p.device)
    torch.cuda.synchronize()
    lib.ref_knn_dist2(C.c_int(p.shape[0]), _p(p), _p(out))
    torch.cuda.synchronize()
    return out


def to_device(scene, cam, device="cuda"):
    """numpy scene / camera dicts (wheat_3dgs_b200.scene) -> dicts of CUDA tensors."""
    g = {k: torch.from_numpy(np.ascontiguousarray(v)).to(device) for k, v in scene.items()
         if k in ("means3D", "scales", "rotations", "opacities", "shs", "colors_precomp", "cov3D_precomp")}
    c = dict(cam)
    for k in ("viewmatrix", "projmatrix", "campos"):
        c[k] = torch.from_numpy(np.ascontiguousarray(cam[k])).to(device)
    return g, c
