"""Helpers of the GPU parity tests: every call goes through the C ABI (wheat_3dgs_b200.ops -> ctypes ->
libwheatsplat.so)."""
import numpy as np
import torch

from wheat_3dgs_b200 import ops

DEV = "cuda"
E = torch.Tensor([])


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def n(x):
    return x.detach().cpu().numpy()


def dev_inputs(g, cam):
    gd = {k: t(v) for k, v in g.items() if k in ("means3D", "scales", "rotations", "opacities", "shs", "colors_precomp",
                                                   "cov3D_precomp")}
    cd = dict(cam)
    for k in ("viewmatrix", "projmatrix", "campos"):
        cd[k] = t(cam[k])
    return gd, cd


def forward(gd, cd, bg, deg, debug=False, scale_modifier=1.0):
    return ops.rasterize_gaussians(bg, gd["means3D"], gd.get("colors_precomp", E), gd["opacities"], gd.get("scales", E),
                                   gd.get("rotations", E), scale_modifier, gd.get("cov3D_precomp", E), cd["viewmatrix"],
                                   cd["projmatrix"], cd["tanfovx"], cd["tanfovy"], cd["image_height"],
                                   cd["image_width"], gd.get("shs", E), deg, cd["campos"], False, debug)


def backward(gd, cd, bg, deg, fw, dc, dd, da, debug=False, scale_modifier=1.0):
    R, color, depth, alpha, radii, geom, binning, image = fw
    return ops.rasterize_gaussians_backward(bg, gd["means3D"], radii, gd.get("colors_precomp", E), gd.get("scales", E),
                                            gd.get("rotations", E), scale_modifier, gd.get("cov3D_precomp", E),
                                            cd["viewmatrix"], cd["projmatrix"], cd["tanfovx"], cd["tanfovy"], dc, dd, da,
                                            gd.get("shs", E), deg, cd["campos"], geom, R, binning, image, alpha, debug)


def fetch(name, fw, P, W, H):
    R, color, depth, alpha, radii, geom, binning, image = fw
    return n(ops.debug_fetch(name, P, W, H, R, geom, binning, image))


def flash_forward(gd, cd, bg, deg, mask, ul, num_obj, used_out=None):
    return ops.flash_rasterize_gaussians(mask, ul if ul is not None else torch.tensor([], dtype=torch.int), bg,
                                         gd["means3D"], E, gd["opacities"], gd["scales"], gd["rotations"], 1.0, E,
                                         cd["viewmatrix"], cd["projmatrix"], cd["tanfovx"], cd["tanfovy"],
                                         cd["image_height"], cd["image_width"], gd["shs"], deg, cd["campos"], False,
                                         num_obj, False, used_count_out=used_out)
