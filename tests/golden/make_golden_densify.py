"""Generates tests/golden/densify.npz in THIS container (CPU) from the UNMODIFIED reference class:
/root/reference/scene/gaussian_model.py is imported as it is (plyfile and simple_knn, absent here and not on this
path, are stubbed in sys.modules), a GaussianModel is filled with the seeded state of cases.densify_inputs(),
`training_setup` builds the reference's own torch.optim.Adam, one optimizer step makes the moments non-zero, and
`densify_and_prune` / `reset_opacity` run unmodified.  The reference hard-codes device="cuda" in its tensor
constructors (:385-389, :395, :399; utils/general_utils.py:83): this script wraps torch.zeros so that the
argument is dropped (no GPU in this container), and wraps torch.normal to record the samples the split drew.

    python tests/golden/make_golden_densify.py
"""
import os
import sys
import types

import numpy as np
import torch
from torch import nn

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import cases  # noqa: E402

REF = "/root/reference"


def import_reference():
    ply = types.ModuleType("plyfile")
    ply.PlyData = ply.PlyElement = object
    knn = types.ModuleType("simple_knn")
    knn_c = types.ModuleType("simple_knn._C")
    knn_c.distCUDA2 = None
    knn._C = knn_c
    sys.modules.update({"plyfile": ply, "simple_knn": knn, "simple_knn._C": knn_c})
    sys.path.insert(0, REF)
    # the module file itself (scene/__init__.py pulls in the dataset readers and ffmpeg, not on this path)
    import importlib.util
    spec = importlib.util.spec_from_file_location("reference_gaussian_model", os.path.join(REF, "scene", "gaussian_model.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.GaussianModel


def on_cpu(fn):
    def wrapped(*a, **k):
        k.pop("device", None)
        return fn(*a, **k)
    return wrapped


RECORDED = {}


def recording_normal(fn):
    def wrapped(*a, **k):
        out = fn(*a, **k)
        RECORDED["samples"] = out.detach().clone()
        return out
    return wrapped


def state_of(gm):
    st = {}
    for group in gm.optimizer.param_groups:
        p = group["params"][0]
        st[group["name"]] = p.detach().numpy().copy()
        s = gm.optimizer.state.get(p)
        st["m_" + group["name"]] = s["exp_avg"].numpy().copy()
        st["v_" + group["name"]] = s["exp_avg_sq"].numpy().copy()
        st["step_" + group["name"]] = np.asarray(float(s["step"]))
    st["which_object"] = gm._which_object.numpy().copy()
    st["xyz_gradient_accum"] = gm.xyz_gradient_accum.numpy().copy()
    st["denom"] = gm.denom.numpy().copy()
    st["max_radii2D"] = gm.max_radii2D.numpy().copy()
    return st


def make_model(GaussianModel, inp):
    gm = GaussianModel(1)
    t = lambda a: torch.from_numpy(a.copy())
    gm._xyz = nn.Parameter(t(inp["xyz"]).requires_grad_(True))
    gm._features_dc = nn.Parameter(t(inp["f_dc"]).requires_grad_(True))
    gm._features_rest = nn.Parameter(t(inp["f_rest"]).requires_grad_(True))
    gm._opacity = nn.Parameter(t(inp["opacity"]).requires_grad_(True))
    gm._scaling = nn.Parameter(t(inp["scaling"]).requires_grad_(True))
    gm._rotation = nn.Parameter(t(inp["rotation"]).requires_grad_(True))
    gm._which_object = t(inp["which_object"])
    gm.spatial_lr_scale = 1.0
    gm.max_radii2D = torch.zeros((inp["xyz"].shape[0],))
    gm.training_setup(types.SimpleNamespace(**cases.DENSIFY_TRAIN_ARGS))
    for group in gm.optimizer.param_groups:   # one step so that exp_avg / exp_avg_sq / step exist and are non-zero
        group["params"][0].grad = t(inp["g_" + group["name"]])
    gm.optimizer.step()
    gm.optimizer.zero_grad(set_to_none=True)
    gm.xyz_gradient_accum = t(inp["xyz_gradient_accum"])
    gm.denom = t(inp["denom"])
    gm.max_radii2D = t(inp["max_radii2D"])
    return gm


if __name__ == "__main__":
    torch.set_num_threads(1)
    GaussianModel = import_reference()
    torch.zeros = on_cpu(torch.zeros)
    torch.normal = recording_normal(torch.normal)
    inp = cases.densify_inputs()
    out = {}
    for tag, kw in cases.DENSIFY_CASES.items():
        torch.manual_seed(cases.DENSIFY_SEED)
        gm = make_model(GaussianModel, inp)
        if tag == list(cases.DENSIFY_CASES)[0]:
            for k, v in state_of(gm).items():
                out["before/" + k] = v
            op = gm._opacity.detach().clone()
        RECORDED.clear()
        gm.densify_and_prune(kw["max_grad"], kw["min_opacity"], kw["extent"], kw["max_screen_size"])
        out[tag + "/samples"] = RECORDED["samples"].numpy()
        for k, v in state_of(gm).items():
            out[tag + "/" + k] = v
        print(tag, "rows", inp["xyz"].shape[0], "->", gm._xyz.shape[0], "split-selected", out[tag + "/samples"].shape[0] // 2)
    gm = make_model(GaussianModel, inp)
    gm.reset_opacity()
    st = state_of(gm)
    out["reset/opacity"], out["reset/m_opacity"], out["reset/v_opacity"] = st["opacity"], st["m_opacity"], st["v_opacity"]
    np.savez_compressed(os.path.join(HERE, "densify.npz"), **out)
    print("wrote densify.npz", os.path.getsize(os.path.join(HERE, "densify.npz")), "bytes")
