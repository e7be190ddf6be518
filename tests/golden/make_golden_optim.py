"""Generates tests/golden/optim.npz in THIS container (CPU):

* the optimiser: torch.optim.Adam configured exactly as GaussianModel.training_setup does
  (scene/gaussian_model.py:172-181: six named groups, lr=0.0 default, eps=1e-15), stepped OPTIM_STEPS
  times on the seeded gradients of cases.optim_inputs() with the xyz learning rate rewritten before every
  step as update_learning_rate does (:188-194).  torch 2.11 is the dependency that holds the arithmetic
  (the reference pins no version); this is its single-tensor CPU path.
* the densification statistics: the reference's own `GaussianModel.add_densification_stats`, cut out of
  /root/reference/scene/gaussian_model.py with `ast` (the module imports plyfile/simple_knn, absent here)
  and executed unmodified, preceded by the max_radii2D statement of train_vanilla_3dgs.py:101.

    python tests/golden/make_golden_optim.py
"""
import ast
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import cases  # noqa: E402

SRC = "/root/reference/scene/gaussian_model.py"


def reference_method(name):
    tree = ast.parse(open(SRC).read())
    cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GaussianModel"][0]
    fn = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == name][0]
    ns = {"torch": torch}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), SRC, "exec"), ns)
    return ns[name]


if __name__ == "__main__":
    torch.set_num_threads(1)
    params, grads, radii, g2d = cases.optim_inputs()
    out = {"input_sha256": np.array(cases.input_hash([params[n] for n, _, _ in cases.OPTIM_GROUPS]))}
    tp = {n: torch.nn.Parameter(torch.from_numpy(params[n].copy())) for n, _, _ in cases.OPTIM_GROUPS}
    opt = torch.optim.Adam([{"params": [tp[n]], "lr": lr, "name": n} for n, _, lr in cases.OPTIM_GROUPS],
                           lr=0.0, eps=1e-15)
    add_stats = reference_method("add_densification_stats")
    P = cases.OPTIM_P
    model = types.SimpleNamespace(xyz_gradient_accum=torch.zeros((P, 1)), denom=torch.zeros((P, 1)),
                                  max_radii2D=torch.zeros((P,)))
    for it in range(cases.OPTIM_STEPS):
        for g in opt.param_groups:
            if g["name"] == "xyz":
                g["lr"] = cases.OPTIM_XYZ_LRS[it]
        for n in tp:
            tp[n].grad = torch.from_numpy(grads[it][n].copy())
        r = torch.from_numpy(radii[it])
        vis = r > 0
        # train_vanilla_3dgs.py:101
        model.max_radii2D[vis] = torch.max(model.max_radii2D[vis], r[vis])
        add_stats(model, types.SimpleNamespace(grad=torch.from_numpy(g2d[it])), vis)
        opt.step()
        opt.zero_grad(set_to_none=True)
        for n in tp:
            st = opt.state[tp[n]]
            out["step%d_%s" % (it, n)] = tp[n].detach().numpy().copy()
            out["step%d_%s_exp_avg" % (it, n)] = st["exp_avg"].numpy().copy()
            out["step%d_%s_exp_avg_sq" % (it, n)] = st["exp_avg_sq"].numpy().copy()
        out["step%d_max_radii2D" % it] = model.max_radii2D.numpy().copy()
        out["step%d_xyz_gradient_accum" % it] = model.xyz_gradient_accum.numpy().copy()
        out["step%d_denom" % it] = model.denom.numpy().copy()
    np.savez_compressed(os.path.join(HERE, "optim.npz"), **out)
    print("torch", torch.__version__, {k: v.shape for k, v in out.items() if k.startswith("step2")})
