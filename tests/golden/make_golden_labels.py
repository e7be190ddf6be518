"""Generates tests/golden/labels.npz in THIS container (CPU) by executing the reference's own
`multi_instance_opt` (run_3d_seg.py:54-72): the function's source is cut out of
/root/reference/run_3d_seg.py with `ast` (the module itself imports packages that are not
installed here) and executed unmodified with torch on CPU.

    python tests/golden/make_golden_labels.py
"""
import ast
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import cases  # noqa: E402

SRC = "/root/reference/run_3d_seg.py"


def reference_function():
    tree = ast.parse(open(SRC).read())
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "multi_instance_opt"][0]
    ns = {"torch": torch, "F": F, "tqdm": lambda it, **kw: it}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), SRC, "exec"), ns)
    return ns["multi_instance_opt"]


if __name__ == "__main__":
    f = reference_function()
    a = cases.votes_inputs()
    out = {"input_sha256": np.array(cases.input_hash([a]))}
    for gamma in (0.0, 0.3, -0.2):
        out["labels_g%+.1f" % gamma] = f(torch.from_numpy(a.copy()), gamma=gamma).numpy()
    np.savez_compressed(os.path.join(HERE, "labels.npz"), **out)
    print({k: (v.shape, int(v.sum())) for k, v in out.items() if k.startswith("labels")})
