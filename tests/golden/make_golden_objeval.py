"""Generates tests/golden/objeval.npz (batched eval_obj_labels) in two stages, because the
reference's rasterizer needs a GPU and the reference's Python is only readable in the build container:

  stage 1 (B200):  gpurun -- python tests/golden/make_golden_objeval.py gpu
      for every non-empty row i >= 1 of cases.objeval_inputs()' all_obj_labels, the UNMODIFIED reference
      flashsplat rasterizer (oracle/_ref/libref_fsr.so) renders the Gaussians of that row as
      flashsplat_render(..., used_mask=row) does (gaussian_renderer/__init__.py:154-187: subset of means3D,
      opacity, scales, rotations, shs; gt_mask None -> ones; obj_num 2) -> gpurun_out/golden/objeval_renders.npz
  stage 2 (here):  python tests/golden/make_golden_objeval.py cpu gpurun_out/golden/objeval_renders.npz
      the reference's own eval_obj_labels, cut out of /root/reference/utils/wheatgs_helper.py with `ast` and
      executed unmodified; its `from gaussian_renderer import flashsplat_render` resolves to a stand-in module
      that hands back the stage-1 alpha/depth of the requested row (rows are passed as small wrappers so that
      the function's `.to("cuda")` works without a GPU) -> tests/golden/objeval.npz
"""
import ast
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import cases  # noqa: E402

SRC = "/root/reference/utils/wheatgs_helper.py"
KEYS = ("means3D", "opacities", "scales", "rotations", "shs")


def stage_gpu():
    import ref as refmod
    sc, cam_np, labels = cases.objeval_inputs()
    out = {}
    rr = refmod.RefRasterizer("fsr")
    bg = torch.from_numpy(cases.BG).cuda()
    for i in range(1, labels.shape[0]):
        if labels[i].sum() == 0:
            continue
        sub = {k: np.ascontiguousarray(sc[k][labels[i]]) for k in KEYS}
        g, cam = refmod.to_device(sub, cam_np, "cuda")
        o = rr.forward(g, cam, bg, 3, gt_mask=None, num_obj=2)
        torch.cuda.synchronize()
        out["alpha_%d" % i] = o["alpha"].cpu().numpy().copy()
        out["depth_%d" % i] = o["depth"].cpu().numpy().copy()
        print(i, int(labels[i].sum()), int((out["alpha_%d" % i] > 0.5).sum()))
    d = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(d, exist_ok=True)
    np.savez_compressed(os.path.join(d, "objeval_renders.npz"), **out)


class _Row:
    """One row of all_obj_labels with the few tensor methods eval_obj_labels calls on it."""

    def __init__(self, idx, row):
        self.idx, self.row = idx, torch.from_numpy(row)

    def bool(self):
        return self

    def sum(self):
        return self.row.sum()

    def to(self, device):
        return self


class _Labels:
    def __init__(self, labels):
        self.labels = labels

    def size(self, dim):
        return self.labels.shape[dim]

    def __getitem__(self, i):
        return _Row(i, self.labels[i])


def stage_cpu(renders_path):
    ren = np.load(renders_path)
    sc, cam_np, labels = cases.objeval_inputs()
    tree = ast.parse(open(SRC).read())
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "eval_obj_labels"][0]
    calls = []

    def flashsplat_render(viewpoint_cam, gaussians, pipe, background, used_mask=None):
        calls.append(used_mask.idx)
        return {"alpha": torch.from_numpy(ren["alpha_%d" % used_mask.idx].copy()),
                "depth": torch.from_numpy(ren["depth_%d" % used_mask.idx].copy())}

    sys.modules["gaussian_renderer"] = types.SimpleNamespace(flashsplat_render=flashsplat_render)
    ns = {"torch": torch}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), SRC, "exec"), ns)
    pred = ns["eval_obj_labels"](_Labels(labels), None, None, None, None)
    out = {"pred_mask": pred.numpy(), "rendered_rows": np.array(calls, np.int32),
           "input_sha256": np.array(cases.input_hash([sc[k] for k in KEYS] + [labels.astype(np.uint8)]))}
    for k in ren.files:
        out[k] = ren[k]
    np.savez_compressed(os.path.join(HERE, "objeval.npz"), **out)
    print("rows rendered", calls, "ids painted", np.unique(out["pred_mask"], return_counts=True))


if __name__ == "__main__":
    if sys.argv[1] == "gpu":
        stage_gpu()
    else:
        stage_cpu(sys.argv[2])
