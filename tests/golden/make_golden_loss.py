"""Generates tests/golden/loss_*.npz in THIS container (CPU) by importing the reference's own
utils/loss_utils.py (pure torch) from /root/reference and running it, with autograd, on the seeded
image pairs of cases.loss_inputs():

    python tests/golden/make_golden_loss.py

Stored per case: loss = (1-l)*l1_loss + l*(1-ssim) (train_vanilla_3dgs.py:78-79), l1, ssim, the gradient
of the loss w.r.t. the image, and for the 4-D case ssim(size_average=False).
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import cases  # noqa: E402

spec = importlib.util.spec_from_file_location("ref_loss_utils", "/root/reference/utils/loss_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

if __name__ == "__main__":
    torch.set_num_threads(8)
    for name in sorted(cases.LOSS_CASES):
        img, gt, lam = cases.loss_inputs(name)
        x = torch.from_numpy(img.copy()).requires_grad_(True)
        y = torch.from_numpy(gt.copy())
        l1 = ref.l1_loss(x, y)
        ss = ref.ssim(x, y)
        loss = (1.0 - lam) * l1 + lam * (1.0 - ss)
        loss.backward()
        out = {"input_sha256": np.array(cases.input_hash([img, gt])), "lambda_dssim": np.float32(lam),
               "loss": loss.detach().numpy(), "l1": l1.detach().numpy(), "ssim": ss.detach().numpy(),
               "grad": x.grad.numpy()}
        if img.ndim == 4:
            out["ssim_per_image"] = ref.ssim(x.detach(), y, size_average=False).numpy()
        np.savez_compressed(os.path.join(HERE, "loss_%s.npz" % name), **out)
        print(name, img.shape, float(loss), float(l1), float(ss), float(np.abs(out["grad"]).max()))
