"""GPU: the fused L1 + SSIM loss (csrc/ws_loss.cu through wheat_3dgs_b200.loss, C ABI ws_l1_ssim_*) against the
reference's own Python (goldens made by tests/golden/make_golden_loss.py from /root/reference/utils/loss_utils.py)
and against the CPU oracle, plus size-independent properties at the training resolution."""
import os

import numpy as np
import pytest
import torch

import cases
import loss_oracle
from util import assert_grad

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from wheat_3dgs_b200 import loss as wloss

LOSS_ATOL = 2e-6   # float32 means of O(1) maps
SSIM_ATOL = 1e-5


def _run(img, gt, lam):
    x = torch.from_numpy(img).cuda().requires_grad_(True)
    y = torch.from_numpy(gt).cuda()
    loss = wloss.l1_ssim_loss(x, y, lam)
    loss.backward()
    torch.cuda.synchronize()
    return float(loss.detach()), x.grad.cpu().numpy()


@pytest.mark.parametrize("name", sorted(cases.LOSS_CASES))
def test_loss_and_gradient_match_reference_and_oracle(golden_dir, name):
    gold = np.load(os.path.join(golden_dir, "loss_%s.npz" % name))
    img, gt, lam = cases.loss_inputs(name)
    loss, grad = _run(img, gt, lam)
    o = loss_oracle.loss_and_grad(img, gt, lam)
    assert abs(loss - float(gold["loss"])) <= LOSS_ATOL
    assert abs(loss - o["loss"]) <= LOSS_ATOL
    assert_grad("dloss/dimage", grad, gold["grad"])
    assert_grad("dloss/dimage vs oracle", grad, o["grad"])
    x, y = torch.from_numpy(img).cuda(), torch.from_numpy(gt).cuda()
    assert abs(float(wloss.l1_loss(x, y)) - float(gold["l1"])) <= LOSS_ATOL
    assert abs(float(wloss.ssim(x, y)) - float(gold["ssim"])) <= SSIM_ATOL
    if "ssim_per_image" in gold.files:
        per = wloss.ssim(x, y, size_average=False).cpu().numpy()
        assert per.shape == gold["ssim_per_image"].shape
        assert np.abs(per - gold["ssim_per_image"]).max() <= SSIM_ATOL


def test_separate_terms_have_their_own_gradients(golden_dir):
    """l1_loss and ssim are differentiable on their own, as in the reference (utils/loss_utils.py)."""
    img, gt, lam = cases.loss_inputs("ragged")
    o0 = loss_oracle.loss_and_grad(img, gt, 0.0)   # pure L1
    o1 = loss_oracle.loss_and_grad(img, gt, 1.0)   # 1 - ssim
    x = torch.from_numpy(img).cuda().requires_grad_(True)
    y = torch.from_numpy(gt).cuda()
    wloss.l1_loss(x, y).backward()
    assert_grad("d l1", x.grad.cpu().numpy(), o0["grad"])
    x.grad = None
    (1.0 - wloss.ssim(x, y)).backward()
    assert_grad("d (1-ssim)", x.grad.cpu().numpy(), o1["grad"])


def test_identical_images_and_errors():
    y = torch.rand(3, 50, 70, device="cuda")
    assert abs(float(wloss.ssim(y, y)) - 1.0) <= 1e-6
    assert float(wloss.l1_loss(y, y)) == 0.0
    with pytest.raises(RuntimeError):
        wloss.ssim(y, y[:, :10])
    with pytest.raises(RuntimeError):
        wloss.ssim(y.cpu(), y.cpu())
    with pytest.raises(RuntimeError):
        wloss.ssim(y, y, window_size=7)


def test_training_resolution_properties():
    """3 x 1536 x 2048 (config 2): deterministic, the gradient is the directional derivative of the loss,
    and it scales linearly with the upstream gradient."""
    g = torch.Generator(device="cuda").manual_seed(3)
    H, W = 1536, 2048
    gt = torch.rand(3, H, W, device="cuda", generator=g)
    base = (gt + 0.1 * torch.randn(3, H, W, device="cuda", generator=g)).clamp(0, 1)
    x = base.clone().requires_grad_(True)
    l0 = wloss.l1_ssim_loss(x, gt, 0.2)
    l0.backward()
    g0 = x.grad.clone()
    x.grad = None
    l1 = wloss.l1_ssim_loss(x, gt, 0.2)
    l1.backward()
    assert float(l0) == float(l1), "loss is not deterministic"
    assert torch.equal(x.grad, g0), "gradient is not deterministic"
    x.grad = None
    (3.0 * wloss.l1_ssim_loss(x, gt, 0.2)).backward()
    assert torch.allclose(x.grad, 3.0 * g0, rtol=1e-5, atol=0.0), "gradient is not linear in the upstream gradient"
    # directional derivative along a smooth direction (float64 reference of the finite difference via the oracle
    # would take minutes at this size: use the kernel's own forward at +-eps and a direction that keeps |x-y| signs)
    d = torch.sign(base - gt) * torch.rand(3, H, W, device="cuda", generator=g)
    eps = 1e-3
    with torch.no_grad():
        lp = float(wloss.l1_ssim_loss(base + eps * d, gt, 0.2))
        lm = float(wloss.l1_ssim_loss(base - eps * d, gt, 0.2))
    fd = (lp - lm) / (2 * eps)
    an = float((g0.double() * d.double()).sum())
    assert abs(fd - an) <= 2e-2 * abs(an) + 1e-6, "directional derivative %g vs gradient %g" % (fd, an)
