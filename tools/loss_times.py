"""Fused L1+SSIM loss (ws_loss.cu) vs the reference's torch implementation (restated inline from
utils/loss_utils.py semantics with F.conv2d) at 3x1536x2048: CUDA-event ms per forward+backward."""
import math
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import loss as wloss  # noqa: E402

H, W = 1536, 2048
gt = torch.rand(3, H, W, device="cuda")
x = (gt + 0.1 * torch.randn_like(gt)).clamp(0, 1).requires_grad_(True)
g1 = torch.tensor([math.exp(-(i - 5) ** 2 / 4.5) for i in range(11)])
g1 = (g1 / g1.sum()).unsqueeze(1)
win = g1.mm(g1.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().cuda()


def torch_loss(a, b, lam=0.2):
    mu1, mu2 = F.conv2d(a, win, padding=5, groups=3), F.conv2d(b, win, padding=5, groups=3)
    s1 = F.conv2d(a * a, win, padding=5, groups=3) - mu1 * mu1
    s2 = F.conv2d(b * b, win, padding=5, groups=3) - mu2 * mu2
    s12 = F.conv2d(a * b, win, padding=5, groups=3) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
    return (1 - lam) * (a - b).abs().mean() + lam * (1 - m.mean())


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def ours():
    x.grad = None
    wloss.l1_ssim_loss(x, gt, 0.2).backward()


def ref():
    x.grad = None
    torch_loss(x, gt).backward()


t_o, t_r = timeit(ours), timeit(ref)
nbytes = 44 * 3 * H * W
print("fused l1+ssim fwd+bwd %.4f ms (%.0f GB/s algorithmic)   torch conv2d version %.4f ms   x%.1f" % (
    t_o, nbytes / t_o / 1e6, t_r, t_r / t_o))
