"""Host mirror of the reference's utils/loss_utils.py (l1_loss :17-18, ssim :35-45, _ssim :47-62) over
libwheatsplat's fused SSIM/L1 kernels (csrc/ws_loss.cu), plus the whole training loss of
train_vanilla_3dgs.py:77-80 in one forward and one backward launch:

    from wheat_3dgs_b200.loss import l1_loss, ssim, l1_ssim_loss        # instead of utils.loss_utils
    loss = l1_ssim_loss(image, gt_image, opt.lambda_dssim)              # == (1-l)*l1_loss + l*(1-ssim)

Same names, argument meaning and return shapes as the reference; CUDA tensors only (no CPU path).
The gradient flows to the first argument (the rendered image) only, as in the reference's use.
"""
import ctypes as C

import torch

from . import lib as _lib


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _planes(img1, img2):
    if img1.shape != img2.shape:
        raise RuntimeError("image shapes differ: %s vs %s" % (tuple(img1.shape), tuple(img2.shape)))
    if img1.dim() not in (3, 4):
        raise RuntimeError("expected [C,H,W] or [N,C,H,W] images")
    if not (img1.is_cuda and img2.is_cuda):
        raise RuntimeError("images must be CUDA tensors (wheat_3dgs_b200 has no CPU path)")
    H, W = int(img1.shape[-2]), int(img1.shape[-1])
    planes = int(img1.numel() // (H * W))
    return planes, H, W


class _SsimL1Sums(torch.autograd.Function):
    """(img, gt) -> float[planes, 2] = per-plane {sum of the SSIM map, sum of |img - gt|}."""

    @staticmethod
    def forward(ctx, img, gt):
        L = _lib.load()
        planes, H, W = _planes(img, gt)
        x = img.detach().float().contiguous()
        y = gt.detach().float().contiguous()
        sums = torch.empty((planes, 2), dtype=torch.float32, device=x.device)
        nbytes = L.ws_loss_bytes(planes, H, W)
        scratch = torch.empty((nbytes,), dtype=torch.uint8, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(L.ws_l1_ssim_forward(planes, H, W, x.data_ptr(), y.data_ptr(), sums.data_ptr(),
                                            scratch.data_ptr(), nbytes, _stream()), "ws_l1_ssim_forward")
        ctx.save_for_backward(x, y, scratch)
        ctx.dims = (planes, H, W)
        ctx.in_shape, ctx.in_dtype = img.shape, img.dtype
        return sums

    @staticmethod
    def backward(ctx, g_sums):
        L = _lib.load()
        x, y, scratch = ctx.saved_tensors
        planes, H, W = ctx.dims
        coef = g_sums.detach().float().contiguous()  # [planes, 2] = {a, b} per plane
        grad = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(L.ws_l1_ssim_backward(planes, H, W, x.data_ptr(), y.data_ptr(), coef.data_ptr(), 1,
                                             scratch.data_ptr(), scratch.numel(), grad.data_ptr(), _stream()),
                       "ws_l1_ssim_backward")
        return grad.view(ctx.in_shape).to(ctx.in_dtype), None


def ssim_l1_sums(img, gt):
    return _SsimL1Sums.apply(img, gt)


_WEIGHTS = {}


def _loss_weights(planes, n, lam, device):
    """[planes, 2] weights of the per-plane {SSIM sum, L1 sum} in the training loss (cached per shape / lambda)."""
    key = (planes, n, float(lam), str(device))
    w = _WEIGHTS.get(key)
    if w is None:
        w = torch.tensor([[-float(lam) / n, (1.0 - float(lam)) / n]] * planes, dtype=torch.float32, device=device)
        _WEIGHTS[key] = w
    return w


class _L1SsimLoss(torch.autograd.Function):
    """The whole loss of train_vanilla_3dgs.py:78-79 as ONE autograd node: the scalar is a dot product of the kernel's
    per-plane sums with cached weights (two small launches), its backward one scale + the fused backward kernel.  Built
    from torch ops on `ssim_l1_sums` the same value costs seventeen one-element kernels per iteration, forward and
    backward (68 us of launch gaps in a 2.7 ms iteration)."""

    @staticmethod
    def forward(ctx, img, gt, lam):
        L = _lib.load()
        planes, H, W = _planes(img, gt)
        x = img.detach().float().contiguous()
        y = gt.detach().float().contiguous()
        sums = torch.empty((planes, 2), dtype=torch.float32, device=x.device)
        nbytes = L.ws_loss_bytes(planes, H, W)
        scratch = torch.empty((nbytes,), dtype=torch.uint8, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(L.ws_l1_ssim_forward(planes, H, W, x.data_ptr(), y.data_ptr(), sums.data_ptr(),
                                            scratch.data_ptr(), nbytes, _stream()), "ws_l1_ssim_forward")
        w = _loss_weights(planes, planes * H * W, lam, x.device)
        ctx.save_for_backward(x, y, scratch, w)
        ctx.dims = (planes, H, W)
        ctx.in_shape, ctx.in_dtype = img.shape, img.dtype
        return torch.dot(sums.view(-1), w.view(-1)) + float(lam)

    @staticmethod
    def backward(ctx, g):
        L = _lib.load()
        x, y, scratch, w = ctx.saved_tensors
        planes, H, W = ctx.dims
        coef = (w * g.detach().float()).contiguous()   # [planes, 2] = {a, b} per plane
        grad = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(L.ws_l1_ssim_backward(planes, H, W, x.data_ptr(), y.data_ptr(), coef.data_ptr(), 1,
                                             scratch.data_ptr(), scratch.numel(), grad.data_ptr(), _stream()),
                       "ws_l1_ssim_backward")
        return grad.view(ctx.in_shape).to(ctx.in_dtype), None, None


def l1_loss(network_output, gt):
    """utils/loss_utils.py:17-18."""
    planes, H, W = _planes(network_output, gt)
    return ssim_l1_sums(network_output, gt)[:, 1].sum() / float(planes * H * W)


def ssim(img1, img2, window_size=11, size_average=True):
    """utils/loss_utils.py:35-62.  Only the reference's window (11, sigma 1.5) exists on the device."""
    if window_size != 11:
        raise RuntimeError("ssim: window_size must be 11 (the only size the reference uses)")
    planes, H, W = _planes(img1, img2)
    s = ssim_l1_sums(img1, img2)[:, 0]
    if size_average:
        return s.sum() / float(planes * H * W)
    if img1.dim() != 4:
        raise RuntimeError("ssim(size_average=False) needs [N,C,H,W] images")  # .mean(1).mean(1).mean(1)
    n, c = int(img1.shape[0]), int(img1.shape[1])
    return s.view(n, c).sum(dim=1) / float(c * H * W)


def l1_ssim_loss(image, gt_image, lambda_dssim=0.2):
    """train_vanilla_3dgs.py:78-79: (1 - lambda) * l1_loss + lambda * (1 - ssim), one forward launch pair
    and one backward launch."""
    return _L1SsimLoss.apply(image, gt_image, float(lambda_dssim))
