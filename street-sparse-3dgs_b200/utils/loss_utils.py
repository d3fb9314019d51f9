"""Drop-in for the reference's ``utils/loss_utils.py`` (:17-63): ``l1_loss``, ``l2_loss``, ``ssim`` with the same
signatures, plus ``fused_l1_ssim_loss`` -- the photometric loss of train_single.py:116-141 / train_post.py:132-137
``(1 - lambda) * L1 + lambda * (1 - SSIM)`` in one forward and one backward kernel.

CUDA fp32 images go through the hand-written kernels of csrc/loss.cu via the C ABI (h3dgs_ssim_forward /
h3dgs_ssim_backward); the second image (ground truth) receives no gradient, as in training.  There is no fallback
for CUDA tensors; CPU tensors are rejected (the reference's torch implementation is what runs on a CPU)."""
import ctypes
from math import exp

import torch

import h3dgs_native as nat


def l1_loss(network_output, gt):
    return torch.abs((network_output - gt)).mean()


def l2_loss(network_output, gt):
    return ((network_output - gt) ** 2).mean()


def gaussian(window_size, sigma):
    # utils/loss_utils.py:23-25, same arithmetic: python doubles -> float32 tensor -> normalised in float32
    gauss = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
    return gauss / gauss.sum()


_WINDOW = None


def _window11():
    global _WINDOW
    if _WINDOW is None:
        w = gaussian(11, 1.5)
        _WINDOW = (ctypes.c_float * 11)(*[float(x) for x in w])
    return _WINDOW


def _check(img1, img2):
    if not (img1.is_cuda and img2.is_cuda):
        raise ValueError("fused SSIM needs CUDA tensors (no CPU fallback)")
    if img1.dtype != torch.float32 or img2.dtype != torch.float32:
        raise TypeError("fused SSIM needs float32 images")
    if img1.shape != img2.shape or img1.dim() not in (3, 4):
        raise RuntimeError("fused SSIM needs two images of the same shape [C,H,W] or [B,C,H,W]")
    return img1.contiguous(), img2.contiguous()


def _forward(img1, img2, want_grad, want_map=False):
    H, W = img1.shape[-2:]
    C = img1.numel() // (H * W)
    sums = torch.zeros(2, dtype=torch.float64, device=img1.device)
    maps = [torch.empty_like(img1) for _ in range(3)] if want_grad else [None, None, None]
    smap = torch.empty_like(img1) if want_map else None
    nat.check(nat.lib().h3dgs_ssim_forward(C, H, W, nat.ptr(img1), nat.ptr(img2), _window11(), nat.ptr(smap),
                                           nat.ptr(maps[0]), nat.ptr(maps[1]), nat.ptr(maps[2]), nat.ptr(sums),
                                           nat.stream_handle()), "h3dgs_ssim_forward")
    return sums, maps, smap


def _backward(img1, img2, maps, upstream, ssim_scale, l1_scale):
    H, W = img1.shape[-2:]
    C = img1.numel() // (H * W)
    grad = torch.empty_like(img1)
    up = upstream.to(torch.float32).contiguous()
    nat.check(nat.lib().h3dgs_ssim_backward(C, H, W, nat.ptr(img1), nat.ptr(img2), _window11(), nat.ptr(maps[0]),
                                            nat.ptr(maps[1]), nat.ptr(maps[2]), None, nat.ptr(up), ssim_scale,
                                            l1_scale, nat.ptr(grad), nat.stream_handle()), "h3dgs_ssim_backward")
    return grad


class _SSIMMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img1, img2):
        img1, img2 = _check(img1, img2)
        want = ctx.needs_input_grad[0]
        sums, maps, _ = _forward(img1, img2, want)
        if want:
            ctx.save_for_backward(img1, img2, *maps)
        return (sums[0] / img1.numel()).to(torch.float32)

    @staticmethod
    def backward(ctx, grad_out):
        img1, img2, m0, m1, m2 = ctx.saved_tensors
        return _backward(img1, img2, (m0, m1, m2), grad_out, 1.0 / img1.numel(), 0.0), None


class _L1SSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim):
        image, gt = _check(image, gt)
        want = ctx.needs_input_grad[0]
        sums, maps, _ = _forward(image, gt, want)
        n = image.numel()
        means = (sums / n).to(torch.float32)
        ll1 = means[1]
        loss = (1.0 - lambda_dssim) * ll1 + lambda_dssim * (1.0 - means[0])
        ctx.lam = float(lambda_dssim)
        if want:
            ctx.save_for_backward(image, gt, *maps)
        ctx.mark_non_differentiable(ll1)
        return loss, ll1

    @staticmethod
    def backward(ctx, grad_loss, _grad_l1):
        image, gt, m0, m1, m2 = ctx.saved_tensors
        n = image.numel()
        return _backward(image, gt, (m0, m1, m2), grad_loss, -ctx.lam / n, (1.0 - ctx.lam) / n), None, None


def ssim(img1, img2, window_size=11, size_average=True):
    """Mean SSIM of two images (utils/loss_utils.py:33-63).  Gradients flow to ``img1``."""
    if window_size != 11:
        raise NotImplementedError("the fused SSIM implements the window the reference training uses (11)")
    if size_average or img1.dim() == 3 or img1.size(0) == 1:
        out = _SSIMMean.apply(img1, img2)
        return out if size_average else out.reshape(1)
    # ssim_map.mean(1).mean(1).mean(1): one value per batch entry
    return torch.stack([_SSIMMean.apply(a, b) for a, b in zip(img1, img2)])


def ssim_map(img1, img2):
    """The per-pixel SSIM map (no gradient): the quantity ``_ssim`` averages."""
    img1, img2 = _check(img1, img2)
    return _forward(img1, img2, False, want_map=True)[2]


def fused_l1_ssim_loss(image, gt, lambda_dssim=0.2):
    """``(loss, Ll1)`` with ``loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))``
    (train_single.py:122-125, train_post.py:132-137); ``Ll1`` is returned for logging and carries no gradient."""
    return _L1SSIM.apply(image, gt, float(lambda_dssim))
