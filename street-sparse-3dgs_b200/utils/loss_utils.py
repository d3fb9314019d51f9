"""Drop-in for the reference's ``utils/loss_utils.py`` (:17-63): ``l1_loss``, ``l2_loss``, ``ssim`` with the same
signatures, plus ``fused_l1_ssim_loss`` -- the photometric loss of train_single.py:116-141 / train_post.py:132-137
``(1 - lambda) * L1 + lambda * (1 - SSIM)`` in one forward and one backward kernel.

CUDA fp32 images with the training window (11) go through the hand-written kernels of csrc/loss.cu via the C ABI
(h3dgs_ssim_forward / h3dgs_ssim_backward); SSIM is symmetric in its two images, so a second image that requires a
gradient gets it from a second pass with the roles swapped.  What the kernels do not cover -- CPU tensors, other
window sizes, and the evaluation-only ``ssim_masked`` (render_hierarchy_final.py:154) -- is evaluated with the same
torch formulas the reference uses (``_ssim_torch``); the training losses (``fused_l1_ssim_loss``, ``masked_l1_loss``)
have no such path and reject anything but CUDA fp32."""
import ctypes
from math import exp

import torch
import torch.nn.functional as F

import h3dgs_native as nat


def l1_loss(network_output, gt):
    return torch.abs((network_output - gt)).mean()


def l2_loss(network_output, gt):
    return ((network_output - gt) ** 2).mean()


def gaussian(window_size, sigma):
    # utils/loss_utils.py:23-25, same arithmetic: python doubles -> float32 tensor -> normalised in float32
    gauss = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
    return gauss / gauss.sum()


def create_window(window_size, channel):
    # utils/loss_utils.py:27-31: separable window as a [channel,1,ws,ws] depthwise conv2d weight
    w = gaussian(window_size, 1.5).unsqueeze(1)
    return w.mm(w.t()).float()[None, None].expand(channel, 1, window_size, window_size).contiguous()


def _moments(x, window, pad, channel):
    return F.conv2d(x, window, padding=pad, groups=channel)


def _ssim_torch(img1, img2, window_size, size_average):
    """utils/loss_utils.py:33-63 in torch ops (depthwise conv2d moments), for inputs the kernels do not take."""
    channel = img1.size(-3)
    window = create_window(window_size, channel).to(device=img1.device, dtype=img1.dtype)
    pad = window_size // 2
    mu1, mu2 = _moments(img1, window, pad, channel), _moments(img2, window, pad, channel)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = _moments(img1 * img1, window, pad, channel) - mu1_sq
    s2 = _moments(img2 * img2, window, pad, channel) - mu2_sq
    s12 = _moments(img1 * img2, window, pad, channel) - mu12
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    smap = ((2 * mu12 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))
    return smap.mean() if size_average else smap.mean(1).mean(1).mean(1)


def ssim_masked(img1, img2, mask, window_size=11, size_average=True):
    """SSIM restricted to ``mask`` (utils/loss_utils.py:65-150; evaluation only, render_hierarchy_final.py:154):
    windowed moments of the masked images renormalised by the windowed mask weight, windows without support and
    degenerate denominators give 0, the mean runs over masked pixels with a valid window."""
    channel = img1.size(-3)
    window = create_window(window_size, channel).to(device=img1.device, dtype=img1.dtype)
    pad = window_size // 2
    if mask.dim() == 3:
        mask = mask.unsqueeze(1)
    if mask.size(1) == 1 and channel > 1:
        mask = mask.expand(-1, channel, -1, -1)
    a, b = img1 * mask, img2 * mask
    wsum = _moments(mask.float(), window, pad, channel)
    valid = wsum > 1e-6
    safe = torch.where(valid, wsum, torch.ones_like(wsum))
    zero = torch.zeros_like(wsum)

    def local(x):
        return torch.where(valid, _moments(x, window, pad, channel) / safe, zero)

    mu1, mu2 = local(a), local(b)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = torch.where(valid, local(a * a) - mu1_sq, zero)
    s2 = torch.where(valid, local(b * b) - mu2_sq, zero)
    s12 = torch.where(valid, local(a * b) - mu12, zero)
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    num = (2 * mu12 + C1) * (2 * s12 + C2)
    den = (mu1_sq + mu2_sq + C1) * (s1 + s2 + C2)
    ok = (den > 1e-10) & valid
    smap = torch.where(ok, num / torch.where(ok, den, torch.ones_like(den)), zero) * mask
    if size_average:
        return smap.sum() / (mask * valid.float()).sum().clamp(min=1)
    return smap


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


def _check_mask(mask, img):
    """The alpha mask of train_single.py:118-120 (``image *= alpha_mask``): [H,W] or [1,H,W], broadcast over the
    channels.  It receives no gradient."""
    if mask is None:
        return None
    if not mask.is_cuda or mask.dtype != torch.float32:
        raise TypeError("fused loss: the mask must be a CUDA float32 tensor")
    H, W = img.shape[-2:]
    if mask.numel() != H * W or img.dim() != 3:
        raise RuntimeError("fused loss: the mask must be [H,W] or [1,H,W] for a [C,H,W] image")
    return mask.detach().contiguous()


def _forward(img1, img2, want_grad, want_map=False, mask=None):
    H, W = img1.shape[-2:]
    C = img1.numel() // (H * W)
    sums = torch.zeros(2, dtype=torch.float64, device=img1.device)
    maps = [torch.empty_like(img1) for _ in range(3)] if want_grad else [None, None, None]
    smap = torch.empty_like(img1) if want_map else None
    nat.check(nat.lib().h3dgs_ssim_forward(C, H, W, nat.ptr(img1), nat.ptr(img2), nat.ptr(mask), _window11(),
                                           nat.ptr(smap),
                                           nat.ptr(maps[0]), nat.ptr(maps[1]), nat.ptr(maps[2]), nat.ptr(sums),
                                           nat.stream_handle()), "h3dgs_ssim_forward")
    return sums, maps, smap


def _backward(img1, img2, maps, upstream, ssim_scale, l1_scale, mask=None):
    H, W = img1.shape[-2:]
    C = img1.numel() // (H * W)
    grad = torch.empty_like(img1)
    up = upstream.to(torch.float32).contiguous()
    nat.check(nat.lib().h3dgs_ssim_backward(C, H, W, nat.ptr(img1), nat.ptr(img2), nat.ptr(mask), _window11(),
                                            nat.ptr(maps[0]),
                                            nat.ptr(maps[1]), nat.ptr(maps[2]), None, nat.ptr(up), ssim_scale,
                                            l1_scale, nat.ptr(grad), nat.stream_handle()), "h3dgs_ssim_backward")
    return grad


class _SSIMMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img1, img2):
        img1, img2 = _check(img1, img2)
        want1, want2 = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        sums, maps, _ = _forward(img1, img2, want1)
        ctx.want = (want1, want2)
        ctx.save_for_backward(img1, img2, *(maps if want1 else ()))
        return (sums[0] / img1.numel()).to(torch.float32)

    @staticmethod
    def backward(ctx, grad_out):
        img1, img2 = ctx.saved_tensors[:2]
        g1 = g2 = None
        if ctx.want[0]:
            g1 = _backward(img1, img2, ctx.saved_tensors[2:], grad_out, 1.0 / img1.numel(), 0.0)
        if ctx.want[1]:
            # SSIM(a, b) = SSIM(b, a): the gradient of the second image is that of the first with the roles swapped
            _, maps, _ = _forward(img2, img1, True)
            g2 = _backward(img2, img1, maps, grad_out, 1.0 / img1.numel(), 0.0)
        return g1, g2


class _L1SSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim, mask):
        image, gt = _check(image, gt)
        mask = _check_mask(mask, image)
        want = ctx.needs_input_grad[0]
        sums, maps, _ = _forward(image, gt, want, mask=mask)
        ctx.mask = mask
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
        return (_backward(image, gt, (m0, m1, m2), grad_loss, -ctx.lam / n, (1.0 - ctx.lam) / n, mask=ctx.mask),
                None, None, None)


def ssim(img1, img2, window_size=11, size_average=True):
    """Mean SSIM of two images (utils/loss_utils.py:33-63)."""
    if window_size != 11 or not (img1.is_cuda and img2.is_cuda) or img1.dtype != torch.float32 \
            or img2.dtype != torch.float32:
        return _ssim_torch(img1, img2, window_size, size_average)
    if size_average or img1.dim() == 3 or img1.size(0) == 1:
        out = _SSIMMean.apply(img1, img2)
        return out if size_average else out.reshape(1)
    # ssim_map.mean(1).mean(1).mean(1): one value per batch entry
    return torch.stack([_SSIMMean.apply(a, b) for a, b in zip(img1, img2)])


def ssim_map(img1, img2):
    """The per-pixel SSIM map (no gradient): the quantity ``_ssim`` averages."""
    img1, img2 = _check(img1, img2)
    return _forward(img1, img2, False, want_map=True)[2]


def fused_l1_ssim_loss(image, gt, lambda_dssim=0.2, mask=None):
    """``(loss, Ll1)`` with ``loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))``
    (train_single.py:122-125, train_post.py:132-137); ``Ll1`` is returned for logging and carries no gradient.
    With ``mask`` (the camera's ``alpha_mask``, [H,W] or [1,H,W]) both terms are evaluated on ``image * mask`` as
    train_single.py:118-120 / train_post.py:134-136 do, without materialising the product."""
    return _L1SSIM.apply(image, gt, float(lambda_dssim), mask)


class _MaskedL1(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target, mask):
        if not (pred.is_cuda and target.is_cuda) or pred.dtype != torch.float32 or target.dtype != torch.float32:
            raise TypeError("masked_l1_loss needs CUDA float32 tensors (no CPU fallback)")
        if pred.shape != target.shape or (mask is not None and mask.numel() != pred.numel()):
            raise RuntimeError("masked_l1_loss: pred, target and mask must have the same number of elements")
        pred, target = pred.contiguous(), target.contiguous()
        if mask is not None:
            if not mask.is_cuda or mask.dtype != torch.float32:
                raise TypeError("masked_l1_loss: the mask must be a CUDA float32 tensor")
            mask = mask.detach().contiguous()
        n = pred.numel()
        s = torch.zeros(1, dtype=torch.float64, device=pred.device)
        nat.check(nat.lib().h3dgs_masked_l1_forward(n, nat.ptr(pred), nat.ptr(target), nat.ptr(mask), nat.ptr(s),
                                                    nat.stream_handle()), "h3dgs_masked_l1_forward")
        ctx.save_for_backward(pred, target)
        ctx.mask = mask
        return (s[0] / max(n, 1)).to(torch.float32)

    @staticmethod
    def backward(ctx, grad_out):
        pred, target = ctx.saved_tensors
        n = pred.numel()
        grad = torch.empty_like(pred)
        up = grad_out.to(torch.float32).contiguous()
        nat.check(nat.lib().h3dgs_masked_l1_backward(n, nat.ptr(pred), nat.ptr(target), nat.ptr(ctx.mask), nat.ptr(up),
                                                     1.0 / max(n, 1), nat.ptr(grad), nat.stream_handle()),
                  "h3dgs_masked_l1_backward")
        return grad, None, None


def masked_l1_loss(pred, target, mask=None):
    """``torch.abs((pred - target) * mask).mean()`` -- the depth loss of train_single.py:133-139
    (``Ll1depth_pure`` with ``pred = invDepth``, ``target = mono_invdepth``, ``mask = depth_mask``) in one forward and
    one backward kernel.  Gradient flows to ``pred`` only."""
    return _MaskedL1.apply(pred, target, mask)
