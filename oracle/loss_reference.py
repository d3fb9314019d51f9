"""TEST INFRASTRUCTURE ONLY -- CPU restatement (plain torch ops on CPU tensors) of the reference's photometric loss,
the checker for csrc/loss.cu.  Follows utils/loss_utils.py:17-18 (l1_loss), :23-31 (gaussian, create_window),
:33-63 (ssim, _ssim) and the loss assembly of train_single.py:122-125 / train_post.py:132-137.
Pinned against outputs of the reference's own module (tests/golden/loss_l1_ssim.npz, made by
tests/golden/make_golden_loss.py, which imports /root/reference/utils/loss_utils.py); the masked cases and the depth
loss (formulas written inline in train_single.py:116-139, not importable functions) are pinned against the same
script's transcription of those lines around the reference's l1_loss / ssim."""
from math import exp

import torch
import torch.nn.functional as F


def gaussian(window_size, sigma):                      # loss_utils.py:23-25
    g = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
    return g / g.sum()


def create_window(window_size, channel):               # loss_utils.py:27-31
    w1 = gaussian(window_size, 1.5).unsqueeze(1)
    w2 = w1.mm(w1.t()).float().unsqueeze(0).unsqueeze(0)
    return w2.expand(channel, 1, window_size, window_size).contiguous()


def ssim_map(img1, img2, window_size=11):              # loss_utils.py:43-58
    channel = img1.size(-3)
    window = create_window(window_size, channel).to(device=img1.device, dtype=img1.dtype)   # :38-40
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))


def ssim(img1, img2, window_size=11):                  # loss_utils.py:33-41,60-61 (size_average=True)
    return ssim_map(img1, img2, window_size).mean()


def l1_loss(a, b):                                     # loss_utils.py:17-18
    return torch.abs(a - b).mean()


def l1_ssim_loss(image, gt, lambda_dssim=0.2, mask=None):   # train_single.py:116-125, train_post.py:132-137
    if mask is not None:                                    # train_single.py:118-120: image *= alpha_mask
        image = image * mask
    ll1 = l1_loss(image, gt)
    return (1.0 - lambda_dssim) * ll1 + lambda_dssim * (1.0 - ssim(image, gt)), ll1


def depth_l1(inv_depth, mono_invdepth, depth_mask=None):    # train_single.py:133-139 (Ll1depth_pure)
    d = inv_depth - mono_invdepth
    if depth_mask is not None:
        d = d * depth_mask
    return torch.abs(d).mean()
