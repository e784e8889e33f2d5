"""CPU/torch restatement of the trainers' image loss (TEST INFRASTRUCTURE; only tests/, smoke() and
bench.py's baseline legs import it).

  l1_loss()     mean |a - b|                                         utils/loss_utils.py:57-58
  window()      normalised 1-D Gaussian (sigma 1.5), outer product   utils/loss_utils.py:66-74
  ssim()        five depthwise 11x11 correlations, zero padding 5    utils/loss_utils.py:76-106
  image_loss()  (1 - lambda) L1 + lambda (1 - SSIM), with the gradient mask hook
                                                                    trainers/train_static_bg.py:98-102

Pinned by tests/golden/loss_*.npz, produced by tests/golden/make_loss_golden.py from the reference's
own utils/loss_utils.py (tests/test_oracle_golden.py::test_loss_restatement_matches_reference).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def l1_loss(a, b):
    return (a - b).abs().mean()


def window(size: int = 11, sigma: float = 1.5, channels: int = 3):
    g = torch.tensor([math.exp(-(x - size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(size)])
    g = (g / g.sum()).unsqueeze(1)
    return (g @ g.t()).float()[None, None].expand(channels, 1, size, size).contiguous()


def ssim(img1, img2, size: int = 11):
    C = img1.size(-3)
    w = window(size, 1.5, C).to(img1)
    conv = lambda t: F.conv2d(t, w, padding=size // 2, groups=C)   # noqa: E731
    mu1, mu2 = conv(img1), conv(img2)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = conv(img1 * img1) - mu1_sq
    s2 = conv(img2 * img2) - mu2_sq
    s12 = conv(img1 * img2) - mu12
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu12 + c1) * (2 * s12 + c2)) / ((mu1_sq + mu2_sq + c1) * (s1 + s2 + c2))).mean()


def image_loss(render, gt, lambda_dssim: float, grad_mask=None):
    """Returns (loss, l1, ssim); `grad_mask` ([H,W] or [1,H,W] int) multiplies the gradient that
    flows into `render` by (1 - mask), like the trainers' register_hook."""
    if grad_mask is not None and render.requires_grad:
        keep = (1 - grad_mask).to(render.dtype)
        render = render * 1.0   # a non-leaf to hook
        render.register_hook(lambda g: g * keep)
    l1 = l1_loss(render, gt)
    s = ssim(render, gt)
    return (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - s), l1, s
