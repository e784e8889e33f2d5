"""Host-side mirror of the reference's image losses (``utils/loss_utils.py:57-106``) on top of the
fused CUDA loss (``egs_image_loss``).

``l1_loss`` and ``ssim`` keep the reference's names and argument meaning. ``image_loss`` is the
combination every trainer evaluates (``trainers/train_static_bg.py:98-102``,
``fine_all.py:99-100``, ``fine_obj.py:142-143``, ``coarse_obj_pose.py:252-253``)::

    render_image.register_hook(lambda grad: grad * (1 - int_mask))
    loss = (1 - lambda_dssim) * l1_loss(render, gt) + lambda_dssim * (1 - ssim(render, gt))

as one call: value and gradient come out of two tiled kernels, and backward is a single multiply
by the incoming gradient. The kernel differentiates with respect to one image. Both terms are
symmetric in their arguments, and three of the five trainers pass the render SECOND
(``fine_all.py:99-100``, ``fine_obj.py:142-147``, ``coarse_obj_pose.py:252-256``:
``l1_loss(gt_image, render_image)``, ``l1_loss(obj_mask, render_alpha)``), so ``l1_loss`` and
``ssim`` hand the kernel whichever argument requires a gradient and refuse the case where both do;
nothing is ever detached silently. ``l2_loss`` (``fine_obj.py:150``, ``coarse_obj_pose.py:259``, a
[1,H,W] alpha against a mask) is two elementwise torch ops on the device, not a fused kernel.

There is no CPU path.
"""
from __future__ import annotations

import torch

from .diff_gaussian_rasterization import _C

__all__ = ["l1_loss", "l2_loss", "ssim", "image_loss"]

_EMPTY = torch.empty(0)


class _ImageLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, target, grad_mask, lambda_dssim):
        want = image.requires_grad
        terms, grad = _C.image_loss(image, target, grad_mask, lambda_dssim, want)
        if want:
            ctx.save_for_backward(grad)
        ctx.mark_non_differentiable(terms)
        return terms[2].clone(), terms

    @staticmethod
    def backward(ctx, g_loss, _g_terms):
        (grad,) = ctx.saved_tensors
        return grad * g_loss, None, None, None


def image_loss(render: torch.Tensor, gt: torch.Tensor, lambda_dssim: float, grad_mask=None):
    """(loss, Ll1, ssim_value) for a [C,H,W] render against ``gt``. ``grad_mask`` ([H,W] or [1,H,W],
    any integer/bool type) multiplies the gradient into ``render`` by ``1 - mask``. ``loss`` is
    differentiable w.r.t. ``render``; the two terms are returned detached for logging."""
    if gt.requires_grad:
        raise NotImplementedError("image_loss: gradients flow to the first argument only")
    mask = _EMPTY if grad_mask is None else grad_mask
    loss, terms = _ImageLoss.apply(render, gt.detach(), mask, float(lambda_dssim))
    return loss, terms[0], terms[1]


def _differentiated_first(a: torch.Tensor, b: torch.Tensor, who: str):
    """Order a symmetric loss's arguments so that the one that requires a gradient comes first."""
    if a.requires_grad and b.requires_grad:
        raise NotImplementedError(f"{who}: both arguments require a gradient; the fused loss differentiates "
                                  "with respect to one image (detach the other explicitly)")
    return (b, a) if b.requires_grad else (a, b)


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """mean |network_output - gt| (utils/loss_utils.py:57-58); the gradient goes to whichever argument
    requires one."""
    x, y = _differentiated_first(network_output, gt, "l1_loss")
    if x.dim() != 3:
        x, y = x.reshape(1, 1, -1), y.reshape(1, 1, -1)
    return _ImageLoss.apply(x, y, _EMPTY, 0.0)[0]


def l2_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """mean (network_output - gt)^2 (utils/loss_utils.py:60-61). Plain torch ops on the device."""
    if not network_output.is_cuda:
        raise RuntimeError("l2_loss: CUDA tensors only (there is no CPU path)")
    return ((network_output - gt) ** 2).mean()


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11, size_average: bool = True) -> torch.Tensor:
    """SSIM with the reference's 11x11 Gaussian window (utils/loss_utils.py:76-106); the gradient goes
    to whichever argument requires one."""
    if window_size != 11 or not size_average or img1.dim() != 3:
        raise NotImplementedError("ssim: only [C,H,W] inputs with window_size=11, size_average=True "
                                  "(what the trainers use) are on the accelerated path")
    x, y = _differentiated_first(img1, img2, "ssim")
    # loss with lambda = 1 is 1 - ssim
    return 1.0 - _ImageLoss.apply(x, y, _EMPTY, 1.0)[0]
