"""GPU parity of the fused image loss (egs_image_loss).

Checked against the golden vectors tests/golden/loss_*.npz (produced by the reference's own
utils/loss_utils.py, tests/golden/make_loss_golden.py) and, at the bench resolution, against the
torch restatement oracle/loss_ref.py run on the same GPU. Tolerances: scalars 2e-6 relative, the
gradient image 1e-4 of its maximum (separable fp32 window sums versus cuDNN's 2-D correlation)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from golden.make_loss_golden import LOSS_CASES, loss_inputs
from test_oracle_golden import assert_loss_close

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"


def run_cabi(x, lam, use_mask):
    from egogaussian_b200 import cabi
    terms, grad = cabi.image_loss(x["render"], x["gt"], lam, x["mask"] if use_mask else None)
    t = terms.cpu().numpy()
    return {"l1": t[0], "ssim": t[1], "loss": t[2], "d_render": grad.cpu().numpy()}


@pytest.mark.parametrize("case", LOSS_CASES, ids=[c["name"] for c in LOSS_CASES])
def test_loss_kernels_against_reference_golden(case, built, cuda_device):
    x = {k: v.to(cuda_device) for k, v in loss_inputs(case).items()}
    gold = np.load(GOLDEN / f"{case['name']}.npz")
    got = run_cabi(x, case["lam"], case["mask"])
    assert_loss_close(got, gold, case["name"])
    again = run_cabi(x, case["lam"], case["mask"])   # fixed summation order: bit-identical reruns
    assert got["loss"] == again["loss"] and np.array_equal(got["d_render"], again["d_render"])


def test_loss_full_resolution_and_python_mirror(built, cuda_device):
    from egogaussian_b200 import loss_utils
    from oracle import loss_ref
    dev = cuda_device
    gen = torch.Generator().manual_seed(11)
    H, W = 720, 1280
    gt = torch.rand(3, H, W, generator=gen).to(dev)
    render = (gt + 0.1 * torch.randn(3, H, W, generator=gen).to(dev)).clamp(0, 1)
    mask = (torch.rand(H, W, generator=gen) < 0.2).to(torch.int32).to(dev)

    ref_leaf = render.clone().requires_grad_(True)
    ref_loss, ref_l1, ref_ssim = loss_ref.image_loss(ref_leaf, gt, 0.2, mask)
    ref_loss.backward()
    gold = {"l1": ref_l1.item(), "ssim": ref_ssim.item(), "loss": ref_loss.item(), "d_render": ref_leaf.grad.cpu().numpy()}

    leaf = render.clone().requires_grad_(True)
    loss, l1, s = loss_utils.image_loss(leaf, gt, 0.2, grad_mask=mask)
    (3.0 * loss).backward()   # the incoming gradient scales the stored one
    got = {"l1": l1.item(), "ssim": s.item(), "loss": loss.item(), "d_render": (leaf.grad / 3.0).cpu().numpy()}
    assert_loss_close(got, gold, "720p", value_rtol=5e-6)
    assert bool((leaf.grad[:, mask.bool()] == 0).all())

    # the reference's separate entry points
    assert abs(loss_utils.l1_loss(render, gt).item() - ref_l1.item()) <= 5e-6 * ref_l1.item()
    assert abs(loss_utils.ssim(render, gt).item() - ref_ssim.item()) <= 5e-6
    a = render[:1].clone().requires_grad_(True)
    loss_utils.l1_loss(gt[:1], a.detach())   # value only
    loss_utils.l1_loss(a, gt[:1]).backward()
    expect = torch.sign(a.detach() - gt[:1]) / a.numel()
    assert torch.allclose(a.grad, expect, rtol=1e-6, atol=0)


def test_loss_argument_order_of_the_trainers(built, cuda_device):
    """fine_all.py:99-100, fine_obj.py:142-147 and coarse_obj_pose.py:252-256 call l1_loss(gt, render),
    ssim(gt, render), l1_loss(mask, alpha) and l2_loss(alpha, mask): the gradient must reach the render
    whichever position it is in, and nothing may be detached silently."""
    from egogaussian_b200 import loss_utils
    dev = cuda_device
    gen = torch.Generator().manual_seed(2)
    gt = torch.rand(3, 96, 128, generator=gen).to(dev)
    base = (gt + 0.05 * torch.randn(3, 96, 128, generator=gen).to(dev)).clamp(0, 1)
    for fn in (loss_utils.l1_loss, loss_utils.ssim):
        a = base.clone().requires_grad_(True)
        b = base.clone().requires_grad_(True)
        va = fn(a, gt)
        vb = fn(gt, b)
        va.backward()
        vb.backward()
        assert va.item() == vb.item()
        assert b.grad is not None and torch.equal(a.grad, b.grad) and float(a.grad.abs().max()) > 0
        with pytest.raises(NotImplementedError, match="both arguments"):
            fn(a, b)
    alpha = torch.rand(1, 96, 128, generator=gen).to(dev).requires_grad_(True)
    mask = (torch.rand(1, 96, 128, generator=gen) < 0.5).float().to(dev)
    loss = loss_utils.l1_loss(mask, alpha) + loss_utils.l2_loss(alpha, mask)
    loss.backward()
    d = alpha.detach() - mask
    expect = (torch.sign(d) + 2 * d) / alpha.numel()
    assert torch.allclose(alpha.grad, expect, rtol=1e-5, atol=1e-9)
