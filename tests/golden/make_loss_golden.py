"""Generate tests/golden/loss_*.npz with the reference's OWN utils/loss_utils.py (pure torch, runs
on the CPU in the build container):   python tests/golden/make_loss_golden.py

Each case is what the trainers evaluate (trainers/train_static_bg.py:98-102): the gradient-mask
hook on the rendered image, l1_loss, ssim, the lambda mix, backward."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).resolve().parent
REFERENCE = "/root/reference"

LOSS_CASES = [
    dict(name="loss_rgb_masked", C=3, H=75, W=118, lam=0.2, mask=True),
    dict(name="loss_rgb_plain", C=3, H=64, W=96, lam=0.2, mask=False),
    dict(name="loss_small", C=3, H=9, W=13, lam=0.5, mask=True),        # smaller than the window
    dict(name="loss_alpha", C=1, H=40, W=33, lam=0.0, mask=False),      # l1_loss(obj_mask, render_alpha)
]


def loss_inputs(case: dict) -> dict:
    gen = torch.Generator().manual_seed(7000 + sum(map(ord, case["name"])))
    C, H, W = case["C"], case["H"], case["W"]
    gt = torch.rand(C, H, W, generator=gen)
    render = (gt + 0.15 * torch.randn(C, H, W, generator=gen)).clamp(0, 1)
    render[:, : H // 4, : W // 4] = gt[:, : H // 4, : W // 4]    # exact ties: sign(0) = 0 in the L1 gradient
    mask = (torch.rand(H, W, generator=gen) < 0.3).to(torch.int32)
    return {"render": render.contiguous(), "gt": gt.contiguous(), "mask": mask}


def run_reference(case: dict) -> dict:
    sys.path.insert(0, REFERENCE)
    from utils.loss_utils import l1_loss, ssim
    x = loss_inputs(case)
    leaf = x["render"].clone().requires_grad_(True)
    render = leaf * 1.0
    if case["mask"]:
        int_mask = x["mask"]
        render.register_hook(lambda grad: grad * (1 - int_mask))
    l1 = l1_loss(render, x["gt"])
    s = ssim(render, x["gt"])
    loss = (1.0 - case["lam"]) * l1 + case["lam"] * (1.0 - s)
    loss.backward()
    return {"l1": l1.detach().numpy(), "ssim": s.detach().numpy(), "loss": loss.detach().numpy(),
            "d_render": leaf.grad.numpy()}


def main():
    for case in LOSS_CASES:
        res = run_reference(case)
        np.savez_compressed(OUT / f"{case['name']}.npz", **res)
        print(case["name"], float(res["l1"]), float(res["ssim"]), float(res["loss"]))


if __name__ == "__main__":
    main()
