"""Shared helpers for the parity tests: run the CUDA path through the C-ABI, compare results."""
from __future__ import annotations

import sys
from pathlib import Path
from typing import Dict, Optional

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

# tolerances stated by BASELINE.json `north_star`
IMAGE_ATOL = 1e-5       # colour / depth / alpha, absolute, fp32
GRAD_RTOL = 1e-4        # parameter gradients, relative to the tensor's max magnitude


def to_np(x):
    if x is None:
        return None
    if hasattr(x, "detach"):
        return x.detach().cpu().numpy()
    return np.asarray(x)


def run_cuda(kw: Dict, ups: Optional[Dict] = None, debug: bool = False):
    """Forward (+ backward) of our CUDA path through the C-ABI (egogaussian_b200.cabi)."""
    from egogaussian_b200 import cabi
    fwd = cabi.forward(debug=debug, **kw)
    grads = None
    if ups is not None:
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        grads = cabi.backward(fwd, debug=debug, dL_dcolor=ups["dL_dcolor"], dL_ddepth=ups["dL_ddepth"],
                              dL_dalpha=ups["dL_dalpha"], **bkw)
    return fwd, grads


def grad_rel_err(a, b) -> float:
    """max |a-b| / max |b|  (per-tensor, L-inf normalised: SURVEY.md 8c pass criteria)."""
    a, b = to_np(a).astype(np.float64), to_np(b).astype(np.float64)
    scale = np.abs(b).max()
    if scale == 0:
        return float(np.abs(a).max())
    return float(np.abs(a - b).max() / scale)


def keys_from(tile_ranges: np.ndarray, values: np.ndarray, depths: np.ndarray) -> np.ndarray:
    """Rebuild the 64-bit sort keys (tile << 32 | depth bits) from ranges, values and depths."""
    keys = np.zeros(values.shape[0], np.uint64)
    dbits = depths.astype(np.float32).view(np.uint32).astype(np.uint64)
    for t, (a, b) in enumerate(tile_ranges):
        if b > a:
            keys[a:b] = (np.uint64(t) << np.uint64(32)) | dbits[values[a:b]]
    return keys


# BASELINE.json configs[0..2] as (name, P, SH degree, W, H, seed, geometry mode, colour mode); seeds
# follow SURVEY.md 8d (1000 * config id + view id). EgoGaussian's production mode is cov3D_precomp
# with SH degree 0 (train.py:49, GaussianModel(sh_degree=0)); the label render uses colors_precomp.
BASELINE_CASES = [
    ("c1_sh0_scale", 10_000, 0, 256, 256, 1001, "scale_rot", "sh"),
    ("c1_sh0_cov", 10_000, 0, 256, 256, 1001, "cov", "sh"),
    ("c2_sh3_scale", 100_000, 3, 1920, 1080, 2001, "scale_rot", "sh"),
    ("c3_sh3_scale", 1_000_000, 3, 1280, 720, 3001, "scale_rot", "sh"),
    ("c3_sh0_cov", 1_000_000, 0, 1280, 720, 3001, "cov", "sh"),
    ("c3_rgb_cov", 1_000_000, 0, 1280, 720, 3001, "cov", "rgb"),
]


def baseline_case(P, deg, W, H, seed, geom, colour, device="cuda"):
    """Inputs (forward kwargs, upstream gradients) of one BASELINE-size case on `device`."""
    import torch
    from egogaussian_b200 import scenes
    sc = scenes.make_scene(P, W, H, sh_degree=deg, seed=seed)
    cam = sc.camera
    kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
              projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
              image_height=H, image_width=W, sh_degree=deg)
    if colour == "sh":
        kw["shs"] = sc.shs
    else:
        kw["colors_precomp"] = sc.colors_precomp
    if geom == "scale_rot":
        kw["scales"], kw["rotations"] = sc.scales, sc.rotations
    else:
        kw["cov3D_precomp"] = sc.cov3D()
    wc, wd, wa = scenes.upstream_weights(W, H, seed)
    kw = {k: (v.to(device).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
    ups = dict(dL_dcolor=wc.to(device), dL_ddepth=wd.to(device), dL_dalpha=wa.to(device))
    return kw, ups
