"""Generate golden vectors by running the UNMODIFIED reference rasterizer (oracle/_ref) on a B200.

Run on the GPU box:  python tests/golden/make_golden.py [--out gpurun_out/golden]
then copy the .npz files into tests/golden/ and commit them. The reference itself holds no
golden vectors or tests, so these outputs of the reference run on the
target hardware are what pins both the CPU oracle and the CUDA path.

Each case is a small seeded scene from egogaussian_b200.scenes (the same generator every test
uses), one file per (scene, geometry mode, colour mode). Stored per case: the inputs' recipe
(seed, sizes, modes -- not the tensors, they are regenerated), all forward outputs, the decoded
intermediates (radii, means2D, conics, rgb, sorted keys/values, tile ranges, n_contrib) and all
gradients for the fixed upstream weight images.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

CASES = [
    # name, P, W, H, sh_degree, max_sh_degree, seed, geom mode, colour mode, bg
    ("g0_sh3_scale", 1500, 80, 64, 3, 3, 11, "scale_rot", "sh", (0.0, 0.0, 0.0)),
    ("g1_sh0_cov", 1500, 80, 64, 0, 0, 12, "cov", "sh", (0.0, 0.0, 0.0)),
    ("g2_rgb_cov_bg", 1200, 70, 50, 0, 0, 13, "cov", "rgb", (0.3, 0.5, 0.7)),
    ("g3_sh1of3_scale", 1000, 64, 64, 1, 3, 14, "scale_rot", "sh", (0.0, 0.0, 0.0)),
    ("g4_sh2_scale_bg", 1000, 48, 80, 2, 2, 15, "scale_rot", "sh", (0.1, 0.2, 0.3)),
]


def case_inputs(case, device="cpu"):
    from egogaussian_b200 import scenes
    name, P, W, H, deg, maxdeg, seed, geom, colour, bg = case
    sc = scenes.make_scene(P, W, H, sh_degree=deg, seed=seed, bg=bg, max_sh_degree=maxdeg)
    # densify the picture a little: small images with few Gaussians would be mostly empty
    sc.scales = (sc.scales * 1.5).contiguous()
    wc, wd, wa = scenes.upstream_weights(W, H, seed)
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
    import torch
    kw = {k: (v.to(device).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
    ups = dict(dL_dcolor=wc.to(device).contiguous(), dL_ddepth=wd.to(device).contiguous(),
               dL_dalpha=wa.to(device).contiguous())
    return kw, ups


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "golden"))
    args = ap.parse_args()
    import torch
    from oracle import ref_dgr
    out_dir = Path(args.out)
    out_dir.mkdir(parents=True, exist_ok=True)
    for case in CASES:
        kw, ups = case_inputs(case, "cuda")
        fwd = ref_dgr.forward(**kw)
        dec = ref_dgr.decode_buffers(fwd)
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        grads = ref_dgr.backward(fwd, **bkw, **ups)
        torch.cuda.synchronize()
        blob = {"num_rendered": np.int64(fwd["num_rendered"])}
        for k in ("color", "depth", "alpha", "radii"):
            blob[k] = fwd[k].cpu().numpy()
        for k, v in dec.items():
            blob["i_" + k] = v.cpu().numpy()
        for k, v in grads.items():
            if v is not None:
                blob["g_" + k] = v.cpu().numpy()
        np.savez_compressed(out_dir / f"{case[0]}.npz", **blob)
        print(case[0], "R =", fwd["num_rendered"], "visible =", int((fwd["radii"] > 0).sum()))


if __name__ == "__main__":
    main()
