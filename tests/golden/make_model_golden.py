"""Generate tests/golden/model_*.npz by running the reference's OWN Python (scene/gaussian_model.py,
utils/general_utils.py, utils/geometry_utils.py) for the activation / covariance glue.

Runs in the build container (CPU, /root/reference mounted), not on the GPU box:
    python tests/golden/make_model_golden.py

The reference modules hard-code device="cuda" in a few torch.zeros calls and import three
packages that are absent here (plyfile, simple_knn, pytorch3d) for code that is not on this path.
Both are worked around from the outside -- empty stub modules and a torch.zeros wrapper that maps
"cuda" to "cpu" -- so the arithmetic that runs is the reference's, unmodified. The vectors pin
oracle/model_ref.py (CPU test) and the CUDA kernels (GPU test).
"""
from __future__ import annotations

import sys
import types
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).resolve().parent
REFERENCE = "/root/reference"

# name, P, M, scale_modifier, objects (fraction | explicit rows), which_object, accum_R?, during_training, mode
MODEL_CASES = [
    dict(name="model_plain", P=3000, M=16, mod=1.0, rot_cov=False),
    dict(name="model_plain_mod", P=1500, M=4, mod=0.6, rot_cov=False),
    dict(name="model_object_fixed", P=3000, M=16, mod=1.0, rot_cov=True, obj_frac=0.1, which=1, accum=True, train=False),
    dict(name="model_object_train", P=2500, M=16, mod=0.8, rot_cov=True, obj_frac=0.15, which=1, accum=True, train=True),
    dict(name="model_train_no_accum", P=1200, M=1, mod=1.0, rot_cov=True, obj_frac=0.2, which=1, accum=False, train=True),
    dict(name="model_all_rows", P=800, M=9, mod=1.0, rot_cov=True, obj_frac=0.3, which=None, accum=True, train=True),
    dict(name="model_no_match", P=900, M=4, mod=1.0, rot_cov=True, obj_frac=0.1, which=2, accum=True, train=True),
    dict(name="model_single_row0", P=700, M=4, mod=1.0, rot_cov=True, obj_rows=[0], which=1, accum=True, train=True),
    dict(name="model_single_row", P=700, M=4, mod=1.0, rot_cov=True, obj_rows=[123], which=1, accum=True, train=False),
]


def rotation_matrix(gen: torch.Generator) -> torch.Tensor:
    """A seeded proper rotation (QR of a normal matrix, sign-fixed)."""
    q, r = torch.linalg.qr(torch.randn(3, 3, generator=gen, dtype=torch.float64))
    q = q * torch.sign(torch.diagonal(r))
    if torch.det(q) < 0:
        q[:, 0] = -q[:, 0]
    return q.to(torch.float32).contiguous()


def model_inputs(case: dict) -> dict:
    """Seeded raw parameters; shared with the tests so that inputs need not be stored twice."""
    gen = torch.Generator().manual_seed(4000 + sum(map(ord, case["name"])))
    P, M = case["P"], case["M"]
    x = {
        "scaling_raw": torch.randn(P, 3, generator=gen) * 0.7 - 3.0,
        "rotation_raw": torch.randn(P, 4, generator=gen) * 1.5,
        "opacity_raw": torch.randn(P, 1, generator=gen) * 2.0,
        "features_dc": torch.randn(P, 1, 3, generator=gen),
        "features_rest": torch.randn(P, M - 1, 3, generator=gen) * 0.2,
        "g_cov": torch.randn(P, 6, generator=gen),
        "g_opacity": torch.randn(P, 1, generator=gen),
        "g_shs": torch.randn(P, M, 3, generator=gen),
    }
    is_object = torch.zeros(P, 1, dtype=torch.int32)
    if "obj_rows" in case:
        is_object[case["obj_rows"]] = 1
    elif case.get("obj_frac"):
        is_object[torch.rand(P, generator=gen) < case["obj_frac"]] = 1
        is_object[0] = 0
    x["is_object"] = is_object
    x["accum_R"] = rotation_matrix(gen)
    # a pose a little away from the identity, as 6-D rotation [3,2]
    x["rot6d"] = (rotation_matrix(gen)[:, :2] + 0.05 * torch.randn(3, 2, generator=gen)).contiguous()
    return x


def _import_reference():
    sys.path.insert(0, REFERENCE)
    for name in ("plyfile", "simple_knn", "simple_knn._C", "pytorch3d", "pytorch3d.transforms"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
    sys.modules["simple_knn._C"].distCUDA2 = None
    sys.modules["pytorch3d.transforms"].euler_angles_to_matrix = None
    real_zeros = torch.zeros

    def zeros_on_cpu(*a, **k):
        if str(k.get("device", "")).startswith("cuda"):
            k["device"] = "cpu"
        return real_zeros(*a, **k)

    torch.zeros = zeros_on_cpu
    from scene.gaussian_model import GaussianModel
    from utils.geometry_utils import ObjectMove
    return GaussianModel, ObjectMove


def run_reference(case: dict, GaussianModel, ObjectMove) -> dict:
    x = model_inputs(case)
    pc = GaussianModel(3)
    leaves = {}
    for k, attr in (("scaling_raw", "_scaling"), ("rotation_raw", "_rotation"), ("opacity_raw", "_opacity"),
                    ("features_dc", "_features_dc"), ("features_rest", "_features_rest")):
        leaves[k] = x[k].clone().requires_grad_(True)
        setattr(pc, attr, leaves[k])
    pc._is_object = x["is_object"]
    pc.trainable_object_move = ObjectMove()
    with torch.no_grad():
        pc.trainable_object_move.obj_rotation_6d.copy_(x["rot6d"])
    rot6d = pc.trainable_object_move.obj_rotation_6d
    if case["rot_cov"]:
        cov = pc.get_rotated_covariance(x["accum_R"] if case["accum"] else None, case["which"], case["train"], case["mod"])
    else:
        cov = pc.get_covariance(case["mod"])
    opacity, shs = pc.get_opacity, pc.get_features
    loss = (cov * x["g_cov"]).sum() + (opacity * x["g_opacity"]).sum() + (shs * x["g_shs"]).sum()
    loss.backward()
    out = {"cov3D": cov, "opacity": opacity, "shs": shs}
    out.update({"d_" + k: v.grad for k, v in leaves.items()})
    out["d_rot6d"] = rot6d.grad if rot6d.grad is not None else torch.zeros(3, 2)
    # get_features and its backward are exact copies; keep them out of the larger fixtures
    assert torch.equal(shs, torch.cat((x["features_dc"], x["features_rest"]), dim=1))
    assert torch.equal(out["d_features_dc"], x["g_shs"][:, :1]) and torch.equal(out["d_features_rest"], x["g_shs"][:, 1:])
    if case["P"] > 1500:
        for k in ("shs", "d_features_dc", "d_features_rest"):
            del out[k]
    return {k: v.detach().numpy() for k, v in out.items()}


def stats_inputs() -> dict:
    gen = torch.Generator().manual_seed(4999)
    P = 5000
    radii = torch.randint(0, 40, (P,), generator=gen, dtype=torch.int32)
    radii[torch.rand(P, generator=gen) < 0.3] = 0
    return {"radii": radii, "grad": torch.randn(P, 3, generator=gen) * 1e-3,
            "max_radii2D": torch.rand(P, generator=gen) * 30, "accum": torch.rand(P, 1, generator=gen),
            "denom": torch.randint(0, 50, (P, 1), generator=gen).float()}


def run_reference_stats(GaussianModel) -> dict:
    """Two iterations of the per-iteration densification bookkeeping:
    GaussianModel.add_densification_stats (scene/gaussian_model.py:735-740) plus the trainer's
    max_radii2D line (trainers/train_static_bg.py:128, restated here: it is not in a function)."""
    x = stats_inputs()
    pc = GaussianModel(3)
    pc.max_radii2D, pc.xyz_gradient_accum, pc.denom = x["max_radii2D"].clone(), x["accum"].clone(), x["denom"].clone()
    for it in range(2):
        vp = torch.zeros(x["grad"].shape)
        vp.grad = x["grad"] * (it + 1)
        vis = x["radii"] > 0
        pc.max_radii2D[vis] = torch.max(pc.max_radii2D[vis], x["radii"][vis])
        pc.add_densification_stats(vp, vis)
    return {"max_radii2D": pc.max_radii2D.numpy(), "accum": pc.xyz_gradient_accum.numpy(), "denom": pc.denom.numpy()}


def main():
    GaussianModel, ObjectMove = _import_reference()
    np.savez_compressed(OUT / "model_stats.npz", **run_reference_stats(GaussianModel))
    for case in MODEL_CASES:
        res = run_reference(case, GaussianModel, ObjectMove)
        np.savez_compressed(OUT / f"{case['name']}.npz", **res)
        print(case["name"], {k: v.shape for k, v in res.items() if k in ("cov3D", "d_rot6d")},
              "max|d_scaling|=%.3g" % np.abs(res["d_scaling_raw"]).max())


if __name__ == "__main__":
    main()
