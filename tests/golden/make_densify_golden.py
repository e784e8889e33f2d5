"""Generate tests/golden/densify_*.npz by running the reference's OWN GaussianModel.densify_and_clone /
densify_and_split / densification_postfix / cat_tensors_to_optimizer / prune_points
(scene/gaussian_model.py:243-272, 506-677) with a real torch.optim.Adam holding state.

Runs in the build container (CPU, /root/reference mounted), not on the GPU box:
    python tests/golden/make_densify_golden.py

Worked around from the outside, as in make_model_golden.py: stub modules for the absent third-party
imports, device="cuda" literals mapped to the CPU, and torch.normal replaced by `mean + std * z` with z
drawn from a seeded generator and recorded -- so that the split's random offsets can be reproduced from
the same standard-normal noise. The arithmetic that runs is the reference's.
"""
from __future__ import annotations

import sys
import types
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).resolve().parent
sys.path.insert(0, str(OUT))
from make_model_golden import REFERENCE  # noqa: E402

NAMES = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation", "label"]
ATTRS = ["_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation", "_label"]
DENSIFY_CASES = [
    dict(name="densify_plain", P=900, M=4, curr_gen=None, which=None, split_prev_gen=True),
    dict(name="densify_object_gen", P=500, M=16, curr_gen=3, which=1, split_prev_gen=True),
    dict(name="densify_only_gen", P=700, M=1, curr_gen=2, which=None, split_prev_gen=False),
]
MAX_GRAD, EXTENT, PERCENT_DENSE = 0.0002, 4.0, 0.01


def densify_inputs(case: dict) -> dict:
    gen = torch.Generator().manual_seed(7000 + sum(map(ord, case["name"])))
    P, M = case["P"], case["M"]
    x = {"xyz": torch.randn(P, 3, generator=gen), "f_dc": torch.randn(P, 1, 3, generator=gen),
         "f_rest": torch.randn(P, M - 1, 3, generator=gen) * 0.2, "opacity": torch.randn(P, 1, generator=gen),
         # log scales around percent_dense * extent = 0.04 so that both sides of the size test are populated
         "scaling": torch.randn(P, 3, generator=gen) * 0.8 + float(np.log(0.03)),
         "rotation": torch.randn(P, 4, generator=gen), "label": torch.randn(P, 1, generator=gen)}
    for n in NAMES:   # Adam state of one earlier step
        x["m_" + n] = torch.randn(x[n].shape, generator=gen) * 1e-3
        x["v_" + n] = torch.rand(x[n].shape, generator=gen) * 1e-6
    x["generation"] = torch.randint(0, 4, (P, 1), generator=gen, dtype=torch.int32)
    x["is_object"] = (torch.rand(P, 1, generator=gen) < 0.3).to(torch.int32)
    x["grad_accum"] = torch.rand(P, 1, generator=gen) * 0.01
    x["denom"] = torch.randint(0, 30, (P, 1), generator=gen).float()      # zeros -> NaN -> 0
    x["grad_accum"][x["denom"] == 0] = 0.0
    x["max_radii2D"] = torch.rand(P, generator=gen) * 30
    return x


def _import_reference():
    sys.path.insert(0, REFERENCE)
    for name in ("plyfile", "simple_knn", "simple_knn._C", "pytorch3d", "pytorch3d.transforms"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
    sys.modules["simple_knn._C"].distCUDA2 = None
    sys.modules["pytorch3d.transforms"].euler_angles_to_matrix = None
    real = {n: getattr(torch, n) for n in ("zeros", "full")}

    def on_cpu(fn):
        def wrapped(*a, **k):
            if str(k.get("device", "")).startswith("cuda"):
                k["device"] = "cpu"
            return fn(*a, **k)
        return wrapped
    torch.zeros, torch.full = on_cpu(real["zeros"]), on_cpu(real["full"])
    from scene.gaussian_model import GaussianModel
    return GaussianModel


def run_reference(case: dict, GaussianModel) -> dict:
    x = densify_inputs(case)
    pc = GaussianModel(3)
    for n, attr in zip(NAMES, ATTRS):
        setattr(pc, attr, torch.nn.Parameter(x[n].clone()))
    pc._generation, pc._is_object = x["generation"].clone(), x["is_object"].clone()
    pc.percent_dense = PERCENT_DENSE
    pc.optimizer = torch.optim.Adam([{"params": [getattr(pc, a)], "lr": 1e-3, "name": n} for n, a in zip(NAMES, ATTRS)],
                                    lr=0.0, eps=1e-15)
    for n, attr in zip(NAMES, ATTRS):
        pc.optimizer.state[getattr(pc, attr)] = {"step": torch.tensor(1.0), "exp_avg": x["m_" + n].clone(),
                                                 "exp_avg_sq": x["v_" + n].clone()}
    pc.xyz_gradient_accum, pc.denom, pc.max_radii2D = x["grad_accum"].clone(), x["denom"].clone(), x["max_radii2D"].clone()
    noise_gen = torch.Generator().manual_seed(99)
    drawn = []
    real_normal = torch.normal

    def recorded_normal(mean=None, std=None, **k):
        z = torch.randn(std.shape, generator=noise_gen)
        drawn.append(z)
        return mean + std * z
    torch.normal = recorded_normal
    out = {}

    def snapshot(tag):
        for n, attr in zip(NAMES, ATTRS):
            p = getattr(pc, attr)
            st = pc.optimizer.state[p]
            out[f"{tag}_{n}"], out[f"{tag}_m_{n}"], out[f"{tag}_v_{n}"] = p.detach().clone(), st["exp_avg"].clone(), st["exp_avg_sq"].clone()
        out[f"{tag}_generation"], out[f"{tag}_is_object"] = pc._generation.clone(), pc._is_object.clone()
    try:
        grads = pc.xyz_gradient_accum / pc.denom
        grads[grads.isnan()] = 0.0
        pc.densify_and_clone(grads, MAX_GRAD, EXTENT, case["curr_gen"], which_object=case["which"])
        snapshot("clone")
        pc.densify_and_split(grads, MAX_GRAD, EXTENT, curr_gen=case["curr_gen"], split_prev_gen=case["split_prev_gen"],
                             which_object=case["which"])
        snapshot("split")
    finally:
        torch.normal = real_normal
    out["noise"] = drawn[0] if drawn else torch.zeros(0, 3)
    return {k: v.numpy() for k, v in out.items()}


def main():
    GaussianModel = _import_reference()
    for case in DENSIFY_CASES:
        res = run_reference(case, GaussianModel)
        np.savez_compressed(OUT / f"{case['name']}.npz", **res)
        print(case["name"], "P", case["P"], "after clone", res["clone_xyz"].shape[0], "after split", res["split_xyz"].shape[0],
              "noise rows", res["noise"].shape[0])


if __name__ == "__main__":
    main()
