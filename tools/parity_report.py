"""GPU-box report: our CUDA path vs the live reference (oracle/_ref) vs the CPU oracle.

  python tools/parity_report.py [--big] [--json gpurun_out/parity.json]

Prints, per case, bit-exactness of radii / sorted keys / values / ranges, image and gradient
errors, and CUDA-event timings of both rasterizers. Development tool; the pass/fail versions of
these checks live in tests/test_parity_gpu.py.
"""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from egogaussian_b200 import cabi, scenes  # noqa: E402
from helpers import BASELINE_CASES, baseline_case, grad_rel_err, run_cuda, to_np  # noqa: E402
from oracle import cpu_oracle, ref_dgr  # noqa: E402
from golden.make_golden import CASES, case_inputs  # noqa: E402


def big_case(name, P, deg, W, H, seed, geom, colour):
    return baseline_case(P, deg, W, H, seed, geom, colour, "cuda")


def time_fn(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def compare(name, kw, ups, with_cpu, timing):
    rep = {"case": name, "P": int(kw["means3D"].shape[0])}
    fwd, grads = run_cuda(kw, ups)
    torch.cuda.synchronize()
    rep["R"] = fwd.num_rendered
    gv, bv, iv = fwd.geom_view(), fwd.binning_view(), fwd.image_view()
    if ref_dgr.available():
        rf = ref_dgr.forward(**kw)
        dec = ref_dgr.decode_buffers(rf)
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        rg = ref_dgr.backward(rf, **bkw, **ups)
        torch.cuda.synchronize()
        vis = rf["radii"] > 0
        rep["ref_R"] = rf["num_rendered"]
        rep["radii_mismatch"] = int((fwd.radii != rf["radii"]).sum())
        rep["vis"] = int(vis.sum())
        if rep["radii_mismatch"] == 0:
            rep["means2D_bitdiff"] = int((gv["means2D"][vis].view(torch.int32) != dec["means2D"][vis].view(torch.int32)).sum())
            rep["depths_bitdiff"] = int((gv["depths"][vis].view(torch.int32) != dec["depths"][vis].view(torch.int32)).sum())
            rep["conic_bitdiff"] = int((gv["conic_opacity"][vis].view(torch.int32) != dec["conic_opacity"][vis].view(torch.int32)).sum())
            if kw.get("colors_precomp") is None:   # with precomputed colours the reference never writes geom.rgb
                rep["rgb_maxabs"] = float((gv["rgb_depth"][vis][:, :3] - dec["rgb"][vis]).abs().max())
                rep["rgb_bitdiff"] = int((gv["rgb_depth"][vis][:, :3].contiguous().view(torch.int32) != dec["rgb"][vis].contiguous().view(torch.int32)).sum())
            rep["tiles_mismatch"] = int((gv["tiles_touched"] != dec["tiles_touched"]).sum())
        if fwd.num_rendered == rf["num_rendered"]:
            rep["keys_mismatch"] = int((bv["keys"] != dec["keys"]).sum())
            rep["values_mismatch"] = int((bv["values"] != dec["values"]).sum())
            rep["ranges_mismatch"] = int((iv["ranges"] != dec["ranges"]).sum())
            rep["n_contrib_mismatch"] = int((iv["n_contrib"] != dec["n_contrib"]).sum())
        for k in ("color", "depth", "alpha"):
            d = (getattr(fwd, k) - rf[k]).abs()
            rep[f"{k}_maxabs"] = float(d.max())
            rep[f"{k}_n_gt_1e-5"] = int((d > 1e-5).sum())
        for k, v in rg.items():
            if v is not None and grads.get(k) is not None:
                rep[f"g_{k}_rel"] = grad_rel_err(grads[k].reshape(v.shape), v)
        # the reference's own run-to-run atomic noise
        rg2 = ref_dgr.backward(rf, **bkw, **ups)
        rep["ref_self_noise_means3D"] = grad_rel_err(rg2["means3D"], rg["means3D"])
        for k, v in rg.items():   # the same figure for every tensor: what two runs of the reference differ by
            if v is not None and rg2.get(k) is not None and k != "means3D":
                rep[f"ref_self_noise_{k}"] = grad_rel_err(rg2[k], v)
        if timing:
            rep["ref_fwd_ms"] = time_fn(lambda: ref_dgr.forward(**kw))
            rep["ref_bwd_ms"] = time_fn(lambda: ref_dgr.backward(rf, **bkw, **ups))
    if timing:
        rep["our_fwd_ms"] = time_fn(lambda: cabi.forward(**kw))
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        rep["our_bwd_ms"] = time_fn(lambda: cabi.backward(fwd, **bkw, **ups))
    if with_cpu:
        ckw = {k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
        t0 = time.time()
        of = cpu_oracle.render_forward(**ckw)
        bk = {k: v for k, v in ckw.items() if k not in ("opacities",)}
        og = cpu_oracle.render_backward(of, **bk, **{k: v.cpu() for k, v in ups.items()})
        rep["cpu_s"] = time.time() - t0
        rep["cpu_radii_mismatch"] = int((to_np(fwd.radii) != of["radii"]).sum())
        for k in ("color", "depth", "alpha"):
            rep[f"cpu_{k}_maxabs"] = float(np.abs(to_np(getattr(fwd, k)) - of[k]).max())
        for k, v in og.items():
            if v is not None and grads.get(k) is not None:
                rep[f"cpu_g_{k}_rel"] = grad_rel_err(grads[k].reshape(v.shape), v)
    return rep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--big", action="store_true")
    ap.add_argument("--json", default=str(ROOT / "gpurun_out" / "parity.json"))
    a = ap.parse_args()
    reports = []
    for case in CASES:
        kw, ups = case_inputs(case, "cuda")
        reports.append(compare(case[0], kw, ups, with_cpu=True, timing=False))
        print(json.dumps(reports[-1]), flush=True)
    big = BASELINE_CASES if a.big else BASELINE_CASES[:2]
    for name, P, deg, W, H, seed, geom, colour in big:
        kw, ups = big_case(name, P, deg, W, H, seed, geom, colour)
        reports.append(compare(name, kw, ups, with_cpu=(P <= 100_000), timing=True))
        print(json.dumps(reports[-1]), flush=True)
    Path(a.json).parent.mkdir(parents=True, exist_ok=True)
    Path(a.json).write_text(json.dumps(reports, indent=1))


if __name__ == "__main__":
    main()
