"""Scale sweep (BASELINE.json configs[4]): fwd+bwd ms of this rasterizer and of the reference CUDA
rasterizer (oracle/_ref) for P = 100k .. 6M Gaussians at 1280x720, plus configs 1-3.

  python tools/sweep.py [--json gpurun_out/sweep.json] [--max-p 6000000]
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

from egogaussian_b200 import cabi  # noqa: E402
from oracle import ref_dgr  # noqa: E402
from parity_report import big_case, time_fn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=str(ROOT / "gpurun_out" / "sweep.json"))
    ap.add_argument("--max-p", type=int, default=6_000_000)
    a = ap.parse_args()
    cases = [("c1", 10_000, 0, 256, 256), ("c2", 100_000, 3, 1920, 1080)]
    for P in (100_000, 300_000, 1_000_000, 2_000_000, 4_000_000, 6_000_000):
        if P <= a.max_p:
            cases.append((f"c5_P{P}", P, 3, 1280, 720))
    out = []
    for name, P, deg, W, H in cases:
        kw, ups = big_case(name, P, deg, W, H, 3001, "scale_rot", "sh")
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        fwd = cabi.forward(**kw)
        rec = {"case": name, "P": P, "W": W, "H": H, "sh_degree": deg, "R": fwd.num_rendered}
        rec["our_fwd_ms"] = time_fn(lambda: cabi.forward(**kw))
        rec["our_bwd_ms"] = time_fn(lambda: cabi.backward(fwd, **bkw, **ups))
        if ref_dgr.available():
            rf = ref_dgr.forward(**kw)
            rec["ref_fwd_ms"] = time_fn(lambda: ref_dgr.forward(**kw))
            rec["ref_bwd_ms"] = time_fn(lambda: ref_dgr.backward(rf, **bkw, **ups))
            rec["speedup"] = (rec["ref_fwd_ms"] + rec["ref_bwd_ms"]) / (rec["our_fwd_ms"] + rec["our_bwd_ms"])
            del rf
        out.append(rec)
        print(json.dumps(rec), flush=True)
        del fwd, kw, ups
        torch.cuda.empty_cache()
    Path(a.json).parent.mkdir(parents=True, exist_ok=True)
    Path(a.json).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
