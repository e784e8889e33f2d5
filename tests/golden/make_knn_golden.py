"""Generate tests/golden/knn_*.npz with the UNMODIFIED reference KNN module (oracle/_ref/ref_knn_C.so,
built from /root/reference/submodules/simple-knn by oracle/build_ref.py). Needs a GPU:
    gpurun -- 'python tests/golden/make_knn_golden.py gpurun_out/golden'
and the files are then copied into tests/golden/."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

KNN_CASES = ["knn_uniform", "knn_clustered_dups", "knn_three", "knn_two"]


def knn_points(name: str) -> np.ndarray:
    gen = np.random.default_rng(9000 + sum(map(ord, name)))
    if name == "knn_uniform":
        return gen.uniform(-3, 3, (5003, 3)).astype(np.float32)
    if name == "knn_clustered_dups":       # tight clusters, exact duplicates, a few far outliers
        c = gen.normal(0, 2.0, (40, 3))
        p = (c[gen.integers(0, 40, 3000)] + gen.normal(0, 0.01, (3000, 3))).astype(np.float32)
        p[100:140] = p[60:100]
        p[:5] *= 500
        return p
    if name == "knn_three":
        return np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0]], np.float32)
    return np.array([[0.5, 0.25, 0], [1, -1, 3]], np.float32)


def main():
    from oracle import knn_ref
    out = Path(sys.argv[1] if len(sys.argv) > 1 else ROOT / "tests" / "golden")
    out.mkdir(parents=True, exist_ok=True)
    dist = knn_ref.reference()
    for name in KNN_CASES:
        p = knn_points(name)
        d = dist(torch.from_numpy(p).cuda()).cpu().numpy()
        np.savez_compressed(out / f"{name}.npz", mean_dist2=d)
        print(name, p.shape, float(d[np.isfinite(d)].mean()) if np.isfinite(d).any() else d)


if __name__ == "__main__":
    main()
