"""CPU restatement and loader of the reference's KNN statistic (TEST INFRASTRUCTURE; only tests/
import it).

  mean_dist2()   brute-force numpy: mean of the three smallest squared distances to other points,
                 self excluded, duplicates included, missing neighbours = FLT_MAX
                 (submodules/simple-knn/simple_knn.cu:130-185; the box pruning there is exact)
  reference()    the UNMODIFIED reference module, compiled by oracle/build_ref.py into
                 oracle/_ref/ref_knn_C.so (needs a GPU)

Pinned by tests/golden/knn_*.npz, produced on a B200 by tests/golden/make_knn_golden.py with the
reference module (tests/test_oracle_golden.py::test_knn_restatement_matches_reference)."""
from __future__ import annotations

import importlib.util
from pathlib import Path

import numpy as np

_REF = Path(__file__).resolve().parent / "_ref" / "ref_knn_C.so"
FLT_MAX = np.float32(3.4028234663852886e38)


def mean_dist2(points: np.ndarray) -> np.ndarray:
    p = np.asarray(points, np.float32)
    n = p.shape[0]
    out = np.empty(n, np.float32)
    for start in range(0, n, 2048):
        q = p[start:start + 2048]
        d = q[:, None, :] - p[None, :, :]
        d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
        d2[np.arange(q.shape[0]), start + np.arange(q.shape[0])] = FLT_MAX
        if n < 4:
            d2 = np.concatenate([d2, np.full((q.shape[0], 3), FLT_MAX, np.float32)], axis=1)
        best = np.sort(np.partition(d2, 2, axis=1)[:, :3], axis=1)
        with np.errstate(over="ignore"):
            out[start:start + 2048] = (best[:, 0] + best[:, 1] + best[:, 2]) / np.float32(3.0)
    return out


def reference_available() -> bool:
    return _REF.exists()


def reference():
    spec = importlib.util.spec_from_file_location("ref_knn_C", _REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.distCUDA2
