import sys, pathlib; sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch, time
from egogaussian_b200 import cabi
from oracle import knn_ref
ref = knn_ref.reference() if knn_ref.reference_available() else None
for P in (100_000, 150_000, 300_000, 1_000_000, 3_000_000):
    g = torch.Generator().manual_seed(1)
    pts = (torch.randn(P, 3, generator=g) * torch.tensor([3.0, 1.0, 2.0])).cuda()
    def t(f, n=3):
        f(); torch.cuda.synchronize()
        a = time.perf_counter()
        for _ in range(n): f()
        torch.cuda.synchronize()
        return (time.perf_counter() - a) / n * 1e3
    row = {"P": P, "grid_ms": round(t(lambda: cabi.knn_mean_dist2(pts)), 3)}
    if P <= 1_000_000: row["sweep_ms"] = round(t(lambda: cabi.knn_mean_dist2(pts, exhaustive=True), 1), 3)
    if ref is not None:
        row["ref_ms"] = round(t(lambda: ref(pts)), 3)
        row["equal"] = bool(torch.equal(cabi.knn_mean_dist2(pts), ref(pts)))
    print(row, flush=True)
