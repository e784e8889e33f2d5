"""GPU parity of egs_knn_mean_dist2 / simple_knn.distCUDA2 against golden
vectors of the reference's simple-knn module, against that module run live, and against the
brute-force restatement. Tolerance 1e-6 relative (same expression, same compiler: in practice the
results are bit-identical; asserted where the live reference is available)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from golden.make_knn_golden import KNN_CASES, knn_points
from test_oracle_golden import assert_knn_close

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"


@pytest.mark.parametrize("name", KNN_CASES)
def test_knn_against_reference_golden(name, built, cuda_device):
    from egogaussian_b200 import simple_knn
    gold = np.load(GOLDEN / f"{name}.npz")["mean_dist2"]
    got = simple_knn.distCUDA2(torch.from_numpy(knn_points(name)).to(cuda_device)).cpu().numpy()
    assert_knn_close(got, gold, name)


def test_knn_against_live_reference_at_sfm_size(built, cuda_device):
    from egogaussian_b200 import simple_knn
    from oracle import knn_ref
    if not knn_ref.reference_available():
        pytest.skip("oracle/_ref/ref_knn_C.so not built")
    gen = torch.Generator().manual_seed(31)
    pts = (torch.randn(120_000, 3, generator=gen) * torch.tensor([4.0, 1.0, 2.5])).to(cuda_device)
    ours = simple_knn.distCUDA2(pts)
    ref = knn_ref.reference()(pts)
    assert torch.equal(ours, ref)
    from egogaussian_b200 import cabi
    assert torch.equal(cabi.knn_mean_dist2(pts, exhaustive=True), ref)   # the sweep, forced at this size
    assert simple_knn.distCUDA2(pts[:0]).numel() == 0
    with pytest.raises(ValueError):
        simple_knn.distCUDA2(pts.cpu())


def _clouds():
    """Large clouds that stress the grid: uniform bulk, tight clusters with far outliers, a plane
    (one flat axis), and exact duplicates."""
    gen = torch.Generator().manual_seed(77)
    n = 220_000
    uniform = torch.rand(n, 3, generator=gen) * torch.tensor([8.0, 3.0, 5.0])
    centres = torch.randn(40, 3, generator=gen) * 5.0
    clustered = centres[torch.randint(0, 40, (n,), generator=gen)] + 0.05 * torch.randn(n, 3, generator=gen)
    clustered[:300] = torch.randn(300, 3, generator=gen) * 400.0           # outliers far beyond 3 sigma
    plane = torch.rand(n, 3, generator=gen)
    plane[:, 2] = 1.25
    dup = uniform.clone()
    dup[1::2] = dup[0::2]                                                    # every point twice
    return {"uniform": uniform, "clustered_outliers": clustered, "plane": plane, "duplicates": dup}


@pytest.mark.parametrize("name", ["uniform", "clustered_outliers", "plane", "duplicates"])
def test_knn_grid_equals_exhaustive_sweep(name, built, cuda_device):
    """From 32768 points on the statistic is searched through a grid; it must be the exact one."""
    from egogaussian_b200 import cabi
    pts = _clouds()[name].to(cuda_device)
    assert cabi.load().egs_knn_workspace_bytes(pts.shape[0]) > 0
    grid = cabi.knn_mean_dist2(pts)
    sweep = cabi.knn_mean_dist2(pts, exhaustive=True)
    assert torch.equal(grid, sweep), (name, (grid != sweep).sum().item())


def test_knn_grid_against_live_reference_at_1m(built, cuda_device):
    from egogaussian_b200 import simple_knn
    from oracle import knn_ref
    if not knn_ref.reference_available():
        pytest.skip("oracle/_ref/ref_knn_C.so not built")
    gen = torch.Generator().manual_seed(5)
    pts = (torch.randn(1_000_000, 3, generator=gen) * torch.tensor([3.0, 1.0, 2.0])).to(cuda_device)
    assert torch.equal(simple_knn.distCUDA2(pts), knn_ref.reference()(pts))


def test_knn_workspace_is_checked(built, cuda_device):
    from egogaussian_b200 import cabi
    lib = cabi.load()
    P = 200_000
    pts = torch.rand(P, 3, device=cuda_device)
    out = torch.empty(P, device=cuda_device)
    ws = torch.empty(1024, dtype=torch.uint8, device=cuda_device)
    assert lib.egs_knn_workspace_bytes(1000) == 0
    rc = lib.egs_knn_mean_dist2(P, pts.data_ptr(), out.data_ptr(), ws.data_ptr(), 1024, None)
    assert rc != 0 and b"workspace" in lib.egs_last_error()
