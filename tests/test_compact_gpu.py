"""Row compaction (egs_compact_plan / egs_compact_rows): identical to
`tensor[mask]`, which is what GaussianModel.prune_points / _prune_optimizer evaluate per tensor
(scene/gaussian_model.py:506-557). Integer/byte work: bit-exact."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("P,frac", [(1, 1.0), (1000, 0.5), (4097, 0.9), (250_003, 0.13), (70_000, 0.0), (5000, 1.0)])
def test_compaction_equals_boolean_indexing(P, frac, built, cuda_device):
    from egogaussian_b200 import cabi
    dev = cuda_device
    gen = torch.Generator().manual_seed(P)
    keep = (torch.rand(P, generator=gen) < frac).to(dev)
    # the tensors a prune touches: parameters, their Adam moments, bookkeeping (float, int, 1-D and 3-D)
    shapes = [(3,), (1, 3), (15, 3), (1,), (3,), (4,), (1,)]
    tensors = []
    for s in shapes:
        for _ in range(3):   # param, exp_avg, exp_avg_sq
            tensors.append(torch.randn((P,) + s, generator=gen).to(dev))
    tensors += [torch.randn(P, generator=gen).to(dev),                               # max_radii2D
                torch.randint(0, 5, (P, 1), generator=gen, dtype=torch.int32).to(dev),   # generation / is_object
                torch.randn(P, 1, generator=gen).to(dev), torch.randn(P, 1, generator=gen).to(dev)]
    for _ in range(12):      # more than one launch's worth of tensors
        tensors.append(torch.randn(P, 2, generator=gen).to(dev))
    assert len(tensors) > 32
    outs = cabi.compact_rows(keep, tensors)
    for t, o in zip(tensors, outs):
        ref = t[keep]
        assert o.shape == ref.shape and o.dtype == ref.dtype
        assert torch.equal(o, ref)
