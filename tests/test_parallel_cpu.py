"""Host-side logic of the view-data-parallel step, on the CPU with gloo and world_size 2: the
flat gradient buffer layout, the single all-reduce, the view partitioning, and that a replicated
Adam step after the exchange leaves every rank with identical parameters."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from egogaussian_b200 import parallel
from oracle import cpu_oracle as orc


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        P, M = 257, 4                      # odd sizes: exercises the 16-byte padding of the views
        fg = parallel.FlatGrads(P, M, "cpu")
        g = torch.Generator().manual_seed(100 + rank)
        for name, v in fg.views.items():
            v.copy_(torch.randn(v.shape, generator=g))
        local = {k: v.clone() for k, v in fg.views.items()}
        parallel.allreduce_gradients(fg.flat)
        # replicated Adam on identical params with the reduced gradient
        p0 = torch.linspace(-1, 1, P * 3).reshape(P, 3).numpy()
        p1, m1, v1 = orc.adam_reference(p0, fg.views["means3D"].numpy(), np.zeros_like(p0), np.zeros_like(p0), 1.6e-4, 1)
        out[rank] = ({k: v.numpy() for k, v in local.items()}, {k: v.clone().numpy() for k, v in fg.views.items()}, p1)
    finally:
        dist.destroy_process_group()


def test_flat_gradient_allreduce_gloo_world2():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    (l0, r0, p0), (l1, r1, p1) = out[0], out[1]
    for k in l0:
        assert np.allclose(r0[k], l0[k] + l1[k], rtol=0, atol=1e-6), k
        assert np.array_equal(r0[k], r1[k]), "every rank holds the same reduced gradient"
    assert np.array_equal(p0, p1), "replicated Adam keeps ranks bit-identical (no broadcast needed)"


def test_flat_layout_and_partition():
    fg = parallel.FlatGrads(1000, 16, "cpu")
    assert fg.views["sh"].shape == (1000, 16, 3) and fg.views["rotations"].shape == (1000, 4)
    # 59 parameter floats per Gaussian at SH3 + means2D 3
    assert fg.flat.numel() >= 1000 * (59 + 3)
    for name, v in fg.views.items():
        assert v.data_ptr() % 16 == 0 and v.is_contiguous(), name
        v.fill_(1.0)
    assert float(fg.flat.sum()) == 1000 * 62
    assert parallel.views_for_rank(8, 1, 4) == [1, 5]
    assert sorted(sum((list(parallel.views_for_rank(8, r, 2)) for r in range(2)), [])) == list(range(8))
    # single process: the exchange is a no-op
    before = fg.flat.clone()
    parallel.allreduce_gradients(fg.flat)
    assert torch.equal(before, fg.flat)


def test_adam_segments_cover_the_parameter_groups_of_the_flat_layout():
    """FlatGrads.adam_segments: the reference's optimiser groups (scene/gaussian_model.py:189-197) as
    4-float-aligned, ascending, disjoint ranges of the flat buffer; means2D stays outside."""
    P, M = 257, 16
    fg = parallel.FlatGrads(P, M, "cpu")
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}
    segs = fg.adam_segments(lrs)
    assert [s[2] for s in segs] == [1.6e-4, 2.5e-3, 0.05, 5e-3, 1e-3]
    prev = 0
    for (b, e, la, lb, per, sp), off in zip(segs, fg.offsets):
        assert b == off and b >= prev and b % 4 == 0 and e % 4 == 0 and e > b
        prev = e
    sh = segs[1]
    assert sh[3] == 1.25e-4 and sh[4] == 3 * M and sh[5] == 3          # first 3 floats of a row: features_dc
    assert segs[-1][1] <= fg.offsets[-1]                                # means2D rides behind the last group
    assert fg.flat.numel() - fg.offsets[-1] >= 3 * P


def test_exchange_shards_tile_the_buffer():
    from egogaussian_b200 import cabi
    lib = cabi.load()
    for n, world in [(0, 1), (4, 8), (1000, 3), (62_000_000, 8), (59_000_004, 5)]:
        shard = lib.egs_exchange_shard_floats(n, world)
        assert shard % 4 == 0 and shard * world >= n and (shard - 4) * world < n or n == 0
    assert lib.egs_exchange_shard_floats(16, 0) == 0
    g = cabi.peer_group(2, 1, [16, 32], [64, 128], grads_multicast=0)
    assert g.world == 2 and g.rank == 1 and g.grads[1] == 32 and g.signals[0] == 64 and not g.grads_multicast


def test_factored_layout_and_view_batch_bookkeeping():
    """FlatGrads(factored=True): the SH gradient leaves the flat buffer (14 floats per Gaussian stay), its
    factors go to a slot of EGS_DRGB_SLOT_FLOATS floats (rows, padding to 16 bytes, camera centre) and the
    rebuilt [P,M,3] tensor is a plain local one; ViewBatch hands every local view its own slot of the gather
    buffer and lists all views' buffers in the global view order (host logic only: no kernel runs here)."""
    from egogaussian_b200 import cabi
    P, M = 1001, 16
    assert cabi.drgb_slot_floats(P) == (3 * P + 3) // 4 * 4 + 4 and cabi.drgb_slot_floats(P) % 4 == 0
    fg = parallel.FlatGrads(P, M, "cpu", factored=True)
    assert [n for n, _ in fg.layout] == ["means3D", "opacity", "scales", "rotations", "means2D"]
    assert fg.views["rgb"].shape == (P, 3) and fg.views["sh"].shape == (P, M, 3)
    assert fg.views["rgb"].data_ptr() == fg.rgb_slot.data_ptr() and fg.rgb_slot.numel() == cabi.drgb_slot_floats(P)
    out = fg.as_backward_out()
    assert out["sh"] is None and out["rgb"] is fg.views["rgb"]
    cov = parallel.FlatGrads(P, M, "cpu", factored=True, geometry="cov")
    assert "cov3D" in cov.views and "scales" not in cov.views and cov.views["cov3D"].shape == (P, 6)
    batch = parallel.ViewBatch(8, P, M, "cpu")
    assert batch.local_views == list(range(8)) and len(batch.layouts) == 8 and batch.result is batch.layouts[0]
    slot = cabi.drgb_slot_floats(P)
    for v, lay in enumerate(batch.layouts):
        assert lay.rgb_slot.data_ptr() == batch.gather.data_ptr() + 4 * v * slot
        assert batch._slot_ptrs[v] == lay.rgb_slot.data_ptr() and batch._view_flat[v] == lay.flat.data_ptr()
        assert batch.backward_args(v)["rgb_remote"] == [] and batch.backward_args(v)["out"]["rgb"] is lay.views["rgb"]
    with pytest.raises(ValueError):
        parallel.ViewBatch(9, P, M, "cpu")
