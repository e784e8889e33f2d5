"""Host-side logic of the view-data-parallel step, on the CPU with gloo and world_size 2: the
flat gradient buffer layout, the single all-reduce, the view partitioning, and that a replicated
Adam step after the exchange leaves every rank with identical parameters."""
import os
import socket

import numpy as np
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
