import sys, torch
sys.path.insert(0, '.')
from egogaussian_b200 import cabi
from egogaussian_b200.parallel import ViewBatch
dev = torch.device('cuda:0')
P, M, deg = 1_000_000, 16, 3
for NV in (1, 8):
    batch = ViewBatch(NV, P, M, dev)
    batch.gather.normal_()
    for lay in batch.layouts:
        lay.flat.normal_()
    means = torch.randn(P, 3, device=dev)
    for ph in (None, "a", "b"):
        cabi.debug_option("exchange_phases", ph)
        for _ in range(3):
            batch.exchange(means, deg)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(10):
            batch.exchange(means, deg)
        e1.record(); torch.cuda.synchronize()
        print(NV, ph, e0.elapsed_time(e1) / 10, "ms")
    cabi.debug_option("exchange_phases", None)
    del batch
