"""Per-rank breakdown of the view-data-parallel step (torchrun): device time of forward+backward and of
the gradient exchange on every rank, gathered on rank 0. The fastest rank's exchange time contains its
wait for the slowest one; the slowest rank's is the transfer alone.

  python -m torch.distributed.run --nproc-per-node N tools/dp_breakdown.py [--steps 30]
"""
import argparse
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    import bench
    from egogaussian_b200 import cabi, parallel
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--full", action="store_true", help="exchange the expanded SH gradient (egs_exchange_sum)")
    ap.add_argument("--no-multicast", action="store_true", help="peer loads and stores even where the switch could reduce")
    x = ap.parse_args()
    sys.argv = sys.argv[:1]
    a = bench.parse()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(device)
    dist.init_process_group("nccl", device_id=device)
    kw, ups, host, cpu_kw, weights = bench.build_inputs(a, rank, device)
    P, M = a.P, (a.sh_degree + 1) ** 2
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    ex = parallel.PeerExchange(device, multicast=False if x.no_multicast else None)
    batch = None if x.full else parallel.ViewBatch(world, P, M, device, exchange=ex)
    flat = parallel.FlatGrads(P, M, device, allocate=ex.allocate) if x.full else batch.result
    bw = dict(out=flat.as_backward_out()) if x.full else batch.backward_args(0)
    scratch = torch.empty((cabi.load().egs_backward_scratch_bytes(P),), dtype=torch.uint8, device=device)
    hint = 0
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(x.steps + 5)]
    for i in range(x.steps + 5):
        ev[i][0].record()
        fwd = cabi.forward(num_rendered_hint=hint, **kw)
        hint = int(fwd.num_rendered * 1.25) + 8192
        cabi.backward(fwd, scratch=scratch, **bw, **bkw, **ups)
        ev[i][1].record()
        if x.full:
            ex.sum(flat.flat)
        else:
            batch.exchange(kw["means3D"], a.sh_degree)
        ev[i][2].record()
    torch.cuda.synchronize()
    # the exchange alone: every rank enters at the same time (barrier), so no waiting for a slower render
    alone = []
    for i in range(10):
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if x.full:
            ex.sum(flat.flat)
        else:
            batch.exchange(kw["means3D"], a.sh_degree)
        e1.record()
        torch.cuda.synchronize()
        alone.append(e0.elapsed_time(e1))
    alone_ms = sorted(alone)[len(alone) // 2]
    phase_ms = {}
    if not x.full:
        for name in ("a", "b"):
            cabi.debug_option("exchange_phases", name)
            ts = []
            for i in range(8):
                torch.cuda.synchronize()
                dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                batch.exchange(kw["means3D"], a.sh_degree)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            phase_ms[name] = sorted(ts)[len(ts) // 2]
        cabi.debug_option("exchange_phases", None)
        if rank == 0:
            print("exchange phases alone on rank 0: slice sum of the small gradients %.3f ms, SH expansion %.3f ms"
                  % (phase_ms["a"], phase_ms["b"]))
    comp = sum(ev[i][0].elapsed_time(ev[i][1]) for i in range(5, x.steps + 5)) / x.steps
    exch = sum(ev[i][1].elapsed_time(ev[i][2]) for i in range(5, x.steps + 5)) / x.steps
    step = ev[5][0].elapsed_time(ev[x.steps + 4][2]) / x.steps
    t = torch.tensor([comp, exch, step, float(fwd.num_rendered), alone_ms], device=device, dtype=torch.float64)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    if rank == 0:
        print("exchange:", "egs_exchange_sum (expanded SH rows)" if x.full else "egs_exchange_sum_factored, small gradients by " + batch.path)
        for r, o in enumerate(out):
            print(f"  rank {r}: fwd+bwd {o[0]:.3f} ms, exchange (incl. waiting) {o[1]:.3f} ms, step {o[2]:.3f} ms, "
                  f"R {int(o[3])}, exchange alone {o[4]:.3f} ms")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
