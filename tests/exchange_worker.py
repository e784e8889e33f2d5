"""torchrun worker of tests/test_exchange_gpu.py: the peer-memory exchange between real GPUs against
NCCL's all-reduce and the single-GPU Adam kernel. Also prints timings (GB/s of the flat buffer)."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main():
    from egogaussian_b200.diff_gaussian_rasterization import _C
    from egogaussian_b200.parallel import FlatGrads, PeerExchange
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(device)
    dist.init_process_group("nccl", device_id=device)
    for multicast in (False, True):
        run(rank, world, device, multicast)
        run_factored(rank, world, device, multicast)
    if rank == 0:
        print("exchange worker ok", flush=True)
    dist.barrier()
    dist.destroy_process_group()


def run(rank, world, device, multicast):
    from egogaussian_b200.diff_gaussian_rasterization import _C
    from egogaussian_b200.parallel import FlatGrads, PeerExchange
    ex = PeerExchange(device, multicast=multicast)
    P, M = int(os.environ.get("EGS_EXCHANGE_P", "200000")), 16
    layout = FlatGrads(P, M, device, allocate=ex.allocate)
    n = layout.flat.numel()
    through_switch = ex.multicast_pointer(layout.flat) != 0
    if multicast and not through_switch:
        if rank == 0:
            print("no multicast mapping on this box: switch path not exercised", flush=True)
        return
    gen = torch.Generator().manual_seed(100 + rank)

    # ---- sum against NCCL (two ranks: a + b in either order; more: within rounding) ----
    for it in range(3):
        layout.flat.copy_(torch.randn(n, generator=gen).to(device))
        ref = layout.flat.clone()
        dist.all_reduce(ref)
        ex.sum(layout.flat)
        assert not ex.failed()
        if world == 2:
            assert torch.equal(layout.flat, ref), (it, (layout.flat != ref).sum().item())
        else:
            assert (layout.flat - ref).abs().max().item() <= 1e-5 * ref.abs().max().item()
        gathered = [torch.empty_like(layout.flat) for _ in range(world)]
        dist.all_gather(gathered, layout.flat)
        assert all(torch.equal(t, gathered[0]) for t in gathered)   # the same bits on every rank

    # ---- fused update against all-reduce + egs_adam_step on the full buffer ----
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}
    segs = layout.adam_segments(lrs)
    params = ex.allocate(n)
    start = torch.randn(n, generator=torch.Generator().manual_seed(7)).to(device)   # same on all ranks
    params.copy_(start)
    m_shard, v_shard = ex.moment_shards(n)
    ref_p, ref_m, ref_v = start.clone(), torch.zeros_like(start), torch.zeros_like(start)
    torch.cuda.synchronize()
    dist.barrier()
    for step in (1, 2, 3):
        layout.flat.copy_(torch.randn(n, generator=gen).to(device))
        summed = layout.flat.clone()
        dist.all_reduce(summed)
        ex_in = layout.flat.clone()
        # reference on the gradient the exchange itself forms (rank order), taken from a sum exchange
        ex.sum(layout.flat)
        summed_ro = layout.flat.clone()
        layout.flat.copy_(ex_in)
        torch.cuda.synchronize()
        dist.barrier()
        ps = [ref_p[b:e] for (b, e, *_r) in segs]
        _C.adam_step(ps, [summed_ro[b:e] for (b, e, *_r) in segs], [ref_m[b:e] for (b, e, *_r) in segs],
                     [ref_v[b:e] for (b, e, *_r) in segs], [s[2] for s in segs], [step] * len(segs), 0.9, 0.999, 1e-15)
        ex.adam_step(layout.flat, params, m_shard, v_shard, segs, step)
        assert not ex.failed()
        assert torch.equal(params, ref_p), (step, (params != ref_p).sum().item())
        tail_b = segs[-1][1]
        assert torch.equal(layout.flat[tail_b:], summed_ro[tail_b:])
        assert (summed_ro - summed).abs().max().item() <= 1e-5 * summed.abs().max().item()

    # ---- timing ----
    def timed(fn, iters=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record(); torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / iters], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()
    step_box = [3]

    def fused():
        step_box[0] += 1
        ex.adam_step(layout.flat, params, m_shard, v_shard, segs, step_box[0])
    t_nccl = timed(lambda: dist.all_reduce(layout.flat))
    t_peer = timed(lambda: ex.sum(layout.flat))
    t_fused = timed(fused)
    assert not ex.failed()
    if rank == 0:
        mb = n * 4 / 1e6
        print(f"exchange timing world={world} buffer={mb:.1f} MB {'switch multicast' if through_switch else 'peer loads/stores'}: "
              f"nccl all_reduce {t_nccl:.3f} ms, sum {t_peer:.3f} ms, sum+adam {t_fused:.3f} ms", flush=True)


def run_factored(rank, world, device, multicast):
    """The factored SH-gradient exchange between real GPUs: rank r renders view r; the rebuilt dL/dsh must
    equal NCCL's all-reduce of the expanded tensors of an ordinary backward within 1e-4 of its maximum,
    and be the same bits on every rank."""
    from egogaussian_b200 import cabi, scenes
    from egogaussian_b200.parallel import FlatGrads, PeerExchange, ViewBatch
    ex = PeerExchange(device, multicast=multicast)
    P, deg, M, W, H = int(os.environ.get("EGS_EXCHANGE_P", "200000")), 3, 16, 640, 360
    sc = scenes.make_scene(P, W, H, sh_degree=deg, seed=71, view=rank).to(device)
    cam = sc.camera
    kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
              projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
              image_height=H, image_width=W, sh_degree=deg, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    wc, wd, wa = [t.to(device) for t in scenes.upstream_weights(W, H, 71 + rank)]
    ups = dict(dL_dcolor=wc, dL_ddepth=wd, dL_dalpha=wa)
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    fwd = cabi.forward(**kw)
    full = cabi.backward(fwd, **bkw, **ups)
    batch = ViewBatch(world, P, M, device, exchange=ex)
    lay = batch.result
    if multicast and not ex.multicast_pointer(lay.flat):
        return
    cabi.backward(fwd, **batch.backward_args(0), **bkw, **ups)
    ref_sh = full["sh"].clone()
    dist.all_reduce(ref_sh)
    ref_small = {k: full[k].clone() for k in ("means3D", "opacity", "scales", "rotations", "means2D")}
    for t in ref_small.values():
        dist.all_reduce(t)
    batch.exchange(sc.means3D, deg)
    assert not ex.failed()
    got = lay.views["sh"]
    scale = ref_sh.abs().max().item()
    # (the expanded rows come from another backward run than the factors: atomic order differs, ~1e-6)
    assert (got - ref_sh).abs().max().item() <= 1e-4 * scale   # north_star's gradient tolerance
    for k, t in ref_small.items():
        assert (lay.views[k] - t.view_as(lay.views[k])).abs().max().item() <= 1e-4 * max(t.abs().max().item(), 1e-30), k
    gathered = [torch.empty_like(got) for _ in range(world)]
    dist.all_gather(gathered, got.contiguous())
    if not multicast:   # rank-order sums: the same bits everywhere
        assert all(torch.equal(t, gathered[0]) for t in gathered)

    def timed(fn, iters=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record(); torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / iters], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()
    t_f = timed(lambda: batch.exchange(sc.means3D, deg))
    assert not ex.failed()
    if rank == 0:
        print(f"factored exchange world={world} P={P} {'switch multicast' if multicast else 'peer loads/stores'}: "
              f"{t_f:.3f} ms ({lay.flat.numel() * 4 / 1e6 + P * 12 / 1e6:.1f} MB per rank instead of "
              f"{(lay.flat.numel() + 3 * M * P) * 4 / 1e6:.1f})", flush=True)


if __name__ == "__main__":
    main()
