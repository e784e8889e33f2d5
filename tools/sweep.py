"""Scale sweep (BASELINE.json configs[4]): fwd+bwd ms of this rasterizer and of the reference CUDA
rasterizer (oracle/_ref) for P = 100k .. 6M Gaussians at 1280x720, plus configs 1-3.

  python tools/sweep.py [--json gpurun_out/sweep.json] [--max-p 6000000]
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/sweep.py --json ...   (configs[4] at N GPUs)

Under torchrun every rank renders its own view of the scene (rank r = view r of the 8-view batch) and a step
also holds the gradient exchange (egs_exchange_sum_factored); times are the maximum over the ranks and the
reference columns (one GPU, no multi-GPU mode) are left out.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

from egogaussian_b200 import cabi  # noqa: E402
from oracle import ref_dgr  # noqa: E402
from parity_report import big_case, time_fn  # noqa: E402


def sweep_dp(a, rank, world):
    """configs[4] on `world` GPUs: fwd + bwd + exchange per step, one view per rank."""
    import os
    import torch.distributed as dist
    from egogaussian_b200 import parallel, scenes
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(device)
    dist.init_process_group("nccl", device_id=device)
    ex = parallel.PeerExchange(device)
    out = []
    for P in (100_000, 300_000, 1_000_000, 2_000_000, 4_000_000, 6_000_000):
        if P > a.max_p:
            continue
        W, H, deg = 1280, 720, 3
        sc = scenes.make_scene(P, W, H, sh_degree=deg, seed=3001, view=rank).to(device)
        cam = sc.camera
        kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
                  projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                  image_height=H, image_width=W, sh_degree=deg, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
        wc, wd, wa = [t.to(device) for t in scenes.upstream_weights(W, H, 3001 + rank)]
        ups = dict(dL_dcolor=wc, dL_ddepth=wd, dL_dalpha=wa)
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        batch = parallel.ViewBatch(world, P, 16, device, exchange=ex)
        scratch = torch.empty((cabi.load().egs_backward_scratch_bytes(P),), dtype=torch.uint8, device=device)
        state = {"hint": 0}

        def step(exchange=True):
            fwd = cabi.forward(num_rendered_hint=state["hint"], **kw)
            state["hint"] = int(fwd.num_rendered * 1.25) + 8192
            cabi.backward(fwd, scratch=scratch, **batch.backward_args(0), **bkw, **ups)
            if exchange:
                batch.exchange(kw["means3D"], deg)
            state["R"] = fwd.num_rendered

        def timed(fn, iters=10):
            for _ in range(3):
                fn()
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / iters], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        rec = {"case": f"c5_P{P}", "P": P, "W": W, "H": H, "sh_degree": deg, "n_gpus": world, "R_rank0": None}
        rec["step_ms"] = timed(step)
        rec["render_only_ms"] = timed(lambda: step(False))
        rec["R_rank0"] = state["R"]
        rec["views_per_s"] = world / rec["step_ms"] * 1e3
        assert not ex.failed()
        out.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
        del batch, sc, kw, ups, bkw, scratch
        torch.cuda.empty_cache()
    if rank == 0:
        Path(a.json).parent.mkdir(parents=True, exist_ok=True)
        Path(a.json).write_text(json.dumps(out, indent=1))
    dist.barrier()
    dist.destroy_process_group()


def main():
    import os
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=str(ROOT / "gpurun_out" / "sweep.json"))
    ap.add_argument("--max-p", type=int, default=6_000_000)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        return sweep_dp(a, int(os.environ["RANK"]), world)
    cases = [("c1", 10_000, 0, 256, 256), ("c2", 100_000, 3, 1920, 1080)]
    for P in (100_000, 300_000, 1_000_000, 2_000_000, 4_000_000, 6_000_000):
        if P <= a.max_p:
            cases.append((f"c5_P{P}", P, 3, 1280, 720))
    out = []
    for name, P, deg, W, H in cases:
        kw, ups = big_case(name, P, deg, W, H, 3001, "scale_rot", "sh")
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        fwd = cabi.forward(**kw)
        rec = {"case": name, "P": P, "W": W, "H": H, "sh_degree": deg, "R": fwd.num_rendered}
        rec["our_fwd_ms"] = time_fn(lambda: cabi.forward(**kw))
        rec["our_bwd_ms"] = time_fn(lambda: cabi.backward(fwd, **bkw, **ups))
        if ref_dgr.available():
            rf = ref_dgr.forward(**kw)
            rec["ref_fwd_ms"] = time_fn(lambda: ref_dgr.forward(**kw))
            rec["ref_bwd_ms"] = time_fn(lambda: ref_dgr.backward(rf, **bkw, **ups))
            rec["speedup"] = (rec["ref_fwd_ms"] + rec["ref_bwd_ms"]) / (rec["our_fwd_ms"] + rec["our_bwd_ms"])
            rec["speedup_fwd"] = rec["ref_fwd_ms"] / rec["our_fwd_ms"]
            del rf
        out.append(rec)
        print(json.dumps(rec), flush=True)
        del fwd, kw, ups
        torch.cuda.empty_cache()
    Path(a.json).parent.mkdir(parents=True, exist_ok=True)
    Path(a.json).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
