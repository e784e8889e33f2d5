"""Minimal fwd+bwd loop of the bench workload for ncu captures (no timing, no threads).

  python tools/profile_step.py [--steps 2] [--warmup 1] [--P 1000000] [--adam]

Used as the command under `ncu` (see /opt/skills/guides/B200_PROFILING.md): first plain, then
`ncu --metrics gpu__time_duration.sum ...` for the launch list, then `ncu --set full -k regex:...`
for the top kernels. Numbers printed under ncu are never bench values.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--P", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=1280)
    ap.add_argument("--height", type=int, default=720)
    ap.add_argument("--sh-degree", type=int, default=3)
    ap.add_argument("--geom", default="scale_rot")
    ap.add_argument("--adam", action="store_true")
    ap.add_argument("--trainer", action="store_true", help="also run two whole training iterations from raw parameters")
    a = ap.parse_args()
    import torch
    import bench
    from egogaussian_b200 import cabi
    dev = torch.device("cuda", 0)
    kw, ups, _host, _cpu, _w = bench.build_inputs(a, None, dev)
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    for _ in range(a.warmup + a.steps):
        fwd = cabi.forward(**kw)
        g = cabi.backward(fwd, **bkw, **ups)
    if a.adam:
        from egogaussian_b200.diff_gaussian_rasterization import _C
        names = [k for k in ("means3D", "sh", "opacity", "scales", "rotations") if g.get(k) is not None]
        params = [torch.randn_like(g[n]) for n in names]
        m = [torch.zeros_like(p) for p in params]
        v = [torch.zeros_like(p) for p in params]
        _C.adam_step(params, [g[n] for n in names], m, v, [1e-3] * len(names), [1] * len(names), 0.9, 0.999, 1e-15)
    if a.trainer:
        # one whole training iteration from the raw parameters: render() in raw-parameter mode, the
        # fused image loss, backward, densification bookkeeping, fused Adam
        from egogaussian_b200 import gaussian_renderer as gr, loss_utils
        from egogaussian_b200.diff_gaussian_rasterization import _C
        pc = bench._TrainerModel(a, dev)
        pc.max_radii2D = torch.zeros(a.P, device=dev)
        pc.xyz_gradient_accum = torch.zeros(a.P, 1, device=dev)
        pc.denom = torch.zeros(a.P, 1, device=dev)
        gt_img, int_mask = bench.trainer_targets(a, dev)
        params = pc.params()
        m = [torch.zeros_like(p) for p in params]
        v = [torch.zeros_like(p) for p in params]
        for it in range(2):
            pkg = gr.render(pc, pc, bench._Pipe(), pc.bg)
            loss_utils.image_loss(pkg["render"], gt_img, 0.2, grad_mask=int_mask)[0].backward()
            gr.add_densification_stats(pc, pkg["viewspace_points"], pkg["radii"])
            with torch.no_grad():
                _C.adam_step(params, [p.grad for p in params], m, v, pc.lrs, [it + 1] * len(params), 0.9, 0.999, 1e-15)
            for p in params:
                p.grad = None
    torch.cuda.synchronize()
    print("profile_step ok: R =", fwd.num_rendered)


if __name__ == "__main__":
    main()
