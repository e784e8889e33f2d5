"""bench.py -- fwd+bwd render iterations/s of the rasterizer hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--P ...]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one forward + one backward of the rasterizer over one camera view of the synthetic
1M-Gaussian, SH-degree-3, 1280x720 scene (BASELINE.json configs[2]), i.e. one
training iteration's render work. Rank r renders view r of the 8-view batch (configs[3]); with
N > 1 the step also contains the data-parallel exchange (one NCCL all-reduce of the flat
gradient buffer). Weak scaling: per-GPU work is fixed, `value` counts the views of all ranks.

JSON keys (one line, rank 0):
  value      device-resident inputs, C-ABI calls, CUDA events, max over ranks
  e2e        same metric through the public Python API (GaussianRasterizer + autograd) with the
             step's host inputs (camera matrices, upstream-gradient images) copied from pinned
             memory and the scalar loss read back, all inside the timed region
  roofline   dominant kernel of the timed region: algorithmic bytes (SURVEY.md 8d formulas,
             DESIGN.md) / its mean duration from the library's built-in CUDA-event stage timers
  cpu_baseline  the CPU oracle (oracle/cpu_rasterizer.c, OpenMP) on the host cores, one iteration
  kernels    every stage: ms per step, achieved GB/s, fraction of the measured HBM peak
--impl reference runs the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built from
/root/reference for sm_100) on the same workload; if that binary did not travel, the CPU oracle.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "fwd+bwd render iters/sec @1M Gaussians SH3 1280x720"
UNIT = "iters/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--P", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=1280)
    ap.add_argument("--height", type=int, default=720)
    ap.add_argument("--sh-degree", type=int, default=3)
    ap.add_argument("--geom", default="scale_rot", choices=["scale_rot", "cov"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-config3", action="store_true", help="skip the configs[3] batch step reported under extra")
    ap.add_argument("--nccl", action="store_true", help="A/B at N>1: NCCL all-reduce instead of the peer-memory exchange kernel")
    ap.add_argument("--full-exchange", action="store_true",
                    help="A/B at N>1: exchange the expanded [P,M,3] SH gradient instead of its factored form")
    ap.add_argument("--no-hint", action="store_true", help="A/B: exact-size forward (host waits for the count mid-way)")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def algorithmic_bytes(P, V, R, N, Tn, M, D, geom, passes):
    """Compulsory global traffic per launch, SURVEY.md 8(d) (restated in DESIGN.md)."""
    c = 12 * (D + 1) ** 2
    g = 28 if geom == "scale_rot" else 24
    cov_store = 24 if geom == "scale_rot" else 0
    # binning (csrc/binning_buckets.cu). Presorted frames (>= 3M expected instances, or the full word does
    # not fit): tile | index words of w bytes (4 when tile and index bits fit 32, else 8); duplicate writes
    # them, one global partition pass reads and writes them, the per-bucket split reads them and writes
    # the 4-byte index list and the tile ranges. Full-word frames: 8-byte tile | depth | index words and
    # one CTA sort per bucket; five global passes or a pair sort when that does not fit.
    ib, tb = max(P - 1, 1).bit_length(), Tn.bit_length()
    packed = ib + 32 + tb <= 64
    presorted = (R >= 3_000_000 or not packed) and P >= 64 and tb - 10 <= 7
    pk_passes = (32 + tb + 8) // 9
    if presorted:
        w = 4 if tb + ib <= 32 else 8
        binning = {"duplicate_keys": 24 * P + w * R, "sort_histogram": 2 * w * R, "sort_passes": w * R + 4 * R + 8 * Tn,
                   # front-to-back presort of the Gaussians: build the (depth key | index) words, three
                   # stable partition passes over them, the last one writing the bare indices
                   "depth_order": (8 + 8) * P + 2 * 16 * P + 12 * P}
    elif packed and R <= 1024 * 12288 * 8 // 10:
        binning = {"duplicate_keys": 24 * P + 8 * R, "sort_histogram": 16 * R, "sort_passes": 8 * R + 4 * R + 8 * Tn}
    else:
        binning = {"duplicate_keys": 20 * P + (8 if packed else 12) * R, "sort_histogram": 8 * R,
                   "sort_passes": (R * 16 * (pk_passes - 1) + R * 20) if packed else R * 24 * passes,
                   "tile_ranges": 8 * R + 16 * Tn}
    return {
        **binning,
        "preprocess_fwd": 20 * P + V * (g + 4 + c + 4 + 8 + 16 + 12 + 3 + cov_store),
        "scan": (12 if presorted else 8) * P,
        "blend_fwd": 44 * R + 24 * N + 8 * Tn,
        "bwd_zero": 48 * P,
        "blend_bwd": 44 * R + 28 * N + 8 * Tn + 40 * R,
        "preprocess_bwd": 4 * P + P * (12 + 24 + 16 + 12 + 12 + 4 + 3 + g + c) + P * (12 + 12 + 4 + (28 if geom == "scale_rot" else 24) + 12 * M),
    }


def build_inputs(a, view, device):
    import torch
    from egogaussian_b200 import scenes
    sc = scenes.make_scene(a.P, a.width, a.height, sh_degree=a.sh_degree, seed=3001, view=view)
    cam = sc.camera
    kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
              projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
              image_height=a.height, image_width=a.width, sh_degree=a.sh_degree, shs=sc.shs)
    if a.geom == "scale_rot":
        kw["scales"], kw["rotations"] = sc.scales, sc.rotations
    else:
        kw["cov3D_precomp"] = sc.cov3D()
    wc, wd, wa = scenes.upstream_weights(a.width, a.height, 3001 + (view or 0))
    # host side of a step: the camera as one 35-float record (view | projection | centre) and the three
    # upstream-gradient images as one [5,H,W] block, both in pinned memory
    host = {"cam": torch.cat((cam.viewmatrix.reshape(-1), cam.projmatrix.reshape(-1), cam.campos.reshape(-1))).float().pin_memory(),
            "w": torch.cat((wc, wd, wa)).contiguous().pin_memory()} if device.type == "cuda" else None
    cpu_kw = dict(kw)
    kw = {k: (v.to(device).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
    ups = dict(dL_dcolor=wc.to(device).contiguous(), dL_ddepth=wd.to(device).contiguous(),
               dL_dalpha=wa.to(device).contiguous())
    return kw, ups, host, cpu_kw, (wc, wd, wa)


class HostLoop:
    """Per-step host<->device traffic of the e2e region, two steps in flight, the way a training loop
    feeds frames and logs its loss: step k's inputs are copied from pinned memory into buffer k % 2 on a
    copy stream (the camera first, then the images), its loss is copied back on a third stream, and the
    host reads the loss of step k - 1 once step k is enqueued. Every step still pays its own copies and
    its own read; they just do not sit between two kernels of the main stream. Used by both arms."""

    def __init__(self, host, device):
        import torch
        self.torch, self.host = torch, host
        self.dv = []
        for _ in range(2):
            cam = torch.empty_like(host["cam"], device=device)
            self.dv.append({"cam": cam, "w": torch.empty_like(host["w"], device=device), "viewmatrix": cam[0:16].view(4, 4),
                            "projmatrix": cam[16:32].view(4, 4), "campos": cam[32:35]})
        self.loss_host = [torch.empty((), dtype=torch.float32).pin_memory() for _ in range(2)]
        self.done = [torch.cuda.Event() for _ in range(2)]
        self.copied = [torch.cuda.Event() for _ in range(2)]
        self.cam_ready = [torch.cuda.Event() for _ in range(2)]
        self.loss_ready = [torch.cuda.Event() for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=device)
        self.back_stream = torch.cuda.Stream(device=device)
        self.k, self.last_loss = 0, None

    def begin(self):
        """Enqueue this step's host->device copies; returns (device buffers, event of the image copy).
        Buffer k % 2 is free: step k - 2 finished before its loss was read at the end of step k - 1."""
        b = self.k % 2
        dv = self.dv[b]
        with self.torch.cuda.stream(self.copy_stream):
            dv["cam"].copy_(self.host["cam"], non_blocking=True)
            self.cam_ready[b].record()
            dv["w"].copy_(self.host["w"], non_blocking=True)
            self.copied[b].record()
        self.torch.cuda.current_stream().wait_event(self.cam_ready[b])
        return dv, self.copied[b]

    @staticmethod
    def loss(color, depth, alpha, w):
        """<(color, depth, alpha), w>: the scalar both arms differentiate."""
        import torch
        return torch.dot(torch.cat((color, depth, alpha)).view(-1), w.view(-1))

    def end(self, loss):
        b = self.k % 2
        self.loss_ready[b].record()
        with self.torch.cuda.stream(self.back_stream):
            self.back_stream.wait_event(self.loss_ready[b])
            loss.record_stream(self.back_stream)
            self.loss_host[b].copy_(loss.detach(), non_blocking=True)
            self.done[b].record()
        if self.k > 0:
            self.done[1 - b].synchronize()                  # the previous step's result is on the host
            self.last_loss = float(self.loss_host[1 - b])
        self.k += 1

    def bytes_in(self):
        return sum(v.numel() * 4 for v in self.host.values())


def timed_region(step, steps, warmup, dist_on, device):
    """W untimed steps, then exactly K steps between barrier+synchronize pairs; CUDA events on the
    launching (current) stream; returns seconds, max over ranks."""
    import torch
    import torch.distributed as dist
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    sec = e0.elapsed_time(e1) / 1e3
    if dist_on:
        t = torch.tensor([sec], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item())
    return sec


class _TrainerModel:
    """Raw optimiser parameters of the bench scene, named like the reference's GaussianModel
    attributes (scene/gaussian_model.py:326-338): what a training iteration really starts from."""

    def __init__(self, a, device):
        import torch
        from egogaussian_b200 import scenes
        sc = scenes.make_scene(a.P, a.width, a.height, sh_degree=a.sh_degree, seed=3001, view=None)
        op = sc.opacities.reshape(-1, 1).clamp(1e-5, 1 - 1e-5)
        raw = {"_xyz": sc.means3D, "_features_dc": sc.shs[:, :1], "_features_rest": sc.shs[:, 1:],
               "_opacity": torch.log(op / (1 - op)), "_scaling": torch.log(sc.scales), "_rotation": sc.rotations}
        for k, v in raw.items():
            setattr(self, k, v.to(device).contiguous().requires_grad_(True))
        self._is_object = torch.zeros(a.P, 1, dtype=torch.int32, device=device)
        self.active_sh_degree = self.max_sh_degree = a.sh_degree
        self.names = list(raw)
        self.lrs = [1.6e-4, 2.5e-3, 2.5e-3 / 20, 0.05, 5e-3, 1e-3]   # arguments/__init__.py defaults
        cam = sc.camera
        self.image_width, self.image_height = a.width, a.height
        self.FoVx, self.FoVy = 2 * math.atan(cam.tanfovx), 2 * math.atan(cam.tanfovy)
        self.world_view_transform = cam.viewmatrix.to(device)
        self.full_proj_transform = cam.projmatrix.to(device)
        self.camera_center = cam.campos.to(device)
        self.bg, self.tanfovx, self.tanfovy = sc.bg.to(device), cam.tanfovx, cam.tanfovy

    def params(self):
        return [getattr(self, k) for k in self.names]


class _Pipe:
    compute_cov3D_python = True    # forced by the reference's trainers (train.py:49)
    convert_SHs_python = False


def eager_glue(pc, scaling_modifier=1.0):
    """What the reference's Python evaluates before every rasterizer call, as the same sequence of
    eager torch ops (get_opacity / get_features / get_covariance: scene/gaussian_model.py:29-33,
    126-153,167-168; zero-filled matrices written entry by entry: utils/general_utils.py:110-158).
    Used only to time the untouched glue beside the fused kernels."""
    import torch
    opacity = torch.sigmoid(pc._opacity)
    shs = torch.cat((pc._features_dc, pc._features_rest), dim=1)
    s = scaling_modifier * torch.exp(pc._scaling)
    r = pc._rotation
    q = r / torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.zeros((q.size(0), 3, 3), device=q.device)
    entries = ((0, 0, 1 - 2 * (y * y + z * z)), (0, 1, 2 * (x * y - w * z)), (0, 2, 2 * (x * z + w * y)),
               (1, 0, 2 * (x * y + w * z)), (1, 1, 1 - 2 * (x * x + z * z)), (1, 2, 2 * (y * z - w * x)),
               (2, 0, 2 * (x * z - w * y)), (2, 1, 2 * (y * z + w * x)), (2, 2, 1 - 2 * (x * x + y * y)))
    for i, j, v in entries:
        R[:, i, j] = v
    S = torch.zeros((s.shape[0], 3, 3), dtype=torch.float, device=s.device)
    for j in range(3):
        S[:, j, j] = s[:, j]
    L = R @ S
    sym = L @ L.transpose(1, 2)
    cov = torch.zeros((sym.shape[0], 6), dtype=torch.float, device=sym.device)
    for k, (i, j) in enumerate(((0, 0), (0, 1), (0, 2), (1, 1), (1, 2), (2, 2))):
        cov[:, k] = sym[:, i, j]
    return cov, opacity, shs


def trainer_targets(a, device):
    """A ground-truth frame and a hand/object mask for the trainers' image loss (seeded noise: the
    loss kernels do not care what the picture shows)."""
    import torch
    gen = torch.Generator().manual_seed(3003)
    gt = torch.rand(3, a.height, a.width, generator=gen).to(device)
    mask = (torch.rand(a.height, a.width, generator=gen) < 0.15).to(torch.int32).to(device)
    return gt, mask


def trainer_iteration_rate(a, device, weights, render_fn, optimizer_step, dist_on, loss_fn=None):
    """iterations/s of: raw parameters -> glue -> rasterizer -> loss -> backward -> Adam. Without
    `loss_fn` the loss is the weighted sum the rasterizer metric uses."""
    wc, wd, wa = (w.to(device).contiguous() for w in weights)

    def it():
        color, depth, alpha = render_fn()
        loss = loss_fn(color) if loss_fn else (color * wc).sum() + (depth * wd).sum() + (alpha * wa).sum()
        loss.backward()
        optimizer_step()

    return a.steps / timed_region(it, a.steps, 3, dist_on, device)


def cpu_baseline(cpu_kw, weights):
    """One fwd+bwd iteration of the same workload on the host cores with the CPU oracle."""
    from oracle import cpu_oracle as orc
    orc.build()
    bk = {k: v for k, v in cpu_kw.items() if k != "opacities"}
    t0, iters = time.time(), 0
    while iters < 20 and (iters == 0 or time.time() - t0 < 10.0):   # bounded: about 10 s of CPU work
        f = orc.render_forward(**cpu_kw)
        orc.render_backward(f, dL_dcolor=weights[0], dL_ddepth=weights[1], dL_dalpha=weights[2], **bk)
        iters += 1
    dt = time.time() - t0
    return {"value": iters / dt, "unit": UNIT, "cores": orc.num_threads(), "kind": "port",
            "sample": "%d full fwd+bwd iterations of the same workload in %.1f s, C/OpenMP oracle "
                      "(oracle/cpu_rasterizer.c)" % (iters, dt)}


# --------------------------------------------------------------------------------------------
def run_config3(a, rank, world, device, ex):
    """BASELINE configs[3], the fine_all dynamic stage (trainers/fine_all.py:88-128) as one optimiser
    step over a FIXED batch of 8 views: rank r renders views r, r + N, ....

        xyz' = R xyz + t on the object rows                    egs_se3_apply     (apply_trans_rot_new)
        cov, opacity, shs = activations, L' = R L on them      egs_model_forward (get_rotated_covariance)
        per local view: egs_forward / egs_backward on (xyz', cov, opacity, shs), factored SH gradient
        ONE exchange over all 8 views in view order            egs_exchange_sum_factored (ViewBatch)
        chain to the raw parameters                            egs_model_backward
        xyz: gradient pulled back through the pose + Adam      egs_adam_se3_step (reverse_trans_rot_new)
        the other five groups: Adam                            egs_adam_step

    Returns timing, and a check of one step's summed gradients against the same batch accumulated on
    ONE GPU (every rank renders all 8 views itself once): the sums run in view order through the same
    kernel, so the two differ only by the atomic order inside each backward."""
    import torch
    import torch.distributed as dist
    from egogaussian_b200 import cabi, parallel, scenes
    from egogaussian_b200.diff_gaussian_rasterization import _C
    NV = 8
    if NV % world:
        return {"skipped": "8 views do not divide over %d ranks" % world}
    P, D = a.P, a.sh_degree
    M = (D + 1) ** 2
    sc = scenes.make_scene(P, a.width, a.height, sh_degree=D, seed=3001, view=None)
    cams = [scenes.make_camera(a.width, a.height, yaw_deg=(k - 3.5) * 2.0, x_shift=(k - 3.5) * 0.05) for k in range(NV)]
    op = sc.opacities.reshape(-1, 1).clamp(1e-5, 1 - 1e-5)
    raw = {"xyz": sc.means3D, "f_dc": sc.shs[:, :1], "f_rest": sc.shs[:, 1:], "opacity": torch.log(op / (1 - op)),
           "scaling": torch.log(sc.scales), "rotation": sc.rotations}
    raw = {k: v.to(device).contiguous() for k, v in raw.items()}
    is_obj = sc.is_object.to(device)
    Rt = scenes.object_pose().to(device)
    R_pose = Rt[:9].reshape(3, 3).contiguous()
    bg = sc.bg.to(device)
    model = dict(scaling_raw=raw["scaling"], rotation_raw=raw["rotation"], opacity_raw=raw["opacity"],
                 features_dc=raw["f_dc"], features_rest=raw["f_rest"], is_object=is_obj, which_object=1, accum_R=R_pose)
    dev_cams = [dict(viewmatrix=c.viewmatrix.to(device), projmatrix=c.projmatrix.to(device), campos=c.campos.to(device),
                     tanfovx=c.tanfovx, tanfovy=c.tanfovy) for c in cams]
    ups = []
    for k in range(NV):
        wc, wd, wa = scenes.upstream_weights(a.width, a.height, 3001 + k)
        ups.append(dict(dL_dcolor=wc.to(device).contiguous(), dL_ddepth=wd.to(device).contiguous(),
                        dL_dalpha=wa.to(device).contiguous()))
    scratch = torch.empty((cabi.load().egs_backward_scratch_bytes(P),), dtype=torch.uint8, device=device)
    names = ["f_dc", "f_rest", "opacity", "scaling", "rotation"]
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}
    moments = {k: (torch.zeros_like(v), torch.zeros_like(v)) for k, v in raw.items()}
    hints = {}

    def batch_gradients(batch, views):
        """Back-propagate `views` (global ids, one per layout of `batch`) and run the batch's one exchange."""
        moved, _ = _C.se3_apply(raw["xyz"], is_obj, 1, Rt, torch.empty(0, device=device))
        act = cabi.model_forward(**model)
        for j, v in enumerate(views):
            kw = dict(means3D=moved, opacities=act["opacity"], bg=bg, image_height=a.height, image_width=a.width,
                      sh_degree=D, shs=act["shs"], cov3D_precomp=act["cov3D"], **dev_cams[v])
            fwd = cabi.forward(num_rendered_hint=hints.get(v, 0), **kw)
            hints[v] = int(fwd.num_rendered * 1.25) + 8192
            bkw = {k: x for k, x in kw.items() if k not in ("opacities", "image_height", "image_width")}
            cabi.backward(fwd, scratch=scratch, **batch.backward_args(j), **bkw, **ups[v])
        res = batch.exchange(moved, D)
        return moved, res

    def optimiser_step(res, step_no):
        g = cabi.model_backward(dL_dcov3D=res.views["cov3D"], dL_dopacity=res.views["opacity"], dL_dshs=res.views["sh"],
                                **model)
        _C.adam_se3_step(raw["xyz"], res.views["means3D"], moments["xyz"][0], moments["xyz"][1], is_obj, 1, Rt,
                         lrs["xyz"], step_no, 0.9, 0.999, 1e-15)
        grads = {"f_dc": g["features_dc"], "f_rest": g["features_rest"], "opacity": g["opacity_raw"],
                 "scaling": g["scaling_raw"], "rotation": g["rotation_raw"]}
        _C.adam_step([raw[k] for k in names], [grads[k] for k in names], [moments[k][0] for k in names],
                     [moments[k][1] for k in names], [lrs[k] for k in names], [step_no] * len(names), 0.9, 0.999, 1e-15)

    batch = parallel.ViewBatch(NV, P, M, device, exchange=ex, geometry="cov")
    my_views = batch.local_views

    # ---- check: the distributed sums against the whole batch on this GPU alone ----
    check = None
    _, res = batch_gradients(batch, my_views)
    torch.cuda.synchronize()
    if world > 1:
        solo = parallel.ViewBatch(NV, P, M, device, exchange=None, geometry="cov")
        _, ref = batch_gradients(solo, list(range(NV)))
        torch.cuda.synchronize()
        worst = 0.0
        for name in ("means3D", "cov3D", "opacity", "means2D", "sh"):
            want, got = ref.views[name], res.views[name]
            worst = max(worst, float((got - want).abs().max() / want.abs().max().clamp_min(1e-30)))
        w = torch.tensor([worst], device=device, dtype=torch.float64)
        dist.all_reduce(w, op=dist.ReduceOp.MAX)
        check = {"max_rel_err_vs_single_gpu_batch": float(w.item()), "tolerance": 1e-4}
        del solo, ref
        torch.cuda.empty_cache()

    stepc = [0]

    def step():
        _, r_ = batch_gradients(batch, my_views)
        stepc[0] += 1
        optimiser_step(r_, stepc[0])

    steps = max(5, a.steps // 2)
    sec = timed_region(step, steps, 3, world > 1, device)
    if ex is not None and ex.failed():
        raise RuntimeError("a rank gave up waiting for its peers inside the exchange kernel")
    return {"workload": "configs[3]: 8-view batch, 10 %% object rows under the App. C pose, one exchange and one "
                        "Adam step (xyz through the pose) per batch; %d view(s) per GPU" % (NV // world),
            "ms_per_batch": sec / steps * 1e3, "views_per_s": NV * steps / sec, "batches_per_s": steps / sec,
            "steps": steps, "check": check}


LEAF_TO_FLAT = {"means3D": "means3D", "opacities": "opacity", "shs": "sh", "scales": "scales", "rotations": "rotations"}


def run_ours(a, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from egogaussian_b200 import cabi, parallel
    from egogaussian_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    dist_on = world > 1
    kw, ups, host, cpu_kw, weights = build_inputs(a, rank if dist_on else None, device)
    P, M = a.P, (a.sh_degree + 1) ** 2
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    # the gradient exchange of view-DP: one kernel per rank over peer-mapped memory (csrc/exchange.cu);
    # NCCL's all-reduce only if the buffers cannot be mapped on this box (or with --nccl)
    ex, exchange = None, "none"
    if dist_on and not a.nccl:
        try:
            ex = parallel.PeerExchange(device)
        except Exception as exc:
            exchange = "nccl all_reduce (peer mapping failed: %s)" % str(exc)[:120]
    elif dist_on:
        exchange = "nccl all_reduce"
    # the SH gradient travels in factored form (3 floats of dL/dRGB per Gaussian and view instead of 3 M;
    # every rank rebuilds the summed [P,M,3] tensor from all ranks' factors: csrc/exchange.cu)
    factored = ex is not None and not a.full_exchange and M > 1
    batch = parallel.ViewBatch(world, P, M, device, exchange=ex) if factored else None
    flat = batch.result if factored else parallel.FlatGrads(P, M, device, allocate=ex.allocate if ex else None)
    if ex:
        exchange = ("egs_exchange_sum_factored, " + batch.path) if factored else \
            ("egs_exchange_sum, " + ("switch multicast" if ex.multicast_pointer(flat.flat) else "peer loads/stores"))
    scratch = torch.empty((cabi.load().egs_backward_scratch_bytes(P),), dtype=torch.uint8, device=device)
    state = {}

    def step(exchange_grads=True):
        # like a training loop (and like the torch binding does by itself): guess this iteration's
        # instance count from the previous one so the forward never stalls the GPU on the host
        fwd = cabi.forward(num_rendered_hint=0 if a.no_hint else state.get("hint", 0), **kw)
        state["hint"] = int(fwd.num_rendered * 1.25) + 8192
        out = flat.as_backward_out()
        if a.geom != "scale_rot":
            out = {"means3D": out["means3D"], "sh": out["sh"], "opacity": out["opacity"], "means2D": out["means2D"],
                   "cov3D": state.setdefault("cov", torch.empty((P, 6), device=device)), "rgb": out.get("rgb")}
        push = {k: v for k, v in batch.backward_args(0).items() if k != "out"} if factored else {}
        cabi.backward(fwd, out=out, scratch=scratch, **push, **bkw, **ups)
        if dist_on and exchange_grads:
            if factored:
                batch.exchange(kw["means3D"], a.sh_degree)
            elif ex:
                ex.sum(flat.flat)
            else:
                parallel.allreduce_gradients(flat.flat)
        state["fwd"] = fwd

    def sh_basis_torch(means, campos, degree):
        """[P,16] values of the real SH basis in the direction campos -> means (the constants and polynomials
        of DGR/cuda_rasterizer/backward.cu:20-139), zero above `degree`: the checker's own restatement."""
        d = means - campos[None, :]
        d = d / d.norm(dim=1, keepdim=True)
        x, y, z = d[:, 0], d[:, 1], d[:, 2]
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        c1, c2, c3 = 0.4886025119029199, [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792,
                                           0.5462742152960396], [-0.5900435899266435, 2.890611442640554, -0.4570457994644658,
                                                                 0.3731763325901154, -0.4570457994644658, 1.445305721320277,
                                                                 -0.5900435899266435]
        cols = [torch.full_like(x, 0.28209479177387814), -c1 * y, c1 * z, -c1 * x,
                c2[0] * xy, c2[1] * yz, c2[2] * (2 * zz - xx - yy), c2[3] * xz, c2[4] * (xx - yy),
                c3[0] * y * (3 * xx - yy), c3[1] * xy * z, c3[2] * y * (4 * zz - xx - yy), c3[3] * z * (2 * zz - 3 * xx - 3 * yy),
                c3[4] * x * (4 * zz - xx - yy), c3[5] * z * (xx - yy), c3[6] * x * (xx - 3 * yy)]
        b = torch.stack(cols, dim=1)
        b[:, (degree + 1) ** 2:] = 0
        return b

    def exchange_check():
        """One more step whose exchanged gradients are checked twice. (1) Against the SAME inputs: the flat
        buffer against NCCL's all-reduce of a copy taken before the exchange, and the rebuilt dL/dsh against
        sum_q basis_q x dL/dRGB_q formed in torch from the all-gathered factors -- what is left is summation
        order. (2) Against NCCL's all-reduce of the expanded gradients of ANOTHER backward run of the same
        step, which also contains the atomic-order noise of two runs (1e-6 .. 3e-5 by tensor, the same size
        as the reference's own run-to-run noise)."""
        step(exchange_grads=False)
        ref = parallel.FlatGrads(P, M, device)
        if factored:   # the expanded rows of this rank's view, from an ordinary backward of the same forward
            cabi.backward(state["fwd"], out=ref.as_backward_out(), scratch=scratch, **bkw, **ups)
            step(exchange_grads=False)   # fresh factors (same inputs) for the exchange under test
        else:
            ref.flat.copy_(flat.flat)
        dist.all_reduce(ref.flat)
        same = {}
        if factored:
            before = flat.flat.clone()
            dist.all_reduce(before)
            slots = [torch.empty_like(flat.rgb_slot) for _ in range(world)]
            dist.all_gather(slots, flat.rgb_slot.contiguous())
            sh_want = torch.zeros((P, M, 3), device=device)
            tail = (3 * P + 3) // 4 * 4
            for q in range(world):
                b = sh_basis_torch(kw["means3D"], slots[q][tail:tail + 3], a.sh_degree)[:, :M]
                sh_want += b[:, :, None] * slots[q][:3 * P].view(P, 3)[:, None, :]
            batch.exchange(kw["means3D"], a.sh_degree)
            same = {"flat": (flat.flat, before), "sh": (flat.views["sh"], sh_want)}
        elif ex:
            before = flat.flat.clone()
            dist.all_reduce(before)
            ex.sum(flat.flat)
            same = {"flat": (flat.flat, before)}
        else:
            parallel.allreduce_gradients(flat.flat)
        torch.cuda.synchronize()

        def rel(got, want):
            return float((got - want.view_as(got)).abs().max() / want.abs().max().clamp_min(1e-30))
        worst_same = max([rel(g, w) for g, w in same.values()] or [0.0])
        worst, sig = 0.0, torch.zeros(1, dtype=torch.int64, device=device)
        for name, want in ref.views.items():
            got = flat.views[name]
            worst = max(worst, rel(got, want))
            sig += got.contiguous().view(torch.int32).to(torch.int64).sum()
        w = torch.tensor([worst_same, worst], device=device, dtype=torch.float64)
        dist.all_reduce(w, op=dist.ReduceOp.MAX)
        sigs = [torch.zeros_like(sig) for _ in range(world)]
        dist.all_gather(sigs, sig)
        failed = bool(ex.failed()) if ex else False
        return {"max_rel_err": float(w[0].item()), "identical_across_ranks": all(int(x) == int(sigs[0]) for x in sigs),
                "against": "the same inputs: NCCL all_reduce of a copy of the flat buffer, and sum_q basis_q x dL/dRGB_q "
                           "formed in torch from the all-gathered factors",
                "max_rel_err_vs_another_backward_run": float(w[1].item()),
                "another_run": "NCCL all_reduce of the expanded gradients of a second backward of the same step "
                               "(adds the atomic-order noise of two runs)", "gave_up": failed}

    clocks = ClockSampler(local_rank) if rank == 0 else None
    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    l0 = cabi.launch_count()
    sec = timed_region(step, a.steps, 0, dist_on, device)
    launches = cabi.launch_count() - l0
    clk = clocks.stop() if clocks else None
    # Per-stage durations: the library's stage timers bracket every stage with a CUDA-event pair on the
    # launching stream; 18 pairs per step cost 0.05 ms (5 %) of device time, so the headline region above
    # runs without them and the same K steps are run once more with them for the table and the roofline.
    cabi.profile_enable(True)
    step()
    torch.cuda.synchronize()
    cabi.profile_read()
    sec_prof = timed_region(step, a.steps, 0, dist_on, device)
    prof = cabi.profile_read()
    cabi.profile_enable(False)
    value = world * a.steps / sec
    xcheck = exchange_check() if dist_on else None

    fwd = state["fwd"]
    R = fwd.num_rendered
    V = int((fwd.radii > 0).sum())
    N = a.width * a.height
    Tn = ((a.width + 15) // 16) * ((a.height + 15) // 16)
    frag = int(fwd.image_view()["n_contrib"].to(torch.int64).sum())
    passes = (32 + Tn.bit_length() + 7) // 8
    alg = algorithmic_bytes(P, V, R, N, Tn, M, a.sh_degree, a.geom, passes)
    peak, peak_src = measured_peak_gbs()
    kernels = {}
    for name, (ms, cnt) in prof.items():
        if cnt == 0 or name not in alg:
            continue
        per = ms / a.steps
        kernels[name] = {"ms_per_step": round(per, 4), "launches_per_step": cnt / a.steps,
                         "alg_MB": round(alg[name] / 1e6, 2), "GBps": round(alg[name] / per / 1e6, 1),
                         "frac_hbm": round(alg[name] / per / 1e6 / peak, 4)}
    top = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
    # per launch, from the committed ncu --set full capture of this command's kernels (tools/ncu_traffic.py ->
    # profiles/traffic.json): DRAM bytes, executed warp instructions, issue-slot and shared-memory utilisation
    ncu = {}
    try:
        tj = json.loads((ROOT / "profiles" / "traffic.json").read_text())["stages"].get(top, {})
        for v in tj.values():
            ncu = v
    except Exception:
        ncu = {}
    sm_count = torch.cuda.get_device_properties(device).multi_processor_count
    sm_mhz = (clk or {}).get("sm_mhz") or (clk or {}).get("sm_max_mhz") or 1965.0
    ms_launch = kernels[top]["ms_per_step"] / max(kernels[top]["launches_per_step"], 1)
    hbm = {"achieved": kernels[top]["GBps"], "peak": peak, "unit": "GB/s", "frac": kernels[top]["frac_hbm"],
           "peak_source": peak_src, "alg_MB": kernels[top]["alg_MB"]}
    inst = ncu.get("warp_instructions_per_launch")
    if top in ("blend_fwd", "blend_bwd") and inst:
        # The tile blends read L2-resident lists (DRAM traffic is a twentieth of the algorithmic bytes), so
        # HBM is not what bounds them: they are bound by instruction issue. Ceiling: 4 schedulers per SM,
        # one warp instruction per cycle each, at the SM clock sampled during the timed region.
        issue_peak = sm_count * 4 * sm_mhz * 1e6 / 1e9                       # G warp-instructions / s
        issue_ach = inst / (ms_launch * 1e-3) / 1e9
        roofline = {"kernel": top, "bound": "issue", "achieved": round(issue_ach, 1), "peak": round(issue_peak, 1),
                    "unit": "Gwarp-inst/s", "frac": round(issue_ach / issue_peak, 4), "traffic": ncu.get("bytes_per_launch"),
                    "ms_per_launch": ms_launch, "warp_instructions_per_launch": inst,
                    "instructions_source": "smsp__inst_executed.sum of the committed ncu capture (profiles/traffic.json); "
                                           "duration live from this run's stage events",
                    "issue_active_pct_ncu": ncu.get("issue_active_pct"), "hbm": hbm,
                    "shared_wavefronts_pct_of_peak_ncu": ncu.get("shared_wavefronts_pct"),
                    "global_red_requests_per_launch_ncu": ncu.get("global_red_requests_per_launch")}
    else:
        roofline = {"kernel": top, "bound": "hbm", **hbm, "traffic": ncu.get("bytes_per_launch"), "ms_per_launch": ms_launch,
                    "issue_active_pct_ncu": ncu.get("issue_active_pct")}
    # fragment-rate ceilings of the two blends: FP32 lanes / ~13 FMA-class operations per
    # fragment, and one MUFU ex2 per fragment on 16 lanes per SM, both at the sampled clock
    frag_ceiling = {"fp32_Gfrag_per_s": round(sm_count * 128 * sm_mhz * 1e6 / 13 / 1e9, 1),
                    "mufu_Gfrag_per_s": round(sm_count * 16 * sm_mhz * 1e6 / 1e9, 1), "sm_count": sm_count, "sm_mhz": sm_mhz,
                    "note": "a fragment = one (pixel, list entry) visit up to the pixel's last contributor; most visits "
                            "are rejected by the footprint bit lists without an instruction, so the rates can pass "
                            "the per-fragment ceilings"}

    # ---- end-to-end through the public Python API with host inputs -------------------------
    e2e = None
    if not a.no_e2e:
        leaves = {k: kw[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "scales", "rotations",
                                                                  "cov3D_precomp") if k in kw}
        loop = HostLoop(host, device)
        h2d = loop.bytes_in()
        # the autograd surface returns the expanded SH gradient: its exchange uses a full-size buffer
        flat_e2e = parallel.FlatGrads(P, M, device, allocate=ex.allocate) if (ex and factored) else flat

        def e2e_step():
            # the host inputs travel on a copy stream: the camera record first (the forward waits for it),
            # then the upstream-gradient images (standing in for the ground-truth frame), which are not
            # needed before the loss
            dv, copied = loop.begin()
            rs = GaussianRasterizationSettings(a.height, a.width, kw["tanfovx"], kw["tanfovy"], kw["bg"], 1.0,
                                               dv["viewmatrix"], dv["projmatrix"], a.sh_degree, dv["campos"], False, False)
            means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
            color, radii, depth, alpha = GaussianRasterizer(rs)(
                means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"], shs=leaves["shs"],
                scales=leaves.get("scales"), rotations=leaves.get("rotations"), cov3D_precomp=leaves.get("cov3D_precomp"))
            torch.cuda.current_stream().wait_event(copied)
            loss = loop.loss(color, depth, alpha, dv["w"])
            loss.backward()
            if dist_on and ex and all(k in LEAF_TO_FLAT for k in leaves):   # gather the autograd gradients into the peer-mapped flat buffer, one exchange
                for k, t in leaves.items():
                    flat_e2e.views[LEAF_TO_FLAT[k]].copy_(t.grad.view_as(flat_e2e.views[LEAF_TO_FLAT[k]]))
                ex.sum(flat_e2e.flat)
            elif dist_on:
                for t in leaves.values():
                    dist.all_reduce(t.grad)
            loop.end(loss)                                  # loss k -> pinned memory; loss k - 1 read on the host
            for t in leaves.values():
                t.grad = None

        sec2 = timed_region(e2e_step, a.steps, a.warmup, dist_on, device)
        e2e = {"value": world * a.steps / sec2, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
               "ms_per_step": sec2 / a.steps * 1e3, "last_loss": loop.last_loss,
               "host_loop": "two steps in flight: every step copies its inputs from pinned memory and its loss back; "
                            "the host reads loss k-1 after enqueueing step k (same loop in the reference arm)"}

    # ---- full training step (adds the fused Adam over all parameter groups), informational ----
    extra = {}
    try:
        from egogaussian_b200.diff_gaussian_rasterization import _C
        names = ["means3D", "sh", "opacity"] + (["scales", "rotations"] if a.geom == "scale_rot" else [])
        params = [torch.randn_like(flat.views[n]) for n in names]
        grads = [flat.views[n] for n in names]
        ms_ = [torch.zeros_like(p) for p in params]
        vs_ = [torch.zeros_like(p) for p in params]
        lrs = [1.6e-4, 2.5e-3, 0.05, 5e-3, 1e-3][:len(names)]
        stepc = [0]

        def train_step():
            step()
            stepc[0] += 1
            _C.adam_step(params, grads, ms_, vs_, lrs, [stepc[0]] * len(params), 0.9, 0.999, 1e-15)

        adam_floats = int(sum(p_.numel() for p_ in params))
        if factored:
            extra["train_step_exchange"] = "egs_exchange_sum_factored, then the replicated egs_adam_step on every rank"
        elif ex:   # exchange and update in one kernel: Adam on the rank's slice, moments sharded by rank
            del params, ms_, vs_
            params_flat = ex.allocate(flat.flat.numel())
            params_flat.normal_()
            m_shard, v_shard = ex.moment_shards(flat.flat.numel())
            segs = flat.adam_segments({"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05,
                                       "scaling": 5e-3, "rotation": 1e-3})
            torch.cuda.synchronize()
            dist.barrier()

            def train_step():   # noqa: F811
                step(exchange_grads=False)
                stepc[0] += 1
                ex.adam_step(flat.flat, params_flat, m_shard, v_shard, segs, stepc[0])
            extra["train_step_exchange"] = "egs_exchange_adam_step"

        sec3 = timed_region(train_step, a.steps, 3, dist_on, device)
        extra["train_step_iters_per_s"] = world * a.steps / sec3
        extra["adam_floats"] = adam_floats
        if ex and ex.failed():
            raise RuntimeError("a rank gave up waiting for its peers inside the exchange kernel")
    except Exception as exc:  # informational only
        extra["train_step_error"] = str(exc)[:200]

    # ---- trainer mode: the iteration starts from the RAW parameters ----------
    # raw -> activations + covariance -> rasterizer -> loss -> backward -> Adam over the six groups,
    # once with the fused model kernels behind gaussian_renderer.render(), once with the reference's
    # eager torch glue in front of the same rasterizer. `--impl reference` reports the same
    # iteration with the reference rasterizer and torch.optim.Adam.
    if world == 1 and not a.no_e2e:
        try:
            from egogaussian_b200.gaussian_renderer import render
            del leaves, loop
            pc = _TrainerModel(a, device)
            params = pc.params()
            ms_ = [torch.zeros_like(p) for p in params]
            vs_ = [torch.zeros_like(p) for p in params]
            stepc = [0]

            def fused_adam():
                stepc[0] += 1
                with torch.no_grad():
                    _C.adam_step(params, [p.grad for p in params], ms_, vs_, pc.lrs, [stepc[0]] * len(params),
                                 0.9, 0.999, 1e-15)
                for p in params:
                    p.grad = None

            def render_fused():
                pkg = render(pc, pc, _Pipe(), pc.bg)
                return pkg["render"], pkg["depth"], pkg["alpha"]

            def render_eager_glue():
                cov, opacity, shs = eager_glue(pc)
                rs = GaussianRasterizationSettings(a.height, a.width, pc.tanfovx, pc.tanfovy, pc.bg, 1.0,
                                                   pc.world_view_transform, pc.full_proj_transform, a.sh_degree,
                                                   pc.camera_center, False, False)
                color, _, depth, alpha = GaussianRasterizer(rs)(
                    means3D=pc._xyz, means2D=torch.zeros_like(pc._xyz, requires_grad=True), opacities=opacity,
                    shs=shs, cov3D_precomp=cov)
                return color, depth, alpha

            extra["trainer_iters_per_s"] = trainer_iteration_rate(a, device, weights, render_fused, fused_adam, False)
            # the same with the trainers' real loss, (1 - l) L1 + l (1 - SSIM) and the gradient mask
            # (trainers/train_static_bg.py:98-102), from the fused loss kernels
            from egogaussian_b200 import loss_utils
            gt_img, int_mask = trainer_targets(a, device)
            cabi.profile_enable(True)
            cabi.profile_read()
            extra["trainer_ssim_iters_per_s"] = trainer_iteration_rate(
                a, device, weights, render_fused, fused_adam, False,
                loss_fn=lambda color: loss_utils.image_loss(color, gt_img, 0.2, grad_mask=int_mask)[0])
            tprof = cabi.profile_read()
            cabi.profile_enable(False)
            # algorithmic bytes of the added stages: raw parameters in, activated
            # tensors out and the reverse; 8 image-sized reads/writes each way for the loss; 28 B per
            # optimised float for Adam
            floats = sum(p.numel() for p in params)
            talg = {"model_fwd": P * (12 + 16 + 4 + 12 * M) + P * (24 + 4 + 12 * M),
                    "model_bwd": P * (12 + 16 + 4 + 24 + 4 + 12 * M) + P * (12 + 16 + 4 + 12 * M),
                    "image_loss": (2 + 3 + 3 + 2 + 1) * 3 * N * 4 + 4 * N, "adam": 28 * floats,
                    # raw-parameter mode of the per-Gaussian kernels (render(fused=True)): raw scale,
                    # quaternion, opacity logit and the split SH rows in; no cov3D / opacity / SH tensors
                    "preprocess_fwd": 20 * P + V * (28 + 4 + 12 * M + 4 + 8 + 16 + 16 + 1),
                    "preprocess_bwd": P * (12 + 4 + 12 * M + 1 + 28 + 4 + 48) + P * (12 + 12 + 4 + 28 + 12 * M)}
            iters = a.steps + 3   # timed + warm-up iterations between the two reads
            extra["trainer_kernels"] = {
                k: {"ms_per_step": round(tprof[k][0] / iters, 4), "alg_MB": round(talg[k] / 1e6, 2),
                    "GBps": round(talg[k] / (tprof[k][0] / iters) / 1e6, 1),
                    "frac_hbm": round(talg[k] / (tprof[k][0] / iters) / 1e6 / peak, 4)}
                for k in talg if k in tprof and tprof[k][1] > 0}
            extra["trainer_stage_ms"] = {k: round(v[0] / iters, 4) for k, v in tprof.items() if v[1] > 0}
            extra["trainer_iters_per_s_eager_glue"] = trainer_iteration_rate(a, device, weights, render_eager_glue,
                                                                             fused_adam, False)
        except Exception as exc:  # informational only
            extra["trainer_error"] = str(exc)[:300]

    if not a.no_config3:
        try:
            torch.cuda.empty_cache()
            extra["config3"] = run_config3(a, rank, world, device, ex)
        except Exception as exc:  # informational only
            extra["config3"] = {"error": str(exc)[:300]}

    if rank != 0:
        return None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": sec / a.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[2]: 1M-Gaussian SH3 scene, 1280x720, one view per GPU per step "
                               "(rank r renders view r of the configs[3] 8-view batch when N>1)",
                   "P": P, "sh_degree": a.sh_degree, "width": a.width, "height": a.height, "geometry": a.geom,
                   "num_rendered": R, "visible": V, "fragments": frag,
                   "parallelism": "view-dp%d" % world if dist_on else "single", "exchange": exchange,
                   "l2": "working set (>300 MB of Gaussian data + sort buffers per step) exceeds the 126 MB L2; no flush"},
        "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "kernels": kernels, "clocks": clk,
        "mfragments_per_s": {"fwd": frag / (kernels["blend_fwd"]["ms_per_step"] * 1e3),
                             "bwd": frag / (kernels["blend_bwd"]["ms_per_step"] * 1e3)},
        "fragment_ceilings": frag_ceiling,
        "kernels_measured_in": "a second pass of the same %d steps with the per-stage event pairs on (%.4f ms per "
                               "step there; the headline region runs without them)" % (a.steps, sec_prof / a.steps * 1e3),
        "extra": extra,
    }
    if xcheck is not None:
        line["exchange_check"] = xcheck
        line["reference_n_gpus"] = 1   # --impl reference always runs on one GPU (the reference has no multi-GPU mode)
    if not a.no_cpu and world == 1:
        line["cpu_baseline"] = cpu_baseline(cpu_kw, weights)
    return line


def run_reference(a, rank, world, local_rank):
    """The unmodified reference on the same workload (rank 0 only)."""
    if rank != 0:
        return None
    from oracle import ref_dgr
    import torch
    if not (ref_dgr.available() and torch.cuda.is_available()):
        # the reference has no CPU implementation; time the CPU restatement instead
        kw, ups, host, cpu_kw, weights = build_inputs(a, None, torch.device("cpu"))
        cb = cpu_baseline(cpu_kw, weights)
        return {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world,
                "steps": 1, "warmup": 0, "ms_per_step": 1e3 / cb["value"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "configs[2] on host cores (oracle port; oracle/_ref missing)"},
                "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0,
                                            "d2h_bytes_per_step": 0}}
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    kw, ups, host, cpu_kw, weights = build_inputs(a, None, device)
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    state = {}

    def step():
        f = ref_dgr.forward(**kw)
        ref_dgr.backward(f, **bkw, **ups)
        state["fwd"] = f

    clocks = ClockSampler(local_rank)
    sec = timed_region(step, a.steps, a.warmup, False, device)
    clk = clocks.stop()
    value = a.steps / sec

    rast = ref_dgr.make_autograd_rasterizer()
    leaves = {k: kw[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "scales", "rotations",
                                                              "cov3D_precomp") if k in kw}
    loop = HostLoop(host, device)
    h2d = loop.bytes_in()

    def e2e_step():
        # same host loop as our arm: images prefetched on a side stream, camera first, loss read one step late
        dv, copied = loop.begin()
        rs = dict(bg=kw["bg"], scale_modifier=1.0, viewmatrix=dv["viewmatrix"], projmatrix=dv["projmatrix"],
                  tanfovx=kw["tanfovx"], tanfovy=kw["tanfovy"], image_height=a.height, image_width=a.width,
                  sh_degree=a.sh_degree, campos=dv["campos"])
        means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
        color, radii, depth, alpha = rast(rs, means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
                                          shs=leaves["shs"], scales=leaves.get("scales"), rotations=leaves.get("rotations"),
                                          cov3D_precomp=leaves.get("cov3D_precomp"))
        torch.cuda.current_stream().wait_event(copied)
        loss = loop.loss(color, depth, alpha, dv["w"])
        loss.backward()
        loop.end(loss)
        for t in leaves.values():
            t.grad = None

    sec2 = timed_region(e2e_step, a.steps, a.warmup, False, device)

    # trainer mode, the reference's way: eager torch glue + its rasterizer + torch.optim.Adam
    # configured as scene/gaussian_model.py:189-198 (one group per tensor, lr per group, eps 1e-15)
    extra = {}
    try:
        del leaves, loop
        pc = _TrainerModel(a, device)
        opt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(pc.params(), pc.lrs)], lr=0.0, eps=1e-15)
        rs = dict(bg=pc.bg, scale_modifier=1.0, viewmatrix=pc.world_view_transform, projmatrix=pc.full_proj_transform,
                  tanfovx=pc.tanfovx, tanfovy=pc.tanfovy, image_height=a.height, image_width=a.width,
                  sh_degree=a.sh_degree, campos=pc.camera_center)

        def render_reference():
            cov, opacity, shs = eager_glue(pc)
            color, _, depth, alpha = rast(rs, means3D=pc._xyz, means2D=torch.zeros_like(pc._xyz, requires_grad=True),
                                          opacities=opacity, shs=shs, scales=None, rotations=None, cov3D_precomp=cov)
            return color, depth, alpha

        def adam():
            opt.step()
            opt.zero_grad(set_to_none=True)

        extra["trainer_iters_per_s"] = trainer_iteration_rate(a, device, weights, render_reference, adam, False)
        from oracle import loss_ref   # the reference's torch loss, restated (utils/loss_utils.py)
        gt_img, int_mask = trainer_targets(a, device)
        extra["trainer_ssim_iters_per_s"] = trainer_iteration_rate(
            a, device, weights, render_reference, adam, False,
            loss_fn=lambda color: loss_ref.image_loss(color, gt_img, 0.2, int_mask)[0])
    except Exception as exc:  # informational only
        extra["trainer_error"] = str(exc)[:300]

    return {"impl": "reference", "extra": extra,
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": sec / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[2]: 1M-Gaussian SH3 scene, 1280x720, unmodified reference CUDA "
                                   "rasterizer built for sm_100 (oracle/_ref)", "P": a.P, "sh_degree": a.sh_degree,
                       "width": a.width, "height": a.height, "geometry": a.geom,
                       "num_rendered": state["fwd"]["num_rendered"]},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 0, "kind": "reference",
                             "sample": "the reference has no CPU implementation: its own CUDA kernels on the same GPU"},
            "e2e": {"value": a.steps / sec2, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": sec2 / a.steps * 1e3, "python_layer": getattr(rast, "python_layer", None)},
            "clocks": clk}


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference" and rank != 0:
        return 0
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner
    # does) are sent to stderr for the whole run and the line is written to the saved descriptor
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    if a.impl == "ours" and not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    dist_on = world > 1 and a.impl == "ours"
    if dist_on:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        line = run_ours(a, rank, world, local_rank) if a.impl == "ours" else run_reference(a, rank, world, local_rank)
        if line is not None:
            os.write(json_fd, (json.dumps(line) + "\n").encode())
    finally:
        if dist_on:
            import torch.distributed as dist
            dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
