"""Scratch: where does the end-to-end step's time go on the host? perf_counter around every phase of the
bench's e2e loop (no synchronisation added), and the device time of the torch glue alone."""
import json
import sys
import time

import torch

sys.path.insert(0, '.')
sys.path.insert(0, 'tests')
import bench  # noqa: E402
from egogaussian_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402
from helpers import baseline_case  # noqa: E402

device = torch.device("cuda:0")
kw, ups = baseline_case(1_000_000, 3, 1280, 720, 3001, "scale_rot", "sh", "cuda")
host = {"w": torch.cat((ups["dL_dcolor"], ups["dL_ddepth"], ups["dL_dalpha"])).cpu().pin_memory(),
        "cam": torch.cat((kw["viewmatrix"].reshape(-1), kw["projmatrix"].reshape(-1), kw["campos"].reshape(-1))).cpu().pin_memory()}
leaves = {k: kw[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
loop = bench.HostLoop(host, device)
acc = {}


def tick(name, t0):
    t = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + (t - t0)
    return t


def e2e_step():
    t = time.perf_counter()
    dv, copied = loop.begin()
    t = tick("begin", t)
    rs = GaussianRasterizationSettings(720, 1280, kw["tanfovx"], kw["tanfovy"], kw["bg"], 1.0,
                                       dv["viewmatrix"], dv["projmatrix"], 3, dv["campos"], False, False)
    means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
    t = tick("settings+zeros", t)
    color, radii, depth, alpha = GaussianRasterizer(rs)(
        means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"], shs=leaves["shs"],
        scales=leaves["scales"], rotations=leaves["rotations"], cov3D_precomp=None)
    t = tick("forward", t)
    torch.cuda.current_stream().wait_event(copied)
    loss = loop.loss(color, depth, alpha, dv["w"])
    t = tick("loss", t)
    loss.backward()
    t = tick("backward", t)
    loop.end(loss)
    t = tick("end(read loss k-1)", t)
    for x in leaves.values():
        x.grad = None
    tick("drop grads", t)


for _ in range(10):
    e2e_step()
torch.cuda.synchronize()
acc.clear()
N = 100
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
w0 = time.perf_counter()
e0.record()
for _ in range(N):
    e2e_step()
e1.record()
torch.cuda.synchronize()
wall = (time.perf_counter() - w0) / N * 1e3
print(json.dumps({"host_ms_per_phase": {k: round(v / N * 1e3, 4) for k, v in acc.items()},
                  "host_sum_ms": round(sum(acc.values()) / N * 1e3, 4), "wall_ms": round(wall, 4),
                  "device_ms": round(e0.elapsed_time(e1) / N, 4)}))

# device time of the glue alone
color = torch.randn(3, 720, 1280, device=device, requires_grad=True)
depth = torch.randn(1, 720, 1280, device=device, requires_grad=True)
alpha = torch.randn(1, 720, 1280, device=device, requires_grad=True)
dv = loop.dv[0]


def glue():
    z = torch.zeros_like(leaves["means3D"])
    loss = loop.loss(color, depth, alpha, dv["w"])
    loss.backward()
    color.grad = depth.grad = alpha.grad = None
    return z


for _ in range(5):
    glue()
torch.cuda.synchronize()
e0.record()
for _ in range(N):
    glue()
e1.record()
torch.cuda.synchronize()
print(json.dumps({"glue_device_or_host_ms": round(e0.elapsed_time(e1) / N, 4)}))

# device timeline of a few steps: kernels and copies in start order, with the idle time before each
if len(sys.argv) > 1 and sys.argv[1] == "timeline":
    from torch.profiler import ProfilerActivity, profile
    for _ in range(5):
        e2e_step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(4):
            e2e_step()
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    t_end, rows = None, []
    for e in evs:
        s, d = e.time_range.start, e.time_range.end - e.time_range.start
        gap = None if t_end is None else s - t_end
        rows.append((round(s - evs[0].time_range.start, 1), round(d, 1), None if gap is None else round(gap, 1), e.name[:90]))
        t_end = max(t_end or 0, e.time_range.end)
    with open("gpurun_out/e2e_timeline.txt", "w") as f:
        for r in rows:
            f.write("%10.1f %8.1f %8s  %s\n" % (r[0], r[1], r[2], r[3]))
    print("timeline rows", len(rows))
