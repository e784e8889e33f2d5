"""GPU tests of the peer-memory gradient exchange (egs_exchange_sum / egs_exchange_adam_step,
SURVEY.md 8e). On one GPU the ranks are emulated: every "rank" owns a buffer and a stream on the same
device, and the kernels of one exchange run side by side (small grids), meeting in the same flag
barriers a real NVLink group uses. With two or more GPUs the real thing runs under torchrun
(tests/exchange_worker.py) against NCCL's all-reduce and the single-GPU Adam kernel."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


class EmulatedGroup:
    def __init__(self, world, n, device, seed=0):
        from egogaussian_b200 import cabi
        self.cabi, self.world, self.n = cabi, world, n
        gen = torch.Generator().manual_seed(seed)
        self.grads = [torch.randn(n, generator=gen).to(device) for _ in range(world)]
        self.signals = [torch.zeros(cabi.EGS_EXCHANGE_SIGNAL_WORDS, dtype=torch.int32, device=device)
                        for _ in range(world)]
        self.streams = [torch.cuda.Stream(device) for _ in range(world)]
        self.groups = [cabi.peer_group(world, r, [g.data_ptr() for g in self.grads],
                                       [s.data_ptr() for s in self.signals]) for r in range(world)]
        self.epoch = 0
        torch.cuda.synchronize(device)

    def each_rank(self, fn):
        self.epoch += 1
        for r in range(self.world):
            with torch.cuda.stream(self.streams[r]):
                fn(r, self.groups[r], self.epoch)
        torch.cuda.synchronize()
        assert all(int(s[-1].item()) == 0 for s in self.signals), "a rank gave up waiting for its peers"


def rank_order_sum(tensors, scale=1.0):
    acc = tensors[0].clone()
    for t in tensors[1:]:
        acc += t
    return acc * scale if scale != 1.0 else acc


@pytest.mark.parametrize("world,n", [(1, 4096), (2, 1 << 20), (3, 1000 * 4), (4, 4), (8, 1 << 18), (5, 0)])
def test_exchange_sum_emulated(world, n, built, cuda_device):
    g = EmulatedGroup(world, n, cuda_device, seed=world)
    for it in range(3):   # epochs advance; the counters reset themselves
        expect = rank_order_sum(g.grads, 0.5 if it == 1 else 1.0)
        g.each_rank(lambda r, grp, e: g.cabi.exchange_sum(grp, n, e, scale=0.5 if it == 1 else 1.0, max_ctas=4))
        for r in range(world):
            assert torch.equal(g.grads[r], expect), (world, n, it, r)   # same bits on every rank
        for t in g.grads:   # new gradients for the next round
            t.add_(torch.randn_like(t))
        torch.cuda.synchronize()


def test_exchange_adam_step_emulated_matches_sum_then_adam(built, cuda_device):
    """Reduce-scatter + Adam + all-gather in one kernel == rank-order sum, then egs_adam_step on
    every group (two learning rates inside the SH rows), bit for bit; moments sharded by rank."""
    from egogaussian_b200 import cabi
    from egogaussian_b200.diff_gaussian_rasterization import _C
    from egogaussian_b200.parallel import FlatGrads
    world, P, M = 3, 2501, 16
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}
    layout = FlatGrads(P, M, cuda_device)
    n = layout.flat.numel()
    segs = layout.adam_segments(lrs)
    g = EmulatedGroup(world, n, cuda_device, seed=9)
    gen = torch.Generator().manual_seed(3)
    start = torch.randn(n, generator=gen).to(cuda_device)
    params = [start.clone() for _ in range(world)]
    shard = int(cabi.load().egs_exchange_shard_floats(n, world))
    moments = [(torch.zeros(shard, device=cuda_device), torch.zeros(shard, device=cuda_device)) for _ in range(world)]

    # expected: dense Adam on the summed gradient, one run per learning rate, selected per element
    exp_p, exp_m, exp_v = start.clone(), torch.zeros_like(start), torch.zeros_like(start)
    lr_of = torch.zeros(n, device=cuda_device, dtype=torch.float64)
    for (b, e, la, lb, per, sp) in segs:
        lr_of[b:e] = la
        if per:
            idx = (torch.arange(e - b, device=cuda_device) % per) >= sp
            lr_of[b:e][idx] = lb
    covered = lr_of > 0
    for step in (1, 2, 3):
        summed = rank_order_sum(g.grads)
        new_p = exp_p.clone()
        new_m = new_v = None
        for lr in sorted(set(lr_of[covered].tolist())):
            p, m, v = exp_p.clone(), exp_m.clone(), exp_v.clone()
            _C.adam_step([p], [summed], [m], [v], [lr], [step], 0.9, 0.999, 1e-15)
            pick = lr_of == lr
            new_p[pick] = p[pick]
            new_m, new_v = m, v   # the moments do not depend on the learning rate
        exp_p = new_p
        exp_m = torch.where(covered, new_m, exp_m)
        exp_v = torch.where(covered, new_v, exp_v)

        before = [t.clone() for t in g.grads]
        g.each_rank(lambda r, grp, e: cabi.exchange_adam_step(
            grp, [p.data_ptr() for p in params], moments[r][0], moments[r][1], segs, n, e, step, max_ctas=4))
        tail = ~covered   # means2D rides along: summed on all ranks, no update
        for r in range(world):
            assert torch.equal(params[r], exp_p), (step, r, (params[r] != exp_p).sum().item())
            assert torch.equal(g.grads[r][tail], summed[tail])
            assert torch.equal(g.grads[r][covered], before[r][covered])
            lo, hi = r * shard, min(n, (r + 1) * shard)
            if hi > lo:
                assert torch.equal(moments[r][0][:hi - lo], exp_m[lo:hi])
                assert torch.equal(moments[r][1][:hi - lo], exp_v[lo:hi])
        for t in g.grads:
            t.copy_(torch.randn(n, generator=gen).to(cuda_device))
        torch.cuda.synchronize()


@pytest.mark.parametrize("world,deg,M", [(2, 3, 16), (3, 1, 4), (8, 3, 16), (4, 2, 16), (1, 0, 1)])
def test_factored_sh_exchange_emulated(world, deg, M, built, cuda_device):
    """egs_exchange_sum_factored against what it replaces. Every emulated rank owns a gather buffer with
    one slot per rank, renders another view of the same scene and back-propagates twice: once the
    ordinary way (dL/dsh as [P,M,3]) and once in factored form, where the backward kernel writes the
    dL/dRGB rows and the camera centre into the rank's slot of EVERY gather buffer (drgb_remote). Two
    backward runs differ in the last bits (atomic order), so the two forms are compared within 2e-5 of
    the tensor maximum. The exchange itself is exact: on every rank it must give the rank-order sum of
    the ranks' own expansions (each taken from a one-rank exchange of the same slot), bit for bit, while
    the flat buffer of the other gradients is summed as by egs_exchange_sum."""
    from helpers import grad_rel_err
    from egogaussian_b200 import cabi, scenes
    from egogaussian_b200.parallel import FlatGrads
    dev = cuda_device
    P, W, H = 5000 + 7, 128, 96
    slot = cabi.drgb_slot_floats(P)
    gathers = [torch.full((world * slot,), float("nan"), device=dev) for _ in range(world)]
    full, layouts = [], []
    means = None
    for r in range(world):
        sc = scenes.make_scene(P, W, H, sh_degree=deg, seed=61, view=r, max_sh_degree={16: 3, 4: 1, 1: 0}[M]).to(dev)
        cam = sc.camera
        means = sc.means3D
        kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
                  projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                  image_height=H, image_width=W, sh_degree=deg, shs=sc.shs,
                  scales=(sc.scales * 2.0).contiguous(), rotations=sc.rotations)
        wc, wd, wa = [t.to(dev) for t in scenes.upstream_weights(W, H, 61 + r)]
        ups = dict(dL_dcolor=wc, dL_ddepth=wd, dL_dalpha=wa)
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        fwd = cabi.forward(**kw)
        g_full = cabi.backward(fwd, **bkw, **ups)
        lay = FlatGrads(P, M, dev, factored=True, rgb_slot=gathers[r][r * slot:(r + 1) * slot])
        for k in ("means3D", "opacity", "scales", "rotations", "means2D"):   # poison (not the padding between views)
            lay.views[k].fill_(float("nan"))
        remote = [gathers[q].data_ptr() + 4 * r * slot for q in range(world) if q != r]
        cabi.backward(fwd, out=lay.as_backward_out(), rgb_remote=remote, **bkw, **ups)
        torch.cuda.synchronize()
        for k in ("means3D", "opacity", "scales", "rotations", "means2D"):
            assert grad_rel_err(lay.views[k], g_full[k].view_as(lay.views[k])) <= 2e-5, (r, k)
        for q in range(world):   # the slot arrived everywhere, rows and camera centre
            got = gathers[q][r * slot:(r + 1) * slot]
            assert torch.equal(got[:3 * P], gathers[r][r * slot:r * slot + 3 * P]) and torch.isfinite(got[:3 * P]).all()
            tail = (3 * P + 3) // 4 * 4
            assert torch.equal(got[tail:tail + 3], cam.campos), (r, q)
        full.append(g_full)
        layouts.append(lay)
    assert sum(float(g["sh"].abs().max()) for g in full) > 0
    # every rank's own expansion: a one-rank exchange of its slot (flat buffer untouched: 0 floats)
    own = []
    for r in range(world):
        sig = torch.zeros(cabi.EGS_EXCHANGE_SIGNAL_WORDS, dtype=torch.int32, device=dev)
        solo = cabi.peer_group(1, 0, [layouts[r].flat.data_ptr()], [sig.data_ptr()])
        out = torch.full((P, M, 3), float("nan"), device=dev)
        cabi.exchange_sum_factored(solo, 0, 1, means3D=means, drgb_ptrs=[gathers[r].data_ptr() + 4 * r * slot],
                                   dL_dsh=out, sh_degree=deg, max_ctas=2)
        torch.cuda.synchronize()
        assert int(sig[-1].item()) == 0
        assert grad_rel_err(out, full[r]["sh"]) <= 2e-5, r     # the expansion is the dL/dsh backward writes
        own.append(out)
    want_sh = rank_order_sum(own)
    want_flat = rank_order_sum([lay.flat for lay in layouts])
    signals = [torch.zeros(cabi.EGS_EXCHANGE_SIGNAL_WORDS, dtype=torch.int32, device=dev) for _ in range(world)]
    streams = [torch.cuda.Stream(dev) for _ in range(world)]
    torch.cuda.synchronize()
    for r in range(world):
        grp = cabi.peer_group(world, r, [lay.flat.data_ptr() for lay in layouts], [s.data_ptr() for s in signals])
        with torch.cuda.stream(streams[r]):
            cabi.exchange_sum_factored(grp, layouts[r].flat.numel(), 1, means3D=means,
                                       drgb_ptrs=[gathers[r].data_ptr() + 4 * q * slot for q in range(world)],
                                       dL_dsh=layouts[r].views["sh"], sh_degree=deg, max_ctas=2)
    torch.cuda.synchronize()
    assert all(int(s[-1].item()) == 0 for s in signals)
    for r in range(world):
        assert torch.equal(layouts[r].flat, want_flat), r
        got = layouts[r].views["sh"]
        assert torch.equal(got, want_sh), (r, float((got - want_sh).abs().max()), float(want_sh.abs().max()))


@pytest.mark.parametrize("geometry", ["scale_rot", "cov"])
def test_view_batch_on_one_gpu(geometry, built, cuda_device):
    """parallel.ViewBatch without a process group: the batch of BASELINE configs[3] accumulated on one GPU
    through the exchange kernel (world = 1, one source per view) against the view-order sum of ordinary
    per-view backward calls."""
    from helpers import grad_rel_err
    from egogaussian_b200 import cabi, scenes
    from egogaussian_b200.parallel import ViewBatch
    dev = cuda_device
    P, W, H, deg, NV = 6001, 160, 96, 2, 4
    M = 9
    sc = scenes.make_scene(P, W, H, sh_degree=deg, seed=81).to(dev)
    cams = [scenes.make_camera(W, H, yaw_deg=(k - 1.5) * 3.0, x_shift=(k - 1.5) * 0.1) for k in range(NV)]
    batch = ViewBatch(NV, P, M, dev, geometry=geometry)
    assert batch.local_views == list(range(NV))
    want = None
    for k, (cam, lay) in enumerate(zip(cams, batch.layouts)):
        kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix.to(dev),
                  projmatrix=cam.projmatrix.to(dev), campos=cam.campos.to(dev), tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                  image_height=H, image_width=W, sh_degree=deg, shs=sc.shs)
        if geometry == "scale_rot":
            kw["scales"], kw["rotations"] = (sc.scales * 2).contiguous(), sc.rotations
        else:
            kw["cov3D_precomp"] = scenes.build_covariance(sc.scales * 2, sc.rotations)
        wc, wd, wa = [t.to(dev) for t in scenes.upstream_weights(W, H, 81 + k)]
        ups = dict(dL_dcolor=wc, dL_ddepth=wd, dL_dalpha=wa)
        bkw = {n: v for n, v in kw.items() if n not in ("opacities", "image_height", "image_width")}
        fwd = cabi.forward(**kw)
        cabi.backward(fwd, **batch.backward_args(k), **bkw, **ups)
        full = cabi.backward(fwd, **bkw, **ups)
        if want is None:
            want = {n: v.clone() for n, v in full.items() if v is not None}
        else:
            for n in want:
                want[n] += full[n]
    res = batch.exchange(sc.means3D, deg)
    torch.cuda.synchronize()
    assert int(batch._signals[-1].item()) == 0
    for n, v in want.items():
        assert grad_rel_err(res.views[n], v.view_as(res.views[n])) <= 2e-5, n
    # a second batch through the same buffers (the epoch advances)
    res2 = batch.exchange(sc.means3D, deg)
    torch.cuda.synchronize()
    assert grad_rel_err(res2.views["sh"], want["sh"]) <= 2e-5


def test_exchange_adam_short_periods(built, cuda_device):
    """Two learning rates alternating with period 1, 2 and 3 inside a segment (a 4-float chunk then
    wraps around the period more than once): against egs_adam_step run per learning rate."""
    from egogaussian_b200 import cabi
    from egogaussian_b200.diff_gaussian_rasterization import _C
    world, n = 2, 4096
    for period, split in ((1, 1), (2, 1), (3, 2)):
        g = EmulatedGroup(world, n, cuda_device, seed=20 + period)
        start = torch.randn(n, generator=torch.Generator().manual_seed(period)).to(cuda_device)
        params = [start.clone() for _ in range(world)]
        shard = int(cabi.load().egs_exchange_shard_floats(n, world))
        moments = [(torch.zeros(shard, device=cuda_device), torch.zeros(shard, device=cuda_device)) for _ in range(world)]
        la, lb = 1e-2, 3e-4
        segs = [(0, n, la, lb, period, split)]
        summed = rank_order_sum(g.grads)
        use_b = (torch.arange(n, device=cuda_device) % period) >= split
        want = start.clone()
        for lr, pick in ((la, ~use_b), (lb, use_b)):
            p, m, v = start.clone(), torch.zeros_like(start), torch.zeros_like(start)
            _C.adam_step([p], [summed], [m], [v], [lr], [1], 0.9, 0.999, 1e-15)
            want[pick] = p[pick]
        g.each_rank(lambda r, grp, e: cabi.exchange_adam_step(
            grp, [p.data_ptr() for p in params], moments[r][0], moments[r][1], segs, n, e, 1, max_ctas=4))
        for r in range(world):
            assert torch.equal(params[r], want), (period, r)


def test_exchange_gives_up_without_touching_data(built, cuda_device):
    """A rank whose peer never arrives must not sum (or step the optimiser on) a stale buffer: after the
    timeout the kernel skips the data phase, leaves gradients and parameters as they were and raises
    the error word."""
    from egogaussian_b200 import cabi
    n = 1 << 14
    g = EmulatedGroup(2, n, cuda_device, seed=31)
    before = [t.clone() for t in g.grads]
    params = [torch.randn(n, device=cuda_device) for _ in range(2)]
    p_before = [t.clone() for t in params]
    m, v = torch.zeros(n // 2, device=cuda_device), torch.zeros(n // 2, device=cuda_device)
    cabi.exchange_set_timeout_ms(200)
    try:
        with torch.cuda.stream(g.streams[0]):   # rank 1 never launches
            cabi.exchange_adam_step(g.groups[0], [p.data_ptr() for p in params], m, v, [(0, n, 1e-2, 0, 0, 0)], n, 1, 1,
                                    max_ctas=4)
        torch.cuda.synchronize()
    finally:
        cabi.exchange_set_timeout_ms(20000)
    assert int(g.signals[0][-1].item()) == 1, "the error word must be set"
    for r in range(2):
        assert torch.equal(g.grads[r], before[r]) and torch.equal(params[r], p_before[r])
    assert float(m.abs().max()) == 0.0 and float(v.abs().max()) == 0.0


def test_exchange_rejects_bad_arguments(built, cuda_device):
    from egogaussian_b200 import cabi
    g = EmulatedGroup(2, 64, cuda_device)
    with pytest.raises(cabi.EgsError):
        cabi.exchange_sum(g.groups[0], 62, 1)            # not a multiple of 4
    bad = cabi.peer_group(2, 0, [t.data_ptr() for t in g.grads], [t.data_ptr() for t in g.signals])
    bad.world = 9
    with pytest.raises(cabi.EgsError):
        cabi.exchange_sum(bad, 64, 1)
    with pytest.raises(cabi.EgsError, match="resident"):  # a grid that could never be co-resident
        cabi.exchange_sum(g.groups[0], 64, 1, max_ctas=1 << 20)
    m = torch.zeros(32, device=cuda_device)
    with pytest.raises(cabi.EgsError):                   # overlapping segments
        cabi.exchange_adam_step(g.groups[0], [t.data_ptr() for t in g.grads], m, m.clone(),
                                [(0, 32, 1e-3, 0, 0, 0), (16, 64, 1e-3, 0, 0, 0)], 64, 1, 1)


@pytest.mark.timeout(600)
def test_exchange_across_gpus_under_torchrun(built):
    """Two real ranks over NVLink peer memory (skipped on a one-GPU box)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29731", str(ROOT / "tests" / "exchange_worker.py")]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=500)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "exchange worker ok" in out.stdout
