"""GPU parity tests: the CUDA path, called through the C-ABI, against
  (a) golden vectors produced by the unmodified reference on a B200 (tests/golden/*.npz),
  (b) the live reference when oracle/_ref/ref_dgr_C.so travelled with the snapshot,
  (c) the CPU oracle on the same seeded inputs,
and size-independent properties at BASELINE.json's full sizes.

Bars (BASELINE.json north_star): radii, tile ranges, sorted keys and values bit-exact; colour,
depth, alpha within 1e-5 absolute; gradients within 1e-4 of the tensor's max magnitude.
"""
from pathlib import Path

import numpy as np
import pytest
import torch

from helpers import BASELINE_CASES, GRAD_RTOL, IMAGE_ATOL, baseline_case, grad_rel_err, run_cuda, to_np
from golden.make_golden import CASES, case_inputs

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"


def _assert_images(ours, ref, atol=IMAGE_ATOL, what=""):
    for k in ("color", "depth", "alpha"):
        a, b = to_np(getattr(ours, k)), to_np(ref[k])
        assert a.shape == b.shape
        err = np.abs(a - b).max()
        assert err <= atol, f"{what}{k}: max abs err {err}"


def _assert_grads(ours, ref, rtol=GRAD_RTOL, what=""):
    checked = 0
    for k in ("means2D", "means3D", "opacity", "colors", "cov3D", "sh", "scales", "rotations"):
        r = ref.get(k) if isinstance(ref, dict) else None
        if r is None:
            continue
        o = ours[k]
        assert o is not None, f"{what}gradient {k} missing"
        e = grad_rel_err(to_np(o).reshape(to_np(r).shape), r)
        assert e <= rtol, f"{what}grad {k}: rel err {e}"
        checked += 1
    assert checked >= 4


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_against_golden_reference_vectors(case, built, cuda_device):
    path = GOLDEN / f"{case[0]}.npz"
    if not path.exists():
        pytest.skip("golden vectors not generated yet")
    gold = np.load(path)
    kw, ups = case_inputs(case, cuda_device)
    fwd, grads = run_cuda(kw, ups)
    torch.cuda.synchronize()
    assert fwd.num_rendered == int(gold["num_rendered"])
    assert np.array_equal(to_np(fwd.radii), gold["radii"]), "radii must be bit-exact"
    bv, iv, gv = fwd.binning_view(), fwd.image_view(), fwd.geom_view()
    assert np.array_equal(to_np(bv["keys"]), gold["i_keys"]), "sorted keys must be bit-exact"
    assert np.array_equal(to_np(bv["values"]), gold["i_values"]), "sorted values must be bit-exact"
    assert np.array_equal(to_np(iv["ranges"]), gold["i_ranges"]), "tile ranges must be bit-exact"
    vis = gold["radii"] > 0
    assert np.array_equal(to_np(gv["tiles_touched"]), gold["i_tiles_touched"])
    assert np.array_equal(to_np(gv["means2D"])[vis].view(np.uint32), gold["i_means2D"][vis].view(np.uint32))
    assert np.array_equal(to_np(gv["depths"])[vis].view(np.uint32), gold["i_depths"][vis].view(np.uint32))
    assert np.abs(to_np(gv["conic_opacity"])[vis] - gold["i_conic_opacity"][vis]).max() <= 1e-6 * np.abs(gold["i_conic_opacity"][vis]).max()
    _assert_images(fwd, gold, what="golden ")
    assert np.array_equal(to_np(iv["n_contrib"]), gold["i_n_contrib"])
    _assert_grads(grads, {k[2:]: gold[k] for k in gold.files if k.startswith("g_")}, what="golden ")


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_against_live_reference(case, built, cuda_device):
    from oracle import ref_dgr
    if not ref_dgr.available():
        pytest.skip("oracle/_ref/ref_dgr_C.so not present")
    kw, ups = case_inputs(case, cuda_device)
    fwd, grads = run_cuda(kw, ups)
    rf = ref_dgr.forward(**kw)
    dec = ref_dgr.decode_buffers(rf)
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    rg = ref_dgr.backward(rf, **bkw, **ups)
    torch.cuda.synchronize()
    assert fwd.num_rendered == rf["num_rendered"]
    assert torch.equal(fwd.radii, rf["radii"])
    bv, iv = fwd.binning_view(), fwd.image_view()
    assert torch.equal(bv["keys"], dec["keys"]) and torch.equal(bv["values"], dec["values"])
    assert torch.equal(iv["ranges"], dec["ranges"])
    _assert_images(fwd, rf, what="live ref ")
    _assert_grads(grads, rg, what="live ref ")


_baseline_cache = {}


def _baseline_inputs(case, dev):
    """One scene per case for the whole session: building 1M Gaussians on the host takes seconds."""
    if case[0] not in _baseline_cache:
        _baseline_cache.clear()   # cases arrive grouped; keep one scene resident at a time
        _baseline_cache[case[0]] = baseline_case(*case[1:], device=dev)
    return _baseline_cache[case[0]]


@pytest.mark.parametrize("hint", [False, True], ids=["exact", "hinted"])
@pytest.mark.parametrize("case", BASELINE_CASES, ids=[c[0] for c in BASELINE_CASES])
def test_against_live_reference_at_baseline_sizes(case, hint, built, cuda_device):
    """BASELINE.json configs[0..2] against the unmodified reference run on the same inputs
    (rasterizer_impl.cu:197-447): radii, sorted keys, sorted values, tile ranges and n_contrib
    identical; colour / depth / alpha within 1e-5; every gradient within 1e-4 of its maximum.
    `hinted` passes num_rendered_hint = 1.2 x the count, the way the torch binding does from the
    second iteration on: at the C3 cases that is the depth presort + one partition + register split,
    the path the benchmark number is measured on; `exact` is the first-iteration path."""
    from egogaussian_b200 import cabi
    from oracle import ref_dgr
    if not ref_dgr.available():
        pytest.skip("oracle/_ref/ref_dgr_C.so not present")
    kw, ups = _baseline_inputs(case, cuda_device)
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    rf = ref_dgr.forward(**kw)
    dec = ref_dgr.decode_buffers(rf)
    # the reference's atomics make two of its own runs differ by up to 6e-5 of the maximum at these sizes
    # (the scale gradient at configs[1]; profiles/r2_parity_vs_reference.json lists it per tensor): the
    # mean of three runs is the target, so that the 1e-4 bound is spent on this implementation
    runs = [ref_dgr.backward(rf, **bkw, **ups) for _ in range(3)]
    rg = {k: (None if v is None else sum(r[k] for r in runs) / len(runs)) for k, v in runs[0].items()}
    R = rf["num_rendered"]
    fwd = cabi.forward(num_rendered_hint=int(R * 1.2) + 1 if hint else 0, **kw)
    grads = cabi.backward(fwd, **bkw, **ups)
    torch.cuda.synchronize()
    assert fwd.num_rendered == R
    assert torch.equal(fwd.radii, rf["radii"]), "radii"
    gv, bv, iv = fwd.geom_view(), fwd.binning_view(), fwd.image_view()
    vis = rf["radii"] > 0
    assert torch.equal(gv["tiles_touched"], dec["tiles_touched"]), "tiles_touched"
    assert torch.equal(gv["depths"][vis].view(torch.int32), dec["depths"][vis].view(torch.int32)), "depths"
    assert torch.equal(gv["means2D"][vis].view(torch.int32), dec["means2D"][vis].view(torch.int32)), "means2D"
    assert torch.equal(bv["keys"], dec["keys"]), "sorted keys"
    assert torch.equal(bv["values"], dec["values"]), "sorted values"
    assert torch.equal(iv["ranges"], dec["ranges"]), "tile ranges"
    assert torch.equal(iv["n_contrib"], dec["n_contrib"]), "n_contrib"
    _assert_images(fwd, rf, what=f"{case[0]} ")
    _assert_grads(grads, rg, what=f"{case[0]} ")


def test_mark_visible_against_live_reference(built, cuda_device):
    """markVisible through the product's own binding and nn.Module (DGR/rasterize_points.cu:210-229,
    DGR/diff_gaussian_rasterization/__init__.py:179-188) against the reference's on 100k points, a
    fifth of them within a few ulp of the z_view = 0.2 plane."""
    from egogaussian_b200 import scenes
    from egogaussian_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _C
    from oracle import ref_dgr
    if not ref_dgr.available():
        pytest.skip("oracle/_ref/ref_dgr_C.so not present")
    dev = cuda_device
    gen = torch.Generator().manual_seed(5)
    P = 100_000
    pts = torch.randn(P, 3, generator=gen) * torch.tensor([3.0, 3.0, 4.0])
    edge = torch.arange(0, P, 5)
    pts[edge, 2] = 0.2 + (torch.randint(-4, 5, (edge.numel(),), generator=gen).float() * 1.5e-8)
    pts[1] = float("nan")
    pts[2] = float("inf")
    pts = pts.to(dev)
    for view_id in (None, 3):   # the camera at the origin (z_view == z exactly) and a yawed, shifted one
        sc = scenes.make_scene(64, 320, 200, sh_degree=0, seed=9, view=view_id)
        cam = sc.camera
        view, proj = cam.viewmatrix.to(dev), cam.projmatrix.to(dev)
        want = ref_dgr.mark_visible(pts, view, proj)
        got = _C.mark_visible(pts, view, proj)
        assert got.dtype == torch.bool and got.shape == (P,)
        assert torch.equal(got, want), view_id
        assert 0.2 < float(want.float().mean()) < 0.8
        rs = GaussianRasterizationSettings(200, 320, cam.tanfovx, cam.tanfovy, sc.bg.to(dev), 1.0, view, proj, 0,
                                           cam.campos.to(dev), False, False)
        assert torch.equal(GaussianRasterizer(rs).markVisible(pts), want), view_id
    # empty input, as the reference binding handles it
    assert _C.mark_visible(torch.empty(0, 3, device=dev), view, proj).numel() == 0


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_against_cpu_oracle(case, built, cuda_device):
    """Stage-wise: the CPU build cannot reproduce nvcc's FMA contraction, so its own radii may
    differ where 3*sqrt(lambda) sits on an integer; the later stages are therefore fed the GPU's
    preprocess output, which makes them exactly comparable."""
    from oracle import cpu_oracle as orc
    kw, ups = case_inputs(case, cuda_device)
    fwd, grads = run_cuda(kw, ups)
    torch.cuda.synchronize()
    ckw = {k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
    own = orc.render_forward(**ckw)
    # (1) preprocess: radii equal except on rounding boundaries; floats to rounding
    r_gpu = to_np(fwd.radii)
    diff = r_gpu != own["radii"]
    frac = np.abs(own["radius_f"] - np.round(own["radius_f"]))
    assert np.all(frac[diff] < 1e-3) and diff.mean() < 2e-3, "radii differ away from a rounding boundary"
    gv = {k: to_np(v) for k, v in fwd.geom_view().items()}
    vis = (r_gpu > 0) & ~diff
    assert np.allclose(gv["means2D"][vis], own["means2D"][vis], rtol=0, atol=2e-3)
    assert np.allclose(gv["conic_opacity"][vis], own["conic_opacity"][vis], rtol=2e-4, atol=1e-6)
    assert np.allclose(gv["rgb_depth"][vis][:, :3], own["rgb"][vis], rtol=0, atol=1e-5)
    # (2) binning from the GPU's preprocess output: bit-exact
    clamped = np.stack([(gv["clamped"] >> c) & 1 for c in range(3)], axis=1).astype(np.uint8)
    pre = dict(radii=r_gpu, means2D=gv["means2D"], depths=gv["depths"], conic_opacity=gv["conic_opacity"],
               rgb=np.ascontiguousarray(gv["rgb_depth"][:, :3]), clamped=clamped,
               tiles_touched=gv["tiles_touched"].astype(np.uint32), cov3D=own["cov3D"], radius_f=own["radius_f"])
    staged = orc.render_forward(preprocessed=pre, **ckw)
    bv, iv = fwd.binning_view(), fwd.image_view()
    assert staged["num_rendered"] == fwd.num_rendered
    assert np.array_equal(to_np(bv["keys"]).view(np.uint64), staged["keys"])
    assert np.array_equal(to_np(bv["values"]).view(np.uint32), staged["values"])
    assert np.array_equal(to_np(iv["ranges"]).view(np.uint32), staged["ranges"])
    # (3) blend forward
    _assert_images(fwd, staged, what="oracle ")
    nc = to_np(iv["n_contrib"]).view(np.uint32)
    assert (nc != staged["n_contrib"]).mean() < 1e-3
    # (4) backward from the same forward state, including the GPU's alpha image: T_final =
    # 1 - alpha amplifies the last bit of alpha (backward.cu:463), see tests/test_oracle_golden.py
    staged["alpha"] = to_np(fwd.alpha)
    staged["n_contrib"] = nc
    bk = {k: v for k, v in ckw.items() if k != "opacities"}
    og = orc.render_backward(staged, **bk, **{k: v.cpu() for k, v in ups.items()})
    _assert_grads(grads, og, what="oracle ")


def _surface_inputs(case, dev):
    kw, ups = case_inputs(case, dev)
    return kw, ups


@pytest.mark.parametrize("case", [CASES[0], CASES[2]], ids=[CASES[0][0], CASES[2][0]])
def test_python_surface_equals_cabi(case, built, cuda_device):
    """GaussianRasterizer (autograd, torch extension) gives the C-ABI's numbers and gradients."""
    from egogaussian_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    kw, ups = case_inputs(case, cuda_device)
    fwd, grads = run_cuda(kw, ups)
    rs = GaussianRasterizationSettings(kw["image_height"], kw["image_width"], kw["tanfovx"], kw["tanfovy"], kw["bg"],
                                       1.0, kw["viewmatrix"], kw["projmatrix"], kw["sh_degree"], kw["campos"],
                                       False, False)
    leaf = {k: kw[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "colors_precomp", "scales",
                                                            "rotations", "cov3D_precomp") if k in kw}
    means2D = torch.zeros_like(leaf["means3D"], requires_grad=True)
    color, radii, depth, alpha = GaussianRasterizer(rs)(
        means3D=leaf["means3D"], means2D=means2D, opacities=leaf["opacities"], shs=leaf.get("shs"),
        colors_precomp=leaf.get("colors_precomp"), scales=leaf.get("scales"), rotations=leaf.get("rotations"),
        cov3D_precomp=leaf.get("cov3D_precomp"))
    assert torch.equal(color, fwd.color) and torch.equal(depth, fwd.depth) and torch.equal(alpha, fwd.alpha)
    assert torch.equal(radii, fwd.radii)
    ((color * ups["dL_dcolor"]).sum() + (depth * ups["dL_ddepth"]).sum() + (alpha * ups["dL_dalpha"]).sum()).backward()
    pairs = {"means3D": "means3D", "opacities": "opacity", "shs": "sh", "colors_precomp": "colors", "scales": "scales",
             "rotations": "rotations", "cov3D_precomp": "cov3D"}
    for name, t in leaf.items():
        assert t.grad is not None, name
        assert grad_rel_err(t.grad.reshape(grads[pairs[name]].shape), grads[pairs[name]]) <= GRAD_RTOL
    assert grad_rel_err(means2D.grad, grads["means2D"]) <= GRAD_RTOL
    assert torch.all(means2D.grad[:, 2] == 0)
    # alpha-only loss (coarse_obj_pose.py:255-260 style): color / depth grads arrive as None
    for t in list(leaf.values()) + [means2D]:
        t.grad = None
    c2, _, _, a2 = GaussianRasterizer(rs)(
        means3D=leaf["means3D"], means2D=means2D, opacities=leaf["opacities"], shs=leaf.get("shs"),
        colors_precomp=leaf.get("colors_precomp"), scales=leaf.get("scales"), rotations=leaf.get("rotations"),
        cov3D_precomp=leaf.get("cov3D_precomp"))
    a2.abs().mean().backward()
    assert leaf["opacities"].grad is not None and torch.isfinite(leaf["opacities"].grad).all()


def test_edge_cases(built, cuda_device):
    from egogaussian_b200 import cabi, scenes
    dev = cuda_device
    cam = scenes.make_camera(37, 21)   # not a multiple of the 16x16 tile
    base = dict(bg=torch.tensor([0.3, 0.5, 0.7], device=dev), viewmatrix=cam.viewmatrix.to(dev),
                projmatrix=cam.projmatrix.to(dev), campos=cam.campos.to(dev), tanfovx=cam.tanfovx,
                tanfovy=cam.tanfovy, image_height=21, image_width=37)
    # P == 0: outputs are zero (rasterize_points.cu:84 skips everything)
    e3 = torch.empty((0, 3), device=dev)
    f = cabi.forward(means3D=e3, opacities=torch.empty((0, 1), device=dev), colors_precomp=e3,
                     cov3D_precomp=torch.empty((0, 6), device=dev), **base)
    assert f.num_rendered == 0 and float(f.color.abs().max()) == 0.0 and f.radii.numel() == 0
    # everything culled: background only, num_rendered == 0, backward gives zeros
    m = torch.tensor([[0.0, 0.0, 0.1], [0.0, 0.0, -3.0]], device=dev)
    kw = dict(means3D=m, opacities=torch.ones(2, 1, device=dev), colors_precomp=torch.ones(2, 3, device=dev),
              scales=torch.full((2, 3), 0.1, device=dev),
              rotations=torch.tensor([[1.0, 0, 0, 0]] * 2, device=dev), **base)
    ups = dict(dL_dcolor=torch.ones(3, 21, 37, device=dev), dL_ddepth=torch.ones(1, 21, 37, device=dev),
               dL_dalpha=torch.ones(1, 21, 37, device=dev))
    f, g = run_cuda(kw, ups)
    assert f.num_rendered == 0 and torch.all(f.radii == 0)
    assert torch.allclose(f.color[:, 3, 5], base["bg"]) and float(f.alpha.abs().max()) == 0.0
    for k in ("means3D", "means2D", "opacity", "colors", "scales", "rotations"):
        assert float(g[k].abs().max()) == 0.0, k
    # one huge opaque Gaussian: every pixel saturates at alpha 0.99 after one contributor
    kw = dict(means3D=torch.tensor([[0.0, 0.0, 2.0]], device=dev), opacities=torch.ones(1, 1, device=dev),
              colors_precomp=torch.tensor([[1.0, 0.0, 0.0]], device=dev), scales=torch.full((1, 3), 500.0, device=dev),
              rotations=torch.tensor([[1.0, 0, 0, 0]], device=dev), **base)
    f, g = run_cuda(kw, ups)
    assert f.num_rendered == 3 * 2
    assert torch.all(f.image_view()["n_contrib"] == 1)
    assert torch.allclose(f.alpha, torch.full_like(f.alpha, 0.99), atol=1e-6)
    assert torch.allclose(f.depth, torch.full_like(f.depth, 1.98), atol=1e-5)
    # mark_visible: z_view > 0.2
    vis = cabi.mark_visible(torch.tensor([[0, 0, 0.2], [0, 0, 0.21], [0, 0, -1.0]], device=dev),
                            base["viewmatrix"], base["projmatrix"])
    assert vis.tolist() == [False, True, False]
    # argument errors come back as EgsError with the library's message
    with pytest.raises(cabi.EgsError, match="exactly one of shs"):
        cabi.forward(means3D=kw["means3D"], opacities=kw["opacities"], scales=kw["scales"],
                     rotations=kw["rotations"], **base)


def test_full_size_properties(built, cuda_device):
    """Config 3 size (1M Gaussians, SH3, 1280x720): properties that need no oracle."""
    from egogaussian_b200 import cabi, scenes
    dev = cuda_device
    sc = scenes.make_scene(1_000_000, 1280, 720, sh_degree=3, seed=3001).to(dev)
    cam = sc.camera
    kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
              projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
              image_height=720, image_width=1280, sh_degree=3, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    f1 = cabi.forward(**kw)
    gv, bv, iv = f1.geom_view(), f1.binning_view(), f1.image_view()
    R = f1.num_rendered
    assert R == int(gv["tiles_touched"].to(torch.int64).sum())
    keys, vals = bv["keys"], bv["values"]
    assert bool((keys[1:] >= keys[:-1]).all()), "sorted"
    ties = keys[1:] == keys[:-1]
    assert bool((vals[1:][ties] > vals[:-1][ties]).all()), "stable"
    # the sorted list is a permutation of what duplication produced: same multiset of values and
    # the depth bits in each key belong to its value
    counts = torch.bincount(vals.to(torch.int64), minlength=sc.P)
    assert torch.equal(counts, gv["tiles_touched"].to(torch.int64))
    assert torch.equal((keys & 0xFFFFFFFF).to(torch.int32), gv["depths"].view(torch.int32)[vals.to(torch.int64)])
    tiles = (keys >> 32)
    rng = iv["ranges"].to(torch.int64)
    tile_counts = torch.bincount(tiles, minlength=rng.shape[0])
    assert torch.equal(rng[:, 1] - rng[:, 0], tile_counts)
    nonempty = tile_counts > 0
    assert torch.equal(rng[nonempty, 0], (torch.cumsum(tile_counts, 0) - tile_counts)[nonempty])
    # image sanity: alpha in [0, 1], colour finite, n_contrib within the tile's list
    assert bool(torch.isfinite(f1.color).all()) and float(f1.alpha.min()) >= 0 and float(f1.alpha.max()) <= 1 + 1e-5
    # idempotence / determinism of the forward
    f2 = cabi.forward(**kw)
    assert torch.equal(f1.color, f2.color) and torch.equal(f1.radii, f2.radii)
    assert torch.equal(f1.binning_view()["values"], f2.binning_view()["values"])
    # linearity of backward in the upstream gradient: g(2w) = 2 g(w)
    wc, wd, wa = [t.to(dev) for t in scenes.upstream_weights(1280, 720, 1)]
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    g1 = cabi.backward(f1, dL_dcolor=wc, dL_ddepth=wd, dL_dalpha=wa, **bkw)
    g2 = cabi.backward(f1, dL_dcolor=2 * wc, dL_ddepth=2 * wd, dL_dalpha=2 * wa, **bkw)
    for k in ("means3D", "sh", "opacity", "scales", "rotations"):
        assert grad_rel_err(g2[k], 2 * g1[k]) <= GRAD_RTOL, k
    assert bool(torch.all(g1["means3D"][f1.radii == 0] == 0))


@pytest.fixture
def debug_options():
    """Library switches (egs_debug_option) are process-wide: hand out the setter, restore the defaults after."""
    from egogaussian_b200 import cabi
    yield cabi.debug_option
    for name in ("sort", "presort", "presort_scramble", "blend_forward", "blend_backward"):
        cabi.debug_option(name, None)


def test_binning_paths_agree(built, cuda_device, debug_options):
    """All binning paths give identical sorted lists, ranges and images: the presorted bucket path
    (depth presort, tile | index words, one global partition, register split), the full-word bucket
    path (presort off: one CTA sorts each bucket), the packed-word LSD sort ("packed"; taken when R is
    too large for the bucket sorter), the (u64 key, u32 value) pair sort ("pairs"; taken when
    tile | depth | index does not fit one 64-bit word and the presort is off) and the plain 8-bit pair
    sort ("pairs8"). A second scene with a few huge Gaussians makes some buckets exceed one chunk of
    the split kernel and the shared memory of the bucket sorter; a third puts 150k Gaussians on sixteen
    tiles, so that nearly all 32 words of a warp share a digit in every round of the partition's ranking
    (its shared-memory counters are then updated by back-to-back atomics on the same address, whose
    order is what keeps the pass stable)."""
    from egogaussian_b200 import cabi, scenes
    dev = cuda_device
    for P, W, H, seed, boost in ((60_000, 640, 360, 77, 1.0), (40_000, 256, 144, 78, 6.0), (150_000, 64, 64, 79, 1.0)):
        sc = scenes.make_scene(P, W, H, sh_degree=1, seed=seed)
        sc.scales = (sc.scales * boost).contiguous()
        sc = sc.to(dev)
        cam = sc.camera
        kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
                  projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                  image_height=H, image_width=W, sh_degree=1, shs=sc.shs, cov3D_precomp=sc.cov3D())
        debug_options("sort", None)
        debug_options("presort", "1")   # these scenes are below the size where it is used by itself
        a = cabi.forward(**kw)
        assert a.num_rendered > 0
        av = a.binning_view()
        keys = av["keys"]
        assert bool((keys[1:] >= keys[:-1]).all())
        ties = keys[1:] == keys[:-1]
        assert bool((av["values"][1:][ties] > av["values"][:-1][ties]).all()), "ties keep index order"
        debug_options("presort", "0")
        for forced in (None, "packed", "pairs", "pairs8"):
            debug_options("sort", forced)
            b = cabi.forward(**kw)
            assert a.num_rendered == b.num_rendered
            bv = b.binning_view()
            assert torch.equal(av["keys"], bv["keys"]) and torch.equal(av["values"], bv["values"]), forced
            assert torch.equal(a.image_view()["ranges"], b.image_view()["ranges"]), forced
            assert torch.equal(a.color, b.color) and torch.equal(a.radii, b.radii), forced
        debug_options("sort", None)
        # the presort's veto (a depth beyond its 30-bit key range): the presorted path skips itself and
        # forward falls back to a path that does not rely on the order; with and without a guess
        debug_options("presort", "1")
        debug_options("presort_scramble", "1")
        for hint in (0, int(a.num_rendered * 1.2)):
            c = cabi.forward(num_rendered_hint=hint, **kw)
            cv = c.binning_view()
            assert torch.equal(av["keys"], cv["keys"]) and torch.equal(av["values"], cv["values"]), ("veto", hint)
            assert torch.equal(a.image_view()["ranges"], c.image_view()["ranges"]) and torch.equal(a.color, c.color)
        debug_options("presort_scramble", None)
        gv = a.geom_view()
        # the scan runs over the Gaussians front to back (presort) or in index order: same total
        assert int(gv["point_offsets"][-1]) == int(gv["tiles_touched"].to(torch.int64).sum())
        if boost > 1.0:   # the stress scene really has buckets beyond one chunk / the shared-memory capacity
            tiles = (keys >> 32)
            n_tiles = ((W + 15) // 16) * ((H + 15) // 16)
            shift = max(n_tiles.bit_length() - 10, 0)
            assert int(torch.bincount(tiles >> shift).max()) > 12288


def test_presorted_path_beyond_a_million_gaussians(built, cuda_device, debug_options):
    """BASELINE configs[4] territory: 2M Gaussians at 1280x720 need 21 index bits, so tile | depth | index
    no longer fits 64 bits. The presorted path (64-bit tile | index words here) must give exactly the
    reference's lists (rasterizer_impl.cu:70-111, 301-319: sorted keys, values, ranges) and images, and
    the same as our pair sort, which is what such a frame gets when the presort is off."""
    from egogaussian_b200 import cabi
    from oracle import ref_dgr
    kw, ups = baseline_case(2_000_000, 0, 1280, 720, 3001, "cov", "sh", cuda_device)
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    f = cabi.forward(**kw)
    R = f.num_rendered
    assert R > 10_000_000   # beyond what the full-word bucket path accepts, too
    fv, iv = f.binning_view(), f.image_view()
    keys, vals = fv["keys"].clone(), fv["values"].clone()
    assert bool((keys[1:] >= keys[:-1]).all())
    ties = keys[1:] == keys[:-1]
    assert bool((vals[1:][ties] > vals[:-1][ties]).all())
    h = cabi.forward(num_rendered_hint=int(R * 1.1), **kw)
    assert torch.equal(h.binning_view()["values"], vals) and torch.equal(h.image_view()["ranges"], iv["ranges"])
    assert torch.equal(h.color, f.color)
    debug_options("presort", "0")
    p = cabi.forward(**kw)
    debug_options("presort", None)
    pv = p.binning_view()
    assert torch.equal(pv["keys"], keys) and torch.equal(pv["values"], vals)
    assert torch.equal(p.image_view()["ranges"], iv["ranges"]) and torch.equal(p.color, f.color)
    del p, h
    if not ref_dgr.available():
        pytest.skip("oracle/_ref/ref_dgr_C.so not present: compared with our pair sort only")
    rf = ref_dgr.forward(**kw)
    dec = ref_dgr.decode_buffers(rf)
    assert rf["num_rendered"] == R and torch.equal(f.radii, rf["radii"])
    assert torch.equal(keys, dec["keys"]), "sorted keys"
    assert torch.equal(vals, dec["values"]), "sorted values"
    assert torch.equal(iv["ranges"], dec["ranges"]), "tile ranges"
    assert torch.equal(iv["n_contrib"], dec["n_contrib"])
    _assert_images(f, rf, what="2M ")
    grads = cabi.backward(f, **bkw, **ups)
    rg = ref_dgr.backward(rf, **bkw, **ups)
    _assert_grads(grads, rg, what="2M ")


def test_wide_images_take_the_generic_split(built, cuda_device, debug_options):
    """More than 8192 tiles leave more than three tile bits below the 10-bit bucket digit: the split
    kernel then ranks with match.any and per-warp counters instead of one ballot per bin. 4096x2304
    (36864 tiles, 6 bits) against the pair sort."""
    from egogaussian_b200 import cabi, scenes
    W, H = 4096, 2304
    sc = scenes.make_scene(150_000, W, H, sh_degree=0, seed=5).to(cuda_device)
    cam = sc.camera
    kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
              projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
              image_height=H, image_width=W, sh_degree=0, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    debug_options("presort", "1")
    a = cabi.forward(**kw)
    debug_options("presort", "0")
    debug_options("sort", "pairs")
    b = cabi.forward(**kw)
    assert a.num_rendered == b.num_rendered > 1_000_000
    av, bv = a.binning_view(), b.binning_view()
    assert torch.equal(av["values"], bv["values"]) and torch.equal(av["keys"], bv["keys"])
    assert torch.equal(a.image_view()["ranges"], b.image_view()["ranges"]) and torch.equal(a.color, b.color)


def test_prefiltered_fault_is_reported(built, cuda_device):
    """prefiltered=True promises that no Gaussian will be culled; the reference prints and traps the
    kernel when one is (auxiliary.h:154-161). Here forward returns EGS_ERR_PREFILTERED with the
    reference's message, and a scene that keeps the promise renders as without the flag."""
    from egogaussian_b200 import cabi
    kw, _ = case_inputs(CASES[0], cuda_device)
    with pytest.raises(cabi.EgsError, match="filtered although prefiltered"):
        cabi.forward(prefiltered=True, **kw)          # the scene has Gaussians behind the near plane
    base = cabi.forward(**kw)
    front = (kw["means3D"][:, 2] > 0.25).nonzero().squeeze(1)
    kw2 = {k: (v[front].contiguous() if isinstance(v, torch.Tensor) and v.shape[:1] == kw["means3D"].shape[:1] else v)
           for k, v in kw.items()}
    a, b = cabi.forward(prefiltered=True, **kw2), cabi.forward(**kw2)
    assert a.num_rendered == b.num_rendered == base.num_rendered
    assert torch.equal(a.color, b.color) and torch.equal(a.color, base.color)


def test_num_rendered_hint_never_changes_results(built, cuda_device):
    """EgsForwardArgs.num_rendered_hint only moves the host wait: a generous guess, a tight one, one
    that is too small (the enqueued work skips itself and the exact path re-runs after a second
    binning_alloc) and no guess at all give identical buffers, images and gradients."""
    from egogaussian_b200 import cabi
    kw, ups = case_inputs(CASES[0], cuda_device)
    base, gbase = run_cuda(kw, ups)
    R = base.num_rendered
    assert R > 4096
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    for hint in (R, R + 1, 2 * R + 12345, R - 1, R // 3, 1):
        f = cabi.forward(num_rendered_hint=hint, **kw)
        assert f.num_rendered == R, hint
        if hint >= R:   # the guessed workspace was kept
            assert f.binning.numel() == cabi.load().egs_binning_bytes(hint, f.W, f.H), hint
        else:
            assert f.binning.numel() == base.binning.numel(), hint
        for k in ("color", "depth", "alpha", "radii"):
            assert torch.equal(getattr(f, k), getattr(base, k)), (hint, k)
        fv, bv = f.binning_view(), base.binning_view()
        assert torch.equal(fv["keys"], bv["keys"]) and torch.equal(fv["values"], bv["values"]), hint
        fi, bi = f.image_view(), base.image_view()
        assert torch.equal(fi["ranges"], bi["ranges"]) and torch.equal(fi["n_contrib"], bi["n_contrib"]), hint
        g = cabi.backward(f, dL_dcolor=ups["dL_dcolor"], dL_ddepth=ups["dL_ddepth"],
                          dL_dalpha=ups["dL_dalpha"], **bkw)
        for k in ("means2D", "means3D", "opacity", "sh", "scales", "rotations"):
            assert grad_rel_err(g[k], gbase[k]) <= GRAD_RTOL, (hint, k)


def test_forward_kernel_variants_agree(built, cuda_device, debug_options):
    """The second forward blend kept as an independent implementation (two pixels per thread, a
    log-domain pre-filter, no bit lists) is bit-identical to the default one."""
    kw, ups = case_inputs(CASES[0], cuda_device)
    f1, _ = run_cuda(kw, None)
    debug_options("blend_forward", "v2")
    f2, _ = run_cuda(kw, None)
    for k in ("color", "depth", "alpha"):
        assert torch.equal(getattr(f1, k), getattr(f2, k)), k
    assert torch.equal(f1.image_view()["n_contrib"], f2.image_view()["n_contrib"])


def test_backward_kernel_variants_agree(built, cuda_device, debug_options):
    """The backward blend with two pixels per thread (four warps per tile) and with four (two warps per
    tile) sums the same fragments in another order: equal within the gradient tolerance, also for the
    colour-only backward of the label render."""
    from egogaussian_b200 import cabi
    for case in (CASES[0], CASES[2]):
        kw, ups = case_inputs(case, cuda_device)
        debug_options("blend_backward", "4")
        _, g4 = run_cuda(kw, ups)
        debug_options("blend_backward", "2")
        _, g2 = run_cuda(kw, ups)
        debug_options("blend_backward", None)
        for k in ("means2D", "means3D", "opacity", "sh", "colors", "scales", "rotations", "cov3D"):
            if g4.get(k) is not None:
                assert grad_rel_err(g2[k], g4[k]) <= GRAD_RTOL, (case[0], k)


@pytest.mark.parametrize("ppl", ["2", "4"])
def test_backward_without_depth_and_alpha_gradients(ppl, built, cuda_device, debug_options):
    """A loss on the colour image only: the C-ABI takes NULL for dL/ddepth and dL/dalpha (the tile kernel
    then leaves their recurrences out) and the autograd surface passes exactly that when autograd hands it
    None. Both must give what zero-filled gradients give (the reference's way), also when only one of
    the two is absent."""
    from egogaussian_b200 import cabi
    from egogaussian_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    debug_options("blend_backward", ppl)
    for case in (CASES[0], CASES[2]):
        kw, ups = case_inputs(case, cuda_device)
        zero = dict(ups, dL_ddepth=torch.zeros_like(ups["dL_ddepth"]), dL_dalpha=torch.zeros_like(ups["dL_dalpha"]))
        _, want = run_cuda(kw, zero)
        fwd = cabi.forward(**kw)
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        got = cabi.backward(fwd, dL_dcolor=ups["dL_dcolor"], dL_ddepth=None, dL_dalpha=None, **bkw)
        for k, v in want.items():
            if v is not None:
                assert grad_rel_err(got[k], v) <= 2e-5, (case[0], k)
        # only the alpha gradient absent
        _, want1 = run_cuda(kw, dict(ups, dL_dalpha=torch.zeros_like(ups["dL_dalpha"])))
        got1 = cabi.backward(fwd, dL_dcolor=ups["dL_dcolor"], dL_ddepth=ups["dL_ddepth"], dL_dalpha=None, **bkw)
        for k, v in want1.items():
            if v is not None:
                assert grad_rel_err(got1[k], v) <= 2e-5, (case[0], k, "alpha absent")
        # through autograd: a loss that touches the colour image only
        leaves = {k: kw[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "colors_precomp", "scales",
                                                                  "rotations", "cov3D_precomp") if kw.get(k) is not None}
        rs = GaussianRasterizationSettings(kw["image_height"], kw["image_width"], kw["tanfovx"], kw["tanfovy"], kw["bg"], 1.0,
                                           kw["viewmatrix"], kw["projmatrix"], kw["sh_degree"], kw["campos"], False, False)
        color, _, _, _ = GaussianRasterizer(rs)(
            means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"], requires_grad=True),
            opacities=leaves["opacities"], shs=leaves.get("shs"), colors_precomp=leaves.get("colors_precomp"),
            scales=leaves.get("scales"), rotations=leaves.get("rotations"), cov3D_precomp=leaves.get("cov3D_precomp"))
        (color * ups["dL_dcolor"]).sum().backward()
        names = {"means3D": "means3D", "opacities": "opacity", "shs": "sh", "colors_precomp": "colors", "scales": "scales",
                 "rotations": "rotations", "cov3D_precomp": "cov3D"}
        for k, t in leaves.items():
            assert grad_rel_err(t.grad.reshape(want[names[k]].shape), want[names[k]]) <= 2e-5, (case[0], k, "autograd")


def test_backward_contributor_set_matches_forward(built, cuda_device):
    """A fragment sitting exactly on the alpha = 1/255 threshold must be in or out in BOTH passes: the
    backward decides in the log domain, the forward with o * expf(power); inside a small band around
    the threshold the backward therefore uses the forward's expression. 256 Gaussians are stacked on
    the optical axis of a 65x65 image, so that their centres hit pixel (32, 32) exactly (power == 0,
    alpha == opacity there), with opacities within +-4 ulp of 1/255; a random scene surrounds them.
    Gradients are compared with the reference (live when present, else the CPU oracle's backward)."""
    import numpy as np
    from egogaussian_b200 import scenes
    dev = cuda_device
    W = H = 65
    sc = scenes.make_scene(1500, W, H, sh_degree=0, seed=321)
    n_axis = 256
    thr = np.float32(1.0) / np.float32(255.0)
    ulps = (np.arange(n_axis) % 9 - 4).astype(np.int32)
    o_axis = (np.full(n_axis, thr, np.float32).view(np.int32) + ulps).view(np.float32)
    means = torch.cat([torch.stack([torch.zeros(n_axis), torch.zeros(n_axis), 2.0 + 0.01 * torch.arange(n_axis)], 1),
                       sc.means3D])
    P = means.shape[0]
    gen = torch.Generator().manual_seed(4)
    cam = sc.camera
    kw = dict(means3D=means.to(dev), opacities=torch.cat([torch.from_numpy(o_axis)[:, None], sc.opacities]).to(dev),
              bg=torch.zeros(3, device=dev), viewmatrix=cam.viewmatrix.to(dev), projmatrix=cam.projmatrix.to(dev),
              campos=cam.campos.to(dev), tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, image_height=H, image_width=W,
              sh_degree=0, colors_precomp=torch.rand(P, 3, generator=gen).to(dev),
              scales=torch.cat([torch.full((n_axis, 3), 0.15), sc.scales]).to(dev),
              rotations=torch.cat([torch.tensor([[1.0, 0, 0, 0]]).repeat(n_axis, 1), sc.rotations]).to(dev))
    wc, wd, wa = [t.to(dev) for t in scenes.upstream_weights(W, H, 5)]
    ups = dict(dL_dcolor=wc, dL_ddepth=wd, dL_dalpha=wa)
    fwd, grads = run_cuda(kw, ups)
    gv = fwd.geom_view()
    assert bool((gv["means2D"][:n_axis] == 32.0).all()), "the axis Gaussians must sit on a pixel centre"
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    from oracle import ref_dgr
    if ref_dgr.available():
        rf = ref_dgr.forward(**kw)
        assert torch.equal(fwd.image_view()["n_contrib"], ref_dgr.decode_buffers(rf)["n_contrib"])
        assert torch.equal(fwd.alpha, rf["alpha"])
        rg = ref_dgr.backward(rf, **bkw, **ups)
    else:
        from oracle import cpu_oracle as orc
        ckw = {k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
        staged = orc.render_forward(**ckw)
        staged["alpha"] = to_np(fwd.alpha)
        rg = orc.render_backward(staged, **{k: v for k, v in ckw.items() if k != "opacities"},
                                 **{k: v.cpu() for k, v in ups.items()})
    _assert_grads(grads, rg, what="threshold ")
    # the axis Gaussians below the threshold took no part at pixel (32, 32): their opacity gradient
    # there is zero in both implementations, those at or above it have one
    g_op = to_np(grads["opacity"])[:n_axis, 0]
    r_op = to_np(rg["opacity"]).reshape(-1)[:n_axis]
    assert np.array_equal(g_op != 0, r_op != 0)


def test_view_batch_step_with_object_pose(built, cuda_device):
    """configs[3] in miniature (SURVEY.md 8e parity definition): object Gaussians are moved by a rigid
    pose, two views are rendered, their gradients accumulated in view order in ONE flat buffer, the
    positions' gradient is pulled back through the pose and one Adam step is taken. Ground truth:
    the reference rasterizer (live, when it travelled) or our own C-ABI per view, torch autograd
    through the pose and torch.optim.Adam."""
    from egogaussian_b200 import cabi, parallel, scenes
    from egogaussian_b200.diff_gaussian_rasterization import _C
    from oracle import ref_dgr
    dev = cuda_device
    P, W, H, deg = 3000, 96, 64, 1
    base = scenes.make_scene(P, W, H, sh_degree=deg, seed=91)
    base.scales = (base.scales * 1.5).contiguous()
    Rt = scenes.object_pose().to(dev)
    R, t = Rt[:9].reshape(3, 3), Rt[9:]
    xyz = base.means3D.to(dev)
    is_obj = base.is_object.to(dev)
    moved, _ = _C.se3_apply(xyz, is_obj, 1, Rt, torch.empty(0, device=dev))
    M = (deg + 1) ** 2
    flat = parallel.FlatGrads(P, M, dev)
    total = torch.zeros_like(flat.flat)
    ref_total = {k: torch.zeros_like(v) for k, v in flat.views.items()}
    for view in (0, 1):
        sc = scenes.make_scene(P, W, H, sh_degree=deg, seed=91, view=view).to(dev)
        cam = sc.camera
        kw = dict(means3D=moved, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
                  projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                  image_height=H, image_width=W, sh_degree=deg, shs=sc.shs,
                  scales=(sc.scales * 1.5).contiguous(), rotations=sc.rotations)
        wc, wd, wa = [x.to(dev) for x in scenes.upstream_weights(W, H, 91 + view)]
        ups = dict(dL_dcolor=wc, dL_ddepth=wd, dL_dalpha=wa)
        bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
        fwd = cabi.forward(**kw)
        cabi.backward(fwd, out=flat.as_backward_out(), **bkw, **ups)
        total += flat.flat                      # accumulation in view order (one GPU, two views)
        if ref_dgr.available():
            rf = ref_dgr.forward(**kw)
            rg = ref_dgr.backward(rf, **bkw, **ups)
        else:
            rg = cabi.backward(fwd, **bkw, **ups)
        for k in ref_total:
            ref_total[k] += rg[k].reshape(ref_total[k].shape)
    flat.flat.copy_(total)
    for k, v in flat.views.items():
        assert grad_rel_err(v, ref_total[k]) <= GRAD_RTOL, k
    # optimizer: canonical positions, gradient pulled back through the pose
    lr = 1.6e-4
    p_ref = xyz.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=lr, eps=1e-15)
    p_ref.grad = torch.where(is_obj[:, None] == 1, ref_total["means3D"] @ R, ref_total["means3D"])
    opt.step()
    p = xyz.clone()
    m1, v1 = torch.zeros_like(p), torch.zeros_like(p)
    _C.adam_se3_step(p, flat.views["means3D"].contiguous(), m1, v1, is_obj, 1, Rt, lr, 1, 0.9, 0.999, 1e-15)
    # Adam's first step is lr * sign(g): compare where the gradient is not at the noise floor
    g = ref_total["means3D"].abs()
    solid = g > 1e-3 * g.max()
    assert torch.allclose(p[solid], p_ref.detach()[solid], rtol=1e-6, atol=4e-6 * lr)
    assert float((p - p_ref.detach()).abs().max()) <= 2.01 * lr


def test_adam_and_se3(built, cuda_device):
    from egogaussian_b200 import scenes
    from egogaussian_b200.diff_gaussian_rasterization import _C
    dev = cuda_device
    torch.manual_seed(0)
    shapes = [(5000, 3), (5000, 1, 3), (5000, 15, 3), (5000, 1), (5000, 3), (5000, 4), (7,)]
    lrs = [1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3, 1e-4]
    ref_p = [torch.randn(s, device=dev).requires_grad_(True) for s in shapes]
    opt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(ref_p, lrs)], lr=0.0, eps=1e-15)
    ours = [p.detach().clone() for p in ref_p]
    m = [torch.zeros_like(p) for p in ours]
    v = [torch.zeros_like(p) for p in ours]
    for step in range(1, 4):
        grads = [torch.randn(s, device=dev) * (10.0 ** -step) for s in shapes]
        for p, g in zip(ref_p, grads):
            p.grad = g.clone()
        opt.step()
        _C.adam_step(ours, grads, m, v, lrs, [step] * len(ours), 0.9, 0.999, 1e-15)
        # each step moves a parameter by about lr; fp32 rounding of that update is ~1e-6 of it
        for a, b, lr in zip(ours, ref_p, lrs):
            assert torch.allclose(a, b.detach(), rtol=1e-6, atol=4e-6 * lr * step), (lr, step)
    # rigid transform and its pull-back
    P = 4000
    xyz = torch.randn(P, 3, device=dev)
    is_obj = (torch.rand(P, device=dev) < 0.1).to(torch.int32)
    Rt = scenes.object_pose().to(dev)
    R, t = Rt[:9].reshape(3, 3), Rt[9:]
    cov = scenes.build_covariance(torch.rand(P, 3) * 0.1 + 0.01, torch.nn.functional.normalize(torch.randn(P, 4), dim=1)).to(dev)
    moved, cov_moved = _C.se3_apply(xyz, is_obj, 1, Rt, cov)
    want = torch.where(is_obj[:, None] == 1, xyz @ R.t() + t, xyz)
    assert torch.allclose(moved, want, atol=1e-6)
    S = torch.zeros(P, 3, 3, device=dev)
    iu = [(0, 0, 0), (0, 1, 1), (0, 2, 2), (1, 1, 3), (1, 2, 4), (2, 2, 5)]
    for a, b, k in iu:
        S[:, a, b] = cov[:, k]
        S[:, b, a] = cov[:, k]
    S2 = R @ S @ R.t()
    want_cov = torch.stack([S2[:, a, b] for a, b, _ in iu], dim=1)
    assert torch.allclose(cov_moved, torch.where(is_obj[:, None] == 1, want_cov, cov), atol=1e-6)
    # reference semantics: un-move then compare (fine_all.py:116 tolerance 1e-3)
    back = torch.where(is_obj[:, None] == 1, (moved - t) @ R, moved)
    assert torch.allclose(back, xyz, atol=1e-3)
    g_moved = torch.randn(P, 3, device=dev)
    p_ref = xyz.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=1.6e-4, eps=1e-15)
    p_ref.grad = torch.where(is_obj[:, None] == 1, g_moved @ R, g_moved)   # R^T g as rows
    opt.step()
    p = xyz.clone()
    m1, v1 = torch.zeros_like(p), torch.zeros_like(p)
    _C.adam_se3_step(p, g_moved, m1, v1, is_obj, 1, Rt, 1.6e-4, 1, 0.9, 0.999, 1e-15)
    assert torch.allclose(p, p_ref.detach(), rtol=1e-6, atol=4e-6 * 1.6e-4)


def test_forward_lists_are_identical_across_calls(built, cuda_device):
    """The sorted instance list is a pure function of the inputs: exact-size and guessed-size calls,
    repeated, give the same count, keys and values (look-back orders, atomics and the depth presort
    must never leak into the result)."""
    from egogaussian_b200 import cabi, scenes
    for P, W, H in ((200_000, 640, 360), (1_000_000, 1280, 720)):
        sc = scenes.make_scene(P, W, H, sh_degree=0, seed=11).to(cuda_device)
        cam = sc.camera
        kw = dict(means3D=sc.means3D, opacities=sc.opacities, bg=sc.bg, viewmatrix=cam.viewmatrix,
                  projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                  image_height=H, image_width=W, sh_degree=0, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
        first = None
        for i in range(12):
            f = cabi.forward(**kw) if i % 2 == 0 else cabi.forward(num_rendered_hint=int(first[0] * 1.2), **kw)
            bv = f.binning_view()
            sig = (f.num_rendered, bv["keys"].clone(), bv["values"].clone())
            if first is None:
                first = sig
            else:
                assert sig[0] == first[0], (P, i)
                assert torch.equal(sig[1], first[1]) and torch.equal(sig[2], first[2]), (P, i)


def test_one_thread_renders_on_two_devices(built):
    """The pinned read-back slot and its event are kept per (thread, device): a thread that renders on
    cuda:0 and then on cuda:1 (single-process multi-GPU, or an evaluation pass on another device) must
    get the same results on both. Skipped on a one-GPU box."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from egogaussian_b200 import cabi
    results = []
    for _ in range(2):
        for d in (0, 1):
            dev = torch.device("cuda", d)
            kw, ups = case_inputs(CASES[0], dev)
            with torch.cuda.device(dev):
                fwd, grads = run_cuda(kw, ups)
                torch.cuda.synchronize(dev)
            results.append((fwd.num_rendered, fwd.color.cpu(), grads["means3D"].cpu()))
    for r in results[1:]:
        assert r[0] == results[0][0] and torch.equal(r[1], results[0][1])
        assert grad_rel_err(r[2], results[0][2]) <= GRAD_RTOL
