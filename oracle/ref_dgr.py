"""Thin caller of the UNMODIFIED reference rasterizer built by oracle/build_ref.py.

TEST INFRASTRUCTURE ONLY (parity checker, golden-vector generator and the `--impl reference`
bench arm). Needs a GPU: the reference has no CPU path.

`forward` / `backward` pass arguments in the order the reference's own Python layer does
(DGR/diff_gaussian_rasterization/__init__.py:63-92, 111-144) and additionally decode the
reference's three opaque byte buffers into named arrays using the `fromChunk` layouts
(DGR/cuda_rasterizer/rasterizer_impl.cu:155-193): every array starts at the
next 128-byte boundary, so all arrays that precede the CUB temporary storage can be located
from (P, R, N) alone.
"""
from __future__ import annotations

import importlib.util
import sys
from pathlib import Path
from typing import Dict

HERE = Path(__file__).resolve().parent
SO = HERE / "_ref" / "ref_dgr_C.so"
_mod = None


def available() -> bool:
    return SO.exists()


def load():
    global _mod
    if _mod is None:
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        if not SO.exists():
            raise ImportError(f"{SO} missing: run `python oracle/build_ref.py` where /root/reference is mounted")
        spec = importlib.util.spec_from_file_location("ref_dgr_C", str(SO))
        _mod = importlib.util.module_from_spec(spec)
        sys.modules["ref_dgr_C"] = _mod
        spec.loader.exec_module(_mod)
    return _mod


def _absent():
    import torch
    return torch.Tensor([])


def _al(x: int, a: int = 128) -> int:
    return (x + a - 1) // a * a


def forward(*, means3D, opacities, bg, viewmatrix, projmatrix, campos, tanfovx, tanfovy, image_height,
            image_width, sh_degree=0, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, scale_modifier=1.0, prefiltered=False, debug=False) -> Dict:
    import torch
    C = load()
    e = _absent()
    args = (bg, means3D, e if colors_precomp is None else colors_precomp, opacities,
            e if scales is None else scales, e if rotations is None else rotations, float(scale_modifier),
            e if cov3D_precomp is None else cov3D_precomp, viewmatrix, projmatrix, float(tanfovx), float(tanfovy),
            int(image_height), int(image_width), e if shs is None else shs, int(sh_degree), campos,
            bool(prefiltered), bool(debug))
    R, color, depth, alpha, radii, geom, binning, img = C.rasterize_gaussians(*args)
    return {"num_rendered": int(R), "color": color, "depth": depth, "alpha": alpha, "radii": radii,
            "geomBuffer": geom, "binningBuffer": binning, "imgBuffer": img,
            "P": int(means3D.shape[0]), "H": int(image_height), "W": int(image_width)}


def backward(fwd: Dict, *, means3D, bg, viewmatrix, projmatrix, campos, tanfovx, tanfovy, dL_dcolor,
             dL_ddepth, dL_dalpha, sh_degree=0, shs=None, colors_precomp=None, scales=None, rotations=None,
             cov3D_precomp=None, scale_modifier=1.0, debug=False, **_ignored) -> Dict:
    C = load()
    e = _absent()
    args = (bg, means3D, fwd["radii"], e if colors_precomp is None else colors_precomp,
            e if scales is None else scales, e if rotations is None else rotations, float(scale_modifier),
            e if cov3D_precomp is None else cov3D_precomp, viewmatrix, projmatrix, float(tanfovx), float(tanfovy),
            dL_dcolor, dL_ddepth, dL_dalpha, e if shs is None else shs, int(sh_degree), campos,
            fwd["geomBuffer"], fwd["num_rendered"], fwd["binningBuffer"], fwd["imgBuffer"], fwd["alpha"], bool(debug))
    g2d, gcol, gop, g3d, gcov, gsh, gsc, grot = C.rasterize_gaussians_backward(*args)
    return {"means2D": g2d, "colors": gcol if colors_precomp is not None else None, "opacity": gop,
            "means3D": g3d, "cov3D": gcov if cov3D_precomp is not None else None,
            "sh": gsh if shs is not None else None, "scales": gsc if scales is not None else None,
            "rotations": grot if rotations is not None else None}


def mark_visible(means3D, viewmatrix, projmatrix):
    return load().mark_visible(means3D, viewmatrix, projmatrix)


def decode_buffers(fwd: Dict) -> Dict:
    """Named views of the reference's intermediates (device tensors; copies, not aliases)."""
    import torch
    P, R, N = fwd["P"], fwd["num_rendered"], fwd["H"] * fwd["W"]
    out = {}
    g = fwd["geomBuffer"]
    base = g.data_ptr()
    off = _al(base) - base

    def take(buf, off, nbytes, dtype, shape):
        t = buf[off:off + nbytes].clone().view(dtype).reshape(shape)
        return t, _al(buf.data_ptr() + off + nbytes) - buf.data_ptr()

    out["depths"], off = take(g, off, 4 * P, torch.float32, (P,))
    out["clamped"], off = take(g, off, 3 * P, torch.uint8, (P, 3))
    _, off = take(g, off, 4 * P, torch.int32, (P,))                       # internal_radii (unused)
    out["means2D"], off = take(g, off, 8 * P, torch.float32, (P, 2))
    out["cov3D"], off = take(g, off, 24 * P, torch.float32, (P, 6))
    out["conic_opacity"], off = take(g, off, 16 * P, torch.float32, (P, 4))
    out["rgb"], off = take(g, off, 12 * P, torch.float32, (P, 3))
    out["tiles_touched"], off = take(g, off, 4 * P, torch.int32, (P,))
    if R > 0:
        b = fwd["binningBuffer"]
        off = _al(b.data_ptr()) - b.data_ptr()
        out["values"], off = take(b, off, 4 * R, torch.int32, (R,))
        _, off = take(b, off, 4 * R, torch.int32, (R,))                   # unsorted values
        out["keys"], off = take(b, off, 8 * R, torch.int64, (R,))
    else:
        out["values"] = torch.empty(0, dtype=torch.int32, device=g.device)
        out["keys"] = torch.empty(0, dtype=torch.int64, device=g.device)
    im = fwd["imgBuffer"]
    off = _al(im.data_ptr()) - im.data_ptr()
    out["n_contrib"], off = take(im, off, 4 * N, torch.int32, (fwd["H"], fwd["W"]))
    tiles = ((fwd["W"] + 15) // 16) * ((fwd["H"] + 15) // 16)
    ranges, off = take(im, off, 8 * N, torch.int32, (N, 2))
    out["ranges"] = ranges[:tiles].clone()
    return out


def load_python_layer():
    """The reference's OWN diff_gaussian_rasterization/__init__.py (staged unmodified into oracle/_ref by
    build_ref.py), imported with the compiled reference standing in for its `_C`. None when the file did
    not travel."""
    py = HERE / "_ref" / "dgr_py" / "diff_gaussian_rasterization" / "__init__.py"
    if not py.exists():
        return None
    name = "ref_dgr_py.diff_gaussian_rasterization"
    if name in sys.modules:
        return sys.modules[name]
    import types
    C = load()
    parent = types.ModuleType("ref_dgr_py")
    parent.__path__ = [str(py.parent.parent)]
    sys.modules["ref_dgr_py"] = parent
    spec = importlib.util.spec_from_file_location(name, str(py), submodule_search_locations=[str(py.parent)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    sys.modules[name + "._C"] = C          # `from . import _C` inside the reference file resolves to the build
    spec.loader.exec_module(mod)
    return mod


def make_autograd_rasterizer():
    """The `--impl reference` bench arm's public call: rasterize(rs, ...) -> (color, radii, depth, alpha),
    differentiable. Goes through the reference's own Python layer (GaussianRasterizationSettings,
    GaussianRasterizer, _RasterizeGaussians of DGR/diff_gaussian_rasterization/__init__.py) whenever
    that file travelled with the snapshot; otherwise through the restatement below, which passes the
    arguments in the same order (DGR/diff_gaussian_rasterization/__init__.py:44-158)."""
    import torch
    C = load()
    ref_py = load_python_layer()
    if ref_py is not None:
        def rasterize_stock(rs, *, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                            rotations=None, cov3D_precomp=None):
            settings = ref_py.GaussianRasterizationSettings(
                image_height=rs["image_height"], image_width=rs["image_width"], tanfovx=rs["tanfovx"], tanfovy=rs["tanfovy"],
                bg=rs["bg"], scale_modifier=rs["scale_modifier"], viewmatrix=rs["viewmatrix"], projmatrix=rs["projmatrix"],
                sh_degree=rs["sh_degree"], campos=rs["campos"], prefiltered=False, debug=False)
            return ref_py.GaussianRasterizer(raster_settings=settings)(
                means3D=means3D, means2D=means2D, opacities=opacities, shs=shs, colors_precomp=colors_precomp,
                scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp)
        rasterize_stock.python_layer = "reference's own diff_gaussian_rasterization/__init__.py"
        return rasterize_stock

    class _Ref(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, means2D, sh, colors, opac, scales, rots, cov, rs):
            out = C.rasterize_gaussians(rs["bg"], means3D, colors, opac, scales, rots, rs["scale_modifier"], cov,
                                        rs["viewmatrix"], rs["projmatrix"], rs["tanfovx"], rs["tanfovy"],
                                        rs["image_height"], rs["image_width"], sh, rs["sh_degree"], rs["campos"],
                                        False, False)
            n, color, depth, alpha, radii, geom, binning, img = out
            ctx.rs, ctx.n = rs, n
            ctx.save_for_backward(colors, means3D, scales, rots, cov, radii, sh, geom, binning, img, alpha)
            return color, radii, depth, alpha

        @staticmethod
        def backward(ctx, g_color, _g_radii, g_depth, g_alpha):
            rs = ctx.rs
            colors, means3D, scales, rots, cov, radii, sh, geom, binning, img, alpha = ctx.saved_tensors
            g2d, gcol, gop, g3d, gcov, gsh, gsc, grot = C.rasterize_gaussians_backward(
                rs["bg"], means3D, radii, colors, scales, rots, rs["scale_modifier"], cov, rs["viewmatrix"],
                rs["projmatrix"], rs["tanfovx"], rs["tanfovy"], g_color, g_depth, g_alpha, sh, rs["sh_degree"],
                rs["campos"], geom, ctx.n, binning, img, alpha, False)
            return g3d, g2d, gsh, gcol, gop, gsc, grot, gcov, None

    def rasterize(rs, *, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                  cov3D_precomp=None):
        e = torch.Tensor([])
        return _Ref.apply(means3D, means2D, e if shs is None else shs, e if colors_precomp is None else colors_precomp,
                          opacities, e if scales is None else scales, e if rotations is None else rotations,
                          e if cov3D_precomp is None else cov3D_precomp, rs)

    rasterize.python_layer = "restatement in oracle/ref_dgr.py (the reference's file did not travel)"
    return rasterize
