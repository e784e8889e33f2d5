"""Drop-in Python surface of the B200-native rasterizer.

Same public names, argument order, return values and error messages as the reference package
(`submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py`, cited as
DGR/py below), written fresh on top of our `_C` extension:

* ``GaussianRasterizationSettings``  -- DGR/py:160-172 (12-field NamedTuple)
* ``GaussianRasterizer``             -- DGR/py:174-223 (``markVisible``, ``forward``)
* ``rasterize_gaussians``            -- DGR/py:21-42
* ``_RasterizeGaussians``            -- DGR/py:44-158 (autograd Function)

``gaussian_renderer/__init__.py:90-98`` and ``gaussian_renderer/render_helper.py:63`` call
``GaussianRasterizer(...)(means3D=..., means2D=..., shs=..., colors_precomp=..., opacities=...,
scales=..., rotations=..., cov3D_precomp=...)`` and unpack ``(color, radii, depth, alpha)``.

There is no CPU path: importing this module without the compiled extension raises.
"""
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

try:
    from . import _C
except ImportError as exc:  # fail loudly: a silent fallback would void every parity claim
    raise ImportError(
        "egogaussian_b200: the CUDA extension diff_gaussian_rasterization/_C.so is missing or "
        "cannot be loaded; build it with `python -m egogaussian_b200.build` "
        "(or `python -c 'import __graft_entry__ as g; g.build()'`)") from exc

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def cpu_deep_copy_tuple(input_tuple):
    """Host snapshot of an argument tuple, used by the ``debug`` dumps (DGR/py:17-19)."""
    return tuple(x.cpu().clone() if isinstance(x, torch.Tensor) else x for x in input_tuple)


def _call_with_snapshot(fn, args, debug, dump_name, what):
    """Run ``fn(*args)``; with ``debug`` set, save the inputs to ``dump_name`` if it throws."""
    if not debug:
        return fn(*args)
    saved = cpu_deep_copy_tuple(args)  # copy first: a faulting kernel may corrupt the inputs
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print("\nAn error occured in %s. Writing %s for debugging.\n" % (what, dump_name))
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered,
                rs.debug)
        (num_rendered, color, depth, alpha, radii, geom_blob, binning_blob, image_blob) = \
            _call_with_snapshot(_C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")

        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom_blob, binning_blob, image_blob, alpha)
        ctx.mark_non_differentiable(radii)
        # backward fills in the missing upstream gradients itself; left on, autograd would also
        # zero-fill a "gradient" for the integer radii on every iteration
        ctx.set_materialize_grads(False)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom_blob,
         binning_blob, image_blob, alpha) = ctx.saved_tensors
        # autograd hands None for outputs that took no part in the loss: the colour gradient is then
        # zeros; depth and alpha go to the binding as empty tensors, and the tile kernel leaves their
        # terms out (the reference zero-fills all three, __init__.py:100-158)
        if grad_color is None:
            grad_color = torch.zeros_like(alpha.expand(3, -1, -1))
        empty = alpha.new_empty(0)
        grad_depth = empty if grad_depth is None else grad_depth
        grad_alpha = empty if grad_alpha is None else grad_alpha

        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color,
                grad_depth, grad_alpha, sh, rs.sh_degree, rs.campos, geom_blob, ctx.num_rendered,
                binning_blob, image_blob, alpha, rs.debug)
        # Only the precomputed colours want a gradient (the trainers' label render,
        # gaussian_renderer/render_helper.py:40-66): both backward kernels skip everything else.
        need = ctx.needs_input_grad
        colors_only = colors_precomp.numel() != 0 and need[3] and not any(need[i] for i in (0, 1, 2, 4, 5, 6, 7))
        fn = _C.rasterize_gaussians_backward_colors if colors_only else _C.rasterize_gaussians_backward
        (g_means2D, g_colors, g_opacities, g_means3D, g_cov3D, g_sh, g_scales, g_rotations) = \
            _call_with_snapshot(fn, args, rs.debug, "snapshot_bw.dump", "backward")
        if colors_only:
            return (None, None, None, g_colors, None, None, None, None, None)

        def present(g, inp):  # absent inputs were empty tensors and get no gradient
            return g if inp.numel() != 0 else None
        return (g_means3D, g_means2D, present(g_sh, sh), present(g_colors, colors_precomp),
                g_opacities, present(g_scales, scales), present(g_rotations, rotations),
                present(g_cov3D, cov3Ds_precomp), None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions: torch.Tensor) -> torch.Tensor:
        """bool[P]: view-space z > 0.2 for the current camera (DGR/py:179-188)."""
        rs = self.raster_settings
        with torch.no_grad():
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs: Optional[torch.Tensor] = None,
                colors_precomp: Optional[torch.Tensor] = None, scales: Optional[torch.Tensor] = None,
                rotations: Optional[torch.Tensor] = None, cov3D_precomp: Optional[torch.Tensor] = None):
        # same checks and messages as DGR/py:194-198 (typo included: callers may match on it)
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_pair = scales is not None and rotations is not None
        have_any = scales is not None or rotations is not None
        if (not have_pair and cov3D_precomp is None) or (have_any and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        absent = torch.Tensor([])
        return rasterize_gaussians(
            means3D, means2D,
            absent if shs is None else shs,
            absent if colors_precomp is None else colors_precomp,
            opacities,
            absent if scales is None else scales,
            absent if rotations is None else rotations,
            absent if cov3D_precomp is None else cov3D_precomp,
            self.raster_settings)
