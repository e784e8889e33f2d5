"""Host-side mirror of the reference's render glue.

``render(viewpoint_camera, pc, pipe, bg_color, ...)`` has the reference's signature, argument
meaning and return dict (``gaussian_renderer/__init__.py:18-107``) and works on any ``pc`` that
carries the GaussianModel attributes it reads (``_xyz, _scaling, _rotation, _opacity,
_features_dc, _features_rest, _is_object, active_sh_degree, max_sh_degree,
trainable_object_move``). What the reference derives from those raw parameters with ~25 eager
torch kernels per iteration and an autograd graph of the same length --

    get_opacity / get_scaling / get_features              scene/gaussian_model.py:126-153
    get_covariance / get_rotated_covariance               scene/gaussian_model.py:29-63,167-171
    build_rotation / build_scaling_rotation / strip_symmetric   utils/general_utils.py:110-158

-- is one CUDA launch each way here (``egs_model_forward`` / ``egs_model_backward`` behind the
C-ABI, exposed as the autograd function ``model_activations``), followed by the rasterizer.

There is no CPU path; the module raises at import when the extension is missing.
"""
from __future__ import annotations

import math
import weakref
from typing import Optional

import torch

from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _C

__all__ = ["render", "rasterize_raw", "model_activations", "get_covariance", "get_rotated_covariance", "rot6d_to_matrix",
           "row0_copies"]

_EMPTY = torch.empty(0)


def rot6d_to_matrix(rot6d: torch.Tensor) -> torch.Tensor:
    """Gram-Schmidt of the [3,2] pose parameter, columns (b1, b2, b1 x b2)
    (utils/geometry_utils.py:69-89). Nine numbers: stays in torch so autograd reaches rot6d."""
    a1, a2 = rot6d[:, 0], rot6d[:, 1]
    b1 = torch.nn.functional.normalize(a1, dim=0)
    b2 = torch.nn.functional.normalize(a2 - torch.dot(b1, a2) * b1, dim=0)
    return torch.stack((b1, b2, torch.linalg.cross(b1, b2)), dim=-1)


_copies_cache: dict = {}


def row0_copies(is_object: Optional[torch.Tensor], which_object: Optional[int], P: int) -> int:
    """Number of entries of the reference's (row, 0) index list that name row 0
    (scene/gaussian_model.py:49-59; see include/egs_rasterizer.h). It only changes when
    ``_is_object`` does (densify / prune), so the one device read it costs is cached on the
    tensor's identity and version."""
    if which_object is None or is_object is None:
        return P + 1 if P > 0 else 0
    key = (id(is_object), int(which_object))
    hit = _copies_cache.get(key)
    # the entry is only valid for the very same live tensor object at the same version (an id or a
    # device address can be reused by another tensor once this one is gone)
    if hit is not None and hit[0]() is is_object and hit[1] == is_object._version:
        return hit[2]
    sel = is_object.reshape(-1) == which_object
    n = int(sel.sum())
    copies = n + int(bool(sel[0])) if n > 0 else 0
    if len(_copies_cache) > 64:
        _copies_cache.clear()
    _copies_cache[key] = (weakref.ref(is_object), is_object._version, copies)
    return copies


class _ModelActivations(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest, scale_modifier,
                is_object, which_object, accum_R, train_R, copies):
        cov3D, opacity, shs = _C.model_forward(scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest,
                                               scale_modifier, is_object, which_object, accum_R, train_R, copies)
        ctx.save_for_backward(scaling_raw, rotation_raw, opacity_raw, is_object, accum_R, train_R)
        ctx.meta = (scale_modifier, which_object, copies, shs.shape[1] if shs.numel() else 0)
        return cov3D, opacity, shs

    @staticmethod
    def backward(ctx, g_cov, g_opacity, g_shs):
        scaling_raw, rotation_raw, opacity_raw, is_object, accum_R, train_R = ctx.saved_tensors
        scale_modifier, which_object, copies, M = ctx.meta
        need = ctx.needs_input_grad

        def grad(g, wanted):
            return g if (g is not None and wanted and g.numel()) else _EMPTY

        ds, dq, dop, ddc, drest, dT = _C.model_backward(
            scaling_raw, rotation_raw, opacity_raw, M, scale_modifier, is_object, which_object, accum_R, train_R,
            copies, grad(g_cov, need[0] or need[1] or need[9]), grad(g_opacity, need[2]),
            grad(g_shs, need[3] or need[4]))

        def out(t, i):
            return t if (need[i] and t.numel()) else None

        return (out(ds, 0), out(dq, 1), out(dop, 2), out(ddc, 3), out(drest, 4), None, None, None, None,
                out(dT, 9), None)


def model_activations(scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest, scale_modifier=1.0,
                      is_object=None, which_object=None, accum_R=None, train_R=None, rotate=False, quirk=True):
    """(cov3D [P,6], opacity [P,1], shs [P,M,3]) from the raw parameters, differentiable.

    ``rotate`` selects get_rotated_covariance semantics: rows with ``is_object == which_object``
    (all rows when ``which_object`` is None) use ``train_R @ (accum_R @ L)``. ``quirk`` keeps the
    reference's treatment of row 0 (see ``row0_copies``)."""
    mask, which, accum_R, train_R, copies = _selection(scaling_raw.shape[0], scaling_raw.device, is_object,
                                                       which_object, accum_R, train_R, rotate, quirk)
    return _ModelActivations.apply(scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest,
                                   float(scale_modifier), mask, which, accum_R, train_R, copies)


def _selection(P, dev, is_object, which_object, accum_R, train_R, rotate, quirk):
    """Arguments of the object-rotation part as the extension wants them:
    (mask int32 or empty, which, accum_R or empty, train_R or empty, row0_copies)."""
    copies = 0
    mask = _EMPTY
    if rotate:
        if accum_R is None and train_R is None:
            # gaussian_model.py:54-55: the identity leaves L unchanged exactly, but the rows still go
            # through the index list, so row 0's gradient multiplicity applies
            accum_R = torch.eye(3, device=dev)
        if which_object is not None:
            mask = is_object
            if mask.dtype != torch.int32:
                mask = mask.to(torch.int32)
        copies = row0_copies(is_object, which_object, P) if quirk else 0
    else:
        accum_R = train_R = None

    def dev_or_empty(t):
        return _EMPTY if t is None else t.to(device=dev, dtype=torch.float32)

    return mask, -1 if which_object is None else int(which_object), dev_or_empty(accum_R), dev_or_empty(train_R), copies


class _RasterizeRaw(torch.autograd.Function):
    """Rasterizer on the raw parameters (egs_forward_raw / egs_backward_raw): activations, covariance
    and the dc/rest split live inside the per-Gaussian kernels, nothing intermediate reaches HBM."""

    @staticmethod
    def forward(ctx, means3D, means2D, colors, scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest,
                mask, which, accum_R, train_R, copies, rs):
        out = _C.rasterize_gaussians_raw(rs.bg, means3D, colors, scaling_raw, rotation_raw, opacity_raw, features_dc,
                                         features_rest, rs.scale_modifier, mask, which, accum_R, train_R, copies,
                                         rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                                         rs.image_width, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        num_rendered, color, depth, alpha, radii, geom, binning, img = out
        ctx.rs, ctx.meta = rs, (num_rendered, which, copies)
        ctx.save_for_backward(means3D, colors, scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest,
                              mask, accum_R, train_R, radii, geom, binning, img, alpha)
        ctx.mark_non_differentiable(radii)
        ctx.set_materialize_grads(False)   # no zero-filled "gradient" for the integer radii
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, g_color, _g_radii, g_depth, g_alpha):
        (means3D, colors, scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest, mask, accum_R, train_R,
         radii, geom, binning, img, alpha) = ctx.saved_tensors
        if g_color is None:
            g_color = torch.zeros_like(alpha.expand(3, -1, -1))
        empty = alpha.new_empty(0)   # "no gradient": the tile kernel then leaves the depth / alpha terms out
        g_depth = empty if g_depth is None else g_depth
        g_alpha = empty if g_alpha is None else g_alpha
        rs = ctx.rs
        num_rendered, which, copies = ctx.meta
        need = ctx.needs_input_grad
        g = _C.rasterize_gaussians_raw_backward(
            rs.bg, means3D, radii, colors, scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest,
            rs.scale_modifier, mask, which, accum_R, train_R, copies, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
            rs.tanfovy, g_color, g_depth, g_alpha, rs.sh_degree, rs.campos, geom, num_rendered, binning, img, alpha,
            bool(need[11]), rs.debug)
        d_means2D, d_means3D, d_colors, ds, dq, dop, ddc, drest, dT = g

        def out(t, i):
            return t if (need[i] and t.numel()) else None

        return (out(d_means3D, 0), out(d_means2D, 1), out(d_colors, 2), out(ds, 3), out(dq, 4), out(dop, 5),
                out(ddc, 6), out(drest, 7), None, None, None, out(dT, 11), None, None)


def rasterize_raw(raster_settings, means3D, means2D, scaling_raw, rotation_raw, opacity_raw, features_dc=None,
                  features_rest=None, colors_precomp=None, is_object=None, which_object=None, accum_R=None,
                  train_R=None, rotate=False, quirk=True):
    """(color, radii, depth, alpha) straight from the raw GaussianModel parameters; the covariance
    semantics are those of ``model_activations`` (get_covariance / get_rotated_covariance)."""
    if (features_dc is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    mask, which, accum_R, train_R, copies = _selection(means3D.shape[0], means3D.device, is_object, which_object,
                                                       accum_R, train_R, rotate, quirk)
    e = _EMPTY
    return _RasterizeRaw.apply(means3D, means2D, e if colors_precomp is None else colors_precomp, scaling_raw,
                               rotation_raw, opacity_raw, e if features_dc is None else features_dc,
                               e if features_rest is None else features_rest, mask, which, accum_R, train_R, copies,
                               raster_settings)


def get_covariance(pc, scaling_modifier: float = 1.0) -> torch.Tensor:
    """GaussianModel.get_covariance (scene/gaussian_model.py:167-168)."""
    return model_activations(pc._scaling, pc._rotation, _EMPTY, _EMPTY, _EMPTY, scaling_modifier)[0]


def get_rotated_covariance(pc, accum_R, which_object, during_training, scaling_modifier: float = 1.0) -> torch.Tensor:
    """GaussianModel.get_rotated_covariance (scene/gaussian_model.py:170-171)."""
    train_R = rot6d_to_matrix(pc.trainable_object_move.obj_rotation_6d) if during_training else None
    return model_activations(pc._scaling, pc._rotation, _EMPTY, _EMPTY, _EMPTY, scaling_modifier,
                             is_object=pc._is_object, which_object=which_object, accum_R=accum_R,
                             train_R=train_R, rotate=True)[0]


def render(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier: float = 1.0,
           override_color=None, rot_cov: bool = False, accum_R=None, which_object=None,
           during_training: bool = False, fused: bool = True):
    """Render one view (gaussian_renderer/__init__.py:18-107). ``bg_color`` must be on the GPU.

    With ``pipe.compute_cov3D_python`` (forced on by the reference's trainers) the activations and
    the covariance build run inside the rasterizer's per-Gaussian kernels (``fused=True``, default)
    or as their own kernel pair in front of it (``fused=False``); both render identically."""
    xyz = pc._xyz
    screenspace_points = torch.zeros_like(xyz, requires_grad=True)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height),
        image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5),
        tanfovy=math.tan(viewpoint_camera.FoVy * 0.5),
        bg=bg_color,
        scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center,
        prefiltered=False,
        debug=False)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)

    if getattr(pipe, "convert_SHs_python", False) and override_color is None:
        raise NotImplementedError("convert_SHs_python=True (eval_sh in torch) is not on the accelerated path; "
                                  "the reference never enables it (arguments/__init__.py:66)")
    want_shs = override_color is None
    dc = pc._features_dc if want_shs else _EMPTY
    rest = pc._features_rest if want_shs else _EMPTY

    scales = rotations = cov3D_precomp = None
    if pipe.compute_cov3D_python and fused:
        train_R = None
        if rot_cov and during_training:
            train_R = rot6d_to_matrix(pc.trainable_object_move.obj_rotation_6d)
        rendered_image, radii, rendered_depth, rendered_alpha = rasterize_raw(
            raster_settings, xyz, screenspace_points, pc._scaling, pc._rotation, pc._opacity,
            features_dc=dc if want_shs else None, features_rest=rest if want_shs else None,
            colors_precomp=override_color, is_object=pc._is_object, which_object=which_object, accum_R=accum_R,
            train_R=train_R, rotate=rot_cov)
        return {"render": rendered_image,
                "viewspace_points": screenspace_points,
                "visibility_filter": radii > 0,
                "radii": radii,
                "depth": rendered_depth,
                "alpha": rendered_alpha}
    if pipe.compute_cov3D_python:
        train_R = None
        if rot_cov and during_training:
            train_R = rot6d_to_matrix(pc.trainable_object_move.obj_rotation_6d)
        cov3D_precomp, opacity, shs = model_activations(
            pc._scaling, pc._rotation, pc._opacity, dc, rest, scaling_modifier, is_object=pc._is_object,
            which_object=which_object, accum_R=accum_R, train_R=train_R, rotate=rot_cov)
    else:
        _, opacity, shs = _activations_only(pc, dc, rest)
        scales = torch.exp(pc._scaling)
        rotations = torch.nn.functional.normalize(pc._rotation)

    rendered_image, radii, rendered_depth, rendered_alpha = rasterizer(
        means3D=xyz,
        means2D=screenspace_points,
        shs=shs if want_shs else None,
        colors_precomp=override_color,
        opacities=opacity,
        scales=scales,
        rotations=rotations,
        cov3D_precomp=cov3D_precomp)

    return {"render": rendered_image,
            "viewspace_points": screenspace_points,
            "visibility_filter": radii > 0,
            "radii": radii,
            "depth": rendered_depth,
            "alpha": rendered_alpha}


def add_densification_stats(pc, viewspace_point_tensor: torch.Tensor, radii: torch.Tensor) -> None:
    """The trainers' per-iteration bookkeeping in one launch and without a host sync
    (trainers/train_static_bg.py:128-129; GaussianModel.add_densification_stats,
    scene/gaussian_model.py:735-740): for the rows with ``radii > 0`` updates ``pc.max_radii2D``,
    ``pc.xyz_gradient_accum`` and ``pc.denom`` in place. ``radii`` is the rasterizer's output;
    the visibility filter is derived from it, as the reference does (``radii > 0``)."""
    with torch.no_grad():
        _C.densify_stats(radii, viewspace_point_tensor.grad, pc.max_radii2D, pc.xyz_gradient_accum, pc.denom)


def _activations_only(pc, dc, rest):
    """opacity and features without the covariance (compute_cov3D_python=False)."""
    return _ModelActivations.apply(_EMPTY, _EMPTY, pc._opacity, dc, rest, 1.0, _EMPTY, -1, _EMPTY, _EMPTY, 0)
