"""Host-side mirror of ``gaussian_renderer/render_helper.py`` (the second caller of the rasterizer:
the per-iteration label render of ``trainers/train_static.py:105``, ``fine_obj.py:174-176`` and
``coarse_obj_pose.py:295-297``).

Same function names, argument meaning and results. ``pc`` is any object with the GaussianModel
attributes read here (``_xyz, _opacity, _scaling, _rotation, _label, active_sh_degree``); the
reference's activated properties (``get_opacity`` = sigmoid, ``get_scaling`` = exp, ``get_rotation``
= normalize, ``scene/gaussian_model.py:126-153``) are evaluated with the same torch ops, and the
rasterizer runs in its scale + rotation / precomputed-colour mode exactly as in the reference
(``render_helper.py:37-66``), so the label image and the gradient that reaches ``_label`` are those
of the drop-in rasterizer.
"""
from __future__ import annotations

import math

import torch

from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer as Renderer

__all__ = ["get_raster_settings", "get_pts_label_as_rgb", "gaussians_to_label_rendervar", "get_render_label"]


def get_raster_settings(viewpoint_camera, pc, bg_color: torch.Tensor, scaling_modifier: float = 1.0):
    """Rasterization settings of a camera without rendering (render_helper.py:7-31)."""
    return GaussianRasterizationSettings(
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


def get_pts_label_as_rgb(gaussians_label: torch.Tensor) -> torch.Tensor:
    """[P,1] label -> [P,3] colour with the label in every channel (render_helper.py:33-38)."""
    return gaussians_label.reshape(-1, 1).float().expand(-1, 3).contiguous()


def gaussians_to_label_rendervar(gaussians) -> dict:
    """Rasterizer inputs of the label render: geometry detached, the label as colour
    (render_helper.py:40-56)."""
    xyz = gaussians._xyz
    return {
        "means3D": xyz.detach(),
        "colors_precomp": get_pts_label_as_rgb(gaussians._label),
        "rotations": torch.nn.functional.normalize(gaussians._rotation).detach(),
        "opacities": torch.sigmoid(gaussians._opacity).detach(),
        "scales": torch.exp(gaussians._scaling).detach(),
        "means2D": torch.zeros_like(xyz, requires_grad=False),
    }


def get_render_label(viewpoint_camera, pc, bg_color: torch.Tensor) -> torch.Tensor:
    """Rendered label image [3,H,W] (render_helper.py:58-66)."""
    label_renderer = Renderer(get_raster_settings(viewpoint_camera, pc, bg_color))
    render_label, _, _, _ = label_renderer(**gaussians_to_label_rendervar(pc))
    return render_label
