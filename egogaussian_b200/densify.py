"""Host-side mirror of the reference's densification (scene/gaussian_model.py:243-272, 506-677) on the
device kernels of csrc/densify.cu and csrc/compact.cu.

The reference keeps its state in a GaussianModel and a torch.optim.Adam; here the state is three plain
dicts, so that any trainer can hold it the way it likes:

    params   name -> tensor      "xyz" [P,3], "f_dc" [P,1,3], "f_rest" [P,M-1,3], "opacity" [P,1],
                                 "scaling" [P,3], "rotation" [P,4], "label" [P,1]   (the optimiser's groups)
    moments  name -> (exp_avg, exp_avg_sq)    the Adam state of the same groups (absent = no state yet)
    extras   "generation" [P,1] int32, "is_object" [P,1] int32

`densify_and_clone` / `densify_and_split` return new dicts (tensors grow, so nothing is updated in place);
the per-iteration bookkeeping tensors (xyz_gradient_accum, denom, max_radii2D) start again from zero
afterwards, as densification_postfix does (:578-580). There is no CPU path.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import cabi

__all__ = ["densify_and_clone", "densify_and_split", "prune_points"]

State = Tuple[Dict[str, torch.Tensor], Dict[str, Tuple[torch.Tensor, torch.Tensor]], Dict[str, torch.Tensor]]


def _grow(selected, params, moments, extras, copies, curr_gen):
    """densification_postfix + cat_tensors_to_optimizer: every tensor gets the selected rows appended
    (`copies` times); moments get zeros (:257-258); generation gets curr_gen when it is an int (:630, :666)."""
    names = list(params)
    tensors = [params[n].contiguous() for n in names]
    fills = ["copy"] * len(names)
    for n in names:
        if n in moments:
            tensors += [moments[n][0].contiguous(), moments[n][1].contiguous()]
            fills += ["zero", "zero"]
    enames = list(extras)
    for n in enames:
        tensors.append(extras[n].contiguous())
        fills.append(int(curr_gen) if n == "generation" and isinstance(curr_gen, int) else "copy")
    outs, count, plan = cabi.append_selected_rows(selected, tensors, fills, copies=copies)
    it = iter(outs)
    new_params = {n: next(it) for n in names}
    new_moments = {}
    for n in names:
        if n in moments:
            new_moments[n] = (next(it), next(it))
    new_extras = {n: next(it) for n in enames}
    return new_params, new_moments, new_extras, count, plan


def densify_and_clone(params, moments, extras, grad_accum, denom, grad_threshold: float, scene_extent: float,
                      percent_dense: float, curr_gen: Optional[int] = None, which_object: Optional[int] = None):
    """GaussianModel.densify_and_clone (:638-678): small Gaussians with a large view-space gradient are
    duplicated in place. Returns (params, moments, extras, n_cloned)."""
    sel = cabi.densify_select(grad_accum=grad_accum.reshape(-1), denom=denom.reshape(-1), scaling_raw=params["scaling"],
                              grad_threshold=grad_threshold, size_threshold=percent_dense * scene_extent, split=False,
                              is_object=extras.get("is_object"), which_object=which_object)
    p, m, e, count, _ = _grow(sel, params, moments, extras, 1, curr_gen)
    return p, m, e, count


def densify_and_split(params, moments, extras, grad_accum, denom, grad_threshold: float, scene_extent: float,
                      percent_dense: float, N: int = 2, curr_gen: Optional[int] = None, split_prev_gen: bool = True,
                      which_object: Optional[int] = None, noise: Optional[torch.Tensor] = None):
    """GaussianModel.densify_and_split (:588-636): large Gaussians with a large gradient are replaced by N
    smaller ones sampled inside them. `noise` [N * n_selected, 3] standard normal (default: torch.randn);
    the reference draws torch.normal(0, scaling), which is scaling * noise. grad_accum / denom may be
    shorter than the model (rows a clone just appended have no statistic: zero, :604-605). Returns
    (params, moments, extras, n_split); the selected originals are removed (:635-636)."""
    P = int(params["xyz"].shape[0])
    sel = cabi.densify_select(grad_accum=grad_accum.reshape(-1), denom=denom.reshape(-1), scaling_raw=params["scaling"],
                              grad_threshold=grad_threshold, size_threshold=percent_dense * scene_extent, split=True,
                              is_object=extras.get("is_object"), which_object=which_object,
                              generation=extras.get("generation"), only_generation=None if split_prev_gen else curr_gen)
    p, m, e, count, plan = _grow(sel, params, moments, extras, N, curr_gen)
    if count:
        if noise is None:
            noise = torch.randn((N * count, 3), device=params["xyz"].device)
        cabi.split_samples(P, plan, count, N, params["xyz"], params["scaling"], params["rotation"], noise.contiguous(),
                           p["xyz"][P:], p["scaling"][P:])
    keep = torch.ones(P + N * count, dtype=torch.uint8, device=sel.device)
    keep[:P] = 1 - sel
    p, m, e = prune_points(keep, p, m, e)
    return p, m, e, count


def prune_points(keep, params, moments, extras):
    """GaussianModel.prune_points / _prune_optimizer (:506-557) on all three dicts: rows with keep != 0 stay."""
    names = list(params)
    tensors = [params[n] for n in names]
    for n in names:
        if n in moments:
            tensors += [moments[n][0], moments[n][1]]
    enames = list(extras)
    tensors += [extras[n] for n in enames]
    outs = cabi.compact_rows(keep, [t.contiguous() for t in tensors])
    it = iter(outs)
    new_params = {n: next(it) for n in names}
    new_moments = {}
    for n in names:
        if n in moments:
            new_moments[n] = (next(it), next(it))
    return new_params, new_moments, {n: next(it) for n in enames}
