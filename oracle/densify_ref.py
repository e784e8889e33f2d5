"""Torch restatement of the reference's densification for the parity tests. TEST INFRASTRUCTURE ONLY.

Follows scene/gaussian_model.py: densify_and_clone :638-678, densify_and_split :588-636,
densification_postfix :559-586, cat_tensors_to_optimizer :243-272, prune_points / _prune_optimizer
:506-557, build_rotation utils/general_utils.py:124-149 -- on plain dicts (see egogaussian_b200/densify.py
for their layout), with the split's normal samples written as scaling * noise for a given standard-normal
`noise`. Pinned against tests/golden/densify_*.npz, which the reference's own code produced
(tests/golden/make_densify_golden.py).
"""
from __future__ import annotations

import torch


def _grads(grad_accum, denom, P):
    g = (grad_accum / denom).reshape(-1)
    g[g.isnan()] = 0.0
    padded = torch.zeros(P, device=g.device)
    padded[:g.shape[0]] = g
    return padded


def _append(params, moments, extras, new, new_gen):
    p = {n: torch.cat((t, new[n]), dim=0) for n, t in params.items()}
    m = {n: (torch.cat((a, torch.zeros_like(new[n])), dim=0), torch.cat((b, torch.zeros_like(new[n])), dim=0))
         for n, (a, b) in moments.items()}
    e = {"generation": torch.cat((extras["generation"], new_gen), dim=0),
         "is_object": torch.cat((extras["is_object"], new["is_object"]), dim=0)}
    return p, m, e


def densify_and_clone(params, moments, extras, grad_accum, denom, grad_threshold, scene_extent, percent_dense,
                      curr_gen=None, which_object=None):
    P = params["xyz"].shape[0]
    g = _grads(grad_accum, denom, P)
    sel = (g.abs() >= grad_threshold) & (torch.exp(params["scaling"]).max(dim=1).values <= percent_dense * scene_extent)
    if which_object is not None:
        sel = sel & (extras["is_object"].reshape(-1) == which_object)
    new = {n: t[sel] for n, t in params.items()}
    new["is_object"] = extras["is_object"][sel]
    gen = extras["generation"][sel]
    if isinstance(curr_gen, int):
        gen = torch.full_like(gen, curr_gen)
    return (*_append(params, moments, extras, new, gen), int(sel.sum()))


def build_rotation(r):
    q = r / r.norm(dim=1, keepdim=True)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.reshape(-1, 3, 3)


def densify_and_split(params, moments, extras, grad_accum, denom, grad_threshold, scene_extent, percent_dense, noise,
                      N=2, curr_gen=None, split_prev_gen=True, which_object=None):
    P = params["xyz"].shape[0]
    g = _grads(grad_accum, denom, P)
    scale = torch.exp(params["scaling"])
    sel = (g >= grad_threshold) & (scale.max(dim=1).values > percent_dense * scene_extent)
    if not split_prev_gen:
        sel = sel & (extras["generation"].reshape(-1) == curr_gen)
    if which_object is not None:
        sel = sel & (extras["is_object"].reshape(-1) == which_object)
    stds = scale[sel].repeat(N, 1)
    samples = stds * noise
    rots = build_rotation(params["rotation"][sel]).repeat(N, 1, 1)
    new = {n: t[sel].repeat(N, *([1] * (t.dim() - 1))) for n, t in params.items()}
    new["xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + params["xyz"][sel].repeat(N, 1)
    new["scaling"] = torch.log(scale[sel].repeat(N, 1) / (0.8 * N))
    new["is_object"] = extras["is_object"][sel].repeat(N, 1)
    gen = extras["generation"][sel].repeat(N, 1)
    if isinstance(curr_gen, int):
        gen = torch.full_like(gen, curr_gen)
    p, m, e = _append(params, moments, extras, new, gen)
    keep = ~torch.cat((sel, torch.zeros(N * int(sel.sum()), dtype=torch.bool, device=sel.device)))
    p = {n: t[keep] for n, t in p.items()}
    m = {n: (a[keep], b[keep]) for n, (a, b) in m.items()}
    e = {n: t[keep] for n, t in e.items()}
    return p, m, e, int(sel.sum())
