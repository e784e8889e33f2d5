"""CPU/torch restatement of the GaussianModel activation + covariance glue (TEST INFRASTRUCTURE).

Only tests/, __graft_entry__.smoke() and bench.py's baseline legs may import this module; nothing
under egogaussian_b200/ does. It restates, op for op in fp32 torch (autograd supplies the
backward), what the reference evaluates every iteration before calling the rasterizer:

  activations()            get_scaling / get_opacity / get_features    scene/gaussian_model.py:126-153
  quaternion_to_rotation() build_rotation                              utils/general_utils.py:124-146
  scaling_rotation()       build_scaling_rotation                      utils/general_utils.py:148-158
  upper_triangle()         strip_lowerdiag / strip_symmetric           utils/general_utils.py:110-122
  covariance()             build_covariance_from_scaling_rotation      scene/gaussian_model.py:29-33
  rotated_covariance()     build_covariance_from_scaling_rotation_w_rot scene/gaussian_model.py:46-63
  rot6d_to_matrix()        rot6d_to_matrix                             utils/geometry_utils.py:69-89

Pinned by tests/golden/model_*.npz, which tests/golden/make_model_golden.py produced by running the
reference's own modules (tests/test_oracle_golden.py::test_model_restatement_matches_reference).

One behaviour worth spelling out, because the restatement writes it explicitly instead of
inheriting it from an indexing idiom: the reference selects rows with
`torch.nonzero(mask).squeeze()` where `mask` is [P,1]. That yields (row, 0) pairs, and indexing the
[P,3,3] factor with an [n,2] index tensor gathers row r AND row 0 for every pair. So whenever any
row is selected, row 0 is rotated too, and the scatter back (`L[indices] = ...`, duplicates
allowed) makes autograd hand row 0's gradient to each of its copies: row 0's raw gradients and its
contribution to the pose gradient are multiplied by the number of index entries naming row 0.
`row0_copies()` computes that count; `rotated_covariance(..., quirk=False)` gives the per-row
behaviour the code presumably intended.
"""
from __future__ import annotations

from typing import Optional

import torch


def activations(scaling_raw, opacity_raw, features_dc, features_rest):
    """(exp(_scaling), sigmoid(_opacity), cat(dc, rest) on dim 1)."""
    return torch.exp(scaling_raw), torch.sigmoid(opacity_raw), torch.cat((features_dc, features_rest), dim=1)


def quaternion_to_rotation(rotation_raw):
    """Normalise (w,x,y,z) and expand to [P,3,3]; every product rounded on its own, as in eager torch."""
    norm = torch.sqrt(rotation_raw[:, 0] * rotation_raw[:, 0] + rotation_raw[:, 1] * rotation_raw[:, 1] +
                      rotation_raw[:, 2] * rotation_raw[:, 2] + rotation_raw[:, 3] * rotation_raw[:, 3])
    q = rotation_raw / norm[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rows = [
        1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y),
    ]
    return torch.stack(rows, dim=1).view(-1, 3, 3)


def scaling_rotation(scales, rotation_raw):
    """L = R diag(s)."""
    return quaternion_to_rotation(rotation_raw) @ torch.diag_embed(scales)


def upper_triangle(sym):
    """[xx, xy, xz, yy, yz, zz] of a batch of 3x3 matrices."""
    return torch.stack((sym[:, 0, 0], sym[:, 0, 1], sym[:, 0, 2], sym[:, 1, 1], sym[:, 1, 2], sym[:, 2, 2]), dim=1)


def covariance(scaling_raw, rotation_raw, scale_modifier: float = 1.0):
    L = scaling_rotation(scale_modifier * torch.exp(scaling_raw), rotation_raw)
    return upper_triangle(L @ L.transpose(1, 2))


def rot6d_to_matrix(rot6d):
    """Gram-Schmidt of the two columns of a [3,2] parameter; columns (b1, b2, b1 x b2)."""
    a1, a2 = rot6d[:, 0], rot6d[:, 1]
    b1 = torch.nn.functional.normalize(a1, dim=0)
    b2 = torch.nn.functional.normalize(a2 - torch.dot(b1, a2) * b1, dim=0)
    b3 = torch.linalg.cross(b1, b2)
    return torch.stack((b1, b2, b3), dim=-1)


def selected_rows(is_object, which_object: Optional[int]):
    """Boolean [P]: rows whose factor is rotated, before the row-0 quirk."""
    flat = is_object.reshape(-1)
    return torch.ones_like(flat, dtype=torch.bool) if which_object is None else flat == which_object


def row0_copies(is_object, which_object: Optional[int]) -> int:
    """How many entries of the reference's (row, 0) index list name row 0 (0 = nothing selected)."""
    sel = selected_rows(is_object, which_object)
    n = int(sel.sum())
    return n + int(bool(sel[0])) if n > 0 else 0


def rotated_covariance(scaling_raw, rotation_raw, scale_modifier, is_object, which_object, accum_R=None,
                       train_R=None, quirk: bool = True):
    """L' = [train_R] (accum_R L) on the selected rows (plus row 0, see the module docstring)."""
    L = scaling_rotation(scale_modifier * torch.exp(scaling_raw), rotation_raw)
    sel = selected_rows(is_object, which_object)
    moved = (accum_R if accum_R is not None else torch.eye(3, dtype=L.dtype, device=L.device)) @ L
    if train_R is not None:
        moved = train_R @ moved
    copies = row0_copies(is_object, which_object) if quirk else 0
    if copies > 0:
        sel = sel.clone()
        sel[0] = True
        # the forward value of row 0 is the moved row; its gradient counts once per index entry
        # (value unchanged exactly: moved - moved.detach() is an exact zero)
        times = torch.ones(L.shape[0], dtype=L.dtype, device=L.device)
        times[0] = float(copies)
        moved = moved.detach() + (moved - moved.detach()) * times[:, None, None]
    Lp = torch.where(sel[:, None, None], moved, L)
    return upper_triangle(Lp @ Lp.transpose(1, 2))
