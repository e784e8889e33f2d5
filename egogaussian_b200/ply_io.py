"""Host-side mirror of GaussianModel.save_ply / load_ply (scene/gaussian_model.py:340-481), SURVEY.md 8f
rank 4. The records are packed and unpacked on the GPU (csrc/ply.cu: egs_ply_pack / egs_ply_unpack) and
cross PCIe as one block; this module only writes and parses the PLY header (binary_little_endian 1.0,
one `vertex` element of float properties, which is what the reference writes through plyfile).

There is no CPU path: the tensors are CUDA tensors, like the reference's (`device="cuda"`, :463-478).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List

import numpy as np
import torch

from . import cabi

_FLOAT_NAMES = {"float", "float32"}


def attribute_names(sh_coeffs: int, with_extras: bool = True) -> List[str]:
    """construct_list_of_attributes (gaussian_model.py:340-359; _og :361-373 without the extras)."""
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)]
    names += [f"f_rest_{i}" for i in range(3 * (sh_coeffs - 1))]
    names += ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
    return names + (["label", "generation", "is_object"] if with_extras else [])


def _header(n: int, names: List[str]) -> bytes:
    lines = ["ply", "format binary_little_endian 1.0", f"element vertex {n}"]
    lines += [f"property float {name}" for name in names] + ["end_header"]
    return ("\n".join(lines) + "\n").encode("ascii")


def _columns(struct, **tensors):
    for k, t in tensors.items():
        setattr(struct, k, t.data_ptr() if t is not None and t.numel() else None)
    return struct


def save_ply(path, *, xyz, features_dc, features_rest, opacity, scaling, rotation, label=None, generation=None,
             is_object=None) -> None:
    """save_ply (gaussian_model.py:375-397). Without label/generation/is_object the original 3DGS
    record (62 floats at SH degree 3) is written."""
    if not xyz.is_cuda:
        raise ValueError("save_ply packs the records on the GPU: CUDA tensors required")
    extras = label is not None
    P, M = int(xyz.shape[0]), 1 + int(features_rest.shape[1])
    lib = cabi.load()
    K = lib.egs_ply_record_floats(M, int(extras))
    f32 = lambda t: t.detach().contiguous().float()   # noqa: E731
    i32 = lambda t: t.detach().contiguous().to(torch.int32)   # noqa: E731
    keep = dict(xyz=f32(xyz), features_dc=f32(features_dc), features_rest=f32(features_rest), opacity=f32(opacity),
                scaling=f32(scaling), rotation=f32(rotation), label=f32(label) if extras else None,
                generation=i32(generation) if extras else None, is_object=i32(is_object) if extras else None)
    rows = torch.empty((P, K), dtype=torch.float32, device=xyz.device)
    with torch.cuda.device(xyz.device):
        cabi._check(lib.egs_ply_pack(P, M, C.byref(_columns(cabi.EgsPlyColumns(), **keep)), int(extras),
                                     rows.data_ptr(), cabi._stream_handle(None)), "egs_ply_pack")
    host = rows.cpu().numpy()
    d = os.path.dirname(str(path))
    if d:
        os.makedirs(d, exist_ok=True)   # mkdir_p(os.path.dirname(path)), :376
    with open(path, "wb") as f:
        f.write(_header(P, attribute_names(M, extras)))
        f.write(host.astype("<f4", copy=False).tobytes())


def _parse_header(f):
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt, count, names, in_vertex = None, None, [], False
    while True:
        line = f.readline()
        if not line:
            raise ValueError("PLY header does not end")
        tok = line.decode("ascii").split()
        if not tok or tok[0] in ("comment", "obj_info"):
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            if count is not None and tok[1] != "vertex":
                in_vertex = False   # later elements (faces) are ignored, like elements[0] in the reference
                continue
            if tok[1] != "vertex":
                raise ValueError("the first PLY element must be `vertex`")
            count, in_vertex = int(tok[2]), True
        elif tok[0] == "property" and in_vertex:
            if tok[1] not in _FLOAT_NAMES:
                raise ValueError(f"vertex property {tok[-1]} is {tok[1]}: only float properties are supported")
            names.append(tok[2])
        elif tok[0] == "end_header":
            break
    if fmt != "binary_little_endian" or count is None:
        raise ValueError("only binary_little_endian PLY files with a vertex element are supported")
    return count, names


def load_ply(path, max_sh_degree: int, is_object: bool = False, force_bg: bool = False, device="cuda") -> Dict[str, torch.Tensor]:
    """load_ply (gaussian_model.py:399-481): the tensors of the model as a dict (xyz, features_dc [P,1,3],
    features_rest [P,M-1,3], opacity, scaling, rotation, label [P,1]; generation, is_object [P,1] int32)."""
    device = torch.device(device)
    if device.type != "cuda":
        raise ValueError("load_ply unpacks the records on the GPU: a CUDA device is required")
    M = (max_sh_degree + 1) ** 2
    with open(path, "rb") as f:
        P, names = _parse_header(f)
        raw = np.fromfile(f, dtype="<f4", count=P * len(names))
    if raw.size != P * len(names):
        raise ValueError("PLY file is shorter than its header says")
    n_rest = sum(1 for n in names if n.startswith("f_rest_"))
    assert n_rest == 3 * M - 3, "the file's SH degree differs from max_sh_degree"   # :420
    where = {n: i for i, n in enumerate(names)}
    canonical = attribute_names(M, True)
    column_of = (C.c_int * len(canonical))(*[where.get(n, -1) for n in canonical])
    rows = torch.from_numpy(raw).to(device)
    out = dict(xyz=torch.empty((P, 3), device=device), features_dc=torch.empty((P, 1, 3), device=device),
               features_rest=torch.empty((P, M - 1, 3), device=device), opacity=torch.empty((P, 1), device=device),
               scaling=torch.empty((P, 3), device=device), rotation=torch.empty((P, 4), device=device),
               label=torch.empty((P, 1), device=device),
               generation=torch.empty((P, 1), dtype=torch.int32, device=device),
               is_object=torch.empty((P, 1), dtype=torch.int32, device=device))
    with torch.cuda.device(device):
        cabi._check(cabi.load().egs_ply_unpack(P, M, rows.data_ptr(), len(names), column_of,
                                               C.byref(_columns(cabi.EgsPlyColumns(), **out)), 0.01,
                                               int(bool(is_object)), cabi._stream_handle(None)), "egs_ply_unpack")
    if force_bg:
        out["is_object"].zero_()   # :461
    return out
