"""Checkpoint records: oracle/ply_ref.py restates the reference's save_ply /
load_ply (parity unpinned: plyfile is not installed, see its header); egs_ply_pack / egs_ply_unpack
through egogaussian_b200/ply_io.py must write the same bytes and read the same tensors."""
import io

import numpy as np
import pytest
import torch

from oracle import ply_ref


def model_arrays(P, M, seed=0):
    rng = np.random.default_rng(seed)
    f = lambda *s: rng.normal(size=s).astype(np.float32)   # noqa: E731
    return dict(xyz=f(P, 3), features_dc=f(P, 1, 3), features_rest=f(P, M - 1, 3), opacity=f(P, 1), scaling=f(P, 3),
                rotation=f(P, 4), label=rng.random((P, 1)).astype(np.float32),
                generation=rng.integers(0, 7, (P, 1)).astype(np.int32), is_object=rng.integers(0, 3, (P, 1)).astype(np.int32))


def test_oracle_layout_and_round_trip():
    a = model_arrays(5, 16)
    blob = ply_ref.save_ply_bytes(**a)
    head = blob[:blob.index(b"end_header\n")].decode().split("\n")
    assert head[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 5"]
    props = [l.split()[-1] for l in head if l.startswith("property float ")]
    assert props[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
    assert props[9] == "f_rest_0" and props[53] == "f_rest_44" and props[54:] == [
        "opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3", "label", "generation", "is_object"]
    rec = np.frombuffer(blob[blob.index(b"end_header\n") + 11:], dtype="<f4").reshape(5, 65)
    assert np.array_equal(rec[:, 3:6], np.zeros((5, 3)))                       # normals
    assert np.array_equal(rec[:, 9 + 15], a["features_rest"][:, 0, 1])          # channel-major: f_rest_15 = coeff 1, green
    assert np.array_equal(rec[:, 63], a["generation"][:, 0].astype(np.float32))
    back = ply_ref.load_ply_bytes(blob, 3)
    assert all(np.array_equal(a[k], back[k]) for k in a)
    og = ply_ref.save_ply_bytes(**{k: v for k, v in a.items() if k not in ("label", "generation", "is_object")})
    back = ply_ref.load_ply_bytes(og, 3, is_object=True)
    assert np.allclose(back["label"], 0.01) and (back["generation"] == 0).all() and (back["is_object"] == 1).all()


def test_header_parser_of_the_host_mirror():
    from egogaussian_b200 import ply_io
    blob = ply_ref.save_ply_bytes(**model_arrays(3, 4))
    n, names = ply_io._parse_header(io.BytesIO(blob))
    assert n == 3 and names == ply_io.attribute_names(4, True) == ply_ref.attribute_names(3, 9, True)
    assert ply_io._header(3, names) == blob[:blob.index(b"end_header\n") + 11]
    with pytest.raises(ValueError):
        ply_io._parse_header(io.BytesIO(b"ply\nformat ascii 1.0\nelement vertex 1\nproperty float x\nend_header\n"))
    with pytest.raises(ValueError):
        ply_io._parse_header(io.BytesIO(b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty uchar red\nend_header\n"))


@pytest.mark.gpu
@pytest.mark.parametrize("P,degree,extras", [(1000, 3, True), (257, 3, False), (33, 0, True), (5000, 1, True), (0, 3, True)])
def test_save_and_load_match_the_restatement(P, degree, extras, built, cuda_device, tmp_path):
    from egogaussian_b200 import ply_io
    M = (degree + 1) ** 2
    a = model_arrays(P, M, seed=P + degree)
    if not extras:
        a = {k: v for k, v in a.items() if k not in ("label", "generation", "is_object")}
    dev = {k: torch.from_numpy(v).to(cuda_device) for k, v in a.items()}
    path = tmp_path / "sub" / "point_cloud.ply"
    ply_io.save_ply(path, **dev)
    blob = path.read_bytes()
    assert blob == ply_ref.save_ply_bytes(**a)                                   # byte for byte
    got = ply_io.load_ply(path, degree, is_object=True, device=cuda_device)
    ref = ply_ref.load_ply_bytes(blob, degree, is_object=True)
    for k, v in ref.items():
        assert got[k].dtype == (torch.int32 if v.dtype == np.int32 else torch.float32)
        assert tuple(got[k].shape) == v.shape, k
        assert np.array_equal(got[k].cpu().numpy(), v), k
    assert (ply_io.load_ply(path, degree, force_bg=True, device=cuda_device)["is_object"] == 0).all()


@pytest.mark.gpu
def test_load_accepts_any_property_order(built, cuda_device, tmp_path):
    """plyfile readers address properties by name; a file written by another tool may order them differently."""
    from egogaussian_b200 import ply_io
    a = model_arrays(300, 16, seed=4)
    blob = ply_ref.save_ply_bytes(**a)
    names = ply_ref.attribute_names(3, 45, True)
    rec = np.frombuffer(blob[blob.index(b"end_header\n") + 11:], dtype="<f4").reshape(300, 65)
    perm = np.random.default_rng(1).permutation(65)
    path = tmp_path / "shuffled.ply"
    path.write_bytes(ply_io._header(300, [names[i] for i in perm]) + np.ascontiguousarray(rec[:, perm]).tobytes())
    got = ply_io.load_ply(path, 3, device=cuda_device)
    assert all(np.array_equal(got[k].cpu().numpy(), a[k]) for k in a)
    with pytest.raises(AssertionError):
        ply_io.load_ply(path, 2, device=cuda_device)                              # wrong SH degree (:420)
    with pytest.raises(ValueError):
        ply_io.load_ply(path, 3, device="cpu")
