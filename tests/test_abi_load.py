"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/egs_rasterizer.h declares; size functions are pure host code and behave; the Python
surface raises the reference's own messages for bad argument combinations."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared_functions():
    text = (ROOT / "include" / "egs_rasterizer.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(egs_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(built):
    from egogaussian_b200 import cabi
    lib = cabi.load()
    names = declared_functions()
    assert len(names) >= 16
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert set(cabi.EXPORTED_SYMBOLS) <= set(names)
    assert lib.egs_abi_version() == 2


def test_workspace_sizes_are_pure_and_monotone(built):
    from egogaussian_b200 import cabi
    lib = cabi.load()
    prev = 0
    for P in (0, 1, 1000, 10**6, 6 * 10**6):
        b = lib.egs_geom_bytes(P)
        assert b == lib.egs_geom_bytes(P) and b >= prev and b % 256 == 0
        prev = b
    # per-Gaussian cost of the geometry blob: 53 B of state (4+8+16+16+1+4+4) + 20 B of scratch for
    # the front-to-back presort (two word buffers and the order) + its 12 KB of status per 4096 Gaussians
    assert 73 <= lib.egs_geom_bytes(10**6) / 10**6 <= 77
    assert lib.egs_binning_bytes(0, 1280, 720) > 0
    # two u64 + two u32 buffers (24 B) plus 3 B of look-back status words per instance
    assert 24 <= lib.egs_binning_bytes(4 * 10**6, 1280, 720) / (4 * 10**6) <= 28
    assert lib.egs_image_bytes(1280, 720) >= 1280 * 720 * 4 + 3600 * 8
    assert lib.egs_backward_scratch_bytes(10**6) >= 48 * 10**6


def test_invalid_arguments_fail_without_touching_the_gpu(built):
    from egogaussian_b200 import cabi
    lib = cabi.load()
    n = ctypes.c_int(0)
    assert lib.egs_forward(None, None, ctypes.byref(n)) == -1
    assert b"NULL" in lib.egs_last_error()
    a = cabi.EgsForwardArgs()
    a.P, a.width, a.height = 5, 0, 16
    assert lib.egs_forward(ctypes.byref(a), None, ctypes.byref(n)) == -1
    assert lib.egs_backward(None, None) == -1
    assert lib.egs_adam_step(None, 99, 0.9, 0.999, 1e-15, None) == -1


def test_python_surface_argument_checks(built):
    import torch
    from egogaussian_b200.diff_gaussian_rasterization import (GaussianRasterizationSettings,
                                                              GaussianRasterizer)
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
        "projmatrix", "sh_degree", "campos", "prefiltered", "debug")
    rs = GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                       torch.zeros(3), False, False)
    r = GaussianRasterizer(rs)
    m, o = torch.zeros(4, 3), torch.zeros(4, 1)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, o, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, o, shs=torch.zeros(4, 1, 3), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, o, colors_precomp=m)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, o, colors_precomp=m, scales=m, rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))
    # there is no CPU path: CPU tensors are rejected loudly, not silently rendered on the host
    with pytest.raises(Exception, match="CUDA"):
        r(m, m, o, colors_precomp=m, cov3D_precomp=torch.zeros(4, 6))


def test_positional_order_matches_reference():
    import inspect
    from egogaussian_b200.diff_gaussian_rasterization import GaussianRasterizer
    params = list(inspect.signature(GaussianRasterizer.forward).parameters)
    assert params == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales",
                      "rotations", "cov3D_precomp"]
