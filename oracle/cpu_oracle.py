"""numpy/ctypes front end of the CPU oracle (oracle/cpu_rasterizer.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs. The product package never imports this module.

The oracle restates the reference rasterizer's CUDA kernels on the CPU (the reference has no CPU
path of its own); see the header of cpu_rasterizer.c for the file:line map. `render_forward` and
`render_backward` chain the stages the way CudaRasterizer::Rasterizer::forward / backward do
(DGR/cuda_rasterizer/rasterizer_impl.cu:197-339, 343-447).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path
from typing import Dict, Optional

import numpy as np

HERE = Path(__file__).resolve().parent
SRC = HERE / "cpu_rasterizer.c"
LIB = HERE / "_build" / "liboracle.so"
_lib = None


def build(force: bool = False) -> Path:
    """gcc -O2 -fopenmp; -ffp-contract=off keeps results identical on every x86 host."""
    if LIB.exists() and not force and LIB.stat().st_mtime >= SRC.stat().st_mtime:
        return LIB
    LIB.parent.mkdir(parents=True, exist_ok=True)
    cmd = ["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared", "-std=gnu11",
           str(SRC), "-o", str(LIB), "-lm"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stderr)
    return LIB


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(LIB))
        _lib.orc_bin.restype = C.c_int64
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(_load().orc_num_threads())


def _p(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a) -> Optional[np.ndarray]:
    if a is None:
        return None
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a, dtype=np.float32)


def render_forward(*, means3D, opacities, bg, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                   image_height, image_width, sh_degree=0, shs=None, colors_precomp=None, scales=None,
                   rotations=None, cov3D_precomp=None, scale_modifier=1.0,
                   preprocessed: Optional[Dict[str, np.ndarray]] = None,
                   binned: Optional[Dict[str, np.ndarray]] = None) -> Dict[str, np.ndarray]:
    """Full forward. `preprocessed` / `binned` let a test substitute another implementation's
    stage outputs (stage-wise parity: feed the GPU's preprocess into the oracle's binning, ...)."""
    lib = _load()
    means3D, opacities = _f32(means3D), _f32(opacities).reshape(-1)
    shs, colors_precomp, scales = _f32(shs), _f32(colors_precomp), _f32(scales)
    rotations, cov3D_precomp = _f32(rotations), _f32(cov3D_precomp)
    bg, view, proj, cam = _f32(bg), _f32(viewmatrix).reshape(-1), _f32(projmatrix).reshape(-1), _f32(campos)
    P, H, W = means3D.shape[0], int(image_height), int(image_width)
    M = 0 if shs is None else shs.shape[1]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    out: Dict[str, np.ndarray] = {}

    if preprocessed is None:
        pre = {
            "radii": np.zeros(P, np.int32), "means2D": np.zeros((P, 2), np.float32),
            "depths": np.zeros(P, np.float32), "cov3D": np.zeros((P, 6), np.float32),
            "conic_opacity": np.zeros((P, 4), np.float32), "rgb": np.zeros((P, 3), np.float32),
            "clamped": np.zeros((P, 3), np.uint8), "tiles_touched": np.zeros(P, np.uint32),
            "radius_f": np.zeros(P, np.float32),
        }
        lib.orc_preprocess(C.c_int(P), C.c_int(int(sh_degree)), C.c_int(M), _p(means3D), _p(scales),
                           C.c_float(scale_modifier), _p(rotations), _p(opacities), _p(shs), _p(cov3D_precomp),
                           _p(colors_precomp), _p(view), _p(proj), _p(cam), C.c_int(W), C.c_int(H),
                           C.c_float(tanfovx), C.c_float(tanfovy), _p(pre["radii"]), _p(pre["means2D"]),
                           _p(pre["depths"]), _p(pre["cov3D"]), _p(pre["conic_opacity"]), _p(pre["rgb"]),
                           _p(pre["clamped"]), _p(pre["tiles_touched"]), _p(pre["radius_f"]))
        if cov3D_precomp is not None:
            pre["cov3D"] = cov3D_precomp
    else:
        pre = {k: np.ascontiguousarray(v) for k, v in preprocessed.items()}
    out.update(pre)

    if binned is None:
        R = int(pre["tiles_touched"].astype(np.int64).sum())
        keys, vals = np.zeros(max(R, 1), np.uint64), np.zeros(max(R, 1), np.uint32)
        tk, tv = np.zeros_like(keys), np.zeros_like(vals)
        ranges = np.zeros((gx * gy, 2), np.uint32)
        n = lib.orc_bin(C.c_int(P), _p(pre["means2D"]), _p(pre["depths"]), _p(pre["radii"]), C.c_int(W),
                        C.c_int(H), _p(keys), _p(vals), _p(tk), _p(tv), _p(ranges))
        assert n == R, (n, R)
        binned = {"keys": keys[:R], "values": vals[:R], "ranges": ranges}
    else:
        binned = {k: np.ascontiguousarray(v) for k, v in binned.items()}
    out.update(binned)
    out["num_rendered"] = int(binned["keys"].shape[0])

    color = np.zeros((3, H, W), np.float32)
    depth = np.zeros((1, H, W), np.float32)
    alpha = np.zeros((1, H, W), np.float32)
    n_contrib = np.zeros((H, W), np.uint32)
    frag = C.c_int64(0)
    vals = np.ascontiguousarray(binned["values"], dtype=np.uint32)
    lib.orc_blend_forward(C.c_int(W), C.c_int(H), _p(np.ascontiguousarray(binned["ranges"], dtype=np.uint32)),
                          _p(vals), _p(pre["means2D"]), _p(pre["rgb"]), _p(pre["depths"]),
                          _p(pre["conic_opacity"]), _p(bg), _p(color), _p(depth), _p(alpha), _p(n_contrib),
                          C.byref(frag))
    out.update(color=color, depth=depth, alpha=alpha, n_contrib=n_contrib, fragments=int(frag.value))
    return out


def render_backward(fwd: Dict[str, np.ndarray], *, means3D, bg, viewmatrix, projmatrix, campos, tanfovx,
                    tanfovy, image_height, image_width, dL_dcolor, dL_ddepth, dL_dalpha, sh_degree=0,
                    shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None,
                    scale_modifier=1.0) -> Dict[str, Optional[np.ndarray]]:
    lib = _load()
    means3D = _f32(means3D)
    shs, scales, rotations = _f32(shs), _f32(scales), _f32(rotations)
    bg, view, proj, cam = _f32(bg), _f32(viewmatrix).reshape(-1), _f32(projmatrix).reshape(-1), _f32(campos)
    P, H, W = means3D.shape[0], int(image_height), int(image_width)
    M = 0 if shs is None else shs.shape[1]
    gc, gd, ga = _f32(dL_dcolor), _f32(dL_ddepth).reshape(H, W), _f32(dL_dalpha).reshape(H, W)

    dmean2D = np.zeros((P, 2), np.float64)
    dconic = np.zeros((P, 3), np.float64)
    dopac = np.zeros(P, np.float64)
    dcol = np.zeros((P, 3), np.float64)
    ddep = np.zeros(P, np.float64)
    lib.orc_blend_backward(C.c_int(W), C.c_int(H), _p(np.ascontiguousarray(fwd["ranges"], dtype=np.uint32)),
                           _p(np.ascontiguousarray(fwd["values"], dtype=np.uint32)), _p(bg),
                           _p(fwd["means2D"]), _p(fwd["conic_opacity"]), _p(fwd["rgb"]), _p(fwd["depths"]),
                           _p(np.ascontiguousarray(fwd["alpha"].reshape(H, W))), _p(fwd["n_contrib"]),
                           _p(gc), _p(gd), _p(ga), _p(dmean2D), _p(dconic), _p(dopac), _p(dcol), _p(ddep))
    dmean2D32, dconic32 = dmean2D.astype(np.float32), dconic.astype(np.float32)
    dcol32, ddep32 = dcol.astype(np.float32), ddep.astype(np.float32)

    dmeans3D = np.zeros((P, 3), np.float32)
    dcov3D = np.zeros((P, 6), np.float32)
    dsh = np.zeros((P, M, 3), np.float32) if M else None
    dscales = np.zeros((P, 3), np.float32) if scales is not None else None
    drots = np.zeros((P, 4), np.float32) if scales is not None else None
    lib.orc_preprocess_backward(C.c_int(P), C.c_int(int(sh_degree)), C.c_int(M), _p(means3D), _p(fwd["radii"]),
                                _p(shs), _p(np.ascontiguousarray(fwd["clamped"], dtype=np.uint8)), _p(scales),
                                _p(rotations), C.c_float(scale_modifier),
                                _p(np.ascontiguousarray(fwd["cov3D"], dtype=np.float32)), _p(view), _p(proj),
                                _p(cam), C.c_int(W), C.c_int(H), C.c_float(tanfovx), C.c_float(tanfovy),
                                _p(dmean2D32), _p(dconic32), _p(dcol32), _p(ddep32), _p(dmeans3D), _p(dcov3D),
                                _p(dsh), _p(dscales), _p(drots))
    means2D_grad = np.zeros((P, 3), np.float32)
    means2D_grad[:, :2] = dmean2D32
    return {
        "means2D": means2D_grad, "opacity": dopac.astype(np.float32)[:, None], "means3D": dmeans3D,
        "colors": dcol32 if colors_precomp is not None else None,
        "cov3D": dcov3D if cov3D_precomp is not None else None,
        "sh": dsh, "scales": dscales, "rotations": drots,
        "conic": dconic32, "depths": ddep32,
    }


def mark_visible(means3D, viewmatrix) -> np.ndarray:
    lib = _load()
    m, v = _f32(means3D), _f32(viewmatrix).reshape(-1)
    out = np.zeros(m.shape[0], np.uint8)
    lib.orc_mark_visible(C.c_int(m.shape[0]), _p(m), _p(v), _p(out))
    return out.astype(bool)


def adam_reference(param, grad, exp_avg, exp_avg_sq, lr, step, beta1=0.9, beta2=0.999, eps=1e-15):
    """One torch.optim.Adam step in numpy fp32 (scene/gaussian_model.py:198 configuration)."""
    p, g = np.asarray(param, np.float32), np.asarray(grad, np.float32)
    m, v = np.asarray(exp_avg, np.float32), np.asarray(exp_avg_sq, np.float32)
    m = m + (g - m) * np.float32(1.0 - beta1)
    v = v * np.float32(beta2) + np.float32(1.0 - beta2) * g * g
    bc1, bc2 = 1.0 - beta1 ** step, 1.0 - beta2 ** step
    denom = np.sqrt(v) / np.float32(np.sqrt(bc2)) + np.float32(eps)
    p = p + np.float32(-(lr / bc1)) * (m / denom)
    return p.astype(np.float32), m.astype(np.float32), v.astype(np.float32)


def scene_kwargs(scene, mode_geom="scale_rot", mode_color="sh") -> dict:
    """Keyword arguments for render_forward / render_backward from an egogaussian_b200.scenes.Scene."""
    cam = scene.camera
    kw = dict(means3D=scene.means3D, bg=scene.bg, viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix,
              campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, image_height=cam.height,
              image_width=cam.width, sh_degree=scene.sh_degree)
    if mode_color == "sh":
        kw["shs"] = scene.shs
    else:
        kw["colors_precomp"] = scene.colors_precomp
    if mode_geom == "scale_rot":
        kw["scales"], kw["rotations"] = scene.scales, scene.rotations
    else:
        kw["cov3D_precomp"] = scene.cov3D()
    return kw
