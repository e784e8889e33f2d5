"""In-tree build of the native code: libegs_rasterizer.so (CUDA, C-ABI) and the torch binding.

Everything is compiled for sm_100a only (`-gencode arch=compute_100a,code=sm_100a`), with
nvcc's default `-fmad=true` and no fast-math, the same floating-point mode the reference builds
with (DGR/setup.py:21-29 passes no arithmetic flags) -- bit-exact radii / sort keys rely on it.
The built `.so` files stay in-tree (git-ignored) so they travel to the GPU box with the snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
OBJ = ROOT / "build" / "obj"
LIB = PKG / "libegs_rasterizer.so"
BINDING = PKG / "diff_gaussian_rasterization" / "_C.so"

CUDA_SOURCES = ["preprocess", "binning", "binning_packed", "binning_buckets", "blend_fwd", "blend_bwd", "preprocess_bwd", "optim", "model", "loss", "compact", "densify", "knn", "exchange", "ply", "abi"]
NVCC_FLAGS = ["-std=c++17", "-O3", *(os.environ.get("EGS_EXTRA_NVCC", "").split()), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: %s\n%s\n%s" % (" ".join(map(str, cmd)), r.stdout, r.stderr))
    return r.stdout + r.stderr


def build_lib(force: bool = False, verbose: bool = False) -> Path:
    """Compile csrc/*.cu and link libegs_rasterizer.so."""
    OBJ.mkdir(parents=True, exist_ok=True)
    headers = [CSRC / "common.cuh", ROOT / "include" / "egs_rasterizer.h"]
    nvcc = _nvcc()
    jobs = []
    for name in CUDA_SOURCES:
        src, obj = CSRC / f"{name}.cu", OBJ / f"{name}.o"
        if force or _stale(obj, [src, *headers]):
            jobs.append([nvcc, *NVCC_FLAGS, *(["-Xptxas", "-v"] if verbose else []), "-c", str(src), "-o", str(obj)])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 1)) as ex:
            for out in ex.map(_run, jobs):
                if verbose and out.strip():
                    print(out)
    objs = [str(OBJ / f"{n}.o") for n in CUDA_SOURCES]
    if force or jobs or _stale(LIB, objs):
        _run([nvcc, "-shared", "-o", str(LIB), *objs, "-gencode", "arch=compute_100a,code=sm_100a",
              "-cudart", "shared"])
    return LIB


def build_binding(force: bool = False) -> Path:
    """Compile the pybind/torch extension that forwards to the C-ABI (host C++ only, no nvcc)."""
    import torch
    from torch.utils import cpp_extension as ce

    src = CSRC / "torch_binding.cpp"
    BINDING.parent.mkdir(parents=True, exist_ok=True)
    if not (force or _stale(BINDING, [src, ROOT / "include" / "egs_rasterizer.h"])):
        return BINDING
    inc = ce.include_paths("cuda")
    inc = list(inc) + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]
    torch_lib = Path(torch.__file__).parent / "lib"
    cmd = ["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=_C",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           *[f"-I{p}" for p in inc], str(src), "-o", str(BINDING),
           f"-L{PKG}", "-legs_rasterizer", f"-L{torch_lib}", "-lc10", "-lc10_cuda", "-ltorch_cpu",
           "-ltorch_cuda", "-ltorch", "-ltorch_python", "-L/usr/local/cuda/lib64", "-lcudart",
           "-Wl,-rpath,$ORIGIN/..", f"-Wl,-rpath,{torch_lib}", "-Wno-deprecated-declarations"]
    _run(cmd)
    return BINDING


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_lib(force=force, verbose=verbose)
    build_binding(force=force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print("built", LIB, BINDING)
