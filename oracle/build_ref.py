"""Build the UNMODIFIED reference rasterizer for sm_100 into oracle/_ref/ (git-ignored).

TEST INFRASTRUCTURE ONLY. Compiles the reference's own five source files where they lie under
/root/reference/submodules/diff-gaussian-rasterization (nothing is copied into the repo) with
nvcc, against the vendored glm headers and this image's torch headers, and writes one pybind
module `oracle/_ref/ref_dgr_C.so` exporting the reference's three functions
(rasterize_gaussians, rasterize_gaussians_backward, mark_visible -- DGR/ext.cpp:15-18).

Same flags as the reference's setup.py (DGR/setup.py:21-29: only the glm include path; nvcc
defaults, i.e. -fmad=true, no fast-math) plus the target architecture. The .so travels to the GPU
box with the gpurun snapshot; /root/reference does not exist there, so this must run here.
"""
from __future__ import annotations

import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
OUT = HERE / "_ref"
REF = Path("/root/reference/submodules/diff-gaussian-rasterization")
SOURCES = ["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu", "cuda_rasterizer/backward.cu",
           "rasterize_points.cu", "ext.cpp"]
MODULE = "ref_dgr_C"
# the reference's second native submodule (init-time KNN, scene/gaussian_model.py:301): three files
KNN_REF = Path("/root/reference/submodules/simple-knn")
KNN_SOURCES = ["simple_knn.cu", "spatial.cu", "ext.cpp"]
KNN_MODULE = "ref_knn_C"


def available() -> bool:
    return (OUT / f"{MODULE}.so").exists()


def knn_available() -> bool:
    return (OUT / f"{KNN_MODULE}.so").exists()


# The reference's own Python layer (autograd Function, settings tuple, nn.Module) rides along as a build
# artefact next to the .so, byte for byte: the `--impl reference` bench arm drives the compiled reference
# through it (with ref_dgr_C standing in for its `_C`), not through a restatement.
PY_LAYER = OUT / "dgr_py" / "diff_gaussian_rasterization" / "__init__.py"


def stage_python_layer(force: bool = False) -> Path | None:
    src = REF / "diff_gaussian_rasterization" / "__init__.py"
    if PY_LAYER.exists() and not force:
        return PY_LAYER
    if not src.exists():
        return None
    PY_LAYER.parent.mkdir(parents=True, exist_ok=True)
    PY_LAYER.write_bytes(src.read_bytes())
    return PY_LAYER


def build(force: bool = False) -> Path | None:
    build_knn(force)
    stage_python_layer(force)
    return _build(MODULE, REF, SOURCES, [str(REF / "third_party/glm"), str(REF)], force)


def build_knn(force: bool = False) -> Path | None:
    """`oracle/_ref/ref_knn_C.so` exporting the reference's distCUDA2 (simple-knn/ext.cpp)."""
    return _build(KNN_MODULE, KNN_REF, KNN_SOURCES, [str(KNN_REF)], force)


def _build(MODULE: str, REF: Path, SOURCES: list, extra_inc: list, force: bool) -> Path | None:
    target = OUT / f"{MODULE}.so"
    if target.exists() and not force:
        return target
    if not REF.exists():
        return None  # e.g. on the GPU box: use the prebuilt file or report it missing
    import torch
    from torch.utils import cpp_extension as ce
    OUT.mkdir(parents=True, exist_ok=True)
    inc = list(ce.include_paths("cuda")) + [sysconfig.get_paths()["include"], *extra_inc]
    common = ["-std=c++17", "-O3", "-gencode", "arch=compute_100,code=sm_100", "-Xcompiler", "-fPIC",
              f"-DTORCH_EXTENSION_NAME={MODULE}", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
              "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__", "--expt-relaxed-constexpr",
              "-w", *[f"-I{p}" for p in inc]]
    objs = []

    def compile_one(rel):
        obj = OUT / (MODULE + "_" + Path(rel).name + ".o")
        cmd = ["nvcc", *common, "-x", "cu", "-c", str(REF / rel), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"reference build failed on {rel}:\n{r.stderr[-4000:]}")
        return str(obj)

    with ThreadPoolExecutor(max_workers=5) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    torch_lib = Path(torch.__file__).parent / "lib"
    link = ["nvcc", "-shared", "-o", str(target), *objs, f"-L{torch_lib}", "-lc10", "-lc10_cuda", "-ltorch_cpu",
            "-ltorch_cuda", "-ltorch", "-ltorch_python", "-Xlinker", f"-rpath={torch_lib}"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("reference link failed:\n" + r.stderr[-4000:])
    for o in objs:
        Path(o).unlink()
    return target


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
