#!/usr/bin/env python
"""Build the UNMODIFIED reference CUDA extensions into oracle/_ref/ (test infrastructure only).

TEST INFRASTRUCTURE - never imported by the product path.

The sources are compiled where they lie under /root/reference (nothing is copied into the
tracked tree); only build outputs land in oracle/_ref/ (git-ignored, NOT gpurun-ignored so the
prebuilt files travel to the GPU box):

  oracle/_ref/diff_gaussian_rasterization/_C.so   <- DGR: cuda_rasterizer/{rasterizer_impl,forward,backward}.cu,
                                                     rasterize_points.cu, ext.cpp   (DGR/setup.py:24-29)
  oracle/_ref/diff_gaussian_rasterization/__init__.py  (installed copy of the reference binding, as
                                                     `pip install --target` would place it)
  oracle/_ref/simple_knn/_C.so                    <- simple-knn: spatial.cu, simple_knn.cu, ext.cpp
                                                     (simple-knn/setup.py:24-28)

We do not run the reference's setup.py/CMake; this is a direct nvcc/g++ recipe.  The two forced
includes (-include cstdint / cfloat) work around headers missing on gcc 13;
no source edits.
"""
import os
import shutil
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("R3DGW_REFERENCE", "/root/reference")
DGR = os.path.join(REF, "submodules", "diff-gaussian-rasterization")
KNN = os.path.join(REF, "submodules", "simple-knn")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")


def _torch_flags():
    import torch
    from torch.utils import cpp_extension as ce
    inc = ce.include_paths() + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]
    lib = ce.library_paths()
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    return inc, lib, abi


def _compile(src, obj, name, inc, abi, extra):
    common = ["-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-O3", "-std=c++17"]
    incs = [f"-I{p}" for p in inc] + extra
    if src.endswith(".cu"):
        cmd = [NVCC, "-c", src, "-o", obj, "-gencode", "arch=compute_100a,code=sm_100a",
               "-include", "cstdint", "-include", "cfloat", "--expt-relaxed-constexpr",
               "-Xcompiler", "-fPIC", "-w"] + common + incs
    else:
        cmd = ["g++", "-c", src, "-o", obj, "-fPIC", "-w"] + common + incs
    subprocess.run(cmd, check=True)
    return obj


def _build_ext(name, root, sources, extra):
    inc, lib, abi = _torch_flags()
    pkg = os.path.join(OUT, name)
    tmp = os.path.join(OUT, "_obj", name)
    os.makedirs(pkg, exist_ok=True)
    os.makedirs(tmp, exist_ok=True)
    jobs = []
    with ThreadPoolExecutor(max_workers=4) as ex:
        for s in sources:
            obj = os.path.join(tmp, os.path.basename(s) + ".o")
            jobs.append(ex.submit(_compile, os.path.join(root, s), obj, name, inc, abi, extra))
        objs = [j.result() for j in jobs]
    so = os.path.join(pkg, "_C.so")
    cmd = ["g++", "-shared", "-o", so] + objs + [f"-L{p}" for p in lib] + \
          ["-L/usr/local/cuda/lib64", "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python",
           "-lc10_cuda", "-ltorch_cuda", "-lcudart"]
    subprocess.run(cmd, check=True)
    return so


def build(force=False):
    """Build both reference extensions; no-op when the sources are absent (GPU box)."""
    if not os.path.isdir(DGR):
        return False
    dgr_so = os.path.join(OUT, "diff_gaussian_rasterization", "_C.so")
    knn_so = os.path.join(OUT, "simple_knn", "_C.so")
    if force or not os.path.exists(dgr_so):
        _build_ext("diff_gaussian_rasterization", DGR,
                   ["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
                    "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp"],
                   [f"-I{os.path.join(DGR, 'third_party', 'glm')}"])
        # "install" the reference's python binding next to its extension (build output, git-ignored)
        shutil.copyfile(os.path.join(DGR, "diff_gaussian_rasterization", "__init__.py"),
                        os.path.join(OUT, "diff_gaussian_rasterization", "__init__.py"))
    if force or not os.path.exists(knn_so):
        _build_ext("simple_knn", KNN, ["spatial.cu", "simple_knn.cu", "ext.cpp"], [])
        open(os.path.join(OUT, "simple_knn", "__init__.py"), "w").close()
    shutil.rmtree(os.path.join(OUT, "_obj"), ignore_errors=True)
    return True


if __name__ == "__main__":
    ok = build(force="--force" in sys.argv)
    print("oracle/_ref built" if ok else "reference sources not found; nothing built")
