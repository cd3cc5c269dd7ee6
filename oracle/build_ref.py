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

  oracle/_ref/py/...                              <- the handful of reference .py files render() and the training
                                                     loss need (gaussian_renderer/__init__.py, scene/gaussian_model.py,
                                                     scene/cameras.py, scene/NVDIFFREC/{light,util}.py, utils/*.py) and
                                                     the split-sum table scene/NVDIFFREC/irrmaps/bsdf_256_256.bin,
                                                     "installed" unmodified so that GPU tests / the reference bench arm
                                                     can import the STOCK caller code on the GPU box
                                                     (oracle/ref_loader.load_ref_python supplies the sys.modules stubs)

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


PY_FILES = [
    "gaussian_renderer/__init__.py", "scene/gaussian_model.py", "scene/cameras.py", "scene/NVDIFFREC/light.py",
    "scene/NVDIFFREC/util.py", "scene/NVDIFFREC/irrmaps/bsdf_256_256.bin", "utils/general_utils.py",
    "utils/graphics_utils.py", "utils/sh_utils.py", "utils/loss_utils.py", "utils/sh_additional_utils.py",
    "utils/spherical_harmonics.py", "utils/camera_utils.py", "utils/system_utils.py", "utils/image_utils.py",
]


def stage_python():
    """Install the reference's caller-side Python (unmodified) under oracle/_ref/py/ - build output, git-ignored."""
    dst_root = os.path.join(OUT, "py")
    for rel in PY_FILES:
        src = os.path.join(REF, rel)
        if not os.path.exists(src):
            continue
        dst = os.path.join(dst_root, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or os.path.getmtime(src) > os.path.getmtime(dst):
            shutil.copyfile(src, dst)
    return dst_root


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
    stage_python()
    return True


if __name__ == "__main__":
    ok = build(force="--force" in sys.argv)
    print("oracle/_ref built" if ok else "reference sources not found; nothing built")
