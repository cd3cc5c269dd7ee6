"""Build + ctypes binding of libr3dgw_b200.so (the C-ABI in include/r3dgw_b200.h).

There is deliberately NO fallback: if the shared library is missing or a symbol is absent the import
fails loudly (the product path is the CUDA path or nothing).
"""
import ctypes
import torch
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(PKG_ROOT, "csrc")
LIB_DIR = os.path.join(PKG_ROOT, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libr3dgw_b200.so")
SOURCES = ["api.cu", "radix_sort.cu", "blend.cu", "knn.cu", "shade.cu", "image_ops.cu", "envmap_sh.cu", "gauss_reg.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]

# every symbol include/r3dgw_b200.h declares
SYMBOLS = [
    "r3dgw_get_geom_layout", "r3dgw_get_binning_layout", "r3dgw_get_image_layout", "r3dgw_version",
    "r3dgw_launch_count", "r3dgw_geometry_forward", "r3dgw_blend_forward", "r3dgw_blend_backward",
    "r3dgw_geometry_backward", "r3dgw_rasterize_forward", "r3dgw_rasterize_backward",
    "r3dgw_mark_visible", "r3dgw_knn_workspace_bytes", "r3dgw_knn_dist2", "r3dgw_shade_forward",
    "r3dgw_shade_backward", "r3dgw_profile_enable", "r3dgw_profile_kernel_count", "r3dgw_profile_kernel_name",
    "r3dgw_profile_read", "r3dgw_image_loss_workspace_bytes", "r3dgw_ssim_l1_forward", "r3dgw_ssim_l1_backward",
    "r3dgw_masked_l1_forward", "r3dgw_masked_l1_backward", "r3dgw_depth_to_normal_forward",
    "r3dgw_depth_to_normal_backward", "r3dgw_alpha_cache_bytes", "r3dgw_blend_forward_cached",
    "r3dgw_blend_backward_color_cached", "r3dgw_materialize_sorted_keys", "r3dgw_envmap_sh_workspace_bytes",
    "r3dgw_envmap_resize_cubic", "r3dgw_envmap_sh_project", "r3dgw_sh_rotate_sweep", "r3dgw_envmap_sh_render",
    "r3dgw_preshade_forward", "r3dgw_preshade_backward", "r3dgw_gaussian_normal_forward",
    "r3dgw_gaussian_normal_backward", "r3dgw_gaussian_reg_workspace_bytes", "r3dgw_gaussian_regularisers_forward",
    "r3dgw_gaussian_regularisers_backward", "r3dgw_normal_tail_forward", "r3dgw_normal_tail_backward",
    "r3dgw_normal_loss_forward", "r3dgw_normal_loss_backward", "r3dgw_blend_backward_geometry_cached",
    "r3dgw_gaussian_regularisers_noalloc_max_gaussians", "r3dgw_gaussian_regularisers_forward_noalloc",
    "r3dgw_blend_backward_geometry_cached_calls",
]


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    deps = list(src_list) + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h", ".cu"))]
    deps.append(os.path.join(os.path.dirname(PKG_ROOT), "include", "r3dgw_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build_native(force=False, verbose=False):
    """Compile csrc/*.cu for sm_100a and link lib/libr3dgw_b200.so (in-tree, travels with gpurun)."""
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    if not force and not _newer(srcs, LIB_PATH):
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    objdir = os.path.join(LIB_DIR, "obj")
    os.makedirs(objdir, exist_ok=True)

    def cc(src):
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        cmd = [NVCC, "-c", src, "-o", obj] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else [])
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(srcs)) as ex:
        objs = list(ex.map(cc, srcs))
    cmd = [NVCC, "-shared", "-o", LIB_PATH] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                       "-Xcompiler", "-fPIC", "-cudart", "static"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB_PATH


ALLOC_FN = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)


class Camera(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int32), ("height", ctypes.c_int32),
                ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float),
                ("scale_modifier", ctypes.c_float), ("prefiltered", ctypes.c_int32),
                ("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p),
                ("cam_pos", ctypes.c_void_p)]


class GeomLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_size_t) for n in
                ("depths", "means2D", "conic_opacity", "cov3D", "tiles_touched", "point_offsets",
                 "rgb", "clamped", "scan_state", "depth_sort", "total")]


class BinningLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_size_t) for n in
                ("point_list", "point_list_keys", "point_list_alt", "point_list_keys_alt", "ranges",
                 "splats", "sort_hist", "total")]


class ImageLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_size_t) for n in ("final_T", "n_contrib", "total")]


class ShadeArgs(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("specular", ctypes.c_int32), ("fix_sky", ctypes.c_int32),
                ("sh_degree", ctypes.c_int32),
                ("positions", ctypes.c_void_p), ("normals", ctypes.c_void_p), ("albedo", ctypes.c_void_p),
                ("roughness", ctypes.c_void_p), ("metalness", ctypes.c_void_p), ("is_sky", ctypes.c_void_p),
                ("fg_index", ctypes.c_void_p), ("cam_pos", ctypes.c_void_p), ("env_sh", ctypes.c_void_p),
                ("sky_sh", ctypes.c_void_p), ("fg_lut", ctypes.c_void_p)]


class PreshadeArgs(ctypes.Structure):
    _fields_ = [("shade", ShadeArgs), ("scales", ctypes.c_void_p), ("rotations", ctypes.c_void_p),
                ("viewmatrix", ctypes.c_void_p)]


_lib = None


def lib():
    """Load (once) and return the ctypes handle; raises if the library is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA library is the product path and there is no fallback. "
            "Run `python -c 'import __graft_entry__ as g; g.build()'` at the repo root.")
    L = ctypes.CDLL(LIB_PATH)
    missing = [s for s in SYMBOLS if not hasattr(L, s)]
    if missing:
        raise ImportError(f"libr3dgw_b200.so lacks symbols: {missing}")
    vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    L.r3dgw_version.restype = ctypes.c_char_p
    L.r3dgw_launch_count.restype = ctypes.c_uint64
    L.r3dgw_get_geom_layout.argtypes = [ci, ci, ctypes.POINTER(GeomLayout)]
    L.r3dgw_get_binning_layout.argtypes = [ci, ci, ci, ctypes.POINTER(BinningLayout)]
    L.r3dgw_get_image_layout.argtypes = [ci, ci, ctypes.POINTER(ImageLayout)]
    cam = ctypes.POINTER(Camera)
    L.r3dgw_geometry_forward.argtypes = [ALLOC_FN, vp, ALLOC_FN, vp, ci, ci, ci, cam,
                                         vp, vp, vp, vp, vp, vp, vp, ctypes.POINTER(ci), vp]
    L.r3dgw_blend_forward.argtypes = [ci, ci, ci, cam, vp, vp, vp, vp, vp, vp, vp]
    L.r3dgw_blend_forward_cached.argtypes = [ci, ci, ci, cam, vp, vp, vp, vp, vp, vp, vp, ci, vp]
    L.r3dgw_blend_backward_color_cached.argtypes = [ci, ci, ci, cam, vp, vp, vp, vp, vp, ci, vp]
    L.r3dgw_blend_backward_geometry_cached.argtypes = [ci, ci, ci, cam] + [vp] * 11
    # (P, R, n_calls, channels*, cam, binning, image, alpha_cache, features**, out_images**, backgrounds**, dL_dpixels**, ...)
    L.r3dgw_blend_backward_geometry_cached_calls.argtypes = [ci, ci, ci, ctypes.POINTER(ci), cam, vp, vp, vp,
                                                             ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(vp),
                                                             ctypes.POINTER(vp), vp, vp, vp, vp]
    L.r3dgw_materialize_sorted_keys.argtypes = [ci, ci, ci, ci, vp, vp, vp]
    L.r3dgw_alpha_cache_bytes.argtypes = [ci]
    L.r3dgw_alpha_cache_bytes.restype = ctypes.c_size_t
    L.r3dgw_blend_backward.argtypes = [ci, ci, ci, cam, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.r3dgw_geometry_backward.argtypes = [ci, ci, ci, cam] + [vp] * 16
    L.r3dgw_rasterize_forward.argtypes = [ALLOC_FN, vp, ALLOC_FN, vp, ALLOC_FN, vp, ci, ci, ci, vp, cam,
                                          vp, vp, vp, vp, vp, vp, vp, vp, vp, ctypes.POINTER(ci), vp]
    L.r3dgw_rasterize_backward.argtypes = [ci, ci, ci, ci, vp, cam] + [vp] * 21
    L.r3dgw_mark_visible.argtypes = [ci, vp, vp, vp, vp, vp]
    L.r3dgw_knn_workspace_bytes.argtypes = [ci]
    L.r3dgw_knn_workspace_bytes.restype = ctypes.c_size_t
    L.r3dgw_knn_dist2.argtypes = [ci, vp, vp, vp, vp]
    sa = ctypes.POINTER(ShadeArgs)
    L.r3dgw_shade_forward.argtypes = [sa, vp, vp, vp, vp]
    L.r3dgw_shade_backward.argtypes = [sa] + [vp] * 11
    pa = ctypes.POINTER(PreshadeArgs)
    L.r3dgw_preshade_forward.argtypes = [pa, ci] + [vp] * 8
    L.r3dgw_preshade_backward.argtypes = [pa, ci] + [vp] * 13
    L.r3dgw_gaussian_normal_forward.argtypes = [ci] + [vp] * 6
    L.r3dgw_gaussian_normal_backward.argtypes = [ci] + [vp] * 7
    sz = ctypes.c_size_t
    L.r3dgw_gaussian_reg_workspace_bytes.argtypes = [ci]
    L.r3dgw_gaussian_reg_workspace_bytes.restype = sz
    L.r3dgw_gaussian_regularisers_forward.argtypes = [ci, vp, vp, vp, vp, vp, ctypes.c_float, vp, vp, vp]
    L.r3dgw_gaussian_regularisers_backward.argtypes = [ci, vp, vp, vp, vp, ctypes.c_float, vp, vp, vp, vp, vp]
    L.r3dgw_gaussian_regularisers_noalloc_max_gaussians.argtypes = []
    L.r3dgw_gaussian_regularisers_forward_noalloc.argtypes = [ci, vp, vp, vp, vp, vp, ctypes.c_float, vp, vp]
    L.r3dgw_envmap_sh_render.argtypes = [ci, vp, ci, vp, vp]
    L.r3dgw_image_loss_workspace_bytes.argtypes = [ci, ci, ci]
    L.r3dgw_image_loss_workspace_bytes.restype = sz
    L.r3dgw_ssim_l1_forward.argtypes = [ci, ci, ci, vp, vp, vp, sz, vp, vp, vp, vp]
    L.r3dgw_ssim_l1_backward.argtypes = [ci, ci, ci, vp, vp, vp, sz, vp, vp, vp, vp, vp]
    L.r3dgw_masked_l1_forward.argtypes = [ci, ci, ci, vp, vp, vp, sz, ci, vp, vp, vp]
    L.r3dgw_masked_l1_backward.argtypes = [ci, ci, ci, vp, vp, vp, sz, ci, vp, vp, vp, vp]
    L.r3dgw_normal_tail_forward.argtypes = [ci, ci, vp, vp, vp, ctypes.c_size_t, vp, vp, vp, vp, vp]
    L.r3dgw_normal_tail_backward.argtypes = [ci, ci, vp, vp, ctypes.c_size_t, vp, vp, vp, vp, vp, vp, vp]
    L.r3dgw_normal_loss_forward.argtypes = [ci, ci] + [vp] * 7
    L.r3dgw_normal_loss_backward.argtypes = [ci, ci] + [vp] * 8
    L.r3dgw_depth_to_normal_forward.argtypes = [ci, ci, vp, vp, vp, vp]
    L.r3dgw_depth_to_normal_backward.argtypes = [ci, ci, vp, vp, vp, vp, vp]
    cf = ctypes.c_float
    L.r3dgw_envmap_sh_workspace_bytes.argtypes = [ci]
    L.r3dgw_envmap_sh_workspace_bytes.restype = sz
    L.r3dgw_envmap_resize_cubic.argtypes = [ci, ci, ci, vp, vp, ci, cf, cf, ci, ci, vp, vp]
    L.r3dgw_envmap_sh_project.argtypes = [ci, ci, ci, vp, vp, ci, cf, cf, ci, vp, vp, vp]
    L.r3dgw_sh_rotate_sweep.argtypes = [ci, ci, vp, vp, ci, vp, vp, vp, vp]
    L.r3dgw_profile_enable.argtypes = [ci]
    L.r3dgw_profile_enable.restype = None
    L.r3dgw_profile_kernel_name.argtypes = [ci]
    L.r3dgw_profile_kernel_name.restype = ctypes.c_char_p
    L.r3dgw_profile_read.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_uint64), ci]
    for name in SYMBOLS:
        fn = getattr(L, name)
        if name not in ("r3dgw_version", "r3dgw_launch_count", "r3dgw_knn_workspace_bytes", "r3dgw_profile_enable",
                        "r3dgw_profile_kernel_name", "r3dgw_image_loss_workspace_bytes", "r3dgw_alpha_cache_bytes",
                        "r3dgw_envmap_sh_workspace_bytes", "r3dgw_gaussian_reg_workspace_bytes"):
            fn.restype = ci
    _lib = L
    return L


def profile_enable(on=True):
    lib().r3dgw_profile_enable(int(bool(on)))


def profile_read(reset=True):
    """{kernel name: (total_ms, launches)} accumulated since the last reset."""
    L = lib()
    n = L.r3dgw_profile_kernel_count()
    ms = (ctypes.c_double * n)()
    cnt = (ctypes.c_uint64 * n)()
    L.r3dgw_profile_read(ms, cnt, int(bool(reset)))
    return {L.r3dgw_profile_kernel_name(i).decode(): (ms[i], int(cnt[i])) for i in range(n)}


class R3dgwError(RuntimeError):
    pass


_ERR = {-1: "bad argument", -2: "unsupported configuration", -3: "allocation callback failed",
        -4: "num_rendered overflows 32-bit int",
        -5: "size exceeds the static scratch of the allocation-free entry point"}


def check(rc, what):
    if rc == 0:
        return
    if rc < 0:
        raise R3dgwError(f"{what}: {_ERR.get(rc, rc)}")
    raise R3dgwError(f"{what}: CUDA error {rc}")


# ---- cheap launch plumbing -------------------------------------------------------------------------
# The training step is ~400 launches and partly launch-bound: torch.cuda.current_stream().cuda_stream builds a Stream
# object and torch.cuda.device(...) resolves indices through several Python layers on every call (~26 of each per step,
# 0.3 ms of host time).  The raw accessors below are single C calls.
try:
    from torch._C import _cuda_getCurrentRawStream as _raw_stream, _cuda_getDevice as _cur_device
except ImportError:   # pragma: no cover - older torch
    _raw_stream = _cur_device = None
if os.environ.get("R3DGW_FAST_PLUMBING", "1") == "0":   # A/B switch: the public torch.cuda accessors
    _raw_stream = _cur_device = None


def raw_stream():
    """cudaStream_t (as int) of torch's current stream on the current device."""
    if _raw_stream is None:
        return torch.cuda.current_stream().cuda_stream
    return _raw_stream(_cur_device())


class _NoGuard:
    __slots__ = ()

    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NO_GUARD = _NoGuard()


def on_device(dev):
    """Context manager making `dev` the current CUDA device; free when it already is."""
    if _cur_device is not None and dev.index is not None and dev.index == _cur_device():
        return _NO_GUARD
    return torch.cuda.device(dev)

