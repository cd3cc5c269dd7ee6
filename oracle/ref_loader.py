"""TEST INFRASTRUCTURE: load the reference extensions built into oracle/_ref/ under alias names.

Only tests/, __graft_entry__.smoke() and bench.py's reference / cpu_baseline legs may import this.
The reference package has the same import name as the drop-in, so it is loaded as `ref_dgr`
(its `from . import _C` resolves inside oracle/_ref/diff_gaussian_rasterization/).
"""
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def available():
    return os.path.exists(os.path.join(REF, "diff_gaussian_rasterization", "_C.so"))


def load_ref_dgr():
    """Reference `diff_gaussian_rasterization` (python binding + _C) as module `ref_dgr`."""
    if "ref_dgr" in sys.modules:
        return sys.modules["ref_dgr"]
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    d = os.path.join(REF, "diff_gaussian_rasterization")
    spec = importlib.util.spec_from_file_location("ref_dgr", os.path.join(d, "__init__.py"),
                                                  submodule_search_locations=[d])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_dgr"] = mod
    spec.loader.exec_module(mod)
    return mod


def load_ref_knn():
    """Reference `simple_knn._C` as module `ref_knn._C`; returns the extension module."""
    if "ref_knn._C" in sys.modules:
        return sys.modules["ref_knn._C"]
    import torch  # noqa: F401
    path = os.path.join(REF, "simple_knn", "_C.so")
    spec = importlib.util.spec_from_file_location("ref_knn._C", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules["ref_knn._C"] = mod
    return mod


def align128(x):
    return (x + 127) // 128 * 128


def parse_ref_buffers(P, R, H, W, geomBuffer, binningBuffer, imgBuffer):
    """Views into the reference's opaque buffers (fromChunk walk, cuda_rasterizer/rasterizer_impl.cu:155-194;
    layout notes in SURVEY.md section 9.2).  Offsets are relative to 128-B-aligned absolute addresses; torch
    allocations are >= 512-B aligned so they equal the null-based offsets."""
    import torch
    out = {}
    o = 0
    def take(buf, nbytes, dtype):
        nonlocal o
        o = align128(o)
        v = buf[o:o + nbytes].view(dtype)
        o += nbytes
        return v
    g = geomBuffer
    out["depths"] = take(g, 4 * P, torch.float32)
    out["clamped"] = take(g, 3 * P, torch.uint8)
    out["internal_radii"] = take(g, 4 * P, torch.int32)
    out["means2D"] = take(g, 8 * P, torch.float32).view(P, 2)
    out["cov3D"] = take(g, 24 * P, torch.float32).view(P, 6)
    out["conic_opacity"] = take(g, 16 * P, torch.float32).view(P, 4)
    out["rgb"] = take(g, 12 * P, torch.float32).view(P, 3)
    out["tiles_touched"] = take(g, 4 * P, torch.int32)
    po = g.numel() - 128 - 4 * P
    out["point_offsets"] = g[po:po + 4 * P].view(torch.int32)
    o = 0
    b = binningBuffer
    out["point_list"] = take(b, 4 * R, torch.int32)
    out["point_list_unsorted"] = take(b, 4 * R, torch.int32)
    out["point_list_keys"] = take(b, 8 * R, torch.int64)
    out["point_list_keys_unsorted"] = take(b, 8 * R, torch.int64)
    o = 0
    im = imgBuffer
    N = H * W
    out["final_T"] = take(im, 4 * N, torch.float32).view(H, W)
    out["n_contrib"] = take(im, 4 * N, torch.int32).view(H, W)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["ranges"] = take(im, 8 * N, torch.int32).view(N, 2)[:tiles]
    return out
