"""ctypes front-end of oracle/raster_oracle.c (CPU restatement of the reference rasterizer + simple-knn).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
reference legs - never by the product path.  `build()` compiles the C file with gcc into
oracle/_build/libraster_oracle.so (git-ignored; travels to the GPU box with the snapshot).
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "raster_oracle.c")
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libraster_oracle.so")

_lib = None


def build(force=False):
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    # -ffp-contract=off: keep every rounding explicit (the GPU contracts to FMA; the mismatch budget in the
    # tests covers that difference)
    subprocess.run(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-ffp-contract=off", "-fopenmp", SRC, "-o", LIB, "-lm"],
                   check=True)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB)
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def set_threads(n):
    lib().oracle_set_threads(ctypes.c_int(int(n)))


def forward(means3D, scales, rotations, opacities, features, viewmatrix, projmatrix, W, H, tanfovx, tanfovy, bg,
            scale_modifier=1.0, cov3D_precomp=None, blend=True):
    """K1..K4 on the CPU.  All inputs numpy fp32 (matrices flat 16, already transposed like the reference's).
    Returns a dict with every intermediate the reference materialises."""
    L = lib()
    means3D, scales, rotations = _f32(means3D), _f32(scales), _f32(rotations)
    opacities, features = _f32(opacities).reshape(-1), _f32(features)
    view, proj, bg = _f32(viewmatrix).reshape(-1), _f32(projmatrix).reshape(-1), _f32(bg)
    cov_pre = _f32(cov3D_precomp)
    P = means3D.shape[0]
    C = features.shape[1] if features is not None else 0
    gx, gy = (W + 15) // 16, (H + 15) // 16
    o = dict(radii=np.zeros(P, np.int32), means2D=np.zeros((P, 2), np.float32), depths=np.zeros(P, np.float32),
             cov3D=np.zeros((P, 6), np.float32), conic_opacity=np.zeros((P, 4), np.float32),
             tiles_touched=np.zeros(P, np.uint32), point_offsets=np.zeros(P, np.uint32))
    L.oracle_preprocess(P, _p(means3D), _p(scales), ctypes.c_float(scale_modifier), _p(rotations), _p(opacities),
                        _p(cov_pre), _p(view), _p(proj), W, H, ctypes.c_float(tanfovx), ctypes.c_float(tanfovy),
                        _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]), _p(o["cov3D"]), _p(o["conic_opacity"]),
                        _p(o["tiles_touched"]))
    R = int(o["tiles_touched"].astype(np.int64).sum())
    o["R"] = R
    o["point_list_keys"] = np.zeros(max(R, 1), np.uint64)[:R]
    o["point_list"] = np.zeros(max(R, 1), np.uint32)[:R]
    o["ranges"] = np.zeros((gx * gy, 2), np.uint32)
    keys = np.zeros(max(R, 1), np.uint64)
    plist = np.zeros(max(R, 1), np.uint32)
    rc = L.oracle_bin(P, W, H, _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]), _p(o["tiles_touched"]),
                      _p(o["point_offsets"]), R, _p(keys), _p(plist), _p(o["ranges"]))
    assert rc == 0
    o["point_list_keys"], o["point_list"] = keys[:R], plist[:R]
    if blend and features is not None:
        o["final_T"] = np.zeros((H, W), np.float32)
        o["n_contrib"] = np.zeros((H, W), np.uint32)
        o["out_color"] = np.zeros((C, H, W), np.float32)
        L.oracle_blend_forward(C, W, H, _p(o["ranges"]), _p(plist), _p(o["means2D"]), _p(o["conic_opacity"]),
                               _p(features), _p(bg), _p(o["final_T"]), _p(o["n_contrib"]), _p(o["out_color"]))
    o["_keep"] = (means3D, scales, rotations, opacities, features, view, proj, bg, cov_pre, plist)
    o["_dims"] = (P, C, W, H, float(tanfovx), float(tanfovy), float(scale_modifier))
    return o


def backward(fw, dL_dpix):
    """K5..K7 on the CPU from a forward() result.  Returns gradients in the reference's layout."""
    L = lib()
    means3D, scales, rotations, opacities, features, view, proj, bg, cov_pre, plist = fw["_keep"]
    P, C, W, H, tfx, tfy, mod = fw["_dims"]
    dL = _f32(dL_dpix)
    g = dict(dL_dmean2D=np.zeros((P, 3), np.float32), dL_dconic=np.zeros((P, 4), np.float32),
             dL_dopacity=np.zeros((P, 1), np.float32), dL_dcolors=np.zeros((P, C), np.float32),
             dL_dmean3D=np.zeros((P, 3), np.float32), dL_dcov3D=np.zeros((P, 6), np.float32),
             dL_dscale=np.zeros((P, 3), np.float32), dL_drot=np.zeros((P, 4), np.float32))
    L.oracle_blend_backward(C, W, H, _p(fw["ranges"]), _p(plist), _p(fw["means2D"]), _p(fw["conic_opacity"]),
                            _p(features), _p(bg), _p(fw["final_T"]), _p(fw["n_contrib"]), _p(dL),
                            _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]), _p(g["dL_dcolors"]))
    cov = cov_pre if cov_pre is not None else fw["cov3D"]
    L.oracle_geometry_backward(P, _p(means3D), _p(fw["radii"]), _p(cov), _p(scales if cov_pre is None else None),
                               ctypes.c_float(mod), _p(rotations), _p(view), _p(proj), W, H, ctypes.c_float(tfx),
                               ctypes.c_float(tfy), _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dmean3D"]),
                               _p(g["dL_dcov3D"]), _p(g["dL_dscale"]), _p(g["dL_drot"]))
    return g


def mark_visible(means3D, viewmatrix):
    means3D, view = _f32(means3D), _f32(viewmatrix).reshape(-1)
    out = np.zeros(means3D.shape[0], np.uint8)
    lib().oracle_mark_visible(means3D.shape[0], _p(means3D), _p(view), _p(out))
    return out.astype(bool)


def knn_dist2(points):
    pts = _f32(points)
    out = np.zeros(pts.shape[0], np.float32)
    lib().oracle_knn_dist2(pts.shape[0], _p(pts), _p(out))
    return out
