"""TEST INFRASTRUCTURE: ctypes front end of oracle/raster_f64.cu (FP64 evaluation of the reference backward pass).

Only tests/ may import this.  build() compiles the .cu with nvcc for sm_100a into oracle/_build/ (git-ignored, travels
with gpurun); backward_f64() runs it on the REFERENCE extension's own forward buffers.
"""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "raster_f64.cu")
OUT_DIR = os.path.join(HERE, "_build")
SO = os.path.join(OUT_DIR, "libraster_f64.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
_lib = None


def build(force=False):
    if not force and os.path.exists(SO) and os.path.getmtime(SO) >= os.path.getmtime(SRC):
        return SO
    os.makedirs(OUT_DIR, exist_ok=True)
    subprocess.run([NVCC, "-shared", "-o", SO, SRC, "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-lineinfo",
                    "-Xcompiler", "-fPIC"], check=True)
    return SO


def available():
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            build()
        L = ctypes.CDLL(SO)
        vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        L.raster_f64_render_backward.argtypes = [ci, ci, ci] + [vp] * 14
        L.raster_f64_geometry_backward.argtypes = [ci, vp, vp, vp, vp, vp, cf, cf, cf, cf, cf] + [vp] * 8
        _lib = L
    return _lib


def backward_f64(rb, sc_means3D, scales, rotations, calls, cam, radii):
    """FP64 gradients of one or several raster calls over ONE geometry (render() blends six feature sets over the
    same Gaussians; the screen-space gradients of all calls add up before the geometry backward, which is linear in
    them).  `rb` = ref_loader.parse_ref_buffers(...) of the reference forward, `radii` its returned radii, `cam` the
    synthetic Camera, `calls` = [(colors [P,C], bg [C], dL_dpix [C,H,W]), ...].  Returns float64 CUDA tensors:
    means3D [P,3], means2D [P,3] (z = 0), opacities [P,1], scales [P,3], rotations [P,4], colors = [one [P,C] per
    call], images = [one FP64 [C,H,W] forward per call]."""
    import torch
    L = lib()
    dev = sc_means3D.device
    P = int(sc_means3D.shape[0])
    H, W = int(cam.image_height), int(cam.image_width)
    z = lambda *s: torch.zeros(*s, dtype=torch.float64, device=dev)   # noqa: E731
    cont = lambda t: t.contiguous()   # noqa: E731
    st = torch.cuda.current_stream().cuda_stream
    d_m2, d_con, d_op = z(P, 2), z(P, 3), z(P)
    d_cols, imgs = [], []
    shared = [cont(rb["ranges"]), cont(rb["point_list"]), cont(rb["means2D"]), cont(rb["conic_opacity"]), cont(rb["n_contrib"])]
    for colors, bg, dL in calls:
        C = int(colors.shape[1])
        d_col, img = z(P, C), z(C, H, W)
        k = [cont(bg.float()), cont(colors.float()), cont(dL.float())]
        rc = L.raster_f64_render_backward(W, H, C, shared[0].data_ptr(), shared[1].data_ptr(), k[0].data_ptr(),
                                          shared[2].data_ptr(), shared[3].data_ptr(), k[1].data_ptr(),
                                          shared[4].data_ptr(), k[2].data_ptr(), d_m2.data_ptr(), d_con.data_ptr(),
                                          d_op.data_ptr(), d_col.data_ptr(), img.data_ptr(), st)
        assert rc == 0, rc
        torch.cuda.synchronize()
        d_cols.append(d_col)
        imgs.append(img)
    d_m3, d_sc, d_rot = z(P, 3), z(P, 3), z(P, 4)
    keep2 = [cont(sc_means3D.float()), cont(radii.int()), cont(rb["cov3D"]), cont(scales.float()), cont(rotations.float())]
    view, proj = cont(cam.world_view_transform.float()), cont(cam.full_proj_transform.float())
    fx, fy = W / (2.0 * cam.tanfovx), H / (2.0 * cam.tanfovy)
    rc = L.raster_f64_geometry_backward(P, *[k.data_ptr() for k in keep2], 1.0, fx, fy, cam.tanfovx, cam.tanfovy,
                                        view.data_ptr(), proj.data_ptr(), d_m2.data_ptr(), d_con.data_ptr(),
                                        d_m3.data_ptr(), d_sc.data_ptr(), d_rot.data_ptr(), st)
    assert rc == 0, rc
    torch.cuda.synchronize()
    m2 = torch.cat([d_m2, z(P, 1)], dim=1)
    return {"means3D": d_m3, "means2D": m2, "colors": d_cols, "opacities": d_op.view(P, 1), "scales": d_sc,
            "rotations": d_rot, "images": imgs}
