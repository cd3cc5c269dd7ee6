#!/usr/bin/env python
"""Generate tests/golden/raster_*.npz from the UNMODIFIED reference CUDA extension (oracle/_ref) on a GPU box.

    gpurun -- 'python tests/golden/make_raster_golden.py gpurun_out/golden'   then copy the .npz into tests/golden/

Records, for the seeded `tiny` scene (r3dgw_b200.synthetic), every intermediate of the reference's forward
(parsed out of its opaque buffers), the image, and the 6 gradients for a seeded dL/dpixel.
These vectors pin oracle/raster_oracle.c (CPU tests) and are a second, box-independent check of the CUDA path.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "relightable3dgaussians-w_b200")]


def main(out_dir):
    import ref_loader
    from r3dgw_b200.synthetic import make_scene
    ref = ref_loader.load_ref_dgr()
    os.makedirs(out_dir, exist_ok=True)
    for name in ("tiny",):
        sc = make_scene(name).to("cuda")
        cam = sc.camera
        H, W, P = cam.image_height, cam.image_width, sc.P
        s = ref.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.world_view_transform,
                                              cam.full_proj_transform, -1, cam.camera_center, False)
        e = torch.Tensor([])
        R, color, radii, gb, bb, ib = ref._C.rasterize_gaussians(
            s.bg, sc.means3D, sc.colors, sc.opacities, sc.scales, sc.rotations, 1.0, e, s.viewmatrix, s.projmatrix,
            s.tanfovx, s.tanfovy, H, W, e, -1, s.campos, False)
        rb = ref_loader.parse_ref_buffers(P, R, H, W, gb, bb, ib)
        gen = torch.Generator().manual_seed(4242)
        dL = torch.randn(3, H, W, generator=gen) * 1e-3
        leaves = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.colors, sc.opacities, sc.scales, sc.rotations)]
        m3, col, op, scl, rot = leaves
        m2 = torch.zeros_like(m3, requires_grad=True)
        img, _ = ref.GaussianRasterizer(s)(means3D=m3, means2D=m2, opacities=op, colors_precomp=col, scales=scl,
                                           rotations=rot)
        img.backward(dL.cuda())
        blob = dict(R=np.array(R), radii=radii, out_color=color, dL_dpix=dL,
                    g_means3D=m3.grad, g_means2D=m2.grad, g_colors=col.grad, g_opacities=op.grad, g_scales=scl.grad,
                    g_rotations=rot.grad)
        for k in ("depths", "means2D", "cov3D", "conic_opacity", "tiles_touched", "point_offsets", "point_list",
                  "point_list_keys", "final_T", "n_contrib", "ranges"):
            blob[k] = rb[k]
        blob = {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else v) for k, v in blob.items()}
        path = os.path.join(out_dir, f"raster_{name}.npz")
        np.savez_compressed(path, **blob)
        print("wrote", path, "R =", R, "bytes =", os.path.getsize(path))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
