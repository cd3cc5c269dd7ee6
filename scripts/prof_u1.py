"""One raster call fwd+bwd (U1) at a named workload - the short command ncu wraps.  usage: prof_u1.py [B] [iters] [C]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import diff_gaussian_rasterization as dgr
from r3dgw_b200.synthetic import make_scene

name = sys.argv[1] if len(sys.argv) > 1 else "B"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
C = int(sys.argv[3]) if len(sys.argv) > 3 else 3
sc = make_scene(name).to("cuda")
cam = sc.camera
s = dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, sc.bg, 1.0,
                                      cam.world_view_transform, cam.full_proj_transform, -1, cam.camera_center, False)
leaves = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.opacities, sc.scales, sc.rotations)]
m3, op, scl, rot = leaves
feat = torch.rand(sc.P, C, device="cuda", requires_grad=True)
dL = torch.full((C, cam.image_height, cam.image_width), 1e-3, device="cuda")
for it in range(iters):
    dgr.clear_geometry_cache()
    m2 = torch.zeros_like(m3, requires_grad=True)
    if C == 3:
        img, _ = dgr.GaussianRasterizer(s)(means3D=m3, means2D=m2, opacities=op, colors_precomp=feat, scales=scl, rotations=rot)
    else:
        img, _ = dgr.rasterize_gaussians_features(m3, m2, feat, op, scl, rot, None, s)
    img.backward(dL)
torch.cuda.synchronize()
print("ok", dgr.last_geometry_state().R)
