"""Every kernel of the library once at a small size - the command compute-sanitizer wraps."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import diff_gaussian_rasterization as dgr
import simple_knn._C as knn
from r3dgw_b200 import render as R
from r3dgw_b200.synthetic import make_scene
import bench

sc = make_scene("tiny", W=173, H=99).to("cuda")
lut = bench.make_lut("cuda")
pc = R.GaussianParams.from_scene(sc, requires_grad=True)
env = sc.env_sh.clone().requires_grad_(True); sky = sc.sky_sh.clone().requires_grad_(True)
bg = torch.zeros(3, device="cuda")
for fn in (lambda: R.render(sc.camera, pc, env, sky, dgr, bg, R.fused_shade_fn(lut), debug=True),
           lambda: R.render_fused(sc.camera, pc, env, sky, bg, R.fused_shade_fn(lut))):
    dgr.clear_geometry_cache()
    out = fn()
    sum(out[k].sum() for k in ("render", "diffuse_color", "specular_color", "depth", "normal")).backward()
s = dgr.GaussianRasterizationSettings(99, 173, sc.camera.tanfovx, sc.camera.tanfovy, bg, 1.0, sc.camera.world_view_transform,
                                      sc.camera.full_proj_transform, 3, sc.camera.camera_center, False)
sh = torch.randn(sc.P, 16, 3, device="cuda", requires_grad=True)
img, _ = dgr.GaussianRasterizer(s)(means3D=pc.means3D, means2D=torch.zeros_like(pc.means3D), opacities=pc.opacity, shs=sh,
                                   scales=pc.scales, rotations=pc.rotations)
img.sum().backward()
dgr.GaussianRasterizer(s).markVisible(pc.means3D)
print("knn", knn.distCUDA2(sc.means3D.contiguous()).sum().item())
torch.cuda.synchronize()
print("sanity ok")
