"""Relative-L2 gradient error new-vs-ref and ref-vs-ref for U1 (one raster call) over a few seeds (GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200"), os.path.join(ROOT, "oracle")]
import torch
import ref_loader
import diff_gaussian_rasterization as dgr
from r3dgw_b200.synthetic import make_scene
ref = ref_loader.load_ref_dgr()

def settings(mod, sc):
    cam = sc.camera
    return mod.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, sc.bg, 1.0,
                                             cam.world_view_transform, cam.full_proj_transform, -1, cam.camera_center, False)

def run(mod, sc, dL):
    leaves = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.colors, sc.opacities, sc.scales, sc.rotations)]
    m3, col, op, scl, rot = leaves
    m2 = torch.zeros_like(m3, requires_grad=True)
    color, _ = mod.GaussianRasterizer(settings(mod, sc))(means3D=m3, means2D=m2, opacities=op, colors_precomp=col, scales=scl, rotations=rot)
    color.backward(dL)
    return [m3.grad, m2.grad, col.grad, op.grad, scl.grad, rot.grad]

names = ("means3D", "means2D", "colors", "opacities", "scales", "rotations")
l2 = lambda a, b: ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()
for scn, seeds in (("A", (None, 1, 2)), ("small", (None, 3, 4, 5))):
    for sd in seeds:
        sc = make_scene(scn, **({} if sd is None else {"seed_offset": sd})).to("cuda")
        H, W = sc.camera.image_height, sc.camera.image_width
        dL = (torch.randn(3, H, W, generator=torch.Generator().manual_seed(1234)) * 1e-3).cuda()
        dgr.clear_geometry_cache()
        g0, g1, g2 = run(ref, sc, dL), run(ref, sc, dL), run(dgr, sc, dL)
        print(scn, sd, " ".join(f"{n}: new {l2(c, a):.1e} ref {l2(b, a):.1e} |" for n, a, b, c in zip(names, g0, g1, g2)))
