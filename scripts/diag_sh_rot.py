"""Run-to-run noise of the rotation gradient on the SH-colour path: ours vs ours, reference vs reference, ours vs reference."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200"), os.path.join(ROOT, "oracle")]
import torch
import ref_loader
import diff_gaussian_rasterization as dgr
from r3dgw_b200.synthetic import make_scene
ref = ref_loader.load_ref_dgr()
for deg in (0, 2):
    sc = make_scene("small", seed_offset=40 + deg).to("cuda")
    H, W = sc.camera.image_height, sc.camera.image_width
    gen = torch.Generator(device="cpu").manual_seed(8)
    shs0 = (torch.randn(sc.P, 16, 3, generator=gen) * 0.3).cuda()
    dL = (torch.randn(3, H, W, generator=gen) * 1e-3).cuda()
    def run(mod):
        cam = sc.camera
        s = mod.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.3, cam.world_view_transform,
                                              cam.full_proj_transform, deg, cam.camera_center, False)
        m3, sh, op, scl, rot = [t.clone().requires_grad_(True) for t in (sc.means3D, shs0, sc.opacities, sc.scales, sc.rotations)]
        m2 = torch.zeros_like(m3, requires_grad=True)
        if mod is dgr: dgr.clear_geometry_cache()
        img, radii = mod.GaussianRasterizer(s)(means3D=m3, means2D=m2, opacities=op, shs=sh, scales=scl, rotations=rot)
        img.backward(dL)
        return rot.grad.double(), scl.grad.double()
    A = [run(dgr) for _ in range(6)]; B = [run(ref) for _ in range(6)]
    l2 = lambda x, y: ((x - y).pow(2).sum().sqrt() / y.pow(2).sum().sqrt()).item()
    for k, nm in ((0, "rot"), (1, "scl")):
        aa = [l2(A[i][k], A[j][k]) for i in range(6) for j in range(i)]
        bb = [l2(B[i][k], B[j][k]) for i in range(6) for j in range(i)]
        ab = [l2(A[i][k], B[j][k]) for i in range(6) for j in range(6)]
        bm = sum(b[k] for b in B) / 6; am = sum(a[k] for a in A) / 6
        print(f"deg{deg} {nm}: ours-ours max {max(aa):.2e} | ref-ref max {max(bb):.2e} min {min(bb):.2e} | ours-ref max {max(ab):.2e} min {min(ab):.2e} | mean-vs-mean {l2(am, bm):.2e}")
