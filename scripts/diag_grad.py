"""Diagnostics (GPU box): gradient error stats new-vs-ref and ref-vs-ref, then timings at config A/B."""
import os, sys, time
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
    rast = mod.GaussianRasterizer(settings(mod, sc))
    color, radii = rast(means3D=m3, means2D=m2, opacities=op, colors_precomp=col, scales=scl, rotations=rot)
    color.backward(dL)
    return color.detach(), [m3.grad, m2.grad, col.grad, op.grad, scl.grad, rot.grad]


names = ("means3D", "means2D", "colors", "opacities", "scales", "rotations")
for scn in sys.argv[1:] or ["tiny", "small"]:
    sc = make_scene(scn).to("cuda")
    H, W = sc.camera.image_height, sc.camera.image_width
    gen = torch.Generator(device="cpu").manual_seed(1234)
    dL = (torch.randn(3, H, W, generator=gen) * 1e-3).cuda()
    dgr.clear_geometry_cache()
    c0, g0 = run(ref, sc, dL)
    c1, g1 = run(ref, sc, dL)
    c2, g2 = run(dgr, sc, dL)
    print(f"== {scn}: P={sc.P} img maxerr {(c0-c2).abs().max().item():.3e}")
    for n, a, b, c in zip(names, g0, g1, g2):
        scale = a.abs().max().item()
        e_rr = (a - b).abs()
        e_nr = (a - c).abs()
        rel_rr = (e_rr / (a.abs() + scale * 1e-5)).max().item()
        rel_nr = (e_nr / (a.abs() + scale * 1e-5)).max().item()
        bad = (e_nr > 1e-3 * a.abs() + scale * 1e-5).float().mean().item()
        bad_rr = (e_rr > 1e-3 * a.abs() + scale * 1e-5).float().mean().item()
        print(f"  {n:10s} scale {scale:.3e} | ref-ref maxabs {e_rr.max().item():.3e} maxrel {rel_rr:.3e} bad {bad_rr:.2e} | new-ref maxabs {e_nr.max().item():.3e} maxrel {rel_nr:.3e} bad {bad:.2e}")
        if bad > 0:
            i = (e_nr / (a.abs() + scale * 1e-5)).flatten().argmax().item()
            print(f"     worst idx {i}: ref {a.flatten()[i].item():.6e} ref2 {b.flatten()[i].item():.6e} new {c.flatten()[i].item():.6e}")


def timeit(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    ev[0].record()
    for i in range(n):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(n))
    return ts[len(ts) // 2]


for scn in ([] if os.environ.get("NO_TIMING") else ["A", "B"]):
    sc = make_scene(scn).to("cuda")
    H, W = sc.camera.image_height, sc.camera.image_width
    dL = (torch.randn(3, H, W) * 1e-3).cuda()
    for nm, mod in (("ref", ref), ("new", dgr), ("new-nocache", dgr)):
        if nm == "new-nocache":
            dgr.set_geometry_cache(False)
        def fwd():
            with torch.no_grad():
                rast = mod.GaussianRasterizer(settings(mod, sc))
                return rast(means3D=sc.means3D, means2D=sc.means3D, opacities=sc.opacities, colors_precomp=sc.colors,
                            scales=sc.scales, rotations=sc.rotations)
        def fwdbwd():
            run(mod, sc, dL)
        t_f = timeit(fwd)
        t_fb = timeit(fwdbwd)
        print(f"{scn} {nm}: fwd {t_f:.3f} ms  fwd+bwd {t_fb:.3f} ms")
        dgr.set_geometry_cache(True)
    fwd()
    st = dgr.last_geometry_state()
    vis = int((st.radii > 0).sum())
    print(f"{scn}: P={sc.P} visible={vis} R={st.R} R/vis={st.R/max(vis,1):.2f}")
