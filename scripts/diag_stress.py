"""Config E (6M Gaussians @ 3840x2160, large splats): U1 fwd / fwd+bwd, ours vs the reference extension, + parity."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200"), os.path.join(ROOT, "oracle")]
import torch
import ref_loader
import diff_gaussian_rasterization as dgr
from r3dgw_b200.synthetic import make_scene
ref = ref_loader.load_ref_dgr()
name = sys.argv[1] if len(sys.argv) > 1 else "E"
Pover = int(sys.argv[2]) if len(sys.argv) > 2 else None
t0 = time.time(); sc = make_scene(name, P=Pover).to("cuda"); print(f"scene {name} built in {time.time()-t0:.1f}s P={sc.P}")
cam = sc.camera
H, W = cam.image_height, cam.image_width
dL = torch.full((3, H, W), 1e-3, device="cuda")

def settings(mod):
    return mod.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.world_view_transform,
                                             cam.full_proj_transform, -1, cam.camera_center, False)

def run(mod, bwd):
    if mod is dgr: dgr.clear_geometry_cache()
    leaves = [t.clone().requires_grad_(bwd) for t in (sc.means3D, sc.colors, sc.opacities, sc.scales, sc.rotations)]
    m3, col, op, scl, rot = leaves
    m2 = torch.zeros_like(m3, requires_grad=bwd)
    with torch.set_grad_enabled(bwd):
        img, radii = mod.GaussianRasterizer(settings(mod))(means3D=m3, means2D=m2, opacities=op, colors_precomp=col, scales=scl, rotations=rot)
    if bwd: img.backward(dL)
    out = img.detach(), radii, [t.grad for t in leaves]
    del img, leaves
    torch.cuda.synchronize(); torch.cuda.empty_cache()
    return out

def timeit(fn, n=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]

i_n, r_n, g_n = run(dgr, False); torch.cuda.synchronize(); print("new fwd ok R=", dgr.last_geometry_state().R, flush=True)
i_n, r_n, g_n = run(dgr, True); torch.cuda.synchronize(); print("new fwd+bwd ok", flush=True)
i_r, r_r, g_r = run(ref, False); torch.cuda.synchronize(); print("ref fwd ok", flush=True)
i_r, r_r, g_r = run(ref, True); torch.cuda.synchronize(); print("ref fwd+bwd ok", flush=True)
st = dgr.last_geometry_state()
print(f"R={st.R} visible={(r_n>0).sum().item()} radii equal={torch.equal(r_r, r_n)} img maxerr={(i_r-i_n).abs().max().item():.2e}")
_, _, g_r2 = run(ref, True)
for nm, a, b, b2 in zip(("means3D","colors","opacities","scales","rotations"), g_n, g_r, g_r2):
    l2 = ((a-b).double().pow(2).sum().sqrt() / b.double().pow(2).sum().sqrt()).item()
    l2r = ((b2-b).double().pow(2).sum().sqrt() / b.double().pow(2).sum().sqrt()).item()
    print(f"  grad {nm}: rel L2 new-ref {l2:.2e}   ref-ref {l2r:.2e}   norms new {a.norm().item():.3e} ref {b.norm().item():.3e} ref2 {b2.norm().item():.3e}")
del i_r, i_n, g_r, g_n, g_r2
torch.cuda.empty_cache()
for nm, mod in (("ref", ref), ("new", dgr)):
    print(f"{name} {nm}: fwd {timeit(lambda: run(mod, False)):.2f} ms  fwd+bwd {timeit(lambda: run(mod, True)):.2f} ms")
print("max mem GB", torch.cuda.max_memory_allocated() / 1e9)
