"""Where does a raster call's wall time go? CPU-side issue time vs GPU time per phase (GPU box)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import diff_gaussian_rasterization as dgr
from r3dgw_b200.synthetic import make_scene
from r3dgw_b200 import native

sc = make_scene(sys.argv[1] if len(sys.argv) > 1 else "B").to("cuda")
cam = sc.camera
s = dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, sc.bg, 1.0,
                                      cam.world_view_transform, cam.full_proj_transform, -1, cam.camera_center, False)
m3, op, scl, rot, col = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.opacities, sc.scales, sc.rotations, sc.colors)]
dL = torch.full((3, cam.image_height, cam.image_width), 1e-3, device="cuda")
rast = dgr.GaussianRasterizer(s)

def sync():
    torch.cuda.synchronize()
    return time.perf_counter()

for it in range(8):
    dgr.clear_geometry_cache()
    m2 = torch.zeros_like(m3, requires_grad=True)
    t0 = sync()
    img, radii = rast(means3D=m3, means2D=m2, opacities=op, colors_precomp=col, scales=scl, rotations=rot)
    t1c = time.perf_counter(); t1 = sync()
    img2, _ = rast(means3D=m3, means2D=m2, opacities=op, colors_precomp=col, scales=scl, rotations=rot)
    t2c = time.perf_counter(); t2 = sync()
    (img + img2).backward(dL)
    t3c = time.perf_counter(); t3 = sync()
    if it >= 3:
        print(f"it{it}: fwd(geom+blend) cpu {1e3*(t1c-t0):.3f} total {1e3*(t1-t0):.3f} | fwd(cached blend) cpu {1e3*(t2c-t1):.3f} total {1e3*(t2-t1):.3f}"
              f" | bwd(2 blend+geom) cpu {1e3*(t3c-t2):.3f} total {1e3*(t3-t2):.3f} ms")
# pure C-ABI cost of a blend launch from python
st = dgr.last_geometry_state()
import ctypes
L = native.lib()
t0 = sync()
for _ in range(100):
    IL = native.ImageLayout(); L.r3dgw_get_image_layout(st.W, st.H, ctypes.byref(IL))
t1 = time.perf_counter()
print(f"ctypes trivial call: {1e6*(t1-t0)/100:.1f} us")
t0 = sync()
for _ in range(20):
    x = torch.empty(3, 1080, 1920, device="cuda"); y = torch.zeros(1000000 * 11, device="cuda")
t1 = time.perf_counter(); t2 = sync()
print(f"torch.empty+zeros: cpu {1e6*(t1-t0)/20:.1f} us total {1e6*(t2-t0)/20:.1f} us")
