"""Per-kernel ms of one U1 raster call (fwd, fwd+bwd) at config E from the library's CUDA-event profile."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import diff_gaussian_rasterization as dgr
from r3dgw_b200 import native
from r3dgw_b200.synthetic import make_scene
name = sys.argv[1] if len(sys.argv) > 1 else "E"
sc = make_scene(name).to("cuda")
cam = sc.camera
H, W = cam.image_height, cam.image_width
dL = torch.full((3, H, W), 1e-3, device="cuda")
s = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.world_view_transform,
                                      cam.full_proj_transform, -1, cam.camera_center, False)
leaves = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.opacities, sc.scales, sc.rotations, sc.colors)]
def call(bwd):
    dgr.clear_geometry_cache()
    m2 = torch.zeros_like(leaves[0], requires_grad=True)
    with torch.set_grad_enabled(bwd):
        img, _ = dgr.GaussianRasterizer(s)(means3D=leaves[0], means2D=m2, opacities=leaves[1], colors_precomp=leaves[4],
                                           scales=leaves[2], rotations=leaves[3])
    if bwd:
        img.backward(dL)
for bwd in (False, True):
    for _ in range(2):
        call(bwd)
    torch.cuda.synchronize()
    native.profile_enable(True)
    n = 4
    for _ in range(n):
        call(bwd)
    torch.cuda.synchronize()
    prof = native.profile_read(reset=True)
    native.profile_enable(False)
    out = {k: round(ms / n, 3) for k, (ms, c) in prof.items() if c}
    out["_sum"] = round(sum(out.values()), 3)
    print("bwd" if bwd else "fwd", "R", dgr.last_geometry_state().R, json.dumps(out))
