"""Where does the scale / rotation gradient error come from?  FP64 truth vs (a) our float screen-space gradients pushed
through the FP64 geometry chain (input error only) and (b) our float end-to-end result.  usage: diag_f64_split.py [A|B]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200"), os.path.join(ROOT, "oracle")]
import torch
import diff_gaussian_rasterization as dgr
import raster_f64, ref_loader
from r3dgw_b200 import native
from r3dgw_b200.synthetic import make_scene

name = sys.argv[1] if len(sys.argv) > 1 else "A"
sc = make_scene(name).to("cuda")
cam = sc.camera
P, H, W = sc.P, cam.image_height, cam.image_width
ref = ref_loader.load_ref_dgr()
e = torch.Tensor([])
gen = torch.Generator().manual_seed(1)
dL = (torch.randn(3, H, W, generator=gen) * 1e-3).cuda()
R, _c, radii, gb, bb, ib = ref._C.rasterize_gaussians(sc.bg, sc.means3D, sc.colors, sc.opacities, sc.scales, sc.rotations, 1.0, e,
                                                     cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy,
                                                     H, W, e, -1, cam.camera_center, False)
rb = ref_loader.parse_ref_buffers(P, R, H, W, gb, bb, ib)
truth = raster_f64.backward_f64(rb, sc.means3D, sc.scales, sc.rotations, [(sc.colors, sc.bg, dL)], cam, radii)

# our blend backward alone (float screen-space gradients), then the FP64 chain on them
L = native.lib()
s = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.world_view_transform,
                                      cam.full_proj_transform, -1, cam.camera_center, False)
dgr.clear_geometry_cache()
leaves = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.colors, sc.opacities, sc.scales, sc.rotations)]
m2 = torch.zeros_like(sc.means3D, requires_grad=True)
img, _ = dgr.GaussianRasterizer(s)(means3D=leaves[0], means2D=m2, opacities=leaves[2], colors_precomp=leaves[1],
                                   scales=leaves[3], rotations=leaves[4])
img.backward(dL)
st = dgr.last_geometry_state()
IL = native.ImageLayout(); L.r3dgw_get_image_layout(W, H, ctypes.byref(IL))
# re-run the blend backward raw to get dL_dconic
imgbuf = torch.empty(IL.total, dtype=torch.uint8, device="cuda")
out = torch.empty(3, H, W, device="cuda")
L.r3dgw_blend_forward(P, st.R, 3, ctypes.byref(st.cam), st.geom.data_ptr(), st.binning.data_ptr(), imgbuf.data_ptr(),
                      sc.colors.data_ptr(), sc.bg.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
z = lambda *s_: torch.zeros(*s_, device="cuda")
d_m2, d_con, d_op, d_col = z(P, 3), z(P, 4), z(P, 1), z(P, 3)
L.r3dgw_blend_backward(P, st.R, 3, ctypes.byref(st.cam), st.geom.data_ptr(), st.binning.data_ptr(), imgbuf.data_ptr(),
                       sc.colors.data_ptr(), sc.bg.data_ptr(), dL.data_ptr(), d_m2.data_ptr(), d_con.data_ptr(),
                       d_op.data_ptr(), d_col.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
zz = lambda *s_: torch.zeros(*s_, dtype=torch.float64, device="cuda")
dm3, dsc, drot = zz(P, 3), zz(P, 3), zz(P, 4)
m2d = d_m2[:, :2].double().contiguous()
cond = torch.stack([d_con[:, 0], d_con[:, 1], d_con[:, 3]], dim=1).double().contiguous()
fx, fy = W / (2.0 * cam.tanfovx), H / (2.0 * cam.tanfovy)
keep = [sc.means3D.contiguous(), radii.int().contiguous(), rb["cov3D"].contiguous(), sc.scales.contiguous(), sc.rotations.contiguous()]
rc = raster_f64.lib().raster_f64_geometry_backward(P, *[k.data_ptr() for k in keep], 1.0, fx, fy, cam.tanfovx, cam.tanfovy,
                                                   cam.world_view_transform.data_ptr(), cam.full_proj_transform.data_ptr(),
                                                   m2d.data_ptr(), cond.data_ptr(), dm3.data_ptr(), dsc.data_ptr(), drot.data_ptr(),
                                                   torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
rel = lambda a, b: ((a.double() - b).norm() / b.norm()).item()
print(f"config {name}: P={P} R={R}")
for nm, chain, full in (("means3D", dm3, leaves[0].grad), ("scales", dsc, leaves[3].grad), ("rotations", drot, leaves[4].grad)):
    print(f"{nm:<10} float screen-space grads -> FP64 chain: {rel(chain, truth[nm]):.3e}   float end to end: {rel(full, truth[nm]):.3e}")
