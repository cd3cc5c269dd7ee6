"""How much of the alpha cache is zeros?  Records the cache for config B (or argv[1]) into a NaN-filled pool and counts,
over the rows the record kernel wrote: non-zero lanes per 128-byte row, all-zero 32-byte sectors (8 lanes = one pixel
row of the warp's 8x4 box) and all-zero 64-byte halves."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import diff_gaussian_rasterization as dgr
from r3dgw_b200 import native
from r3dgw_b200.synthetic import make_scene

name = sys.argv[1] if len(sys.argv) > 1 else "B"
sc = make_scene(name).to("cuda")
cam = sc.camera
s = dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, sc.bg, 1.0,
                                      cam.world_view_transform, cam.full_proj_transform, -1, cam.camera_center, False)
with torch.no_grad():
    dgr.GaussianRasterizer(s)(means3D=sc.means3D, means2D=torch.zeros_like(sc.means3D), opacities=sc.opacities,
                              colors_precomp=sc.colors, scales=sc.scales, rotations=sc.rotations)
st = dgr.last_geometry_state()
L = native.lib()
nbytes = int(L.r3dgw_alpha_cache_bytes(st.R))
pool = torch.full((nbytes // 4,), float("nan"), device="cuda")
IL = native.ImageLayout(); L.r3dgw_get_image_layout(st.W, st.H, ctypes.byref(IL))
img = torch.zeros(IL.total, dtype=torch.uint8, device="cuda")
out = torch.zeros(3, st.H, st.W, device="cuda")
rc = L.r3dgw_blend_forward_cached(st.P, st.R, 3, ctypes.byref(st.cam), st.geom.data_ptr(), st.binning.data_ptr(), img.data_ptr(),
                                  sc.colors.data_ptr(), sc.bg.data_ptr(), out.data_ptr(), pool.data_ptr(), 1,
                                  torch.cuda.current_stream().cuda_stream)
assert rc == 0
torch.cuda.synchronize()
rows_total = nbytes // 128
tot = dict(rows=0, lanes=0, sect_dead=0, half_dead=0, hist=torch.zeros(33, dtype=torch.long, device="cuda"),
           sect_hist=torch.zeros(5, dtype=torch.long, device="cuda"))
chunk = 1 << 22
for r0 in range(0, rows_total, chunk):
    blk = pool[r0 * 32:(min(r0 + chunk, rows_total)) * 32].view(-1, 32)
    written = ~torch.isnan(blk[:, 0])
    w = blk[written]
    if w.numel() == 0:
        continue
    nz = w != 0
    tot["rows"] += int(w.shape[0])
    tot["lanes"] += int(nz.sum())
    sec = nz.view(-1, 4, 8).any(dim=2)
    tot["sect_dead"] += int((~sec).sum())
    tot["half_dead"] += int((~nz.view(-1, 2, 16).any(dim=2)).sum())
    tot["hist"] += torch.bincount(nz.sum(dim=1), minlength=33)
    tot["sect_hist"] += torch.bincount(sec.sum(dim=1), minlength=5)
R = tot["rows"]
print(f"config {name}: R={st.R} rows written {R} ({R * 128 / 1e6:.0f} MB)  non-zero lanes/row {tot['lanes'] / R:.2f}")
print(f"all-zero 32-B sectors: {tot['sect_dead'] / (4 * R):.3f} of sectors   all-zero 64-B halves: {tot['half_dead'] / (2 * R):.3f}")
print("live sectors per row histogram (0..4):", (tot["sect_hist"].float() / R).tolist())
print("rows with 0 live lanes:", int(tot["hist"][0]) / R)
