"""Auxiliary timings on the GPU box: distCUDA2 and the fused shading vs the reference implementations."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200"), os.path.join(ROOT, "oracle")]
import torch
import ref_loader, shade_oracle as so
import simple_knn._C as knn
from r3dgw_b200.shading import shade_gaussians
from r3dgw_b200.synthetic import make_scene
import bench

def timeit(fn, n=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]

ref_knn = ref_loader.load_ref_knn()
g = torch.Generator().manual_seed(1)
for name, pts in (("uniform 1M", torch.rand(1_000_000, 3, generator=g)),
                  ("scene B positions 1M", make_scene("B").means3D),
                  ("gaussian clusters 1M", torch.randn(1_000_000, 3, generator=g) * 0.05 + torch.randint(0, 50, (1_000_000, 1), generator=g).float())):
    p = pts.cuda().contiguous()
    a, b = knn.distCUDA2(p), ref_knn.distCUDA2(p)
    print(f"distCUDA2 {name}: bit-equal {torch.equal(a.view(torch.int32), b.view(torch.int32))}  new {timeit(lambda: knn.distCUDA2(p)):.3f} ms  reference {timeit(lambda: ref_knn.distCUDA2(p)):.3f} ms")

sc = make_scene("B").to("cuda")
lut = bench.make_lut("cuda")
nrm = torch.nn.functional.normalize(torch.randn(sc.P, 3, device="cuda"), dim=1)
names = ("positions", "normals", "albedo", "roughness", "metalness", "env_sh", "sky_sh")
vals = (sc.means3D, nrm, sc.albedo, sc.roughness, sc.metalness, sc.env_sh, sc.sky_sh)
ups = [torch.randn(sc.P, 3, device="cuda") for _ in range(3)]
def new(bwd):
    L = [v.clone().requires_grad_(bwd) for v in vals]
    o = shade_gaussians(L[0], L[1], L[2], L[3], L[4], L[5], sc.camera.camera_center, lut, 4, specular=True, sky_sh=L[6], is_sky=sc.is_sky)
    if bwd: sum((x * u).sum() for x, u in zip(o, ups)).backward()
def ref(bwd):
    L = [v.clone().requires_grad_(bwd) for v in vals]
    o = so.shade_all(L[5], 4, lut, L[0], L[1], L[2], L[3], L[4], sc.is_sky, sc.camera.camera_center, L[6], True, False)
    if bwd: sum((x * u).sum() for x, u in zip(o, ups)).backward()
print(f"shading 1M (specular, sky): new fwd {timeit(lambda: new(False)):.3f} ms fwd+bwd {timeit(lambda: new(True)):.3f} ms | reference torch graph on the same GPU fwd {timeit(lambda: ref(False)):.3f} ms fwd+bwd {timeit(lambda: ref(True)):.3f} ms  (both include cloning the 7 inputs)")
