"""GPU parity: fused relit shading (csrc/shade.cu) and distCUDA2 (csrc/knn.cu) through the C-ABI.

Shading is floating point: tolerance 1e-5 relative/2e-6 absolute on values, 1e-3 relative on gradients (same bar
as the rasterizer gradients), against (i) the golden vectors recorded from the reference's own Python and
(ii) the torch oracle on larger seeded inputs.  distCUDA2 must equal the reference extension bit for bit.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _close(a, b, name, rtol, atol_scale):
    a, b = a.detach().cpu().double(), torch.as_tensor(b).double()
    scale = max(b.abs().max().item(), 1e-12)
    err = (a - b).abs()
    ok = err <= rtol * b.abs() + atol_scale * scale
    assert bool(ok.all()), f"{name}: max err {err.max().item():.3e} scale {scale:.3e}"


@pytest.mark.parametrize("tag", ["spec_km", "spec_nokm", "diffuse"])
def test_shade_kernel_vs_reference_python_golden(native_lib, tag):
    import shade_oracle as so
    from r3dgw_b200.shading import EnvironmentLight, shade_gaussians
    z = np.load(os.path.join(GOLD, f"shade_{tag}.npz"))
    t = {k[3:]: torch.from_numpy(z[k]).cuda() for k in z.files if k.startswith("in_")}
    specular, use_km = bool(z["specular"]), bool(z["use_km"])
    lut = so.synthetic_fg_lut().cuda()
    leaves = {k: t[k].clone().requires_grad_(True) for k in ("pos", "nrm", "albedo", "kr", "km", "env")}
    N = t["pos"].shape[0]
    env = EnvironmentLight(leaves["env"], 4, fg_lut=lut)
    view_pos = t["cam"].reshape(1, 3).repeat(N, 1)
    km = leaves["km"][None, None] if use_km else None
    rgb, ex = env.shade(gb_pos=leaves["pos"][None, None], gb_normal=leaves["nrm"][None, None],
                        albedo=leaves["albedo"][None, None], view_pos=view_pos[None, None], kr=leaves["kr"][None, None],
                        km=km, specular=specular)
    assert rgb.shape == (1, 1, N, 3)
    rgb, dif, spc = rgb.squeeze(), ex["diffuse"].squeeze(), ex["specular"].squeeze()
    loss = (rgb * t["g_rgb"]).sum() + (dif * t["g_dif"]).sum()
    if specular:
        loss = loss + (spc * t["g_spc"]).sum()
    loss.backward()
    for nm, v in (("rgb", rgb), ("dif", dif), ("spc", spc)):
        _close(v, z["ref_" + nm], nm, 1e-5, 2e-6)
    # sky gradient contributions in the fixture come from a separate term; compare the shade-only leaves
    for k in ("nrm", "albedo", "kr", "km", "env"):
        if k == "km" and not use_km:
            continue
        g = leaves[k].grad if leaves[k].grad is not None else torch.zeros_like(leaves[k])
        _close(g, z["ref_d_" + k], "d_" + k, 1e-3, 1e-5)
    _close(leaves["pos"].grad if leaves["pos"].grad is not None else torch.zeros_like(t["pos"]), z["ref_d_pos"],
           "d_pos", 1e-3, 1e-5)
    # sky colour: all Gaussians marked sky, directions = positions - cam
    sky_leaf = t["sky_sh"].clone().requires_grad_(True)
    pos_sky = (t["cam"].reshape(1, 3) + t["sky_dirs"]).contiguous()
    is_sky = torch.ones(N, dtype=torch.bool, device="cuda")
    c, _, _ = shade_gaussians(pos_sky, t["nrm"], t["albedo"], t["kr"], None, t["env"], t["cam"], lut, 4,
                              specular=False, sky_sh=sky_leaf, is_sky=is_sky)
    _close(c, z["ref_sky"], "sky", 1e-5, 2e-6)


@pytest.mark.parametrize("specular,fix_sky", [(True, False), (False, False), (True, True)])
def test_shade_all_vs_oracle_scene(native_lib, specular, fix_sky):
    """The whole colour block of render() on a seeded scene with 5 % sky Gaussians vs the torch oracle + autograd."""
    import shade_oracle as so
    from r3dgw_b200.shading import shade_gaussians
    from r3dgw_b200.synthetic import make_scene
    sc = make_scene("small")
    lut = so.synthetic_fg_lut()
    g = torch.Generator().manual_seed(17)
    normals = torch.nn.functional.normalize(torch.randn(sc.P, 3, generator=g), dim=1)
    ups = [torch.randn(sc.P, 3, generator=g) for _ in range(3)]
    names = ("positions", "normals", "albedo", "roughness", "metalness", "env_sh", "sky_sh")
    vals = (sc.means3D, normals, sc.albedo, sc.roughness, sc.metalness, sc.env_sh, sc.sky_sh)

    def run(dev, fn):
        L = {n: v.clone().to(dev).requires_grad_(True) for n, v in zip(names, vals)}
        outs = fn(L, dev)
        loss = sum((o * u.to(dev)).sum() for o, u in zip(outs, ups))
        loss.backward()
        return [o.detach().cpu() for o in outs], {n: (L[n].grad.cpu() if L[n].grad is not None else torch.zeros_like(vals[i]))
                                                  for i, n in enumerate(names)}

    o_ref, g_ref = run("cpu", lambda L, d: so.shade_all(L["env_sh"], 4, lut, L["positions"], L["normals"], L["albedo"],
                                                         L["roughness"], L["metalness"], sc.is_sky,
                                                         sc.camera.camera_center, L["sky_sh"], specular, fix_sky))
    o_new, g_new = run("cuda", lambda L, d: shade_gaussians(L["positions"], L["normals"], L["albedo"], L["roughness"],
                                                            L["metalness"], L["env_sh"], sc.camera.camera_center.cuda(),
                                                            lut.cuda(), 4, specular=specular, sky_sh=L["sky_sh"],
                                                            is_sky=sc.is_sky.cuda(), fix_sky=fix_sky))
    for nm, a, b in zip(("rgb", "diffuse", "specular"), o_new, o_ref):
        _close(a, b, nm, 1e-5, 2e-6)
    for n in names:
        _close(g_new[n], g_ref[n], "d_" + n, 1e-3, 2e-5)


@pytest.mark.parametrize("P,kind", [(5, "rand"), (1000, "rand"), (20000, "rand"), (4099, "clustered"), (3000, "dup")])
def test_knn_bit_exact_vs_reference(native_lib, P, kind):
    import ref_loader
    if not ref_loader.available():
        pytest.skip("oracle/_ref not built")
    ref_knn = ref_loader.load_ref_knn()
    import simple_knn._C as knn
    g = torch.Generator().manual_seed(100 + P)
    pts = torch.randn(P, 3, generator=g)
    if kind == "clustered":
        pts = pts * 0.01 + torch.randint(0, 4, (P, 1), generator=g).float() * 5.0
    if kind == "dup":
        pts[P // 2:] = pts[:P - P // 2]          # exact duplicates -> zero distances
    pts = pts.cuda().contiguous()
    a = knn.distCUDA2(pts)
    b = ref_knn.distCUDA2(pts)
    assert torch.equal(a.view(torch.int32), b.view(torch.int32)), f"max diff {(a - b).abs().max().item():.3e}"


def test_knn_vs_cpu_oracle_and_empty(native_lib):
    import raster_oracle as ro
    import simple_knn._C as knn
    g = torch.Generator().manual_seed(9)
    pts = torch.rand(2500, 3, generator=g)
    a = knn.distCUDA2(pts.cuda()).cpu().numpy()
    np.testing.assert_allclose(a, ro.knn_dist2(pts.numpy()), rtol=1e-6, atol=0)
    assert knn.distCUDA2(torch.zeros(0, 3, device="cuda")).numel() == 0
