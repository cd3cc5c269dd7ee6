"""GPU parity: image-space post-ops (csrc/image_ops.cu) through the C-ABI.

Floating point. Tolerances (against the golden vectors recorded from the reference's own Python, and against the
torch oracle run on the GPU on larger seeded inputs):
  loss values              1e-5 relative (SSIM / L1 means)
  loss gradients           every entry <= 1e-3*|ref| + 2e-5*max|ref|  (same bar as the rasterizer gradients)
  depth_to_normal normals  2e-5 absolute;  its depth gradient: relative L2 <= 1e-3 and entries <= 1e-3*|ref| + 1e-3*scale
    (the reference subtracts back-projected POINTS ~10 units from the origin to get ~1e-2 differences, so its own fp32
     rounding is ~1e-4 relative on the cross product - the tolerance is the reference's noise floor, not ours)
Full-size (1080p) cases use size-independent properties: ssim(x, x) == 1 with zero gradient, L1 scaling linearity,
bit-reproducibility of the fixed-order reduction, and agreement with the oracle on the GPU.
"""
import math
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _close(a, b, name, rtol, atol_scale):
    a, b = a.detach().cpu().double(), torch.as_tensor(np.asarray(b)).double()
    scale = max(b.abs().max().item(), 1e-12)
    err = (a - b).abs()
    ok = err <= rtol * b.abs() + atol_scale * scale
    assert bool(ok.all()), f"{name}: max err {err.max().item():.3e} scale {scale:.3e} bad {int((~ok).sum())}"


class _View:
    def __init__(self, wvt, W, H, fovx, fovy):
        self.world_view_transform, self.image_width, self.image_height, self.FoVx, self.FoVy = wvt, W, H, fovx, fovy


def test_ssim_l1_vs_reference_python_golden(native_lib):
    from r3dgw_b200 import image_ops as I
    z = np.load(os.path.join(GOLD, "image_ops.npz"))
    t = {k[3:]: torch.from_numpy(z[k]).cuda() for k in z.files if k.startswith("in_")}
    H, W = t["img"].shape[-2:]
    m3 = t["occl"].expand(3, H, W)                 # expanded view, channel stride 0, as train.py:64-65 builds it
    lam = float(z["lam"])
    launches0 = native_lib.r3dgw_launch_count()
    x = t["img"].clone().requires_grad_(True)
    rec, l1, d = I.photometric_loss(x, t["gt"], m3, lam)
    rec.backward()
    assert native_lib.r3dgw_launch_count() - launches0 == 3      # fwd + finalize + bwd
    np.testing.assert_allclose(l1.item(), z["ref_l1"], rtol=1e-5)
    np.testing.assert_allclose(1.0 - d.item(), z["ref_ssim"], rtol=1e-5)
    np.testing.assert_allclose(rec.item(), z["ref_rec"], rtol=1e-5)
    _close(x.grad, z["ref_d_img"], "d_img", 1e-3, 2e-5)
    # drop-in functions, separately
    x = t["img"].clone().requires_grad_(True)
    s = I.ssim(x, t["gt"], mask=m3)
    s.backward()
    np.testing.assert_allclose(s.item(), z["ref_ssim"], rtol=1e-5)
    _close(x.grad, z["ref_d_ssim"], "d_ssim", 1e-3, 2e-5)
    x = t["img"].clone().requires_grad_(True)
    l = I.l1_loss(x, t["gt"], mask=m3.contiguous())                # materialised [3,H,W] mask: channel stride H*W
    l.backward()
    np.testing.assert_allclose(l.item(), z["ref_l1"], rtol=1e-5)
    # sky-BRDF regulariser: zeros target, mask = 1 - sky_mask (both spellings)
    x = t["img"].clone().requires_grad_(True)
    lz = I.l1_loss(x, torch.zeros_like(x), mask=1 - t["sky"].expand(3, H, W))
    lz.backward()
    np.testing.assert_allclose(lz.item(), z["ref_l1_sky"], rtol=1e-5)
    _close(x.grad, z["ref_d_l1_sky"], "d_l1_sky", 1e-6, 1e-7)
    lz2 = I.l1_loss(t["img"], None, mask=t["sky"], complement_mask=True)
    assert lz2.item() == lz.item()
    # empty mask: the reference returns the constants 1 / 0
    zero = torch.zeros(3, H, W, device="cuda")
    x = t["img"].clone().requires_grad_(True)
    s0, l0 = I.ssim(x, t["gt"], mask=zero), I.l1_loss(x, t["gt"], mask=zero)
    assert s0.item() == float(z["ref_ssim_empty"]) == 1.0 and l0.item() == float(z["ref_l1_empty"]) == 0.0
    (s0 + l0).backward()
    assert float(x.grad.abs().max()) == 0.0


def test_depth_to_normal_vs_reference_python_golden(native_lib):
    from r3dgw_b200 import image_ops as I
    z = np.load(os.path.join(GOLD, "image_ops.npz"))
    depth, sky = torch.from_numpy(z["in_depth"]).cuda(), torch.from_numpy(z["in_sky"]).cuda()
    H, W = depth.shape
    fovx, fovy = (float(v) for v in z["in_fov"])
    view = _View(torch.from_numpy(z["in_wvt"]).cuda(), W, H, fovx, fovy)
    d = (depth * sky).unsqueeze(0).clone().requires_grad_(True)
    n = I.depth_to_normal(view, d)
    assert n.shape == (H, W, 3)
    (n * torch.from_numpy(z["in_gn"]).cuda()).sum().backward()
    assert float((n.detach().cpu() - torch.from_numpy(z["ref_normal"])).abs().max()) <= 2e-5
    assert float(n[0].abs().max()) == 0 and float(n[:, -1].abs().max()) == 0            # zero border
    ref = torch.from_numpy(z["ref_d_depth"]).double()
    got = d.grad.cpu().double()
    assert float((got - ref).norm() / ref.norm()) <= 1e-3
    _close(d.grad, z["ref_d_depth"], "d_depth", 1e-3, 1e-3)


@pytest.mark.parametrize("shape", [(3, 7, 9), (3, 33, 31), (1, 64, 64), (3, 270, 480)])
def test_ssim_l1_vs_oracle_on_gpu_ragged_sizes(native_lib, shape):
    """Sizes smaller than the window, one pixel past a tile edge, single channel; non-binary mask values."""
    import image_oracle as io
    from r3dgw_b200 import image_ops as I
    C, H, W = shape
    g = torch.Generator().manual_seed(C * 1000 + H)
    gt = torch.rand(C, H, W, generator=g).cuda()
    img = (gt + 0.2 * torch.randn(C, H, W, generator=g).cuda()).clamp(0, 1)
    mask = (torch.rand(C, H, W, generator=g) > 0.3).float().cuda()
    mask[:, : H // 3] *= 0.5                                     # fractional entries weigh the sum but are not counted
    for lam in (0.2, 1.0):
        xa = img.clone().requires_grad_(True)
        xb = img.clone().requires_grad_(True)
        ra = I.photometric_loss(xa, gt, mask, lam)[0]
        rb = io.photometric(xb, gt, mask, lam)[0]
        ra.backward(); rb.backward()
        np.testing.assert_allclose(ra.item(), rb.item(), rtol=1e-5)
        _close(xa.grad, xb.grad.cpu().numpy(), f"d_img {shape}", 1e-3, 2e-5)


def test_ssim_l1_full_size_properties(native_lib):
    from r3dgw_b200 import image_ops as I
    import image_oracle as io
    g = torch.Generator().manual_seed(4)
    C, H, W = 3, 1080, 1920
    gt = torch.rand(C, H, W, generator=g).cuda()
    img = (gt + 0.1 * torch.randn(C, H, W, generator=g).cuda()).clamp(0, 1)
    mask = torch.ones(1, H, W, device="cuda")
    mask[:, :216] = 0
    m3 = mask.expand(C, H, W)
    # identity: ssim(x, x) == 1, l1 == 0, gradient of the SSIM term vanishes
    x = gt.clone().requires_grad_(True)
    rec, l1, d = I.photometric_loss(x, gt, m3, 0.2)
    rec.backward()
    assert abs(d.item()) <= 1e-6 and l1.item() == 0.0
    assert float(x.grad.abs().max()) <= 1e-9
    # bit-reproducible (fixed-order reduction, no float atomics)
    vals = set()
    for _ in range(3):
        xx = img.clone().requires_grad_(True)
        r = I.photometric_loss(xx, gt, m3, 0.2)[0]
        r.backward()
        vals.add((r.item(), float(xx.grad.double().sum())))
    assert len(vals) == 1
    # L1 is positively homogeneous: l1(a*x, a*gt) == a * l1(x, gt)
    la, lb = I.l1_loss(img, gt, mask=m3).item(), I.l1_loss(2 * img, 2 * gt, mask=m3).item()
    np.testing.assert_allclose(lb, 2 * la, rtol=1e-6)
    # and the oracle on the GPU at full size
    xa = img.clone().requires_grad_(True)
    xb = img.clone().requires_grad_(True)
    ra = I.photometric_loss(xa, gt, m3, 0.2)[0]
    rb = io.photometric(xb, gt, m3, 0.2)[0]
    ra.backward(); rb.backward()
    np.testing.assert_allclose(ra.item(), rb.item(), rtol=1e-5)
    _close(xa.grad, xb.grad.cpu().numpy(), "d_img 1080p", 1e-3, 2e-5)


def test_depth_to_normal_vs_oracle_full_size(native_lib):
    import image_oracle as io
    from r3dgw_b200 import image_ops as I
    from r3dgw_b200.synthetic import arc_cameras
    H, W = 1080, 1920
    cam = arc_cameras(W, H, 3, centroid_z=8.0, step_deg=10.0)[2].to("cuda")
    yy, xx = torch.meshgrid(torch.arange(H, device="cuda").float(), torch.arange(W, device="cuda").float(), indexing="ij")
    depth = 5.0 + 0.002 * xx + 0.001 * yy + 0.5 * torch.sin(xx / 37.0) * torch.cos(yy / 23.0)
    g = torch.Generator().manual_seed(9)
    gn = torch.randn(H, W, 3, generator=g).cuda()
    da = depth.unsqueeze(0).clone().requires_grad_(True)
    db = depth.unsqueeze(0).clone().requires_grad_(True)
    na = I.depth_to_normal(cam, da)
    nb = io.depth_to_normal(cam.world_view_transform, W, H, cam.FoVx, cam.FoVy, db)
    (na * gn).sum().backward(); (nb * gn).sum().backward()
    # unit length inside, zero on the border
    inner = na[1:-1, 1:-1].norm(dim=-1)
    assert float((inner - 1).abs().max()) <= 1e-5
    assert float(na[0].abs().max()) == 0 and float(na[-1].abs().max()) == 0
    # the oracle's fp32 point subtraction is the noisy side (see module docstring): compare in aggregate and entrywise
    assert float((na - nb).abs().max()) <= 2e-3
    assert float((na - nb).abs().mean()) <= 2e-5
    rel = float((da.grad - db.grad).double().norm() / db.grad.double().norm())
    assert rel <= 2e-3, rel
    # against float64 evaluation of the same formulas (the true value): the kernel is the accurate side
    dd = depth.double().unsqueeze(0).clone().requires_grad_(True)
    nd = io.depth_to_normal(cam.world_view_transform.double(), W, H, cam.FoVx, cam.FoVy, dd)
    (nd * gn.double()).sum().backward()
    err_kernel = float((na.double() - nd).abs().max())
    err_oracle = float((nb.double() - nd).abs().max())
    assert err_kernel <= max(2 * err_oracle, 1e-5), (err_kernel, err_oracle)


def test_u2_training_step_native_vs_reference_arm(native_lib):
    """The whole U2 step of bench.py (render + image-space tail + train.py:78-112 loss + backward): the native
    drop-in arm, the native fused arm and the reference arm (reference CUDA extension + the reference's torch graphs
    for shading / SSIM / L1 / depth_to_normal) must give the same loss and the same parameter gradients."""
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import bench
    import ref_loader
    if not ref_loader.available():
        pytest.skip("oracle/_ref not built")
    import diff_gaussian_rasterization as dgr
    from r3dgw_b200.synthetic import make_scene
    sc = make_scene("small", seed_offset=12)
    cam = sc.camera.to("cuda")
    H, W = cam.image_height, cam.image_width
    gt = torch.rand(3, H, W, generator=torch.Generator().manual_seed(3)).cuda()
    res = {}
    ref_runs = []     # the reference's own atomic-order noise (heavy-tailed on rotations / scales) sets the bar
    for impl, mode in (("reference", "dropin"), ("reference", "dropin"), ("reference", "dropin"), ("native", "dropin"),
                       ("native", "fused")):
        dgr.clear_geometry_cache()
        st = bench.Stepper(impl, mode, sc, cam, torch.device("cuda"))
        loss = st.step(gt)
        res[(impl, mode)] = (loss.item(), [None if p.grad is None else p.grad.detach().clone() for p in st.params])
        if impl == "reference":
            ref_runs.append(res[(impl, mode)][1])
    l_ref, g_ref = res[("reference", "dropin")]
    rel2 = lambda x, y: float((x.double() - y.double()).norm() / y.double().norm().clamp_min(1e-30))   # noqa: E731
    names = ("means3D", "opacity", "scales", "rotations", "albedo", "roughness", "metalness", "env_sh", "sky_sh")
    for key in (("native", "dropin"), ("native", "fused")):
        l, g = res[key]
        np.testing.assert_allclose(l, l_ref, rtol=2e-5, err_msg=str(key))
        for k, (nm, a, b) in enumerate(zip(names, g, g_ref)):
            assert (a is None) == (b is None), nm
            if b is None:
                continue
            refs = [r[k] for r in ref_runs]
            rel = min(rel2(a, r) for r in refs)
            noise = max(rel2(refs[i], refs[j]) for i in range(len(refs)) for j in range(i))
            bar = max(2e-3, 3.0 * noise)
            assert rel <= bar, f"{key} d_{nm}: relative L2 {rel:.3e} (bar {bar:.3e}, reference run-to-run {noise:.3e})"


def test_fused_normal_tail_and_normal_loss_vs_reference_expressions(native_lib):
    """normal_tail / normal_consistency_loss (one kernel each way) against the reference's chain of torch ops around the
    pinned depth_to_normal kernel: gaussian_renderer/__init__.py:231-237, 266-271 and train.py:92-96."""
    import math
    from types import SimpleNamespace
    from r3dgw_b200 import image_ops
    from r3dgw_b200.synthetic import make_camera
    dev = "cuda"
    H, W = 270, 480
    cam = make_camera(W, H).to(dev)
    view = SimpleNamespace(world_view_transform=cam.world_view_transform, image_width=W, image_height=H, FoVx=cam.FoVx,
                           FoVy=cam.FoVy)
    g = torch.Generator().manual_seed(12)
    yy, xx = torch.meshgrid(torch.linspace(0, 1, H), torch.linspace(0, 1, W), indexing="ij")
    depth0 = (4.0 + 2.0 * torch.sin(6 * xx) * torch.cos(4 * yy) + 0.05 * torch.rand(H, W, generator=g))
    base = {"n01": torch.rand(3, H, W, generator=g), "depth": depth0.unsqueeze(0).repeat(3, 1, 1).contiguous()}
    alpha = torch.rand(3, H, W, generator=g).to(dev)
    sky = torch.ones(H, W)
    sky[: H // 5] = 0
    sky[:, 100:140] = 0.5                      # a soft mask value exercises the products
    sky = sky.to(dev)
    occl = (torch.rand(1, H, W, generator=g) > 0.1).float().to(dev)
    w_n = torch.randn(3, H, W, generator=g).to(dev)
    w_r = torch.randn(3, H, W, generator=g).to(dev)

    def leaves():
        return {k: v.clone().to(dev).requires_grad_(True) for k, v in base.items()}

    a = leaves()
    normal, nref = image_ops.normal_tail(view, a["n01"], a["depth"], alpha, sky)
    loss_a = image_ops.normal_consistency_loss(normal, nref, occl, sky)
    ((normal * w_n).sum() + (nref * w_r).sum() + 3.0 * loss_a).backward()

    b = leaves()
    n = (b["n01"] - 0.5) * 2.0
    n = n * sky + torch.ones_like(n) * (1 - sky)
    r = image_ops.depth_to_normal(view, (b["depth"][0] * sky).unsqueeze(0)).permute(2, 0, 1)
    r = r * alpha.detach()
    r = r + torch.ones_like(r) * (1 - sky)
    occl3, sky3 = occl.expand(3, H, W), sky.unsqueeze(0).expand(3, H, W)
    rn, rs = n * occl3 * sky3, r * occl3 * sky3
    loss_b = (1 - (rn * rs).sum(dim=0))[None].mean()
    ((n * w_n).sum() + (r * w_r).sum() + 3.0 * loss_b).backward()

    assert torch.equal(normal, n.detach()), (normal - n).abs().max().item()
    assert (nref - r).abs().max().item() <= 1e-6
    assert abs(float(loss_a) - float(loss_b)) <= 2e-6 * max(1.0, abs(float(loss_b)))
    for k in base:
        ga, gb = a[k].grad, b[k].grad
        scale = gb.abs().max().item()
        assert (ga - gb).abs().max().item() <= 1e-5 * scale + 1e-9, (k, (ga - gb).abs().max().item(), scale)
    # bit-reproducible loss (fixed-order reduction)
    assert torch.equal(image_ops.normal_consistency_loss(normal.detach(), nref.detach(), occl, sky),
                       image_ops.normal_consistency_loss(normal.detach(), nref.detach(), occl, sky))


def test_training_loss_node_equals_the_separate_dropins(native_lib):
    """r3dgw_b200.losses.training_loss (one autograd node) against the sum of the separately pinned drop-ins
    (l1_loss, ssim, normal_consistency_loss, min_scale_loss / depth_loss_gaussians) with train.py's lambdas."""
    from types import SimpleNamespace
    from r3dgw_b200 import image_ops, regularisers
    from r3dgw_b200.losses import training_loss
    dev = "cuda"
    H, W, P = 135, 240, 5000
    g = torch.Generator().manual_seed(21)
    lam = {"dssim": 0.2, "sky_brdf": 0.5, "normal": 0.05, "scale": 100.0, "sky_gauss": 0.05}
    base = {k: torch.rand(3, H, W, generator=g) for k in ("render", "diffuse_color", "specular_color")}
    base["normal"] = torch.randn(3, H, W, generator=g)
    base["normal_ref"] = torch.randn(3, H, W, generator=g)
    base["scales"] = torch.rand(P, 3, generator=g) * 0.1 + 0.01
    base["means3D"] = torch.randn(P, 3, generator=g) * 3 + torch.tensor([0.0, 0.0, 8.0])
    gt = torch.rand(3, H, W, generator=g).to(dev)
    occl = (torch.rand(1, H, W, generator=g) > 0.1).float().to(dev)
    sky = torch.ones(H, W)
    sky[: H // 5] = 0
    sky = sky.to(dev)
    radii = ((torch.rand(P, generator=g) < 0.7).int() * 5).to(dev)
    is_sky = (torch.rand(P, generator=g) < 0.1).to(dev)
    wvt = torch.eye(4, device=dev)
    wvt[3, 2] = 0.5

    def leaves():
        return {k: v.clone().to(dev).requires_grad_(True) for k, v in base.items()}

    a = leaves()
    pkg = {k: a[k] for k in ("render", "diffuse_color", "specular_color", "normal", "normal_ref")}
    pkg["radii"] = radii
    loss_a, parts = training_loss(pkg, gt, occl, sky, a["scales"], a["means3D"], is_sky, wvt, lam)
    (loss_a * 1.7).backward()

    b = leaves()
    occl3, sky3 = occl.expand(3, H, W), sky.unsqueeze(0).expand(3, H, W)
    l1 = image_ops.l1_loss(b["render"], gt, mask=occl3)
    ss = 1.0 - image_ops.ssim(b["render"], gt, mask=occl3)
    loss_b = l1 * (1 - lam["dssim"]) + lam["dssim"] * ss
    loss_b = loss_b + lam["sky_brdf"] * (image_ops.l1_loss(b["diffuse_color"], torch.zeros_like(b["diffuse_color"]), mask=1 - sky3)
                                         + image_ops.l1_loss(b["specular_color"], torch.zeros_like(b["specular_color"]), mask=1 - sky3))
    loss_b = loss_b + lam["normal"] * image_ops.normal_consistency_loss(b["normal"], b["normal_ref"], occl, sky)
    ls, ld, _ = regularisers.gaussian_regularisers(b["scales"], b["means3D"], radii, is_sky, wvt, 0.02)
    loss_b = loss_b + lam["scale"] * ls + lam["sky_gauss"] * ld
    (loss_b * 1.7).backward()

    assert abs(float(loss_a) - float(loss_b)) <= 2e-6 * abs(float(loss_b)), (float(loss_a), float(loss_b))
    assert abs(float(parts[1]) - float(l1)) <= 1e-7 and abs(float(parts[16]) - float(ls)) <= 1e-7
    for k in base:
        ga, gb = a[k].grad, b[k].grad
        assert ga is not None and gb is not None, k
        scale = gb.abs().max().item()
        assert (ga - gb).abs().max().item() <= 2e-6 * scale + 1e-12, (k, (ga - gb).abs().max().item(), scale)
