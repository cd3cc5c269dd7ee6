"""Gradient parity adjudicated against a deterministic FP64 evaluation of the reference backward pass.

The reference accumulates its gradients with float atomics, so it is a NOISY oracle: its own scale / rotation
gradients differ run to run by up to 1e-2 relative L2 (profiles/r01_grad_noise*.txt).  oracle/raster_f64.cu restates
backward.cu:144-557 with every decision taken in float exactly as the reference takes it and the arithmetic in
double, evaluated over the reference's OWN forward buffers.  Enforced per gradient tensor, at configs small, A and
B (the config every bench number is quoted on), for the single U1 call and for render()'s six-call pattern:

    ||new - f64||_2 <= 1e-4 ||f64||_2                  (north_star asks <= 1e-3 relative; measured: <= 2.2e-5)
    ||new - f64||_2 <= k ||ref - f64||_2 + 1e-6 ||f64||_2    (k = 1.5 for the back-to-front geometry pass; k = 4 with the
                                                               default front-to-back cached pass, see REF_FACTOR below)
    99.99 % of entries: |new - f64| <= 1e-3 |f64| + 1e-5 max|f64|
(The reference itself sits at 6e-4 .. 1e-3 on rotations and 1e-4 .. 2e-4 on scales at A / B: its float evaluation of
the cov2D -> cov3D -> quaternion chain cancels heavily.  The new fused geometry backward evaluates that chain in FP64
- the kernel is HBM-bound - and lands 40-60x closer to the FP64 truth; screen-space gradients, whose error is the
float accumulation order, are level with the reference.)

The measured errors of both implementations are printed (pytest -s) and kept in profiles/r02_grad_f64.txt.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
NAMES = ("means3D", "means2D", "opacities", "scales", "rotations")


def _settings(mod, sc, bg):
    cam = sc.camera
    return mod.GaussianRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=bg,
        scale_modifier=1.0, viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform, sh_degree=-1,
        campos=cam.camera_center, prefiltered=False)


@pytest.fixture(scope="module")
def mods(native_lib):
    import ref_loader
    import raster_f64
    if not ref_loader.available() or not raster_f64.available():
        pytest.skip("oracle/_ref or oracle/_build/libraster_f64.so not built")
    import diff_gaussian_rasterization as dgr
    return dgr, ref_loader.load_ref_dgr(), ref_loader, raster_f64


def _run(mod, sc, feats, bgs, dLs):
    """render()'s call pattern: every feature set blended over the same leaves, one backward of the summed loss."""
    m3, op, scl, rot = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.opacities, sc.scales, sc.rotations)]
    cols = [f.clone().requires_grad_(True) for f in feats]
    m2 = torch.zeros_like(m3, requires_grad=True)
    loss = 0
    for c, bg, d in zip(cols, bgs, dLs):
        rast = mod.GaussianRasterizer(_settings(mod, sc, bg))
        img, radii = rast(means3D=m3, means2D=m2, opacities=op, colors_precomp=c, scales=scl, rotations=rot)
        loss = loss + (img * d).sum()
    loss.backward()
    g = {"means3D": m3.grad, "means2D": m2.grad, "opacities": op.grad, "scales": scl.grad, "rotations": rot.grad}
    return g, [c.grad for c in cols]


# R3DGW_GEOM_BWD_F32=1 selects the float instantiation of the geometry chain: as accurate as the reference (north_star's
# 1e-3), not as the FP64 default
import os
F32_CHAIN = os.environ.get("R3DGW_GEOM_BWD_F32", "0") == "1"
L2_BAR = 1.5e-3 if F32_CHAIN else 1e-4
# The default geometry pass of a multi-blend frame walks the alpha cache front to back and forms the reference's suffix
# sums as <forward pixel, dL/dpixel> minus a running prefix (csrc/blend.cu, blend_backward_geometry_cached_kernel): that
# subtraction costs a factor 2-3 on the SCREEN-SPACE gradients' error (3e-6 instead of 1.2e-6 relative L2 at config B -
# both 30x inside the 1e-4 bar) for a pass that is 26 % faster.  R3DGW_BWDG_CACHED=0 selects the back-to-front pass,
# which is held to 1.5x the reference's own error.
BWDG_CACHED = os.environ.get("R3DGW_BWDG_CACHED", "1") != "0"
REF_FACTOR = 2.5 if F32_CHAIN else (4.0 if BWDG_CACHED else 1.5)
STRICT_FRAC = 0.999 if F32_CHAIN else 0.9999


def _check(name, new, ref, f64, report):
    new, ref = new.double(), ref.double()
    nf = f64.norm().item()
    e_new, e_ref = (new - f64).norm().item(), (ref - f64).norm().item()
    scale = f64.abs().max().item()
    strict = ((new - f64).abs() <= 1e-3 * f64.abs() + 1e-5 * scale).double().mean().item()
    report.append(f"{name:<12} ||f64|| {nf:.4e}  new-f64 {e_new / nf:.3e}  ref-f64 {e_ref / nf:.3e}  "
                  f"new/ref {e_new / max(e_ref, 1e-300):.2f}  entries within strict bar {strict:.6f}")
    bad = []
    if not e_new <= L2_BAR * nf:
        bad.append(f"{name}: relative L2 error vs FP64 {e_new / nf:.3e} > {L2_BAR}")
    if not e_new <= REF_FACTOR * e_ref + 1e-6 * nf:
        bad.append(f"{name}: new {e_new / nf:.3e} vs reference {e_ref / nf:.3e} (FP64 oracle)")
    if not strict >= STRICT_FRAC:
        bad.append(f"{name}: only {strict:.5f} of entries within 1e-3 relative + 1e-5 of scale")
    return bad


@pytest.mark.parametrize("name,n_calls", [("small", 1), ("small", 4), ("A", 1), ("A", 5), ("B", 1), ("B", 5)])
def test_gradients_vs_fp64_evaluation_of_the_reference_backward(mods, name, n_calls):
    dgr, ref, rl, f64 = mods
    from r3dgw_b200.synthetic import make_scene
    sc = make_scene(name).to("cuda")
    P, H, W = sc.P, sc.camera.image_height, sc.camera.image_width
    gen = torch.Generator().manual_seed(4321 + n_calls)
    feats = [sc.colors] + [torch.rand(P, 3, generator=gen).cuda() for _ in range(n_calls - 1)]
    bgs = [torch.full((3,), 0.25 * (k % 2), device="cuda") for k in range(n_calls)]
    dLs = [(torch.randn(3, H, W, generator=gen) * 1e-3).cuda() for _ in range(n_calls)]

    e = torch.Tensor([])
    s = _settings(ref, sc, bgs[0])
    R, _c, radii, gb, bb, ib = ref._C.rasterize_gaussians(
        s.bg, sc.means3D, feats[0], sc.opacities, sc.scales, sc.rotations, s.scale_modifier, e, s.viewmatrix,
        s.projmatrix, s.tanfovx, s.tanfovy, s.image_height, s.image_width, e, s.sh_degree, s.campos, s.prefiltered)
    rb = rl.parse_ref_buffers(P, R, H, W, gb, bb, ib)
    truth = f64.backward_f64(rb, sc.means3D, sc.scales, sc.rotations, list(zip(feats, bgs, dLs)), sc.camera, radii)

    g_ref, c_ref = _run(ref, sc, feats, bgs, dLs)
    dgr.clear_geometry_cache()
    g_new, c_new = _run(dgr, sc, feats, bgs, dLs)
    report = [f"config {name}, {n_calls} raster call(s), P={P} R={R}"]
    bad = []
    for nm in NAMES:
        bad += _check(nm, g_new[nm], g_ref[nm], truth[nm], report)
    for k in range(n_calls):
        bad += _check(f"colors[{k}]", c_new[k], c_ref[k], truth["colors"][k], report)
    print("\n".join(report))
    assert not bad, "\n".join(bad + report)
