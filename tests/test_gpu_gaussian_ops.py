"""GPU parity of the per-Gaussian caller graph moved into kernels against goldens produced by
the reference's OWN Python (tests/golden/make_gaussian_golden.py runs scene/gaussian_model.py, utils/general_utils.py,
utils/loss_utils.py and scene/NVDIFFREC/light.py unmodified on CPU tensors):

  gaussian_normals            <- GaussianModel.get_normal (build_rotation / get_minimum_axis / flip_align_view)
  preshade_gaussians          <- get_normal + dir_pp + get_depth + the depth / normal colours + shade, one kernel each way
  gaussian_regularisers       <- min_scale_loss + depth_loss_gaussians
  EnvironmentLight.{get_diffuse_irradiance, get_specular_light_sh, sample_illumination, render_sh}

Tolerances (floating point): values 1e-5 relative + 2e-6 of the tensor scale, gradients 1e-3 relative + 1e-5 of scale.
The one discrete decision - the flip sign of a normal whose dot product with the view direction is ~0 - is excluded
where |n . viewdir| < 1e-5.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _close(a, b, name, rtol, atol_scale, rows=None):
    a, b = a.detach().cpu().double(), torch.as_tensor(np.asarray(b)).double()
    assert a.shape == b.shape, f"{name}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    if rows is not None:
        a, b = a[rows], b[rows]
    scale = max(b.abs().max().item(), 1e-12)
    err = (a - b).abs()
    ok = err <= rtol * b.abs() + atol_scale * scale
    assert bool(ok.all()), f"{name}: max err {err.max().item():.3e} scale {scale:.3e} bad {int((~ok).sum())}"


@pytest.fixture(scope="module")
def gold():
    z = np.load(os.path.join(GOLD, "gaussian_ops.npz"))
    return {k: z[k] for k in z.files}


def _stable_rows(gold):
    """rows whose flip decision is not a rounding coin toss"""
    d = (gold["normal_noflip"] * -gold["dir_pp_normalized"]).sum(-1)
    return torch.from_numpy(np.abs(d) > 1e-5)


def test_gaussian_normals_vs_reference_get_normal(native_lib, gold):
    from r3dgw_b200.shading import gaussian_normals
    sc = torch.from_numpy(gold["scales"]).cuda()
    rot = torch.from_numpy(gold["rotations"]).cuda().requires_grad_(True)
    xyz = torch.from_numpy(gold["xyz"]).cuda()
    cam = torch.from_numpy(gold["campos"]).cuda()
    n0 = gaussian_normals(sc, rot)
    _close(n0, gold["normal_noflip"], "normal (no flip)", 1e-5, 2e-6)
    rows = _stable_rows(gold)
    n = gaussian_normals(sc, rot, xyz, cam)
    _close(n, gold["normal"], "normal", 1e-5, 2e-6, rows)
    (n * torch.from_numpy(gold["w_normal"]).cuda()).sum().backward()
    _close(rot.grad, gold["g_rotations"], "d_rotations", 1e-3, 1e-5, rows)
    assert float((n.norm(dim=1) - 1).abs().max()) < 1e-5


def test_preshade_matches_reference_graph_and_separate_ops(native_lib, gold):
    """Values of the fused kernel against (i) the reference goldens for normal / depth and (ii) the already pinned
    shade_gaussians kernel fed with those normals; gradients against the reference's autograd (rotations, xyz via
    depth) and against shade_gaussians' backward."""
    import shade_oracle as so
    from r3dgw_b200.shading import gaussian_normals, preshade_gaussians, shade_gaussians
    from r3dgw_b200.synthetic import make_env_sh
    dev = "cuda"
    N = gold["xyz"].shape[0]
    g = torch.Generator().manual_seed(3)
    is_sky = torch.from_numpy(gold["is_sky"]).to(dev)
    nfg = int((~is_sky).sum())
    base = {"xyz": torch.from_numpy(gold["xyz"]), "scales": torch.from_numpy(gold["scales"]),
            "rot": torch.from_numpy(gold["rotations"]), "alb": torch.rand(nfg, 3, generator=g),
            "kr": torch.rand(nfg, 1, generator=g), "km": torch.rand(nfg, 1, generator=g),
            "env": make_env_sh(g)[0], "sky": torch.randn(4, 3, generator=g) * 0.3}
    cam = torch.from_numpy(gold["campos"]).to(dev)
    wvt = torch.from_numpy(gold["wvt"]).to(dev)
    lut = so.synthetic_fg_lut().to(dev)
    W = [torch.randn(N, 3, generator=g).to(dev) for _ in range(5)]

    def leaves():
        return {k: v.clone().to(dev).requires_grad_(True) for k, v in base.items()}

    a = leaves()
    outs = preshade_gaussians(a["xyz"], a["scales"], a["rot"], a["alb"], a["kr"], a["km"], a["env"], cam, wvt, lut, 4,
                              specular=True, sky_sh=a["sky"], is_sky=is_sky)
    assert len(outs) == 5
    sum((o * w).sum() for o, w in zip(outs, W)).backward()

    b = leaves()
    nrm = gaussian_normals(b["scales"], b["rot"], b["xyz"], cam)
    col = shade_gaussians(b["xyz"], nrm, b["alb"], b["kr"], b["km"], b["env"], cam, lut, 4, specular=True,
                          sky_sh=b["sky"], is_sky=is_sky)
    depth = (b["xyz"] @ wvt[:3, 2:3] + wvt[3, 2]).repeat(1, 3)
    ref_outs = list(col) + [depth, 0.5 * nrm + 0.5]
    sum((o * w).sum() for o, w in zip(ref_outs, W)).backward()

    rows = _stable_rows(gold)
    _close(outs[3][:, :1], gold["depth"], "depth vs reference get_depth", 1e-5, 2e-6)
    _close((outs[4] - 0.5) * 2.0, gold["normal"], "normal colour vs reference get_normal", 1e-5, 4e-6, rows)
    for nm, o, r in zip(("rgb", "diffuse", "specular", "depth3", "normal01"), outs, ref_outs):
        _close(o, r.detach().cpu(), nm, 1e-5, 2e-6)
    for k in base:
        ga = a[k].grad if a[k].grad is not None else torch.zeros_like(a[k])
        gb = b[k].grad if b[k].grad is not None else torch.zeros_like(b[k])
        _close(ga, gb.cpu(), "d_" + k, 1e-3, 1e-5)

    # packed [P,16] form: same numbers in the columns, ones in column 15
    c = leaves()
    feats = preshade_gaussians(c["xyz"], c["scales"], c["rot"], c["alb"], c["kr"], c["km"], c["env"], cam, wvt, lut, 4,
                               specular=True, sky_sh=c["sky"], is_sky=is_sky, packed=True)
    assert feats.shape == (N, 16)
    assert torch.equal(feats[:, :15], torch.cat(outs, dim=1))
    assert bool((feats[:, 15] == 1).all())
    gw = torch.cat(W + [torch.zeros(N, 1, device=dev)], dim=1)
    (feats * gw).sum().backward()
    for k in base:
        ga = a[k].grad if a[k].grad is not None else torch.zeros_like(a[k])
        gc = c[k].grad if c[k].grad is not None else torch.zeros_like(c[k])
        _close(gc, ga.cpu(), "packed d_" + k, 1e-5, 1e-6)


def test_preshade_gradients_vs_reference_autograd(native_lib, gold):
    """d(sum w_n . normal + sum w_d . depth) w.r.t. rotations / xyz against the reference's own autograd through
    get_normal and get_depth (golden g_rotations / g_xyz)."""
    import shade_oracle as so
    from r3dgw_b200.shading import preshade_gaussians
    from r3dgw_b200.synthetic import make_env_sh
    dev = "cuda"
    N = gold["xyz"].shape[0]
    g = torch.Generator().manual_seed(3)
    xyz = torch.from_numpy(gold["xyz"]).to(dev).requires_grad_(True)
    rot = torch.from_numpy(gold["rotations"]).to(dev).requires_grad_(True)
    sc = torch.from_numpy(gold["scales"]).to(dev).requires_grad_(True)
    outs = preshade_gaussians(xyz, sc, rot, torch.rand(N, 3, generator=g).to(dev), torch.rand(N, 1, generator=g).to(dev),
                              None, make_env_sh(g)[0].to(dev), torch.from_numpy(gold["campos"]).to(dev),
                              torch.from_numpy(gold["wvt"]).to(dev), so.synthetic_fg_lut().to(dev), 4, specular=False)
    wn, wd = torch.from_numpy(gold["w_normal"]).to(dev), torch.from_numpy(gold["w_depth"]).to(dev)
    # normal01 = 0.5 n + 0.5 -> weight 2 w_n; depth3 = 3 copies -> weight w_d on the first copy
    ((outs[4] * 2.0 * wn).sum() + (outs[3][:, :1] * wd).sum()).backward()
    rows = _stable_rows(gold)
    _close(rot.grad, gold["g_rotations"], "d_rotations", 1e-3, 1e-5, rows)
    _close(xyz.grad, gold["g_xyz"], "d_xyz (depth)", 1e-3, 1e-5)
    assert sc.grad is None or float(sc.grad.abs().max()) == 0.0


def test_regularisers_vs_reference_losses(native_lib, gold):
    from r3dgw_b200.regularisers import gaussian_regularisers
    dev = "cuda"
    raw = torch.from_numpy(gold["scaling_raw"]).to(dev).requires_grad_(True)
    xyz = torch.from_numpy(gold["xyz"]).to(dev).requires_grad_(True)
    radii = torch.from_numpy(gold["radii"]).to(dev)
    is_sky = torch.from_numpy(gold["is_sky"]).to(dev)
    wvt = torch.from_numpy(gold["wvt"]).to(dev)
    ls, ld, stats = gaussian_regularisers(torch.exp(raw), xyz, radii, is_sky, wvt, 0.02)
    (3.0 * ls + 0.7 * ld).backward()
    _close(ls, gold["min_scale_loss"], "min_scale_loss", 2e-6, 0)
    _close(ld, gold["depth_loss"], "depth_loss_gaussians", 2e-6, 0)
    _close(raw.grad, gold["g_reg_scaling_raw"], "d_scaling", 1e-3, 1e-6)
    _close(xyz.grad, gold["g_reg_xyz"], "d_xyz", 1e-3, 1e-6)
    vis = gold["radii"] > 0
    assert int(stats[2]) == int((vis & ~gold["is_sky"]).sum()) and int(stats[3]) == int((vis & gold["is_sky"]).sum())
    # bit-reproducible (fixed-order reduction)
    ls2, ld2, _ = gaussian_regularisers(torch.exp(raw), xyz, radii, is_sky, wvt, 0.02)
    assert torch.equal(ls, ls2) and torch.equal(ld, ld2)
    # edge cases: nothing visible -> 0 (the reference returns None there); no sky Gaussian -> NaN depth term like
    # torch.mean of an empty selection
    z = torch.zeros_like(radii)
    l0, _, _ = gaussian_regularisers(torch.exp(raw), xyz, z, is_sky, wvt)
    assert float(l0) == 0.0
    _, d_nosky, _ = gaussian_regularisers(torch.exp(raw), xyz, radii, torch.zeros_like(is_sky), wvt)
    assert bool(torch.isnan(d_nosky))


def test_environment_light_is_a_superset_of_the_reference_class(native_lib, gold):
    """Every public attribute of scene/NVDIFFREC/light.py's class exists on the drop-in, and the four non-hot methods
    reproduce the reference's Python (goldens)."""
    from r3dgw_b200.shading import EnvironmentLight
    import shade_oracle as so
    ref_public = ["clone", "get_shdim", "get_shdegree", "get_base", "set_base", "get_diffuse_irradiance",
                  "get_specular_light_sh", "sample_illumination", "shade", "render_sh"]       # light.py:14-205
    for name in ref_public:
        assert hasattr(EnvironmentLight, name), name
    base = torch.from_numpy(gold["env_base"]).cuda()
    env = EnvironmentLight(base, 4, fg_lut=so.synthetic_fg_lut().cuda())
    nrm = torch.from_numpy(gold["env_normals"]).cuda()
    kr = torch.from_numpy(gold["env_kr"]).cuda()
    pos = torch.from_numpy(gold["env_pos"]).cuda()
    cam = torch.from_numpy(gold["campos"]).cuda()
    _close(env.get_diffuse_irradiance(nrm), gold["diffuse_irradiance"], "get_diffuse_irradiance", 1e-5, 2e-6)
    sl = env.get_specular_light_sh(kr)
    assert sl.shape == (257, 25, 3)
    _close(sl, gold["specular_light_sh"], "get_specular_light_sh", 1e-5, 2e-6)
    _close(env.sample_illumination(pos, cam.repeat(257, 1)), gold["sample_illumination"], "sample_illumination", 1e-5, 4e-6)
    img = env.render_sh(width=64)
    assert img.device.type == "cpu" and img.dtype == torch.float64 and tuple(img.shape) == (32, 64, 3)
    _close(img, gold["render_sh_64"], "render_sh", 1e-6, 1e-7)
    assert tuple(env.render_sh().shape) == (300, 600, 3)
    c = env.clone()
    assert isinstance(c, EnvironmentLight) and torch.equal(c.get_base, base)
