"""TEST INFRASTRUCTURE: run the reference's UNMODIFIED caller code (gaussian_renderer.render, GaussianModel,
EnvironmentLight, utils.loss_utils) from the copies oracle/build_ref.py installs under oracle/_ref/py/.

Only tests/ and bench.py's reference / stock legs may import this.  Nothing here is product code: it supplies the
sys.modules stubs the reference needs on a box without its pip dependencies and builds a reference
GaussianModel / camera from a synthetic scene, so that
  * `--impl reference` times the STOCK path: reference render() + reference extension + reference torch shading/loss;
  * GPU tests prove that the same unmodified render() works on top of the drop-in modules.

Stubs: imageio / skimage / plyfile / matplotlib / hydra / omegaconf (unused by the math), `scene` and
`scene.NVDIFFREC` as bare namespace packages (their __init__ imports the dataset readers),
`scene.NVDIFFREC.renderutils` (dead code w.r.t. this path; compiles a plugin at import), and
`nvdiffrast.torch.texture` = F.grid_sample(bilinear, border, align_corners=False), the documented equivalent of
filter_mode='linear', boundary_mode='clamp' (nvdiffrast is an un-vendored submodule: parity for that one call is
unpinned).
"""
import contextlib
import importlib
import importlib.util
import math
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
PY = os.path.join(HERE, "_ref", "py")
_loaded = {}


def available():
    return os.path.exists(os.path.join(PY, "gaussian_renderer", "__init__.py"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _texture(tex, uv, filter_mode="linear", boundary_mode="clamp"):
    import torch.nn.functional as F
    assert filter_mode == "linear" and boundary_mode == "clamp"
    out = F.grid_sample(tex.permute(0, 3, 1, 2), uv * 2.0 - 1.0, mode="bilinear", padding_mode="border",
                        align_corners=False)
    return out.permute(0, 2, 3, 1)


def load_ref_python():
    """Import the staged reference modules (once).  Needs a CUDA device (utils/general_utils.py:295 builds a CUDA
    tensor at import time).  Returns a namespace with gaussian_model, loss_utils, light, graphics_utils, sh_utils."""
    if "ns" in _loaded:
        return _loaded["ns"]
    if not available():
        raise ImportError("oracle/_ref/py is absent: run oracle/build_ref.py in the build container")
    if PY not in sys.path:
        sys.path.insert(0, PY)
    for name in ("imageio", "imageio.v3", "skimage", "skimage.measure", "plyfile", "matplotlib", "matplotlib.pyplot",
                 "hydra", "omegaconf", "lpips", "PIL", "PIL.Image", "cv2", "scipy", "scipy.ndimage"):
        try:
            importlib.import_module(name)
        except Exception:   # noqa: BLE001
            _stub(name)
    sys.modules["plyfile"].__dict__.setdefault("PlyData", object)
    sys.modules["plyfile"].__dict__.setdefault("PlyElement", object)
    if "nvdiffrast.torch" not in sys.modules:
        nd = _stub("nvdiffrast")
        nd.torch = _stub("nvdiffrast.torch", texture=_texture)
    scene = _stub("scene")
    scene.__path__ = [os.path.join(PY, "scene")]
    nv = _stub("scene.NVDIFFREC")
    nv.__path__ = [os.path.join(PY, "scene", "NVDIFFREC")]
    _stub("scene.NVDIFFREC.renderutils")
    ns = types.SimpleNamespace()
    ns.general_utils = importlib.import_module("utils.general_utils")
    ns.graphics_utils = importlib.import_module("utils.graphics_utils")
    ns.sh_utils = importlib.import_module("utils.sh_utils")
    ns.loss_utils = importlib.import_module("utils.loss_utils")
    ns.light = importlib.import_module("scene.NVDIFFREC.light")
    ns.util = importlib.import_module("scene.NVDIFFREC.util")
    ns.gaussian_model = importlib.import_module("scene.gaussian_model")
    _loaded["ns"] = ns
    return ns


def load_stock_renderer(dgr_module, alias):
    """The UNMODIFIED gaussian_renderer/__init__.py executed as module `alias`, with `diff_gaussian_rasterization`
    resolving to `dgr_module` (the reference extension's binding or the drop-in) while it is imported."""
    if alias in _loaded:
        return _loaded[alias]
    load_ref_python()
    prev = sys.modules.get("diff_gaussian_rasterization")
    sys.modules["diff_gaussian_rasterization"] = dgr_module
    try:
        spec = importlib.util.spec_from_file_location(alias, os.path.join(PY, "gaussian_renderer", "__init__.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[alias] = mod
        spec.loader.exec_module(mod)
    finally:
        if prev is not None:
            sys.modules["diff_gaussian_rasterization"] = prev
        else:
            sys.modules.pop("diff_gaussian_rasterization", None)
    _loaded[alias] = mod
    return mod


@contextlib.contextmanager
def _cwd(path):
    old = os.getcwd()
    os.chdir(path)
    try:
        yield
    finally:
        os.chdir(old)


def real_fg_lut(device):
    """The reference's split-sum table scene/NVDIFFREC/irrmaps/bsdf_256_256.bin (installed copy), [1,256,256,2]."""
    import numpy as np
    import torch
    path = os.path.join(PY, "scene", "NVDIFFREC", "irrmaps", "bsdf_256_256.bin")
    return torch.as_tensor(np.fromfile(path, dtype=np.float32).reshape(1, 256, 256, 2), dtype=torch.float32, device=device)


def make_ref_envlight(base, sh_degree=4):
    """Reference EnvironmentLight (its __init__ reads the LUT from a cwd-relative path, light.py:41)."""
    ns = load_ref_python()
    with _cwd(PY):
        return ns.light.EnvironmentLight(base, sh_degree)


def make_ref_model(sc, device, requires_grad=True):
    """Reference GaussianModel holding the synthetic scene `sc` (r3dgw_b200.synthetic.Scene): raw parameters are the
    inverse activations of the scene's values; sky Gaussians are stored as (theta, phi) on the radius-50 dome
    (scene/gaussian_model.py:94-103).  get_sky_angles clamps phi to [-pi/2, pi/2], so sky Gaussians behind the z = 0
    plane land on it - they are culled either way (z <= 0.2), the rendered scene is unchanged."""
    import torch
    ns = load_ref_python()
    gm = ns.gaussian_model.GaussianModel()
    sky = sc.is_sky.reshape(-1).to(device)
    m = sc.means3D.to(device)

    def leaf(t):
        return t.detach().clone().contiguous().requires_grad_(requires_grad)

    def logit(x):
        x = x.clamp(1e-6, 1 - 1e-6)
        return torch.log(x / (1 - x))
    gm._xyz = leaf(m[~sky])
    gm._scaling = leaf(torch.log(sc.scales.to(device)))
    gm._rotation = leaf(sc.rotations.to(device))
    gm._opacity = leaf(logit(sc.opacities.to(device)))
    gm._albedo = leaf(logit(sc.albedo.to(device)))
    gm._roughness = leaf(logit(sc.roughness.to(device)))
    gm._metalness = leaf(logit(sc.metalness.to(device)))
    gm._is_sky = sky.reshape(-1, 1)
    ms = m[sky]
    r = 50.0
    theta = torch.acos((-ms[:, 1] / r).clamp(-1, 1))
    phi = torch.atan2(ms[:, 0], ms[:, 2])
    gm._sky_angles = leaf(torch.stack([theta, phi], dim=1))
    gm._sky_radius = torch.tensor(r, dtype=torch.float32, device=device)
    gm._sky_gauss_center = torch.zeros(1, 3, dtype=torch.float32, device=device)
    gm.max_radii2D = torch.zeros(m.shape[0], device=device)
    gm.xyz_gradient_accum = torch.zeros(m.shape[0], 1, device=device)
    gm.denom = torch.zeros(m.shape[0], 1, device=device)
    return gm


def model_leaves(gm):
    return [gm._xyz, gm._scaling, gm._rotation, gm._opacity, gm._albedo, gm._roughness, gm._metalness, gm._sky_angles]


def make_view(cam, sky_mask, occluders_mask=None, gt=None):
    """The attributes render() / depth_to_normal / the loss read from a scene.cameras.Camera."""
    return types.SimpleNamespace(
        FoVx=cam.FoVx, FoVy=cam.FoVy, image_width=int(cam.image_width), image_height=int(cam.image_height),
        world_view_transform=cam.world_view_transform, full_proj_transform=cam.full_proj_transform,
        camera_center=cam.camera_center, sky_mask=sky_mask, occluders_mask=occluders_mask, original_image=gt, uid=0)


PIPE = types.SimpleNamespace(compute_cov3D_python=False, convert_SHs_python=False, debug=False)


def stock_render(renderer, view, gm, envlight, sky_sh, bg, debug=False, specular=True, fix_sky=False):
    """render(...) exactly as train.py:74 calls it (sky_sh [1,4,3], sky_sh_degree 1)."""
    return renderer.render(view, gm, envlight, sky_sh, 1, PIPE, bg, debug=debug, fix_sky=fix_sky, specular=specular)


def train_loss(render_pkg, view, gm, fns, lam):
    """train.py:78-112 in the post-15 000 regime with the function set `fns` (l1_loss, ssim, min_scale_loss,
    depth_loss_gaussians): the reference's utils.loss_utils or drop-ins with the same signatures.  The env-SH
    positivity term (:99-102, 10 random directions, stochastic) is left out, as in every arm of bench.py."""
    import torch
    gt_image = view.original_image
    sky_mask = view.sky_mask.expand_as(gt_image)
    occluders_mask = view.occluders_mask.expand_as(gt_image)
    image, radii = render_pkg["render"], render_pkg["radii"]
    diff_col, spec_col = render_pkg["diffuse_color"], render_pkg["specular_color"]
    Ll1 = fns.l1_loss(image, gt_image, mask=occluders_mask)
    Ssim = 1.0 - fns.ssim(image, gt_image, mask=occluders_mask)
    loss = Ll1 * (1 - lam["dssim"]) + lam["dssim"] * Ssim
    loss_sky_brdf = fns.l1_loss(diff_col, torch.zeros_like(diff_col), mask=1 - sky_mask) + \
        fns.l1_loss(spec_col, torch.zeros_like(spec_col), mask=1 - sky_mask)
    loss = loss + lam["sky_brdf"] * loss_sky_brdf
    rendered_normal = render_pkg["normal"] * occluders_mask * sky_mask
    rendered_surf_normal = render_pkg["normal_ref"] * occluders_mask * sky_mask
    ncl = (1 - (rendered_normal * rendered_surf_normal).sum(dim=0))[None]
    loss = loss + lam["normal"] * ncl.mean()
    loss = loss + lam["scale"] * fns.min_scale_loss(radii, gm)
    loss = loss + lam["sky_gauss"] * fns.depth_loss_gaussians(gm, view, render_pkg["visibility_filter"])
    return loss


def fov_from_tan(t):
    return 2.0 * math.atan(t)
