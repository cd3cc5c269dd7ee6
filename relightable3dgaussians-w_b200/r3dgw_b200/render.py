"""Host-side mirror of the reference's render() call pattern (gaussian_renderer/__init__.py:69-274).

The reference's render() is caller code that must keep working unchanged on top of the drop-ins; it cannot
travel to the GPU box (it imports the whole scene/ package), so tests and bench.py drive the drop-ins through
this mirror, which issues the same sequence: per-Gaussian colours -> 6 raster calls on identical geometry
(render, diffuse_color, specular_color, depth, normal, alpha) [+4 in debug].  `raster` is any module with the
reference binding's API (ours, or the reference extension for the baseline arm); `shade_fn` produces
(colors_precomp, diffuse_color, specular_color).

render_fused() is the single-pass alternative: same return dict, one K4/K5 over all channels.
"""
import math
from dataclasses import dataclass

import torch

from .shading import preshade_gaussians, shade_gaussians, sky_index_of


@dataclass
class GaussianParams:
    """Post-activation Gaussian parameters, laid out as the reference's getters return them
    (scene/gaussian_model.py:73-179): material arrays are foreground-only."""
    means3D: torch.Tensor      # [P,3]
    opacity: torch.Tensor      # [P,1]
    scales: torch.Tensor       # [P,3]
    rotations: torch.Tensor    # [P,4] normalised
    albedo: torch.Tensor       # [Pfg,3]
    roughness: torch.Tensor    # [Pfg,1]
    metalness: torch.Tensor    # [Pfg,1]
    is_sky: torch.Tensor       # [P] bool

    _sky_index = None

    def sky_index(self):
        """(is_sky u8, foreground rank) for the fused kernels; constant until the Gaussian set changes."""
        if self._sky_index is None:
            self._sky_index = sky_index_of(self.is_sky)
        return self._sky_index

    def leaves(self):
        return [self.means3D, self.opacity, self.scales, self.rotations, self.albedo, self.roughness, self.metalness]

    @staticmethod
    def from_scene(sc, requires_grad=False):
        t = [x.clone().requires_grad_(requires_grad) for x in
             (sc.means3D, sc.opacities, sc.scales, sc.rotations, sc.albedo, sc.roughness, sc.metalness)]
        return GaussianParams(*t, is_sky=sc.is_sky)


def gaussian_normals(scales, rotations, dir_pp_normalized):
    """GaussianModel.get_normal (scene/gaussian_model.py:115-122): column of R(q) at argmin(scale), flipped towards
    the camera (utils/general_utils.py:98-119, 158-170).  Caller-side torch, as in the reference."""
    q = rotations / rotations.norm(dim=1, keepdim=True)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)
    idx = scales.min(dim=-1)[1][..., None, None].expand(-1, 3, -1)
    n = R.gather(2, idx).squeeze(2)
    dotprod = torch.sum(n * -dir_pp_normalized, dim=-1, keepdim=True)
    return n * torch.where(dotprod >= 0, 1.0, -1.0)


def fused_shade_fn(fg_lut, sh_degree=4, preshade=True):
    """Shading callback of the native arm.  With preshade=True the whole per-Gaussian graph in front of the rasterizer
    (get_normal, dir_pp, get_depth, the depth / normal colours, shading) is ONE kernel; with False only
    the shading is fused and the rest stays caller-side torch, as in round 1."""
    def fn(pc, normals, env_sh, sky_sh, cam_pos, specular, fix_sky):
        return shade_gaussians(pc.means3D, normals, pc.albedo, pc.roughness, pc.metalness, env_sh, cam_pos, fg_lut,
                               sh_degree, specular=specular, sky_sh=sky_sh, is_sky=pc.is_sky, fix_sky=fix_sky)

    def prepare(camera, pc, env_sh, sky_sh, specular, fix_sky, packed=False):
        return preshade_gaussians(pc.means3D, pc.scales, pc.rotations, pc.albedo, pc.roughness, pc.metalness, env_sh,
                                  camera.camera_center, camera.world_view_transform, fg_lut, sh_degree,
                                  specular=specular, sky_sh=sky_sh, fix_sky=fix_sky, packed=packed,
                                  sky_index=pc.sky_index())
    if preshade:
        fn.fused_prepare = prepare
    return fn


def _prepare(camera, pc, env_sh, sky_sh, shade_fn, specular, fix_sky):
    fused = getattr(shade_fn, "fused_prepare", None)
    if fused is not None:
        return fused(camera, pc, env_sh, sky_sh, specular, fix_sky)
    cam_pos = camera.camera_center
    dir_pp = pc.means3D - cam_pos.reshape(1, 3)
    dir_pp_n = dir_pp / torch.sqrt(torch.clamp((dir_pp * dir_pp).sum(-1, keepdim=True), min=1e-20))
    normals = gaussian_normals(pc.scales, pc.rotations, dir_pp_n)
    colors, diffuse, spec = shade_fn(pc, normals, env_sh, sky_sh, cam_pos, specular, fix_sky)
    # get_depth (scene/gaussian_model.py:125-130): z of the view-space position, repeated over 3 channels
    wvt = camera.world_view_transform
    depth = (pc.means3D @ wvt[:3, 2:3] + wvt[3, 2]).repeat(1, 3)
    return colors, diffuse, spec, depth, 0.5 * normals + 0.5


def image_post_ops_fused(out, camera, sky_mask, n01_image):
    """The same tail as image_post_ops in ONE kernel (r3dgw_b200.image_ops.normal_tail): `n01_image` is the rasterized
    normal-colour image before the (x - 0.5) * 2 step."""
    from .image_ops import normal_tail
    out["normal"], out["normal_ref"] = normal_tail(camera, n01_image, out["depth"], out["alpha"], sky_mask)
    return out


def image_post_ops(out, camera, sky_mask, depth_to_normal):
    """The image-space tail of the reference's render() (gaussian_renderer/__init__.py:233-237, 266-272), same torch
    expressions: the rendered normal composited over the sky, and `normal_ref` from the masked depth map.
    `depth_to_normal(view, depth[1,H,W]) -> [H,W,3]` is the reference's function or its drop-in.
    (normalize_normal_inplace, :54-57, rebinds a local and has no effect; it is not mirrored.)"""
    n = out["normal"]
    out["normal"] = n * sky_mask + torch.ones_like(n) * (1 - sky_mask)
    nref = depth_to_normal(camera, (out["depth"][0] * sky_mask).unsqueeze(0)).permute(2, 0, 1)
    nref = nref * out["alpha"].detach()
    out["normal_ref"] = nref + torch.ones_like(nref) * (1 - sky_mask)
    return out


def render(camera, pc, env_sh, sky_sh, raster, bg_color, shade_fn, scaling_modifier=1.0, debug=False, specular=True,
           fix_sky=False, sky_mask=None, depth_to_normal=None):
    """Same raster-call sequence as the reference's render(); returns its dict.  With `sky_mask` [H,W] and a
    `depth_to_normal` callable the image-space tail (normal compositing, normal_ref) is appended as in the
    reference."""
    means3D = pc.means3D
    screenspace_points = torch.zeros_like(means3D, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:   # noqa: BLE001
        pass
    tanfovx, tanfovy = math.tan(camera.FoVx * 0.5), math.tan(camera.FoVy * 0.5)

    def settings(bg):
        return raster.GaussianRasterizationSettings(
            image_height=int(camera.image_height), image_width=int(camera.image_width), tanfovx=tanfovx, tanfovy=tanfovy,
            bg=bg, scale_modifier=scaling_modifier, viewmatrix=camera.world_view_transform,
            projmatrix=camera.full_proj_transform, sh_degree=-1, campos=camera.camera_center, prefiltered=False)

    rasterizer = raster.GaussianRasterizer(raster_settings=settings(bg_color))
    colors, diffuse, spec, depth, normal01 = _prepare(camera, pc, env_sh, sky_sh, shade_fn, specular, fix_sky)

    def call(r, feats):
        return r(means3D=means3D, means2D=screenspace_points, shs=None, colors_precomp=feats, opacities=pc.opacity,
                 scales=pc.scales, rotations=pc.rotations, cov3D_precomp=None)

    rendered, radii = call(rasterizer, colors)
    out = {"render": rendered, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii}
    extras = {"diffuse_color": diffuse, "specular_color": spec, "depth": depth, "normal": normal01}
    if debug:
        P = means3D.shape[0]
        fg = ~pc.is_sky
        rough_all = torch.zeros(P, 1, device=means3D.device)
        rough_all[fg] = pc.roughness
        metal_all = torch.zeros(P, 1, device=means3D.device)
        metal_all[fg] = pc.metalness
        alb_all = torch.ones_like(means3D)
        alb_all[fg] = pc.albedo
        sky_color = torch.zeros_like(colors)
        sky_color[pc.is_sky] = colors[pc.is_sky]
        extras.update({"sky_color": sky_color, "roughness": rough_all.repeat(1, 3), "metalness": metal_all.repeat(1, 3),
                       "albedo": alb_all})
    fuse_tail = sky_mask is not None and depth_to_normal is not None and getattr(depth_to_normal, "fused_tail", False)
    n01_img = None
    for k, f in extras.items():
        img = call(rasterizer, f)[0]
        if k == "normal":
            n01_img = img
            if not fuse_tail:
                img = (img - 0.5) * 2.0
        out[k] = img
    rasterizer_alpha = raster.GaussianRasterizer(raster_settings=settings(torch.zeros(3, device=means3D.device)))
    out["alpha"] = call(rasterizer_alpha, torch.ones_like(means3D))[0]
    if fuse_tail:
        image_post_ops_fused(out, camera, sky_mask, n01_img)
    elif sky_mask is not None and depth_to_normal is not None:
        image_post_ops(out, camera, sky_mask, depth_to_normal)
    return out


def render_fused(camera, pc, env_sh, sky_sh, bg_color, shade_fn, scaling_modifier=1.0, specular=True, fix_sky=False,
                 sky_mask=None, depth_to_normal=None):
    """One geometry pass + ONE 16-channel blend (render 3, diffuse 3, specular 3, depth 3, normal 3, alpha 1) instead
    of six 3-channel raster calls.  Returns the same keys as render(debug=False)."""
    import diff_gaussian_rasterization as dgr
    means3D = pc.means3D
    screenspace_points = torch.zeros_like(means3D, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:   # noqa: BLE001
        pass
    s = dgr.GaussianRasterizationSettings(
        image_height=int(camera.image_height), image_width=int(camera.image_width),
        tanfovx=math.tan(camera.FoVx * 0.5), tanfovy=math.tan(camera.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=camera.world_view_transform,
        projmatrix=camera.full_proj_transform, sh_degree=-1, campos=camera.camera_center, prefiltered=False)
    fused = getattr(shade_fn, "fused_prepare", None)
    if fused is not None:
        feats = fused(camera, pc, env_sh, sky_sh, specular, fix_sky, packed=True)                         # [P,16]
    else:
        colors, diffuse, spec, depth, normal01 = _prepare(camera, pc, env_sh, sky_sh, shade_fn, specular, fix_sky)
        feats = torch.cat([colors, diffuse, spec, depth, normal01, torch.ones_like(means3D[:, :1])], dim=1)
    bg16 = torch.cat([bg_color.reshape(3).repeat(5), torch.zeros(1, device=means3D.device)])
    img, radii = dgr.rasterize_gaussians_features(means3D, screenspace_points, feats, pc.opacity, pc.scales,
                                                  pc.rotations, None, s, bg=bg16)
    out = {"render": img[0:3], "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii,
           "diffuse_color": img[3:6], "specular_color": img[6:9], "depth": img[9:12],
           "alpha": img[15:16].expand(3, -1, -1)}
    if sky_mask is not None and depth_to_normal is not None and getattr(depth_to_normal, "fused_tail", False):
        image_post_ops_fused(out, camera, sky_mask, img[12:15])
        return out
    out["normal"] = (img[12:15] - 0.5) * 2.0
    if sky_mask is not None and depth_to_normal is not None:
        image_post_ops(out, camera, sky_mask, depth_to_normal)
    return out
