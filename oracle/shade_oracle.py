"""CPU restatement (torch, fp32) of the reference's per-Gaussian relit shading.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may import this; the
product path (relightable3dgaussians-w_b200/) never does.

Every function cites the reference lines it follows (paths relative to /root/reference/).  It is written as
the same chain of torch ops as the reference (same association order), so that autograd through it is the
gradient oracle for csrc/shade.cu as well.

Parity status: PINNED against the reference's own Python (scene/NVDIFFREC/light.py + utils/sh_utils.py,
imported in the build container by tests/golden/make_shade_golden.py, fixture tests/golden/shade_*.npz)
for everything except the FG-LUT fetch: `nvdiffrast.torch.texture` is an un-vendored submodule
(nvdiffrast @ 729261dc, absent from /root/reference), so the bilinear/clamp lookup below is a restatement of
its published semantics (texel centres at (i+0.5)/res, clamp-to-edge) and is "parity unpinned" for that
one call (light.py:170).
"""
import torch

# utils/sh_utils.py:32-75
C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435]
C4 = [2.5033429417967046, -1.7701307697799304, 0.9461746957575601, -0.6690465435572892, 0.10578554691520431,
      -0.6690465435572892, 0.47308734787878004, -1.7701307697799304, 0.6258357354491761]


def eval_sh(deg, sh, dirs):
    """utils/sh_utils.py:81-151 (degrees 0..4).  sh: [..., C, (deg+1)^2], dirs: [..., 3] -> [..., C]."""
    assert 0 <= deg <= 4
    result = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        result = (result - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3])
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            result = (result + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5] +
                      C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + C2[3] * xz * sh[..., 7] +
                      C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                result = (result + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10] +
                          C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] +
                          C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                          C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14] +
                          C3[6] * x * (xx - 3 * yy) * sh[..., 15])
                if deg > 3:
                    result = (result + C4[0] * xy * (xx - yy) * sh[..., 16] + C4[1] * yz * (3 * xx - yy) * sh[..., 17] +
                              C4[2] * xy * (7 * zz - 1) * sh[..., 18] + C4[3] * yz * (7 * zz - 3) * sh[..., 19] +
                              C4[4] * (zz * (35 * zz - 30) + 3) * sh[..., 20] + C4[5] * xz * (7 * zz - 3) * sh[..., 21] +
                              C4[6] * (xx - yy) * (7 * zz - 1) * sh[..., 22] + C4[7] * xz * (xx - 3 * yy) * sh[..., 23] +
                              C4[8] * (xx * (xx - 3 * yy) - yy * (3 * xx - yy)) * sh[..., 24])
    return result


def gauss_kernel(roughness, sh_degree):
    """utils/sh_utils.py:162-181: per-band Gauss-Weierstrass factors exp(-l(l+1) * 0.3 * roughness) -> [N, (deg+1)^2]."""
    l_idxs = torch.arange(sh_degree + 1, dtype=torch.float32, device=roughness.device).view(1, -1)
    per_band = torch.exp(-l_idxs * (l_idxs + 1) * (0.3 * roughness))
    cols = []
    for l in range(sh_degree + 1):
        cols.append(per_band[..., l:l + 1].expand(roughness.shape[0], 2 * l + 1))
    return torch.cat(cols, dim=-1)


def gamma_correction(rgb, gamma=2.2):
    """utils/sh_utils.py:184-187 == scene/NVDIFFREC/util.py:523-526."""
    return (rgb.clamp(min=0.0, max=1.0) + 1e-4).pow(1.0 / gamma)


# scene/NVDIFFREC/util.py:21-31
def dot(x, y):
    return torch.sum(x * y, -1, keepdim=True)


def reflect(x, n):
    return 2 * dot(x, n) * n - x


def length(x, eps=1e-20):
    return torch.sqrt(torch.clamp(dot(x, x), min=eps))


def safe_normalize(x, eps=1e-20):
    return x / length(x, eps)


def texture_linear_clamp(lut, uv):
    """Restatement of nvdiffrast.torch.texture(lut[1,H,W,C], uv[...,2], filter_mode='linear',
    boundary_mode='clamp') (call site light.py:170).  Texel centres at (i+0.5)/res, clamp-to-edge.
    PARITY UNPINNED: nvdiffrast is not in /root/reference."""
    H, W = lut.shape[1], lut.shape[2]
    u = uv[..., 0].clamp(0.0, 1.0)
    v = uv[..., 1].clamp(0.0, 1.0)
    x = u * W - 0.5
    y = v * H - 0.5
    x0f, y0f = torch.floor(x), torch.floor(y)
    fx, fy = (x - x0f).unsqueeze(-1), (y - y0f).unsqueeze(-1)
    x0 = x0f.long().clamp(0, W - 1)
    x1 = (x0f.long() + 1).clamp(0, W - 1)
    y0 = y0f.long().clamp(0, H - 1)
    y1 = (y0f.long() + 1).clamp(0, H - 1)
    t = lut[0]
    t00, t01, t10, t11 = t[y0, x0], t[y0, x1], t[y1, x0], t[y1, x1]
    return (1 - fy) * ((1 - fx) * t00 + fx * t01) + fy * ((1 - fx) * t10 + fx * t11)


IRR_C1, IRR_C2, IRR_C3, IRR_C4, IRR_C5 = 0.429043, 0.511664, 0.743125, 0.886227, 0.247708


def diffuse_irradiance(base, normal):
    """scene/NVDIFFREC/light.py:65-94 (order-2 Ramamoorthi-Hanrahan irradiance; base [25,3] or [9,3])."""
    x, y, z = normal[..., 0, None], normal[..., 1, None], normal[..., 2, None]
    return (IRR_C1 * base[8, :] * (x ** 2 - y ** 2) + IRR_C3 * base[6, :] * (z ** 2) + IRR_C4 * base[0, :] -
            IRR_C5 * base[6, :] + 2 * IRR_C1 * base[4, :] * x * y + 2 * IRR_C1 * base[7, :] * x * z +
            2 * IRR_C1 * base[5, :] * y * z + 2 * IRR_C2 * base[3, :] * x + 2 * IRR_C2 * base[1, :] * y +
            2 * IRR_C2 * base[2, :] * z)


def shade(base, sh_degree, fg_lut, pos, normal, albedo, view_pos, kr=None, km=None, specular=True):
    """scene/NVDIFFREC/light.py:131-193 on flat [N,*] tensors.  Returns (rgb, diffuse_ldr, specular_ldr)."""
    irr = torch.clamp_min(diffuse_irradiance(base, normal), 1e-4)
    diffuse_hdr = albedo * irr
    diffuse_ldr = gamma_correction(diffuse_hdr)
    if not specular:
        return diffuse_ldr, diffuse_ldr, torch.zeros_like(diffuse_ldr)
    wo = safe_normalize(view_pos - pos)
    reflvec = safe_normalize(reflect(wo, normal))
    NdotV = torch.clamp(dot(wo, normal), min=1e-4)
    fg_uv = torch.cat((NdotV, kr), dim=-1)
    fg = texture_linear_clamp(fg_lut, fg_uv)
    spec_light = gauss_kernel(kr, sh_degree).unsqueeze(-1) * base.unsqueeze(0)      # light.py:97-119
    spec_irr = torch.clamp_min(eval_sh(sh_degree, spec_light.transpose(1, 2), reflvec), 1e-4)
    if km is None:
        F0 = torch.ones_like(albedo) * 0.04
    else:
        F0 = (1.0 - km) * 0.04 + albedo * km
    reflectivity = F0 * fg[..., 0:1] + fg[..., 1:2]
    spec_hdr = spec_irr * reflectivity
    shaded = diffuse_hdr + spec_hdr if km is None else (1 - km) * diffuse_hdr + spec_hdr
    return gamma_correction(shaded), diffuse_ldr, gamma_correction(spec_hdr)


def sky_color(sky_sh, dirs, fix_sky=False):
    """gaussian_renderer/__init__.py:143-148.  sky_sh [1,4,3] or [4,3]; dirs [Ns,3] unit."""
    if fix_sky:
        return torch.ones_like(dirs)
    s = sky_sh if sky_sh.dim() == 3 else sky_sh.unsqueeze(0)
    return torch.clamp_min(eval_sh(1, s.transpose(1, 2), dirs) + 0.5, 0.0)


def shade_all(base, sh_degree, fg_lut, positions, normals, albedo, roughness, metalness, is_sky, cam_pos, sky_sh,
              specular=True, fix_sky=False):
    """The per-Gaussian colour block of render(): gaussian_renderer/__init__.py:119-156.
    positions/normals are [P,3] for ALL Gaussians; albedo/roughness/metalness are foreground-only.
    Returns (colors_precomp, diffuse_color, specular_color) each [P,3]."""
    P = positions.shape[0]
    fg = ~is_sky
    view_pos = cam_pos.reshape(1, 3).repeat(P, 1)
    dir_pp = safe_normalize(positions - view_pos)
    rgb, dif, spc = shade(base, sh_degree, fg_lut, positions[fg], normals[fg], albedo, view_pos[fg], roughness,
                          metalness, specular)
    colors = torch.zeros(P, 3, dtype=positions.dtype, device=positions.device)
    diffuse = torch.zeros(P, 3, dtype=positions.dtype, device=positions.device)
    spec = torch.zeros(P, 3, dtype=positions.dtype, device=positions.device)
    colors[fg] = rgb
    colors[is_sky] = sky_color(sky_sh, dir_pp[is_sky], fix_sky)
    diffuse[fg] = dif
    spec[fg] = spc
    return colors, diffuse, spec


def synthetic_fg_lut(res=256):
    """A smooth stand-in for irrmaps/bsdf_256_256.bin (the split-sum FG table is a data asset of the reference and
    is not copied).  Same shape/dtype [1,res,res,2]; analytic approximation (Karis 2014 mobile fit)."""
    u = (torch.arange(res, dtype=torch.float32) + 0.5) / res
    ndv = u.view(1, res).expand(res, res)
    rough = u.view(res, 1).expand(res, res)
    c0 = torch.tensor([-1.0, -0.0275, -0.572, 0.022])
    c1 = torch.tensor([1.0, 0.0425, 1.04, -0.04])
    r = rough.unsqueeze(-1) * c0 + c1
    a004 = torch.minimum(r[..., 0] * r[..., 0], torch.exp2(-9.28 * ndv)) * r[..., 0] + r[..., 1]
    A = -1.04 * a004 + r[..., 2]
    B = 1.04 * a004 + r[..., 3]
    return torch.stack([A, B], dim=-1).unsqueeze(0).contiguous()
