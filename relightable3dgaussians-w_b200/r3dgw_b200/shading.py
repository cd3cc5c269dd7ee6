"""Fused per-Gaussian relit shading: autograd op over r3dgw_shade_forward/backward + a drop-in EnvironmentLight.

Reference being replaced (a torch graph of dozens of elementwise kernels and an [N,25,3] temporary):
  scene/NVDIFFREC/light.py:14-205   EnvironmentLight (same class name, constructor and method signatures here)
  gaussian_renderer/__init__.py:60-66, 119-156   get_shaded_colors + sky colour + masked scatters
There is no CPU path: tensors must be CUDA tensors, otherwise the call raises.
"""
import ctypes

import numpy as np
import torch

from . import native as _nat


def _c(t):
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _ptr(t):
    return None if t is None else t.data_ptr()


class _ShadeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, positions, normals, albedo, roughness, metalness, env_sh, sky_sh, cam_pos, fg_lut, is_sky,
                fg_index, sh_degree, specular, fix_sky):
        if positions.device.type != "cuda":
            raise RuntimeError("shade_gaussians (B200-native): tensors must live on a CUDA device; there is no CPU path")
        dev = positions.device
        P = int(positions.shape[0])
        pos, nrm, alb = _c(positions), _c(normals), _c(albedo)
        kr = _c(roughness) if roughness is not None else None
        km = _c(metalness) if metalness is not None else None
        env = _c(env_sh.reshape(-1, 3))
        sky = _c(sky_sh.reshape(-1, 3)) if sky_sh is not None else None
        cam = _c(cam_pos.reshape(-1)[:3])
        lut = _c(fg_lut) if fg_lut is not None else None
        args = _nat.ShadeArgs(P, int(bool(specular)), int(bool(fix_sky)), int(sh_degree), _ptr(pos), _ptr(nrm), _ptr(alb),
                              _ptr(kr), _ptr(km), _ptr(is_sky), _ptr(fg_index), _ptr(cam), _ptr(env), _ptr(sky), _ptr(lut))
        out = torch.empty(3, P, 3, dtype=torch.float32, device=dev)
        with _nat.on_device(dev):
            rc = _nat.lib().r3dgw_shade_forward(ctypes.byref(args), out[0].data_ptr(), out[1].data_ptr(),
                                                out[2].data_ptr(), _nat.raw_stream())
        _nat.check(rc, "r3dgw_shade_forward")
        ctx.keep = (pos, nrm, alb, kr, km, env, sky, cam, lut, is_sky, fg_index)
        ctx.cfg = (P, int(sh_degree), bool(specular), bool(fix_sky))
        ctx.shapes = (positions.shape, normals.shape, albedo.shape, None if roughness is None else roughness.shape,
                      None if metalness is None else metalness.shape, env_sh.shape,
                      None if sky_sh is None else sky_sh.shape)
        return out[0], out[1], out[2]

    @staticmethod
    def backward(ctx, g_rgb, g_dif, g_spc):
        pos, nrm, alb, kr, km, env, sky, cam, lut, is_sky, fg_index = ctx.keep
        P, deg, specular, fix_sky = ctx.cfg
        dev = pos.device
        args = _nat.ShadeArgs(P, int(specular), int(fix_sky), deg, _ptr(pos), _ptr(nrm), _ptr(alb), _ptr(kr), _ptr(km),
                              _ptr(is_sky), _ptr(fg_index), _ptr(cam), _ptr(env), _ptr(sky), _ptr(lut))
        g_rgb, g_dif, g_spc = _c(g_rgb), _c(g_dif), _c(g_spc)
        # sky rows of the material gradients and culled rows stay zero -> zero-initialised outputs
        d_pos = torch.zeros_like(pos)
        d_nrm = torch.zeros_like(nrm)
        d_alb = torch.zeros_like(alb)
        d_kr = torch.zeros_like(kr) if kr is not None else None
        d_km = torch.zeros_like(km) if km is not None else None
        d_env = torch.zeros_like(env)
        d_sky = torch.zeros_like(sky) if sky is not None else None
        with _nat.on_device(dev):
            rc = _nat.lib().r3dgw_shade_backward(ctypes.byref(args), _ptr(g_rgb), _ptr(g_dif), _ptr(g_spc), _ptr(d_pos),
                                                 _ptr(d_nrm), _ptr(d_alb), _ptr(d_kr), _ptr(d_km), _ptr(d_env),
                                                 _ptr(d_sky), _nat.raw_stream())
        _nat.check(rc, "r3dgw_shade_backward")
        s = ctx.shapes
        return (d_pos.reshape(s[0]), d_nrm.reshape(s[1]), d_alb.reshape(s[2]),
                None if d_kr is None else d_kr.reshape(s[3]), None if d_km is None else d_km.reshape(s[4]),
                d_env.reshape(s[5]), None if d_sky is None else d_sky.reshape(s[6]),
                None, None, None, None, None, None, None)


def shade_gaussians(positions, normals, albedo, roughness, metalness, env_sh, cam_pos, fg_lut, sh_degree=4,
                    specular=True, sky_sh=None, is_sky=None, fix_sky=False):
    """One fused pass for ALL Gaussians (foreground + sky): returns (colors_precomp, diffuse_color, specular_color),
    each [P,3] - the three tensors render() assembles with masked scatters (gaussian_renderer/__init__.py:132-156).

    positions/normals: [P,3] for all Gaussians.  albedo [Pfg,3], roughness [Pfg,1], metalness [Pfg,1] or None are
    foreground-only, like the reference's parameter store.  is_sky: bool [P] or None (no sky Gaussians)."""
    sky_u8, fg_index = _sky_inputs(is_sky)
    return _ShadeFn.apply(positions, normals, albedo, roughness, metalness, env_sh, sky_sh, cam_pos, fg_lut, sky_u8,
                          fg_index, sh_degree, specular, fix_sky)


def _sky_inputs(is_sky):
    if is_sky is None:
        return None, None
    sky_b = is_sky.reshape(-1).bool()
    return sky_b.to(torch.uint8).contiguous(), (torch.cumsum((~sky_b).to(torch.int32), 0, dtype=torch.int32) - 1).contiguous()


class _PreshadeFn(torch.autograd.Function):
    """r3dgw_preshade_forward/backward: get_normal + dir_pp + get_depth + the normal / depth "colours" + shading in ONE
    kernel each way.  packed=True returns the [P,16] feature matrix of the single-pass blend."""

    @staticmethod
    def forward(ctx, positions, scales, rotations, albedo, roughness, metalness, env_sh, sky_sh, cam_pos, viewmatrix,
                fg_lut, sky_u8, fg_index, sh_degree, specular, fix_sky, packed):
        if positions.device.type != "cuda":
            raise RuntimeError("preshade_gaussians (B200-native): tensors must live on a CUDA device; there is no CPU path")
        dev = positions.device
        P = int(positions.shape[0])
        pos, scl, rot, alb = _c(positions), _c(scales), _c(rotations), _c(albedo)
        kr = _c(roughness) if roughness is not None else None
        km = _c(metalness) if metalness is not None else None
        env = _c(env_sh.reshape(-1, 3))
        sky = _c(sky_sh.reshape(-1, 3)) if sky_sh is not None else None
        cam = _c(cam_pos.reshape(-1)[:3])
        vm = _c(viewmatrix)
        lut = _c(fg_lut) if fg_lut is not None else None
        sh = _nat.ShadeArgs(P, int(bool(specular)), int(bool(fix_sky)), int(sh_degree), _ptr(pos), None, _ptr(alb),
                            _ptr(kr), _ptr(km), _ptr(sky_u8), _ptr(fg_index), _ptr(cam), _ptr(env), _ptr(sky), _ptr(lut))
        args = _nat.PreshadeArgs(sh, _ptr(scl), _ptr(rot), _ptr(vm))
        if packed:
            out = torch.empty(P, 16, dtype=torch.float32, device=dev)
            b, stride = out.data_ptr(), 16
            ptrs = [b, b + 12, b + 24, b + 36, b + 48, b + 60]
        else:
            out = torch.empty(5, P, 3, dtype=torch.float32, device=dev)
            stride = 3
            ptrs = [out[k].data_ptr() for k in range(5)] + [None]
        with _nat.on_device(dev):
            rc = _nat.lib().r3dgw_preshade_forward(ctypes.byref(args), stride, *ptrs, None, _nat.raw_stream())
        _nat.check(rc, "r3dgw_preshade_forward")
        ctx.keep = (pos, scl, rot, alb, kr, km, env, sky, cam, vm, lut, sky_u8, fg_index)
        ctx.cfg = (P, int(sh_degree), bool(specular), bool(fix_sky), bool(packed))
        ctx.shapes = (positions.shape, rotations.shape, albedo.shape, None if roughness is None else roughness.shape,
                      None if metalness is None else metalness.shape, env_sh.shape,
                      None if sky_sh is None else sky_sh.shape)
        if packed:
            return out
        return out[0], out[1], out[2], out[3], out[4]

    @staticmethod
    def backward(ctx, *grads):
        pos, scl, rot, alb, kr, km, env, sky, cam, vm, lut, sky_u8, fg_index = ctx.keep
        P, deg, specular, fix_sky, packed = ctx.cfg
        dev = pos.device
        sh = _nat.ShadeArgs(P, int(specular), int(fix_sky), deg, _ptr(pos), None, _ptr(alb), _ptr(kr), _ptr(km),
                            _ptr(sky_u8), _ptr(fg_index), _ptr(cam), _ptr(env), _ptr(sky), _ptr(lut))
        args = _nat.PreshadeArgs(sh, _ptr(scl), _ptr(rot), _ptr(vm))
        if packed:
            g = _c(grads[0])
            b, stride = g.data_ptr(), 16
            gptrs = [b, b + 12, b + 24, b + 36, b + 48]
            keep = g
        else:
            # the blend backward hands out its 16-byte padded gradient rows as [P,3] views of [P,4] buffers: read them
            # in place through the row stride (five .contiguous() copies per step otherwise)
            live = [x for x in grads if x is not None]
            stride = int(live[0].stride(0)) if live and live[0].dim() == 2 else 3
            if not (stride >= 3 and all(x.dim() == 2 and x.dtype == torch.float32 and x.stride(1) == 1 and
                                        x.stride(0) == stride and x.shape == (P, 3) for x in live)):
                stride = 3
            keep = [torch.zeros(P, stride, dtype=torch.float32, device=dev) if (x is None and k == 0) else
                    (None if x is None else (x if stride != 3 else _c(x))) for k, x in enumerate(grads)]
            gptrs = [_ptr(x) for x in keep]
        # one zero-filled block for every per-Gaussian gradient + the SH gradients (accumulated with atomics)
        n_fg = int(alb.shape[0])
        sizes = [3 * P, 4 * P, 3 * n_fg, n_fg if kr is not None else 0, n_fg if km is not None else 0, env.numel(),
                 sky.numel() if sky is not None else 0]
        flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
        parts, o = [], 0
        for n in sizes:
            parts.append(flat[o:o + n] if n else None)
            o += n
        d_pos, d_rot, d_alb, d_kr, d_km, d_env, d_sky = parts
        with _nat.on_device(dev):
            rc = _nat.lib().r3dgw_preshade_backward(ctypes.byref(args), stride, *gptrs, _ptr(d_pos), _ptr(d_rot),
                                                    _ptr(d_alb), _ptr(d_kr), _ptr(d_km), _ptr(d_env), _ptr(d_sky),
                                                    _nat.raw_stream())
        _nat.check(rc, "r3dgw_preshade_backward")
        del keep
        s = ctx.shapes
        return (d_pos.view(s[0]), None, d_rot.view(s[1]), d_alb.view(s[2]),
                None if d_kr is None else d_kr.view(s[3]), None if d_km is None else d_km.view(s[4]),
                d_env.view(s[5]), None if d_sky is None else d_sky.view(s[6]),
                None, None, None, None, None, None, None, None, None)


def preshade_gaussians(positions, scales, rotations, albedo, roughness, metalness, env_sh, cam_pos, viewmatrix, fg_lut,
                       sh_degree=4, specular=True, sky_sh=None, is_sky=None, fix_sky=False, packed=False,
                       sky_index=None):
    """Everything render() computes per Gaussian in front of the rasterizer, in one kernel (and one for the backward):
    flipped min-axis normal (GaussianModel.get_normal), shaded / diffuse / specular colours (EnvironmentLight.shade + the
    sky colour), the depth "colour" (get_depth repeated over 3 channels) and the normal "colour" 0.5 n + 0.5.
    Returns (colors, diffuse, specular, depth3, normal01), each [P,3] - or, packed=True, the [P,16] feature matrix
    (those five + a column of ones for alpha) the single-pass rasterize_gaussians_features call takes.
    `viewmatrix` is world_view_transform as the camera stores it (transposed).  `sky_index` = (is_sky u8, fg_index i32)
    precomputed by sky_index_of(is_sky) saves two small kernels per call."""
    sky_u8, fg_index = sky_index if sky_index is not None else _sky_inputs(is_sky)
    return _PreshadeFn.apply(positions, scales, rotations, albedo, roughness, metalness, env_sh, sky_sh, cam_pos,
                             viewmatrix, fg_lut, sky_u8, fg_index, sh_degree, specular, fix_sky, packed)


def sky_index_of(is_sky):
    """(is_sky as u8, foreground rank of every Gaussian) - constant between densifications; pass as sky_index=."""
    return _sky_inputs(is_sky)


class _NormalFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scales, rotations, positions, cam_pos):
        if rotations.device.type != "cuda":
            raise RuntimeError("gaussian_normals (B200-native): tensors must live on a CUDA device; there is no CPU path")
        scl, rot = _c(scales), _c(rotations)
        pos = _c(positions) if cam_pos is not None else None
        cam = _c(cam_pos.reshape(-1)[:3]) if cam_pos is not None else None
        P = int(rot.shape[0])
        out = torch.empty(P, 3, dtype=torch.float32, device=rot.device)
        with _nat.on_device(rot.device):
            rc = _nat.lib().r3dgw_gaussian_normal_forward(P, _ptr(scl), _ptr(rot), _ptr(pos), _ptr(cam), out.data_ptr(),
                                                          _nat.raw_stream())
        _nat.check(rc, "r3dgw_gaussian_normal_forward")
        ctx.keep = (scl, rot, pos, cam)
        ctx.rot_shape = rotations.shape
        return out

    @staticmethod
    def backward(ctx, g):
        scl, rot, pos, cam = ctx.keep
        P = int(rot.shape[0])
        d_rot = torch.empty(P, 4, dtype=torch.float32, device=rot.device)
        g = _c(g)
        with _nat.on_device(rot.device):
            rc = _nat.lib().r3dgw_gaussian_normal_backward(P, _ptr(scl), _ptr(rot), _ptr(pos), _ptr(cam), g.data_ptr(),
                                                           d_rot.data_ptr(), _nat.raw_stream())
        _nat.check(rc, "r3dgw_gaussian_normal_backward")
        return None, d_rot.view(ctx.rot_shape), None, None


def gaussian_normals(scales, rotations, positions=None, cam_pos=None):
    """GaussianModel.get_normal (scene/gaussian_model.py:115-122): column of R(q / |q|) at argmin(scale); with positions
    and cam_pos, flipped towards the camera (dir_pp_normalized = safe_normalize(positions - cam_pos))."""
    if cam_pos is not None and positions is None:
        raise ValueError("cam_pos needs positions")
    return _NormalFn.apply(scales, rotations, positions, cam_pos)


# band index of each SH coefficient up to degree 5 and the (l, m)-ordered polynomial basis on torch tensors: used by the
# non-hot methods of EnvironmentLight below (the hot path is the fused kernel)
def _sh_basis_torch(deg, d):
    x, y, z = d[..., 0], d[..., 1], d[..., 2]
    one = torch.ones_like(x)
    cols = [0.28209479177387814 * one]
    if deg > 0:
        cols += [-0.4886025119029199 * y, 0.4886025119029199 * z, -0.4886025119029199 * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        cols += [1.0925484305920792 * xy, -1.0925484305920792 * yz, 0.31539156525252005 * (2.0 * zz - xx - yy),
                 -1.0925484305920792 * xz, 0.5462742152960396 * (xx - yy)]
    if deg > 2:
        cols += [-0.5900435899266435 * y * (3 * xx - yy), 2.890611442640554 * xy * z,
                 -0.4570457994644658 * y * (4 * zz - xx - yy), 0.3731763325901154 * z * (2 * zz - 3 * xx - 3 * yy),
                 -0.4570457994644658 * x * (4 * zz - xx - yy), 1.445305721320277 * z * (xx - yy),
                 -0.5900435899266435 * x * (xx - 3 * yy)]
    if deg > 3:
        cols += [2.5033429417967046 * xy * (xx - yy), -1.7701307697799304 * yz * (3 * xx - yy),
                 0.9461746957575601 * xy * (7 * zz - 1), -0.6690465435572892 * yz * (7 * zz - 3),
                 0.10578554691520431 * (zz * (35 * zz - 30) + 3), -0.6690465435572892 * xz * (7 * zz - 3),
                 0.47308734787878004 * (xx - yy) * (7 * zz - 1), -1.7701307697799304 * xz * (xx - 3 * yy),
                 0.6258357354491761 * (xx * (xx - 3 * yy) - yy * (3 * xx - yy))]
    if deg > 4:
        raise NotImplementedError("SH degree 5 evaluation")
    return torch.stack(cols, dim=-1)


class EnvironmentLight(torch.nn.Module):
    """Drop-in for scene/NVDIFFREC/light.py:14-205 (same constructor, properties and `shade` signature/shapes)."""

    LUT_PATH = "scene/NVDIFFREC/irrmaps/bsdf_256_256.bin"   # cwd-relative, as in light.py:41

    def __init__(self, base: torch.Tensor, sh_degree: int = 4, fg_lut: torch.Tensor = None):
        super().__init__()
        if sh_degree > 5:
            raise NotImplementedError
        self.sh_degree = sh_degree
        self.sh_dim = (sh_degree + 1) ** 2
        self.base = base.squeeze()
        self.NUM_CHANNELS = 3
        self.C1, self.C2, self.C3, self.C4, self.C5 = 0.429043, 0.511664, 0.743125, 0.886227, 0.247708
        if fg_lut is None:
            fg_lut = torch.as_tensor(np.fromfile(self.LUT_PATH, dtype=np.float32).reshape(1, 256, 256, 2),
                                     dtype=torch.float32, device="cuda")
        self._FG_LUT = fg_lut

    def clone(self):
        return EnvironmentLight(self.base.clone().detach(), self.sh_degree, self._FG_LUT)

    @property
    def get_shdim(self):
        return self.sh_dim

    @property
    def get_shdegree(self):
        return self.sh_degree

    @property
    def get_base(self):
        return self.base

    def set_base(self, base: torch.Tensor):
        assert base.squeeze().shape[0] == self.sh_dim, f"The number of SH coefficients must be {self.sh_dim}"
        self.base = base.squeeze()

    def get_diffuse_irradiance(self, normal):
        """light.py:65-94: order-2 irradiance polynomial at `normal` [...,3] -> [...,3] (same expression, torch ops on
        the caller's tensors; the training path evaluates it inside the fused kernel)."""
        x, y, z = normal[..., 0, None], normal[..., 1, None], normal[..., 2, None]
        b = self.base
        return (self.C1 * b[8, :] * (x ** 2 - y ** 2) + self.C3 * b[6, :] * (z ** 2) + self.C4 * b[0, :]
                - self.C5 * b[6, :] + 2 * self.C1 * b[4, :] * x * y + 2 * self.C1 * b[7, :] * x * z
                + 2 * self.C1 * b[5, :] * y * z + 2 * self.C2 * b[3, :] * x + 2 * self.C2 * b[1, :] * y
                + 2 * self.C2 * b[2, :] * z)

    def get_specular_light_sh(self, kr):
        """light.py:97-119 (+ gauss_kernel, utils/sh_utils.py:162-181): kr [N,1] -> [N, sh_dim, 3], the environment SH
        attenuated per band by exp(-l (l+1) 0.3 kr)."""
        l_of = torch.cat([torch.full((2 * l + 1,), float(l)) for l in range(self.sh_degree + 1)]).to(kr.device)
        lobe = torch.exp(-l_of.view(1, -1) * (l_of.view(1, -1) + 1) * (0.3 * kr.reshape(-1, 1)))       # [N, sh_dim]
        return lobe.unsqueeze(-1) * self.base.unsqueeze(0)

    def sample_illumination(self, gb_pos, view_pos):
        """light.py:122-128: gamma(relu(eval_sh(sh_degree, base, normalize(gb_pos - view_pos))))."""
        v = gb_pos - view_pos
        d = (v / torch.sqrt(torch.clamp((v * v).sum(-1, keepdim=True), min=1e-20))).squeeze()
        hdr = torch.relu(_sh_basis_torch(self.sh_degree, d) @ self.base[: self.sh_dim])
        return (hdr.clamp(min=0.0, max=1.0) + 1e-4).pow(1.0 / 2.2)

    def render_sh(self, width: int = 600):
        """light.py:196-205: the band-weighted SH reconstruction (sh_render, utils/sh_additional_utils.py:288-299) as an
        equirectangular [width/2, width, 3] image, gamma corrected; a CPU float64 tensor like the reference's."""
        from . import envmap_sh
        self.base = self.base.squeeze()
        img = envmap_sh.sh_render_device(self.base, width)
        return (img.clamp(min=0.0, max=1.0) + 1e-4).pow(1.0 / 2.2).cpu()

    def shade(self, gb_pos, gb_normal, albedo, view_pos, kr=None, km=None, specular=True):
        """Same call shape as the reference: inputs [1,1,N,*]; view_pos is the camera centre repeated N times.
        Returns (rgb [1,1,N,3], {"diffuse": [1,1,N,3], "specular": [1,1,N,3]})."""
        if self.sh_degree > 4 or self.sh_degree < 2:
            raise NotImplementedError("fused shading supports env SH degree 2..4")
        lead = gb_pos.shape[:-1]
        pos, nrm, alb = gb_pos.reshape(-1, 3), gb_normal.reshape(-1, 3), albedo.reshape(-1, 3)
        cam = view_pos.reshape(-1, 3)[0]
        krf = kr.reshape(-1, 1) if kr is not None else None
        kmf = km.reshape(-1, 1) if km is not None else None
        rgb, dif, spc = shade_gaussians(pos, nrm, alb, krf, kmf, self.base, cam, self._FG_LUT, self.sh_degree,
                                        specular=specular)
        shp = tuple(lead) + (3,)
        return rgb.reshape(shp), {"diffuse": dif.reshape(shp), "specular": spc.reshape(shp)}
