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
    fg_index = None
    sky_u8 = None
    if is_sky is not None:
        sky_b = is_sky.reshape(-1).bool()
        sky_u8 = sky_b.to(torch.uint8).contiguous()
        fg_index = (torch.cumsum((~sky_b).to(torch.int32), 0, dtype=torch.int32) - 1).contiguous()
    return _ShadeFn.apply(positions, normals, albedo, roughness, metalness, env_sh, sky_sh, cam_pos, fg_lut, sky_u8,
                          fg_index, sh_degree, specular, fix_sky)


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
