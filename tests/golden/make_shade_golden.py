#!/usr/bin/env python
"""Generate tests/golden/shade_*.npz by running the REFERENCE's own Python shading on CPU tensors.

Run in the build container (needs /root/reference; nothing under tests/ or bench.py reads it at run time):
    python tests/golden/make_shade_golden.py

What is executed is the unmodified reference code:
    scene/NVDIFFREC/light.py  EnvironmentLight.shade / get_diffuse_irradiance / get_specular_light_sh
    utils/sh_utils.py         eval_sh / gauss_kernel
    scene/NVDIFFREC/util.py   safe_normalize / reflect / dot / gamma_correction
Import shims (no reference file is edited or copied):
  * missing pip deps unused by the math are stubbed in sys.modules (imageio, skimage, nvdiffrast, ...);
  * `scene` / `scene.NVDIFFREC` are registered as bare namespace packages so scene/__init__.py (dataset
    loaders, CUDA-at-import) does not run; `utils.general_utils` (CUDA tensor as a default argument at import
    time, general_utils.py:295) is stubbed - light.py only imports `get_homogeneous` from it and never calls it;
  * `Tensor.cuda()` is an identity here (gauss_kernel hard-codes .cuda(), sh_utils.py:173-175);
  * EnvironmentLight.__init__ loads its LUT onto 'cuda' from a cwd-relative path (light.py:41), so the object is
    created with __new__ and the same attributes are set by hand; the LUT is a synthetic table (oracle
    shade_oracle.synthetic_fg_lut) because bsdf_256_256.bin is a reference data asset that is not copied;
  * nvdiffrast.torch.texture (un-vendored submodule) is provided as F.grid_sample(bilinear, border,
    align_corners=False), the documented equivalent of filter_mode='linear', boundary_mode='clamp'.
"""
import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("R3DGW_REFERENCE", "/root/reference")
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "relightable3dgaussians-w_b200")]


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _texture(tex, uv, filter_mode="linear", boundary_mode="clamp"):
    assert filter_mode == "linear" and boundary_mode == "clamp"
    # tex [1,H,W,C], uv [1,h,w,2] in [0,1] -> [1,h,w,C]
    grid = uv * 2.0 - 1.0
    out = F.grid_sample(tex.permute(0, 3, 1, 2), grid, mode="bilinear", padding_mode="border", align_corners=False)
    return out.permute(0, 2, 3, 1)


def load_reference():
    sys.path.insert(0, REF)
    for name in ("imageio", "imageio.v3", "skimage", "skimage.measure", "plyfile"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:   # noqa: BLE001
                _stub(name)
    nd = _stub("nvdiffrast")
    ndt = _stub("nvdiffrast.torch", texture=_texture)
    nd.torch = ndt
    _stub("utils.general_utils", get_homogeneous=lambda p: torch.cat([p, torch.ones_like(p[..., :1])], -1))
    scene = _stub("scene")
    scene.__path__ = [os.path.join(REF, "scene")]
    nv = _stub("scene.NVDIFFREC")
    nv.__path__ = [os.path.join(REF, "scene", "NVDIFFREC")]
    torch.Tensor.cuda = lambda self, *a, **k: self
    import importlib
    light = importlib.import_module("scene.NVDIFFREC.light")
    sh_utils = importlib.import_module("utils.sh_utils")
    return light, sh_utils


def make_inputs(N, seed):
    g = torch.Generator().manual_seed(seed)
    from r3dgw_b200.synthetic import make_env_sh
    pos = torch.randn(N, 3, generator=g) * 4.0
    nrm = F.normalize(torch.randn(N, 3, generator=g), dim=1)
    cam = torch.tensor([0.3, -0.2, -6.0])
    albedo = torch.sigmoid(torch.randn(N, 3, generator=g))
    kr = torch.sigmoid(torch.randn(N, 1, generator=g))
    km = torch.sigmoid(torch.randn(N, 1, generator=g))
    env = make_env_sh(g, 1)[0]
    sky_sh = 0.3 * torch.randn(1, 4, 3, generator=g)
    sky_dirs = F.normalize(torch.randn(N, 3, generator=g), dim=1)
    gr, gd, gs = (torch.randn(N, 3, generator=g) for _ in range(3))
    return dict(pos=pos, nrm=nrm, cam=cam, albedo=albedo, kr=kr, km=km, env=env, sky_sh=sky_sh, sky_dirs=sky_dirs,
                g_rgb=gr, g_dif=gd, g_spc=gs)


def run_reference(light, sh_utils, lut, inp, specular, use_km):
    env = light.EnvironmentLight.__new__(light.EnvironmentLight)
    torch.nn.Module.__init__(env)
    env.sh_degree, env.sh_dim, env.NUM_CHANNELS = 4, 25, 3
    env.C1, env.C2, env.C3, env.C4, env.C5 = 0.429043, 0.511664, 0.743125, 0.886227, 0.247708
    env._FG_LUT = lut
    leaves = {k: inp[k].clone().requires_grad_(True) for k in ("pos", "nrm", "albedo", "kr", "km", "env", "sky_sh")}
    env.base = leaves["env"]
    N = inp["pos"].shape[0]
    view_pos = inp["cam"].reshape(1, 3).repeat(N, 1)
    # call shape exactly as get_shaded_colors does (gaussian_renderer/__init__.py:60-66)
    km = leaves["km"][None, None, ...] if use_km else None
    rgb, extras = env.shade(gb_pos=leaves["pos"][None, None, ...], gb_normal=leaves["nrm"][None, None, ...],
                            albedo=leaves["albedo"][None, None, ...], view_pos=view_pos[None, None, ...],
                            kr=leaves["kr"][None, None, ...], km=km, specular=specular)
    rgb, dif, spc = rgb.squeeze(), extras["diffuse"].squeeze(), extras["specular"].squeeze()
    # sky colour as in render() (gaussian_renderer/__init__.py:146-147)
    sky = torch.clamp_min(sh_utils.eval_sh(1, leaves["sky_sh"].transpose(1, 2), inp["sky_dirs"]) + 0.5, 0.0)
    loss = (rgb * inp["g_rgb"]).sum() + (dif * inp["g_dif"]).sum() + (sky * inp["g_rgb"]).sum()
    if specular:
        loss = loss + (spc * inp["g_spc"]).sum()
    loss.backward()
    out = dict(rgb=rgb, dif=dif, spc=spc, sky=sky)
    for k, v in leaves.items():
        out["d_" + k] = v.grad if v.grad is not None else torch.zeros_like(v)
    return {k: v.detach().numpy() for k, v in out.items()}


def main():
    light, sh_utils = load_reference()
    import shade_oracle
    lut = shade_oracle.synthetic_fg_lut()
    for tag, N, seed, specular, use_km in (("spec_km", 96, 1, True, True), ("spec_nokm", 64, 2, True, False),
                                           ("diffuse", 64, 3, False, True)):
        inp = make_inputs(N, 1000 + seed)
        out = run_reference(light, sh_utils, lut, inp, specular, use_km)
        blob = {"in_" + k: v.numpy() for k, v in inp.items()}
        blob.update({"ref_" + k: v for k, v in out.items()})
        blob["specular"] = np.array(int(specular))
        blob["use_km"] = np.array(int(use_km))
        path = os.path.join(HERE, f"shade_{tag}.npz")
        np.savez_compressed(path, **blob)
        print("wrote", path, {k: v.shape for k, v in out.items() if k in ("rgb", "sky")})


if __name__ == "__main__":
    main()
