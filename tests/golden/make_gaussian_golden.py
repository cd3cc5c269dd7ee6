#!/usr/bin/env python
"""Generate tests/golden/gaussian_ops.npz by running the REFERENCE's own Python on CPU tensors:

    GaussianModel.get_normal / get_depth / get_scaling / get_rotation   scene/gaussian_model.py:73-130
    build_rotation / get_minimum_axis / flip_align_view                 utils/general_utils.py:98-119, 158-170
    safe_normalize (dir_pp_normalized, gaussian_renderer/__init__.py:127-129)   scene/NVDIFFREC/util.py:30-31
    min_scale_loss / depth_loss_gaussians                               utils/loss_utils.py:210-220, 140-148
    EnvironmentLight.get_diffuse_irradiance / get_specular_light_sh / sample_illumination / render_sh
                                                                        scene/NVDIFFREC/light.py:65-128, 196-205
and their autograd gradients.  Run in the build container (needs /root/reference):
    python tests/golden/make_gaussian_golden.py

Import shims (no reference file is edited or copied): as make_shade_golden.py, plus - because
utils/general_utils.py builds CUDA tensors at import time (:295) and inside build_rotation (:103) - the torch
factory functions drop a device='cuda' argument while the reference code runs; simple_knn / plyfile / the dataset
readers are stubbed (not used by these methods).  The GaussianModel instance is the reference class with its
parameter tensors assigned directly (no checkpoint / COLMAP input).
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_shade_golden as msg   # noqa: E402  (the shared stubs)

REF = msg.REF


def _cpu_factories():
    for name in ("zeros", "ones", "empty", "tensor", "arange", "full", "eye", "rand", "randn", "linspace",
                 "zeros_like", "ones_like", "empty_like"):
        fn = getattr(torch, name)

        def wrap(*a, __fn=fn, **k):
            if str(k.get("device", "")).startswith("cuda"):
                k.pop("device")
            return __fn(*a, **k)
        setattr(torch, name, wrap)
    torch.cuda.set_device = lambda *a, **k: None
    torch.Tensor.cuda = lambda self, *a, **k: self


def load_reference():
    sys.path.insert(0, REF)
    _cpu_factories()
    for name in ("imageio", "imageio.v3", "skimage", "skimage.measure", "plyfile", "simple_knn", "simple_knn._C",
                 "matplotlib", "matplotlib.pyplot", "cv2", "PIL", "PIL.Image", "hydra", "omegaconf", "tqdm", "lpips"):
        try:
            __import__(name)
        except Exception:   # noqa: BLE001
            msg._stub(name)
    sys.modules["plyfile"].__dict__.setdefault("PlyData", object)
    sys.modules["plyfile"].__dict__.setdefault("PlyElement", object)
    sys.modules["simple_knn._C"].__dict__.setdefault("distCUDA2", None)
    nd = msg._stub("nvdiffrast")
    nd.torch = msg._stub("nvdiffrast.torch", texture=msg._texture)
    scene = msg._stub("scene")
    scene.__path__ = [os.path.join(REF, "scene")]
    nv = msg._stub("scene.NVDIFFREC")
    nv.__path__ = [os.path.join(REF, "scene", "NVDIFFREC")]
    sys.modules.pop("utils.general_utils", None)
    gm = importlib.import_module("scene.gaussian_model")
    lu = importlib.import_module("utils.loss_utils")
    light = importlib.import_module("scene.NVDIFFREC.light")
    util = importlib.import_module("scene.NVDIFFREC.util")
    return gm, lu, light, util


def make_model(gm, N, seed):
    """Reference GaussianModel with synthetic parameters: 10 % sky Gaussians (no sky_angles: _xyz holds every
    position, so the model is built with an all-False mask for get_xyz and the sky mask is swapped in after)."""
    g = torch.Generator().manual_seed(seed)
    m = gm.GaussianModel()
    m._xyz = (torch.randn(N, 3, generator=g) * 3.0 + torch.tensor([0.0, 0.0, 6.0])).requires_grad_(True)
    m._scaling = (torch.randn(N, 3, generator=g) * 0.7 - 3.0).requires_grad_(True)
    m._rotation = (torch.randn(N, 4, generator=g)).requires_grad_(True)
    m._is_sky = torch.zeros(N, 1, dtype=torch.bool)
    return m, (torch.rand(N, generator=g) < 0.1)


def main():
    gm, lu, light, util = load_reference()
    N = 4096
    m, sky = make_model(gm, N, 11)
    g = torch.Generator().manual_seed(5)
    campos = torch.tensor([0.4, -0.3, -1.0])
    # a world->view transform stored transposed, like scene/cameras.py:77
    A = torch.linalg.qr(torch.randn(3, 3, generator=g))[0]
    wvt = torch.eye(4)
    wvt[:3, :3] = A
    wvt[3, :3] = torch.tensor([0.1, -0.2, 0.5])
    cam = types.SimpleNamespace(world_view_transform=wvt, camera_center=campos)

    out = {"xyz": m._xyz.detach().numpy(), "scaling_raw": m._scaling.detach().numpy(),
           "rotation_raw": m._rotation.detach().numpy(), "is_sky": sky.numpy(), "campos": campos.numpy(),
           "wvt": wvt.numpy()}
    scales, rots = m.get_scaling, m.get_rotation
    out["scales"], out["rotations"] = scales.detach().numpy(), rots.detach().numpy()
    view_pos = campos.repeat(N, 1)
    dir_pp_n = util.safe_normalize(m.get_xyz - view_pos)
    normal = m.get_normal(dir_pp_normalized=dir_pp_n)
    normal_noflip = m.get_normal()
    depth = m.get_depth(cam)
    out.update(dir_pp_normalized=dir_pp_n.detach().numpy(), normal=normal.detach().numpy(),
               normal_noflip=normal_noflip.detach().numpy(), depth=depth.detach().numpy())
    # gradients of a fixed linear functional of (normal, depth) w.r.t. the RAW parameters and the activations
    wn = torch.randn(N, 3, generator=g)
    wd = torch.randn(N, 1, generator=g)
    act_s, act_r = scales.detach().clone().requires_grad_(True), rots.detach().clone().requires_grad_(True)
    m2 = gm.GaussianModel()
    m2._xyz, m2._is_sky = m._xyz, m._is_sky
    m2.scaling_activation = lambda x: x        # parameters ARE the activations: gradients w.r.t. get_scaling / get_rotation
    m2.rotation_activation = lambda x: x
    m2._scaling, m2._rotation = act_s, act_r
    n2 = m2.get_normal(dir_pp_normalized=dir_pp_n.detach())
    ((n2 * wn).sum() + (m2.get_depth(cam) * wd).sum()).backward()
    out.update(w_normal=wn.numpy(), w_depth=wd.numpy(), g_rotations=act_r.grad.numpy(),
               g_scales=np.zeros((N, 3), np.float32) if act_s.grad is None else act_s.grad.numpy(),
               g_xyz=m._xyz.grad.numpy())
    m._xyz.grad = None

    # regularisers: radii > 0 for ~70 %, sky mask as the model's is_sky
    radii = (torch.rand(N, generator=g) < 0.7).int() * torch.randint(1, 40, (N,), generator=g).int()
    m3 = gm.GaussianModel()
    m3._xyz = m._xyz.detach().clone().requires_grad_(True)
    m3._scaling = m._scaling.detach().clone().requires_grad_(True)
    m3._rotation = m._rotation
    m3._is_sky = torch.zeros(N, 1, dtype=torch.bool)        # get_xyz returns _xyz unchanged
    xyz_all = m3.get_xyz
    m3.__class__ = type("GM", (gm.GaussianModel,), {"get_is_sky": property(lambda self: sky.unsqueeze(1)),
                                                    "get_xyz": property(lambda self: self._xyz)})
    ls = lu.min_scale_loss(radii, m3)
    ld = lu.depth_loss_gaussians(m3, cam, radii > 0)
    (3.0 * ls + 0.7 * ld).backward()
    out.update(radii=radii.numpy(), min_scale_loss=ls.detach().numpy(), depth_loss=ld.detach().numpy(),
               g_reg_scaling_raw=m3._scaling.grad.numpy(), g_reg_xyz=m3._xyz.grad.numpy())

    # EnvironmentLight's four small methods
    from r3dgw_b200.synthetic import make_env_sh
    base = make_env_sh(torch.Generator().manual_seed(3))[0].clone()
    env = light.EnvironmentLight.__new__(light.EnvironmentLight)
    torch.nn.Module.__init__(env)
    env.sh_degree, env.sh_dim, env.base = 4, 25, base
    env.NUM_CHANNELS = 3
    env.C1, env.C2, env.C3, env.C4, env.C5 = 0.429043, 0.511664, 0.743125, 0.886227, 0.247708
    nrm = torch.nn.functional.normalize(torch.randn(257, 3, generator=g), dim=1)
    kr = torch.rand(257, 1, generator=g)
    pos = torch.randn(257, 3, generator=g) * 4.0
    out.update(env_base=base.numpy(), env_normals=nrm.numpy(), env_kr=kr.numpy(), env_pos=pos.numpy(),
               diffuse_irradiance=env.get_diffuse_irradiance(nrm).numpy(),
               specular_light_sh=env.get_specular_light_sh(kr).numpy(),
               sample_illumination=env.sample_illumination(pos, campos.repeat(257, 1)).numpy(),
               render_sh_64=env.render_sh(width=64).numpy())
    path = os.path.join(HERE, "gaussian_ops.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
