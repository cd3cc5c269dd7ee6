#!/usr/bin/env python
"""Generate tests/golden/image_ops.npz by running the REFERENCE's own Python image-space ops on CPU tensors.

Run in the build container (needs /root/reference; nothing under tests/ or bench.py reads it at run time):
    python tests/golden/make_image_golden.py

What is executed is the unmodified reference code:
    utils/loss_utils.py      l1_loss, ssim (create_window / _ssim)
    utils/graphics_utils.py  depths_to_points, depth_to_normal
Import shims (no reference file is edited or copied):
  * `utils.general_utils` evaluates a CUDA tensor as a default argument at import time (general_utils.py:295) and
    loss_utils only imports `rand_hemisphere_dir` from it (unused here) -> stubbed module;
  * missing pip deps unused by the math (imageio, skimage) are stubbed in sys.modules;
  * depths_to_points hard-codes `.cuda()` / `device='cuda'` (graphics_utils.py:150-152): `Tensor.cuda()` is an
    identity here and `torch.arange` drops the device keyword while the reference function runs.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("R3DGW_REFERENCE", "/root/reference")
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "relightable3dgaussians-w_b200")]


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def load_reference():
    sys.path.insert(0, REF)
    for name in ("imageio", "imageio.v3", "skimage", "skimage.measure", "plyfile"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:   # noqa: BLE001
                _stub(name)
    _stub("utils.general_utils", rand_hemisphere_dir=None)
    torch.Tensor.cuda = lambda self, *a, **k: self
    real_arange = torch.arange

    def arange_cpu(*a, **k):
        k.pop("device", None)
        return real_arange(*a, **k)
    torch.arange = arange_cpu
    import importlib
    return importlib.import_module("utils.loss_utils"), importlib.import_module("utils.graphics_utils")


class View:
    def __init__(self, cam):
        self.world_view_transform = cam.world_view_transform
        self.image_width, self.image_height = cam.image_width, cam.image_height
        self.FoVx, self.FoVy = cam.FoVx, cam.FoVy


def main():
    loss_utils, gfx = load_reference()
    from r3dgw_b200.synthetic import arc_cameras
    H, W = 70, 93                      # not multiples of the 32-pixel kernel tile; larger than the 11x11 window
    g = torch.Generator().manual_seed(20261104)
    gt = torch.rand(3, H, W, generator=g)
    img = (gt + 0.15 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
    yy, xx = torch.meshgrid(torch.arange(H).float(), torch.arange(W).float(), indexing="ij")
    occl = ((xx - 40) ** 2 + (yy - 30) ** 2 > 15 ** 2).float()          # occluders_mask: a disc is masked out
    sky = (yy > 0.2 * H).float()                                         # sky_mask: 1 = not sky (top band is sky)
    out = {"in_img": img, "in_gt": gt, "in_occl": occl, "in_sky": sky}

    x = img.clone().requires_grad_(True)
    m3 = occl.expand(3, H, W)
    l1 = loss_utils.l1_loss(x, gt, mask=m3)
    s = loss_utils.ssim(x, gt, mask=m3)
    lam = 0.2
    rec = l1 * (1 - lam) + lam * (1.0 - s)
    rec.backward()
    out.update(ref_l1=l1.detach(), ref_ssim=s.detach(), ref_rec=rec.detach(), ref_d_img=x.grad.clone(), lam=torch.tensor(lam))
    # SSIM alone (gradient of the mean SSIM), and the sky-BRDF masked L1 against zeros with mask = 1 - sky_mask
    x = img.clone().requires_grad_(True)
    loss_utils.ssim(x, gt, mask=m3).backward()
    out["ref_d_ssim"] = x.grad.clone()
    x = img.clone().requires_grad_(True)
    lz = loss_utils.l1_loss(x, torch.zeros_like(x), mask=1 - sky.expand(3, H, W))
    lz.backward()
    out.update(ref_l1_sky=lz.detach(), ref_d_l1_sky=x.grad.clone())
    # empty mask: the reference returns the Python constants
    z = torch.zeros(3, H, W)
    out["ref_l1_empty"] = torch.tensor(float(loss_utils.l1_loss(img, gt, mask=z)))
    out["ref_ssim_empty"] = torch.tensor(float(loss_utils.ssim(img, gt, mask=z)))

    # depth_to_normal on a smooth depth map with a step edge, camera off the origin (second arc camera)
    cam = arc_cameras(W, H, 3, centroid_z=8.0, step_deg=10.0)[1]
    depth = 6.0 + 0.02 * xx + 0.015 * yy + 0.4 * torch.sin(xx / 7.0) * torch.cos(yy / 5.0)
    depth = depth + 1.5 * (xx > 60).float()
    d = (depth * sky).unsqueeze(0).clone().requires_grad_(True)
    n = gfx.depth_to_normal(View(cam), d)
    # upstream gradient: zero where the sky mask is zero, as the normal regulariser's masks make it (train.py:91-92);
    # (zero-depth pixels have a zero cross product there, and F.normalize's eps branch scales gradients by 1e12)
    gn = torch.randn(H, W, 3, generator=g) * sky.unsqueeze(-1)
    (n * gn).sum().backward()
    out.update(in_depth=depth, in_wvt=cam.world_view_transform, in_fov=torch.tensor([cam.FoVx, cam.FoVy]),
               in_gn=gn, ref_normal=n.detach(), ref_d_depth=d.grad.clone())
    np.savez_compressed(os.path.join(HERE, "image_ops.npz"), **{k: v.numpy() for k, v in out.items()})
    print("wrote image_ops.npz:", {k: tuple(v.shape) for k, v in out.items()})
    print("l1", float(l1), "ssim", float(s), "rec", float(rec))


if __name__ == "__main__":
    main()
