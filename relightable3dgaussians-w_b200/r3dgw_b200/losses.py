"""The training loss of train.py:78-112 (post-15 000 regime) as ONE autograd node over the fused kernels.

    rec   = (1 - l_dssim) * l1_loss(image, gt, occl) + l_dssim * (1 - ssim(image, gt, occl))          train.py:78-81
    sky   = l1_loss(diffuse, 0, 1 - sky) + l1_loss(specular, 0, 1 - sky)                              train.py:86-87
    nrm   = mean(1 - sum_c (normal occl sky)_c (normal_ref occl sky)_c)                               train.py:92-96
    scale = min_scale_loss(radii, gaussians)                                                          train.py:105-107
    depth = depth_loss_gaussians(gaussians, camera, visibility_filter)                                train.py:110-112
    loss  = rec + l_sky_brdf * sky + l_normal * nrm + l_scale * scale + l_sky_gauss * depth

Every term is one of the kernels behind r3dgw_b200.image_ops / regularisers (same values, same gradients); what this
node removes is the host side: eight autograd Function calls and ~25 scalar torch kernels per step (the loss section
of the step was launch-bound: its kernels take ~0.3 ms of GPU time, its Python ~0.6 ms).  All partial results land in
one device buffer; the weighted sum is a dot product; the backward coefficients are one multiply.
(The env-SH positivity term, train.py:99-102, is stochastic and stays with the caller.)
There is no CPU path.
"""
import torch

from . import native as _nat
from .image_ops import _f32c, _mask_args, _require_cuda, _stream

_W_CACHE = {}


def _weights(lam, device):
    key = (tuple(sorted(lam.items())), str(device))
    w = _W_CACHE.get(key)
    if w is None:
        ld = float(lam["dssim"])
        # slots of the result buffer that enter the loss: ssim mean (the loss has + ld * (1 - ssim): the constant ld is
        # added apart), masked L1 mean of the render, L1 of diffuse / specular under 1 - sky, normal-consistency mean,
        # min_scale_loss, depth_loss_gaussians.  Only these are gathered: a NaN statistic in another slot (e.g. the mean
        # sky depth when no sky Gaussian is visible) must not leak in through 0 * NaN
        idx = torch.tensor([0, 1, 5, 9, 12, 16, 17])
        v = torch.tensor([-ld, 1.0 - ld, lam["sky_brdf"], lam["sky_brdf"], lam["normal"], lam["scale"], lam["sky_gauss"]],
                         dtype=torch.float32)
        v = (idx.to(device), v.to(device))
        coef = torch.tensor([-ld, 1.0 - ld, lam["sky_brdf"], lam["sky_brdf"], lam["normal"], lam["scale"], lam["sky_gauss"], 0.0],
                            dtype=torch.float32)
        w = (v, coef.to(device), ld)
        _W_CACHE[key] = w
    return w


class _TrainLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, diffuse, specular, normal, normal_ref, scales, means3D, gt, occl, sky, radii, is_sky_u8,
                viewmatrix, wvec, cvec, const, gamma):
        _require_cuda(image, "training_loss")
        dev = image.device
        L = _nat.lib()
        x, y = _f32c(image), _f32c(gt)
        C, H, W = (int(s) for s in x.shape)
        dif, spc, nrm, nrf = _f32c(diffuse), _f32c(specular), _f32c(normal), _f32c(normal_ref)
        occ_k, occ_p, occ_s = _mask_args(occl, C, H, W)
        sky_k, sky_p, sky_s = _mask_args(sky, C, H, W)
        scl, m3, vm = _f32c(scales), _f32c(means3D), _f32c(viewmatrix)
        rad = radii if radii.dtype == torch.int32 else radii.to(torch.int32)
        P = int(scl.shape[0])
        res = torch.empty(24, dtype=torch.float32, device=dev)
        wsz = max(int(L.r3dgw_image_loss_workspace_bytes(C, H, W)), int(L.r3dgw_gaussian_reg_workspace_bytes(P)))
        ws = torch.empty(5, wsz, dtype=torch.uint8, device=dev)
        dmap = torch.empty(3, C, H, W, dtype=torch.float32, device=dev)
        rp, st = res.data_ptr(), _stream()
        sky1 = sky_k if sky_s == 0 else sky_k[0]          # [H,W] plane for the kernels that take one
        occ1 = occ_k if occ_s == 0 else occ_k[0]
        with _nat.on_device(dev):
            rc = L.r3dgw_ssim_l1_forward(C, H, W, x.data_ptr(), y.data_ptr(), occ_p, occ_s, dmap.data_ptr(),
                                         ws[0].data_ptr(), rp, st)
            rc = rc or L.r3dgw_masked_l1_forward(C, H, W, dif.data_ptr(), None, sky_p, sky_s, 1, ws[1].data_ptr(), rp + 16, st)
            rc = rc or L.r3dgw_masked_l1_forward(C, H, W, spc.data_ptr(), None, sky_p, sky_s, 1, ws[2].data_ptr(), rp + 32, st)
            rc = rc or L.r3dgw_normal_loss_forward(H, W, nrm.data_ptr(), nrf.data_ptr(), occ1.data_ptr(), sky1.data_ptr(),
                                                   ws[3].data_ptr(), rp + 48, st)
            rc = rc or L.r3dgw_gaussian_regularisers_forward(P, rad.data_ptr(), None if is_sky_u8 is None else is_sky_u8.data_ptr(),
                                                             scl.data_ptr(), m3.data_ptr(), vm.data_ptr(), float(gamma),
                                                             ws[4].data_ptr(), rp + 64, st)
        _nat.check(rc, "training_loss forward")
        ctx.keep = (x, y, dif, spc, nrm, nrf, occ_k, sky_k, scl, rad, is_sky_u8, vm, dmap, res, cvec, occ1, sky1)
        ctx.cfg = (C, H, W, P, occ_p, occ_s, sky_p, sky_s, float(gamma))
        ctx.shapes = (image.shape, diffuse.shape, specular.shape, normal.shape, normal_ref.shape, scales.shape, means3D.shape)
        ctx.mark_non_differentiable(res)
        return torch.dot(res.index_select(0, wvec[0]), wvec[1]) + const, res

    @staticmethod
    def backward(ctx, g, _g_res):
        x, y, dif, spc, nrm, nrf, occ_k, sky_k, scl, rad, sky_u8, vm, dmap, res, cvec, occ1, sky1 = ctx.keep
        C, H, W, P, occ_p, occ_s, sky_p, sky_s, gamma = ctx.cfg
        dev = x.device
        L = _nat.lib()
        coef = (cvec * g.reshape(())).contiguous()       # [ssim, l1, sky(dif), sky(spc), normal, scale, depth, 0]
        cp, rp, st = coef.data_ptr(), res.data_ptr(), _stream()
        gi = torch.empty(5, C, H, W, dtype=torch.float32, device=dev)
        gp = torch.empty(2, P, 3, dtype=torch.float32, device=dev)
        with _nat.on_device(dev):
            rc = L.r3dgw_ssim_l1_backward(C, H, W, x.data_ptr(), y.data_ptr(), occ_p, occ_s, dmap.data_ptr(), rp, cp,
                                          gi[0].data_ptr(), st)
            rc = rc or L.r3dgw_masked_l1_backward(C, H, W, dif.data_ptr(), None, sky_p, sky_s, 1, rp + 16, cp + 8,
                                                  gi[1].data_ptr(), st)
            rc = rc or L.r3dgw_masked_l1_backward(C, H, W, spc.data_ptr(), None, sky_p, sky_s, 1, rp + 32, cp + 12,
                                                  gi[2].data_ptr(), st)
            rc = rc or L.r3dgw_normal_loss_backward(H, W, nrm.data_ptr(), nrf.data_ptr(), occ1.data_ptr(), sky1.data_ptr(),
                                                    cp + 16, gi[3].data_ptr(), gi[4].data_ptr(), st)
            rc = rc or L.r3dgw_gaussian_regularisers_backward(P, rad.data_ptr(), None if sky_u8 is None else sky_u8.data_ptr(),
                                                              scl.data_ptr(), vm.data_ptr(), gamma, rp + 64, cp + 20,
                                                              gp[0].data_ptr(), gp[1].data_ptr(), st)
        _nat.check(rc, "training_loss backward")
        s = ctx.shapes
        return (gi[0].view(s[0]), gi[1].view(s[1]), gi[2].view(s[2]), gi[3].view(s[3]), gi[4].view(s[4]),
                gp[0].view(s[5]), gp[1].view(s[6])) + (None,) * 10


def training_loss(render_pkg, gt_image, occluders_mask, sky_mask, scales, means3D, is_sky, viewmatrix, lambdas, gamma=0.02):
    """-> (loss, parts).  `render_pkg`: render()'s dict (render, diffuse_color, specular_color, normal, normal_ref,
    radii); masks [H,W] / [1,H,W] / expanded [3,H,W]; `scales` / `means3D` the activated Gaussian parameters; `is_sky`
    bool / u8 [P] or None; `lambdas` = {dssim, sky_brdf, normal, scale, sky_gauss}.  `parts` is the raw f32[24] result
    buffer (detached): [0] ssim mean, [1] L1 mean, [5] sky L1 diffuse, [9] sky L1 specular, [12] normal term,
    [16] min-scale term, [17] sky-depth term."""
    dev = render_pkg["render"].device
    wvec, cvec, const = _weights(lambdas, dev)
    sky_u8 = None
    if is_sky is not None:
        sky_u8 = is_sky.reshape(-1)
        sky_u8 = (sky_u8 if sky_u8.dtype == torch.uint8 else sky_u8.to(torch.uint8)).contiguous()
    loss, parts = _TrainLossFn.apply(render_pkg["render"], render_pkg["diffuse_color"], render_pkg["specular_color"],
                                     render_pkg["normal"], render_pkg["normal_ref"], scales, means3D, gt_image,
                                     occluders_mask, sky_mask, render_pkg["radii"], sky_u8, viewmatrix, wvec, cvec, const,
                                     gamma)
    return loss, parts
