"""Host side of the eval grid on the device over the C-ABI in include/r3dgw_b200.h.

Drop-ins with the reference's names and argument meaning (there is no CPU path; numpy / CPU inputs are uploaded):
  get_coefficients_from_image(ibl, l_max=2, resize_width=None)   utils/sh_additional_utils.py:176-206
  process_environment_map_image(img, scale_high, threshold)      eval_with_gt_envmaps.py:42-49
  rotate_sh(F_nm, yaw, pitch, roll, sh_type='real')              spaudiopy.sph.rotate_sh (spaudiopy==0.1.6)
and the batched form the sun-angle sweep needs (eval_with_gt_envmaps.py:88-100), ONE launch for all 51 angles:
  sun_sweep_envmaps(coeffs, init_rot_zyx, sun_angles) -> CUDA f32 [B, n, 3], entry b ready for envlight.set_base

The *_device variants return CUDA tensors and never synchronise; the reference-named functions return numpy like
the reference does (one small device -> host read).
"""
import math

import numpy as np
import torch

from . import native as _nat

MAX_L = 4


def _stream():
    return _nat.raw_stream()


def _device(device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("envmap_sh (B200-native): a CUDA device is required; there is no CPU path")
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def _upload_image(ibl, device):
    """-> (tensor kept alive [H,W,S], f32 ptr or None, u8 ptr or None, H, W, S)."""
    t = torch.from_numpy(np.ascontiguousarray(ibl)) if isinstance(ibl, np.ndarray) else ibl
    if t.dim() != 3 or t.shape[2] < 3:
        raise ValueError("environment map must be [H, W, >=3]")
    if t.dtype != torch.uint8 and t.dtype != torch.float32:
        t = t.float()       # the reference's f64 product of an f32 image; f64 images are not used by any caller
    t = t.to(device, non_blocking=True).contiguous()
    H, W, S = (int(s) for s in t.shape)
    if t.dtype == torch.uint8:
        return t, None, t.data_ptr(), H, W, S
    return t, t.data_ptr(), None, H, W, S


def _target_size(W, resize_width):
    """utils/sh_additional_utils.py:178-184."""
    if resize_width is not None:
        return int(resize_width), int(resize_width / 2)
    if W > 1000:
        return 1000, 500
    return None


def get_coefficients_from_image_device(ibl, l_max=2, resize_width=None, filder_amount=None, *, boost=None, device=None):
    """-> CUDA f64 [(l_max+1)^2, 3].  boost=(scale_high, threshold) fuses process_environment_map_image's u8 -> f32/255
    + bright-pixel scaling into the load (u8 input only)."""
    if filder_amount is not None:
        raise NotImplementedError("blur_ibl pre-filtering (never enabled by the reference's callers)")
    if not 0 <= int(l_max) <= MAX_L:
        raise NotImplementedError(f"l_max must be in 0..{MAX_L} (the environment light holds 25 coefficients)")
    dev = _device(device)
    L = _nat.lib()
    t, f32, u8, H, W, S = _upload_image(ibl, dev)
    if boost is not None and u8 is None:
        raise ValueError("boost=(scale_high, threshold) applies to uint8 images only")
    bo, scale, thr = (1, float(boost[0]), float(boost[1])) if boost is not None else (0, 1.0, 0.0)
    with _nat.on_device(dev):
        size = _target_size(W, resize_width)
        if size is not None:
            Wd, Hd = size
            small = torch.empty(Hd, Wd, 3, dtype=torch.float32, device=dev)
            rc = L.r3dgw_envmap_resize_cubic(H, W, S, f32, u8, bo, thr, scale, Hd, Wd, small.data_ptr(), _stream())
            _nat.check(rc, "r3dgw_envmap_resize_cubic")
            t, f32, u8, H, W, S, bo = small, small.data_ptr(), None, Hd, Wd, 3, 0
        if H != W // 2:
            raise ValueError(f"lat-long environment map must be [W/2, W] (got {H}x{W})")
        n = (int(l_max) + 1) ** 2
        ws = torch.empty(L.r3dgw_envmap_sh_workspace_bytes(H), dtype=torch.uint8, device=dev)
        coeffs = torch.empty(n, 3, dtype=torch.float64, device=dev)
        rc = L.r3dgw_envmap_sh_project(H, W, S, f32, u8, bo, thr, scale, int(l_max), ws.data_ptr(), coeffs.data_ptr(),
                                       _stream())
        _nat.check(rc, "r3dgw_envmap_sh_project")
    return coeffs


def get_coefficients_from_image(ibl, l_max=2, resize_width=None, filder_amount=None):
    """utils/sh_additional_utils.py:176-206 -> numpy f64 [(l_max+1)^2, 3]."""
    return get_coefficients_from_image_device(ibl, l_max, resize_width, filder_amount).cpu().numpy()


def process_environment_map_image(img, scale_high, threshold, *, device=None, as_numpy=True):
    """eval_with_gt_envmaps.py:42-49.  `img`: the decoded uint8 [H,W,3] image (numpy or tensor); decoding the file
    (plt.imread in the reference) stays with the caller."""
    if isinstance(img, (str, bytes)):
        raise TypeError("pass the decoded uint8 image; file decoding is outside the accelerated path")
    c = get_coefficients_from_image_device(img, 4, boost=(scale_high, threshold), device=device)
    return c.cpu().numpy() if as_numpy else c


def sh_render_device(coeffs, width=600, *, device=None):
    """sh_render (utils/sh_additional_utils.py:288-299) on the device: coeffs [(l_max+1)^2, 3] -> CUDA f64
    [width/2, width, 3], linear."""
    dev = _device(device if device is not None else (coeffs.device if isinstance(coeffs, torch.Tensor)
                                                       and coeffs.is_cuda else None))
    c = _as_f64_device(coeffs, dev)
    n = int(c.shape[0])
    l_max = int(math.sqrt(n) - 1)             # sh_l_max_from_terms, :390-391
    if l_max > MAX_L or c.dim() != 2 or c.shape[1] != 3:
        raise NotImplementedError("sh_render_device: [(l_max+1)^2, 3] coefficients with l_max <= 4")
    W = int(width)
    out = torch.empty(W // 2, W, 3, dtype=torch.float64, device=dev)
    with _nat.on_device(dev):
        rc = _nat.lib().r3dgw_envmap_sh_render(l_max, c.data_ptr(), W, out.data_ptr(), _stream())
    _nat.check(rc, "r3dgw_envmap_sh_render")
    return out


def _as_f64_device(x, dev):
    if isinstance(x, torch.Tensor):
        return x.to(device=dev, dtype=torch.float64).contiguous()
    return torch.as_tensor(np.asarray(x, dtype=np.float64), device=dev).contiguous()


def sh_rotate_sweep(coeffs, angles_ypr, first_ypr=None, *, out_dtype=torch.float32, device=None):
    """coeffs [n, C] (n = (l_max+1)^2 <= 25, C <= 4) -> [B, n, C]: every entry rotated by first_ypr (optional) and then
    by angles_ypr[b], each (yaw, pitch, roll) of a z-y-x Euler rotation."""
    dev = _device(device if device is not None else (coeffs.device if isinstance(coeffs, torch.Tensor)
                                                       and coeffs.is_cuda else None))
    F = _as_f64_device(coeffs, dev)
    n, C = (int(s) for s in F.shape)
    l_max = int(round(math.sqrt(n))) - 1
    if (l_max + 1) ** 2 != n or l_max > MAX_L or C > 4:
        raise NotImplementedError("sh_rotate_sweep: (l_max+1)^2 coefficients with l_max <= 4 and at most 4 channels")
    A = _as_f64_device(angles_ypr, dev).reshape(-1, 3)
    first = None if first_ypr is None else _as_f64_device(first_ypr, dev).reshape(3)
    B = int(A.shape[0])
    out = torch.empty(B, n, C, dtype=out_dtype, device=dev)
    o32, o64 = (out.data_ptr(), None) if out_dtype == torch.float32 else (None, out.data_ptr())
    if out_dtype not in (torch.float32, torch.float64):
        raise ValueError("out_dtype must be float32 or float64")
    with _nat.on_device(dev):
        rc = _nat.lib().r3dgw_sh_rotate_sweep(l_max, C, F.data_ptr(), None if first is None else first.data_ptr(), B,
                                              A.data_ptr(), o32, o64, _stream())
    _nat.check(rc, "r3dgw_sh_rotate_sweep")
    return out


def rotate_sh(F_nm, yaw, pitch, roll, sh_type="real"):
    """spaudiopy.sph.rotate_sh: F_nm [..., (N+1)^2] rows of real SH coefficients -> rotated rows (same container
    kind as the input: numpy in -> numpy f64 out, CUDA tensor in -> CUDA f64 tensor out)."""
    if sh_type != "real":
        raise NotImplementedError("only real SH rotation (the reference never asks for another)")
    is_np = not isinstance(F_nm, torch.Tensor)
    F = torch.as_tensor(np.asarray(F_nm, dtype=np.float64)) if is_np else F_nm.double()
    lead = F.shape[:-1]
    rows = F.reshape(-1, F.shape[-1])
    out = []
    for r0 in range(0, rows.shape[0], 4):       # the kernel takes up to 4 channels (rows) per call
        blk = rows[r0:r0 + 4]
        o = sh_rotate_sweep(blk.T.contiguous(), [[float(yaw), float(pitch), float(roll)]], out_dtype=torch.float64)
        out.append(o[0].T)
    res = torch.cat(out, 0).reshape(*lead, F.shape[-1])
    return res.cpu().numpy() if is_np else res


def sun_sweep_envmaps(coeffs, init_rot_zyx, sun_angles, *, device=None):
    """eval_with_gt_envmaps.py:88-100 for all sun angles at once.
    coeffs [n,3] (get_coefficients_from_image's output, numpy or CUDA f64); init_rot_zyx = (init_rot_z, init_rot_y,
    init_rot_x); sun_angles [B,3] rows `angle` used as rotate_sh(., angle[2], angle[0], angle[1]) -> CUDA f32 [B,n,3]."""
    a = torch.as_tensor(np.asarray(sun_angles, dtype=np.float64) if not isinstance(sun_angles, torch.Tensor)
                        else sun_angles).double().reshape(-1, 3)
    ypr = torch.stack([a[:, 2], a[:, 0], a[:, 1]], dim=1)
    return sh_rotate_sweep(coeffs, ypr, first_ypr=[float(v) for v in init_rot_zyx], device=device)
