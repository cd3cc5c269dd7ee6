"""Per-Gaussian regularisers of the training step as one fused, sync-free op.

Reference being replaced (boolean-mask gathers, a [P_vis,3] torch.sort, `visibility_filter.sum() > 0` host syncs):
  min_scale_loss(radii, gaussians)                         utils/loss_utils.py:210-220
  depth_loss_gaussians(gaussians, camera, visibility)      utils/loss_utils.py:140-148
There is no CPU path: tensors must be CUDA tensors.
"""
import torch

from . import native as _nat


def _c(t, dtype=torch.float32):
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


class _RegFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scales, means3D, radii, is_sky_u8, viewmatrix, gamma):
        if scales.device.type != "cuda":
            raise RuntimeError("gaussian_regularisers (B200-native): tensors must live on a CUDA device; there is no CPU path")
        dev = scales.device
        P = int(scales.shape[0])
        scl, m3, vm = _c(scales), _c(means3D), _c(viewmatrix)
        rad = _c(radii, torch.int32)
        L = _nat.lib()
        ws = torch.empty(L.r3dgw_gaussian_reg_workspace_bytes(P), dtype=torch.uint8, device=dev)
        out = torch.empty(8, dtype=torch.float32, device=dev)
        with _nat.on_device(dev):
            rc = L.r3dgw_gaussian_regularisers_forward(P, rad.data_ptr(), None if is_sky_u8 is None else is_sky_u8.data_ptr(),
                                                       scl.data_ptr(), m3.data_ptr(), vm.data_ptr(), float(gamma),
                                                       ws.data_ptr(), out.data_ptr(), _nat.raw_stream())
        _nat.check(rc, "r3dgw_gaussian_regularisers_forward")
        ctx.keep = (scl, rad, is_sky_u8, vm, out)
        ctx.cfg = (P, float(gamma), scales.shape, means3D.shape)
        ctx.mark_non_differentiable(out)
        return out[0], out[1], out

    @staticmethod
    def backward(ctx, g_scale, g_depth, _g_out):
        scl, rad, sky, vm, out = ctx.keep
        P, gamma, s_shape, m_shape = ctx.cfg
        dev = scl.device
        z = torch.zeros((), dtype=torch.float32, device=dev)
        g = torch.stack([z if g_scale is None else g_scale.reshape(()).float(),
                         z if g_depth is None else g_depth.reshape(()).float()])
        d = torch.empty(2, P, 3, dtype=torch.float32, device=dev)
        with _nat.on_device(dev):
            rc = _nat.lib().r3dgw_gaussian_regularisers_backward(P, rad.data_ptr(), None if sky is None else sky.data_ptr(),
                                                                 scl.data_ptr(), vm.data_ptr(), gamma, out.data_ptr(),
                                                                 g.data_ptr(), d[0].data_ptr(), d[1].data_ptr(),
                                                                 _nat.raw_stream())
        _nat.check(rc, "r3dgw_gaussian_regularisers_backward")
        return d[0].view(s_shape), d[1].view(m_shape), None, None, None, None


def gaussian_regularisers(scales, means3D, radii, is_sky, viewmatrix, gamma=0.02):
    """-> (min_scale_loss, depth_loss_gaussians, stats): two 0-d CUDA tensors with gradients to `scales` / `means3D`
    and the raw f32[8] statistics {scale_loss, depth_loss, count fg, count sky, mean depth sky, mean depth fg,
    count visible, 0}.  `radii` int32 [P] from the rasterizer (visible = radii > 0); `is_sky` bool / u8 [P] or None;
    `viewmatrix` = camera.world_view_transform as stored (transposed)."""
    sky = None
    if is_sky is not None:
        sky = is_sky.reshape(-1)
        sky = sky if sky.dtype == torch.uint8 else sky.to(torch.uint8)
        sky = sky.contiguous()
    return _RegFn.apply(scales, means3D, radii, sky, viewmatrix, gamma)


def min_scale_loss(radii, gaussians):
    """Drop-in for utils/loss_utils.py:210-220, same signature: `gaussians` is any object with the reference
    GaussianModel's `get_scaling` [P,3] and `get_is_sky` [P,1] getters."""
    scales = gaussians.get_scaling
    vm = torch.eye(4, dtype=torch.float32, device=scales.device)
    return gaussian_regularisers(scales, torch.zeros_like(scales), radii, gaussians.get_is_sky, vm)[0]


def depth_loss_gaussians(gaussians, camera, visibility_filter, gamma=0.02):
    """Drop-in for utils/loss_utils.py:140-148, same signature (`gaussians.get_xyz`, `gaussians.get_is_sky`,
    `camera.world_view_transform`)."""
    xyz = gaussians.get_xyz
    return gaussian_regularisers(torch.ones_like(xyz), xyz, visibility_filter.to(torch.int32), gaussians.get_is_sky,
                                 camera.world_view_transform, gamma)[1]


def gaussian_regularisers_forward_noalloc(scales, means3D, radii, is_sky_u8, viewmatrix, gamma, out):
    """Allocation-free forward (no counterpart in the reference; for callers that cannot call cudaMalloc, e.g. under
    CUDA graph capture): fills the caller's f32[8] `out` with the statistics of `gaussian_regularisers` and returns it.
    Nothing is converted or copied here (that would allocate): every tensor must already be a contiguous CUDA tensor
    of the dtype the C-ABI takes (f32 scales / means3D / viewmatrix / out, int32 radii, uint8 is_sky or None).  The
    partial sums live in a static device array shared by all calls on a GPU: calls that overlap on the GPU race.
    P above `r3dgw_gaussian_regularisers_noalloc_max_gaussians()` (2^24) raises."""
    f32 = torch.float32
    for name, t, dt in (("scales", scales, f32), ("means3D", means3D, f32), ("radii", radii, torch.int32),
                        ("viewmatrix", viewmatrix, f32), ("out", out, f32)) + \
            ((("is_sky_u8", is_sky_u8, torch.uint8),) if is_sky_u8 is not None else ()):
        if t.device.type != "cuda" or t.dtype != dt or not t.is_contiguous():
            raise RuntimeError(f"gaussian_regularisers_forward_noalloc: `{name}` must be a contiguous {dt} CUDA tensor")
    if out.numel() < 8:
        raise RuntimeError("gaussian_regularisers_forward_noalloc: `out` needs 8 floats")
    P = int(scales.shape[0])
    with _nat.on_device(scales.device):
        rc = _nat.lib().r3dgw_gaussian_regularisers_forward_noalloc(
            P, radii.data_ptr(), None if is_sky_u8 is None else is_sky_u8.data_ptr(), scales.data_ptr(),
            means3D.data_ptr(), viewmatrix.data_ptr(), float(gamma), out.data_ptr(), _nat.raw_stream())
    _nat.check(rc, "r3dgw_gaussian_regularisers_forward_noalloc")
    return out
