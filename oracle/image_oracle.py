"""TEST INFRASTRUCTURE - torch restatement of the reference's image-space post-ops.

Only tests/, __graft_entry__.smoke() and bench.py's reference / cpu_baseline legs may import this module; the
product path (r3dgw_b200.image_ops) never does.  Device-agnostic (CPU here, the GPU on the box).

Pinned: tests/golden/image_ops.npz holds inputs, outputs and gradients produced by the reference's own Python
(utils/loss_utils.py, utils/graphics_utils.py imported unmodified by tests/golden/make_image_golden.py);
tests/test_cpu_oracle_and_abi.py checks this restatement against it.

  masked_l1        utils/loss_utils.py:27-36
  gaussian_window  utils/loss_utils.py:41-51
  ssim             utils/loss_utils.py:54-99
  depth_points     utils/graphics_utils.py:141-155
  depth_to_normal  utils/graphics_utils.py:157-169
  normal_composite gaussian_renderer/__init__.py:233-237, 268-271
"""
import math

import torch
import torch.nn.functional as F


def masked_l1(x, gt, mask=None):
    """loss_utils.py:27-36: with a mask the sum of |x*mask - gt*mask| is divided by the number of mask entries equal
    to one; an all-zero mask returns the constant 0."""
    if mask is None:
        return (x - gt).abs().mean()
    if float(mask.sum()) == 0:
        return 0
    assert mask.shape[0] == x.shape[0]
    return (x * mask - gt * mask).abs().sum() / (mask == 1).sum()


def gaussian_window(size=11, sigma=1.5, channels=3):
    """loss_utils.py:41-51: normalised 1-D Gaussian (Python doubles -> f32), outer product, one copy per channel."""
    g = torch.Tensor([math.exp(-(i - size // 2) ** 2 / float(2 * sigma ** 2)) for i in range(size)])
    g = (g / g.sum()).unsqueeze(1)
    w2d = g.mm(g.t()).float()
    return w2d.expand(channels, 1, size, size).contiguous()


def ssim(img1, img2, mask, size=11):
    """loss_utils.py:54-99 with size_average=True and a mask (the only form train.py:79 uses)."""
    if float(mask.sum()) == 0:
        return 1
    C = img1.shape[-3]
    win = gaussian_window(size, 1.5, C).to(img1)
    a, b, m = img1.reshape(1, C, *img1.shape[-2:]), img2.reshape(1, C, *img2.shape[-2:]), mask[None]
    blur = lambda t: F.conv2d(t, win, padding=size // 2, groups=C)   # noqa: E731
    mu_a, mu_b = blur(a), blur(b)
    var_a = blur(a * a) - mu_a.pow(2)
    var_b = blur(b * b) - mu_b.pow(2)
    cov = blur(a * b) - mu_a * mu_b
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    smap = ((2 * mu_a * mu_b + c1) * (2 * cov + c2)) / ((mu_a.pow(2) + mu_b.pow(2) + c1) * (var_a + var_b + c2))
    return (smap * m).sum() / (m == 1).sum()


def photometric(image, gt, mask, lambda_dssim):
    """train.py:78-81."""
    l1 = masked_l1(image, gt, mask)
    d = 1.0 - ssim(image, gt, mask)
    return l1 * (1 - lambda_dssim) + lambda_dssim * d, l1, d


def depth_points(world_view_transform, W, H, fovx, fovy, depth):
    """graphics_utils.py:141-155: back-project every pixel centre index (x, y, 1) through K^-1 and the
    camera-to-world rotation, scale by depth, add the camera origin."""
    dev = depth.device
    c2w = world_view_transform.T.inverse()
    fx = W / (2 * math.tan(fovx / 2.))
    fy = H / (2 * math.tan(fovy / 2.))
    dt = depth.dtype      # f32 in the reference (.float()); f64 only when a test asks for the exact value
    K = torch.tensor([[fx, 0., W / 2.], [0., fy, H / 2.], [0., 0., 1.0]]).to(dt).to(dev)
    gx, gy = torch.meshgrid(torch.arange(W, device=dev).to(dt), torch.arange(H, device=dev).to(dt), indexing="xy")
    pix = torch.stack([gx, gy, torch.ones_like(gx)], dim=-1).reshape(-1, 3)
    rays_d = pix @ K.inverse().T @ c2w[:3, :3].T
    return depth.reshape(-1, 1) * rays_d + c2w[:3, 3]


def depth_to_normal(world_view_transform, W, H, fovx, fovy, depth):
    """graphics_utils.py:157-169: central differences of the back-projected points (rows first, then columns),
    cross product, F.normalize, zero border."""
    pts = depth_points(world_view_transform, W, H, fovx, fovy, depth).reshape(*depth.shape[1:], 3)
    out = torch.zeros_like(pts)
    d_rows = pts[2:, 1:-1] - pts[:-2, 1:-1]
    d_cols = pts[1:-1, 2:] - pts[1:-1, :-2]
    out[1:-1, 1:-1, :] = F.normalize(torch.cross(d_rows, d_cols, dim=-1), dim=-1)
    return out


def normal_composite(normal_img, depth_img, alpha_img, sky_mask, world_view_transform, W, H, fovx, fovy):
    """gaussian_renderer/__init__.py:233-237 and :268-271 - the rendered normal mapped back to (-1, 1) and
    composited over the sky, and the depth-derived reference normal weighted by the detached alpha.
    sky_mask is [H,W] (viewpoint_camera.sky_mask.squeeze(), :119)."""
    n = (normal_img - 0.5) * 2.
    n = n * sky_mask + torch.ones_like(n) * (1 - sky_mask)
    n_ref = depth_to_normal(world_view_transform, W, H, fovx, fovy, (depth_img[0] * sky_mask).unsqueeze(0))
    n_ref = n_ref.permute(2, 0, 1) * alpha_img.detach()
    n_ref = n_ref + torch.ones_like(n_ref) * (1 - sky_mask)
    return n, n_ref
