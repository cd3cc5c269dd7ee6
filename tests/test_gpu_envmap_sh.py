"""GPU parity: host side of the eval grid on the device (csrc/envmap_sh.cu) through the C-ABI.

Floating point, FP64 on the device like numpy on the host.  Tolerances:
  projected coefficients  1e-11 * max|coeff| absolute against the golden vectors recorded from the reference's own
                          get_coefficients_from_image (summation order differs; the arithmetic is f64 in both)
  cubic resize pixels     2e-6 * max|pixel| (f32 arithmetic; OpenCV's SIMD accumulation order is not specified);
                          coefficients after a resize 1e-6 * max|coeff|
  SH rotation             1e-12 (f64 outputs) / f32 rounding of the same value (f32 outputs) against the oracle's
                          restatement of spaudiopy's algorithm (parity with spaudiopy itself is unpinned: absent here)
"""
import math
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_projection_vs_reference_python_golden(native_lib):
    from r3dgw_b200 import envmap_sh as E
    z = np.load(os.path.join(GOLD, "envmap_sh.npz"))
    for name in "abcde":
        ref = z[name + "_coeffs"]
        got = E.get_coefficients_from_image(z[name + "_img"], int(z[name + "_lmax"]))
        assert got.dtype == np.float64 and got.shape == ref.shape
        np.testing.assert_allclose(got, ref, rtol=0, atol=1e-11 * np.abs(ref).max(), err_msg=name)
    # u8 path with the bright-pixel boost fused into the load (process_environment_map_image)
    got = E.process_environment_map_image(z["u8_img"], 10, 0.999)
    np.testing.assert_allclose(got, z["u8_coeffs"], rtol=0, atol=1e-11 * np.abs(z["u8_coeffs"]).max())
    # CUDA tensor input, RGBA stride, device output without a sync
    rgba = torch.cat([torch.from_numpy(z["a_img"]), torch.ones(32, 64, 1)], dim=2).cuda()
    dev = E.get_coefficients_from_image_device(rgba, 4)
    assert dev.is_cuda and dev.dtype == torch.float64
    np.testing.assert_allclose(dev.cpu().numpy(), z["a_coeffs"], rtol=0, atol=1e-11 * np.abs(z["a_coeffs"]).max())


def test_resize_and_projection_vs_reference_python_golden(native_lib):
    import sh_oracle as so
    from r3dgw_b200 import envmap_sh as E, native
    z = np.load(os.path.join(GOLD, "envmap_sh.npz"))
    for key, w, ph, l, rw in (("big", 1200, 0.1, 4, None), ("rw", 300, 0.2, 3, 128)):
        got = E.get_coefficients_from_image(so.synth_envmap(w, ph), l, rw)
        np.testing.assert_allclose(got, z[key + "_coeffs"], rtol=0, atol=1e-6 * np.abs(z[key + "_coeffs"]).max())
    # the resize kernel on its own against cv2's output
    src = torch.from_numpy(so.synth_envmap(96, 0.3)).cuda()
    dst = torch.empty(20, 40, 3, device="cuda")
    rc = native.lib().r3dgw_envmap_resize_cubic(48, 96, 3, src.data_ptr(), None, 0, 0.0, 1.0, 20, 40, dst.data_ptr(),
                                                torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    ref = z["resized_pixels"]
    np.testing.assert_allclose(dst.cpu().numpy(), ref, rtol=0, atol=2e-6 * np.abs(ref).max())
    np.testing.assert_allclose(dst.cpu().numpy(), so.resize_cubic(src.cpu().numpy(), 40, 20), rtol=0,
                               atol=1e-6 * np.abs(ref).max())
    # u8 + boost + resize (the path a real 2048-px JPEG takes) against the oracle
    rng = np.random.default_rng(3)
    u8 = rng.integers(0, 256, size=(600, 1200, 3), dtype=np.uint8)
    u8[100:120, 300:330] = 255
    got = E.process_environment_map_image(u8, 10, 0.999)
    ref = so.coefficients_from_image(so.scale_bright(u8, 10, 0.999), 4)
    np.testing.assert_allclose(got, ref, rtol=0, atol=1e-6 * np.abs(ref).max())


def test_projection_properties_full_size(native_lib):
    """Size-independent checks at the reference's working size (1000 x 500): linearity, a constant map projects onto
    Y_00 only (the quadrature's DC response), bit-reproducibility of the fixed-order reduction."""
    import sh_oracle as so
    from r3dgw_b200 import envmap_sh as E
    a = torch.from_numpy(so.synth_envmap(1000, 0.0)).cuda()
    b = torch.from_numpy(so.synth_envmap(1000, 0.37)).cuda()
    ca, cb = E.get_coefficients_from_image_device(a, 4), E.get_coefficients_from_image_device(b, 4)
    cs = E.get_coefficients_from_image_device(a * 2 + b, 4)        # exact in f32 only up to rounding of 2a+b
    ref = so.project((a * 2 + b).cpu().numpy(), 4)
    np.testing.assert_allclose(cs.cpu().numpy(), ref, rtol=0, atol=1e-11 * np.abs(ref).max())
    np.testing.assert_allclose(cs.cpu().numpy(), (2 * ca + cb).cpu().numpy(), rtol=0, atol=1e-5 * np.abs(ref).max())
    assert torch.equal(ca, E.get_coefficients_from_image_device(a, 4))
    const = E.get_coefficients_from_image_device(torch.ones(500, 1000, 3, device="cuda"), 4).cpu().numpy()
    assert abs(const[0, 0] - math.sqrt(4 * math.pi)) < 1e-4 * math.sqrt(4 * math.pi)
    assert np.abs(const[[1, 3, 4, 5, 7, 8], :]).max() < 1e-9      # every m != 0 term integrates to zero over phi
    # a lower l_max returns the leading block of the same coefficients
    c2 = E.get_coefficients_from_image_device(a, 2)
    np.testing.assert_allclose(c2.cpu().numpy(), ca[:9].cpu().numpy(), rtol=0, atol=1e-12)
    with pytest.raises(NotImplementedError):
        E.get_coefficients_from_image(a, 5)
    with pytest.raises(ValueError):
        E.get_coefficients_from_image(torch.ones(30, 50, 3), 2)     # not 2:1


def test_sh_rotation_sweep_vs_oracle(native_lib):
    import sh_oracle as so
    from r3dgw_b200 import envmap_sh as E
    rng = np.random.default_rng(11)
    coeffs = rng.normal(size=(25, 3))
    init = (0.0, 0.0, -math.pi / 2)                                 # eval_with_gt_envmaps.py:150-152
    angles = np.stack([np.linspace(0, math.pi, 51), np.zeros(51), np.zeros(51)], axis=1)
    got = E.sun_sweep_envmaps(coeffs, init, angles)
    ref = so.eval_grid_rotations(coeffs, init, angles)
    assert got.shape == (51, 25, 3) and got.dtype == torch.float32 and got.is_cuda
    np.testing.assert_allclose(got.cpu().numpy(), ref, rtol=0, atol=2e-7 * np.abs(ref).max())
    # general Euler angles, f64 output, every l_max
    for l_max in range(5):
        n = (l_max + 1) ** 2
        F = rng.normal(size=(n, 4))
        ypr = rng.uniform(-3, 3, size=(7, 3))
        got = E.sh_rotate_sweep(F, ypr, first_ypr=(0.3, -0.2, 0.1), out_dtype=torch.float64).cpu().numpy()
        base = so.rotate_sh(F.T, 0.3, -0.2, 0.1)
        for b in range(7):
            np.testing.assert_allclose(got[b], so.rotate_sh(base, *ypr[b]).T, rtol=0, atol=1e-12)
    # the spaudiopy-named single call: numpy in -> numpy out, CUDA in -> CUDA out, any number of rows
    F = rng.normal(size=(5, 25))
    r = E.rotate_sh(F, 0.7, -0.4, 1.9)
    assert isinstance(r, np.ndarray)
    np.testing.assert_allclose(r, so.rotate_sh(F, 0.7, -0.4, 1.9), rtol=0, atol=1e-12)
    rc = E.rotate_sh(torch.from_numpy(F).cuda(), 0.7, -0.4, 1.9)
    assert rc.is_cuda and torch.allclose(rc.cpu(), torch.from_numpy(r), atol=1e-12, rtol=0)
    # properties at any size: identity, norm preservation per band, inverse rotation undoes it (pure yaw)
    np.testing.assert_allclose(E.rotate_sh(F, 0, 0, 0), F, rtol=0, atol=0)
    back = E.rotate_sh(E.rotate_sh(F, 0.9, 0, 0), -0.9, 0, 0)
    np.testing.assert_allclose(back, F, rtol=0, atol=1e-13)
    for l in range(5):
        np.testing.assert_allclose(np.linalg.norm(r[:, l * l:(l + 1) ** 2], axis=1),
                                   np.linalg.norm(F[:, l * l:(l + 1) ** 2], axis=1), rtol=1e-12)


def test_sweep_feeds_the_relit_shading(native_lib):
    """End of the N4 chain: projected + rotated coefficients go straight into the fused shading op as env_sh
    (eval_with_gt_envmaps.py:100-103 set_base + render) without leaving the device."""
    import sh_oracle as so
    from r3dgw_b200 import envmap_sh as E
    img = so.synth_envmap(256, 0.0)
    c = E.get_coefficients_from_image_device(img, 4)
    sweep = E.sun_sweep_envmaps(c, (0.0, 0.0, -math.pi / 2), [[a, 0, 0] for a in np.linspace(0, math.pi, 5)])
    ref = so.eval_grid_rotations(so.project(img, 4), (0.0, 0.0, -math.pi / 2),
                                 [[a, 0, 0] for a in np.linspace(0, math.pi, 5)])
    np.testing.assert_allclose(sweep.cpu().numpy(), ref, rtol=0, atol=2e-7 * np.abs(ref).max())
    assert sweep[0].shape == (25, 3) and sweep.is_contiguous()
