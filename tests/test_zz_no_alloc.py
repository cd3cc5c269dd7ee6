"""Allocation-free C-ABI entry point (r3dgw_gaussian_regularisers_forward_noalloc) against the workspace entry point.

The entry point has no counterpart in the reference (added for callers that
cannot call cudaMalloc): its scratch is a file-scope static __device__ array resolved with cudaGetSymbolAddress.
Collected LAST on purpose; passes alone in a fresh process and after the other tests.
Tolerance: bit-equal - both entry points run the same kernels with the same fixed-order reduction.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(torch, P, seed, sky_fraction=0.1):
    g = torch.Generator().manual_seed(seed)
    scales = (torch.rand(P, 3, generator=g) * 0.05 + 1e-3).cuda()
    means3D = (torch.randn(P, 3, generator=g) * 3.0).cuda()
    radii = (torch.rand(P, generator=g) < 0.7).to(torch.int32).mul_(7).cuda()
    is_sky = (torch.rand(P, generator=g) < sky_fraction).to(torch.uint8).cuda()
    vm = torch.eye(4)
    vm[3, :3] = torch.tensor([0.3, -0.2, 5.0])      # stored transposed, like camera.world_view_transform
    vm[:3, :3] = torch.linalg.qr(torch.randn(3, 3, generator=g))[0]
    return scales, means3D, radii, is_sky, vm.contiguous().cuda()


def _with_workspace(torch, L, P, scales, means3D, radii, is_sky, vm, gamma):
    ws = torch.empty(L.r3dgw_gaussian_reg_workspace_bytes(P), dtype=torch.uint8, device="cuda")
    out = torch.full((8,), -7.0, dtype=torch.float32, device="cuda")
    rc = L.r3dgw_gaussian_regularisers_forward(P, radii.data_ptr(), is_sky.data_ptr(), scales.data_ptr(),
                                               means3D.data_ptr(), vm.data_ptr(), gamma, ws.data_ptr(), out.data_ptr(),
                                               torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    return out


def test_noalloc_regularisers_match_the_workspace_entry_point_twice_in_a_row(native_lib):
    import torch
    L = native_lib
    gamma = 0.02
    stream = torch.cuda.current_stream().cuda_stream
    assert L.r3dgw_gaussian_regularisers_noalloc_max_gaussians() == 1 << 24
    # two calls in a row on the shared static scratch: a large P first (every per-CTA row of the array written), then
    # a small one (3 rows; the rest of the array still holds the first call's sums and must not be read), then a
    # third with another seed
    cases = [(300_000, 11), (700, 12), (65_537, 13)]
    prepared = [(P, _inputs(torch, P, seed)) for P, seed in cases]
    expected = [_with_workspace(torch, L, P, *inp, gamma) for P, inp in prepared]
    outs = [torch.full((8,), -3.0, dtype=torch.float32, device="cuda") for _ in cases]
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    alloc0 = torch.cuda.memory_allocated()
    for (P, (scales, means3D, radii, is_sky, vm)), out in zip(prepared, outs):
        rc = L.r3dgw_gaussian_regularisers_forward_noalloc(P, radii.data_ptr(), is_sky.data_ptr(), scales.data_ptr(),
                                                           means3D.data_ptr(), vm.data_ptr(), gamma, out.data_ptr(),
                                                           stream)
        assert rc == 0, rc
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == alloc0
    assert torch.cuda.mem_get_info()[0] >= free0 - (1 << 20)      # nothing was allocated on the device either
    for (P, _), out, exp in zip(prepared, outs, expected):
        a, b = out.cpu().numpy(), exp.cpu().numpy()
        assert np.isfinite(b[:2]).all() and b[6] > 0, (P, b)
        assert np.array_equal(a, b), (P, a, b)
    # the Python wrapper (no conversions, no allocations) goes through the same entry point
    from r3dgw_b200.regularisers import gaussian_regularisers_forward_noalloc, gaussian_regularisers
    P, (scales, means3D, radii, is_sky, vm) = prepared[1]
    out = torch.empty(8, dtype=torch.float32, device="cuda")
    gaussian_regularisers_forward_noalloc(scales, means3D, radii, is_sky, vm, gamma, out)
    ls, ld, stats = gaussian_regularisers(scales, means3D, radii, is_sky, vm, gamma)
    assert torch.equal(out, stats) and float(ls) == float(out[0]) and float(ld) == float(out[1])
    with pytest.raises(RuntimeError):
        gaussian_regularisers_forward_noalloc(scales.double(), means3D, radii, is_sky, vm, gamma, out)


def test_noalloc_regularisers_edge_cases_and_limit(native_lib, capfd):
    import torch
    from r3dgw_b200 import native
    L = native_lib
    stream = torch.cuda.current_stream().cuda_stream
    P = 1000
    scales, means3D, radii, is_sky, vm = _inputs(torch, P, 21)
    out = torch.full((8,), -3.0, dtype=torch.float32, device="cuda")
    # beyond the stated maximum: a clear error, nothing launched, `out` untouched
    launches = L.r3dgw_launch_count()
    too_many = L.r3dgw_gaussian_regularisers_noalloc_max_gaussians() + 1
    rc = L.r3dgw_gaussian_regularisers_forward_noalloc(too_many, radii.data_ptr(), is_sky.data_ptr(), scales.data_ptr(),
                                                       means3D.data_ptr(), vm.data_ptr(), 0.02, out.data_ptr(), stream)
    assert rc == -5
    assert L.r3dgw_launch_count() == launches
    assert "exceeds the maximum" in capfd.readouterr().err
    with pytest.raises(native.R3dgwError, match="static scratch"):
        native.check(rc, "r3dgw_gaussian_regularisers_forward_noalloc")
    torch.cuda.synchronize()
    assert float(out[0]) == -3.0
    # bad arguments
    assert L.r3dgw_gaussian_regularisers_forward_noalloc(P, radii.data_ptr(), None, scales.data_ptr(),
                                                         means3D.data_ptr(), vm.data_ptr(), 0.02, None, stream) == -1
    assert L.r3dgw_gaussian_regularisers_forward_noalloc(-1, radii.data_ptr(), None, scales.data_ptr(),
                                                         means3D.data_ptr(), vm.data_ptr(), 0.02, out.data_ptr(), stream) == -1
    # P = 0, no sky array, nothing visible: same answers as the workspace entry point
    zero_sky = torch.zeros_like(is_sky)
    for p_, rad, sky in ((0, radii, is_sky), (P, radii, None), (P, torch.zeros_like(radii), is_sky)):
        ws = torch.empty(L.r3dgw_gaussian_reg_workspace_bytes(p_), dtype=torch.uint8, device="cuda")
        exp = torch.empty(8, dtype=torch.float32, device="cuda")
        sky_ptr = None if sky is None else sky.data_ptr()
        assert L.r3dgw_gaussian_regularisers_forward(p_, rad.data_ptr(), sky_ptr, scales.data_ptr(), means3D.data_ptr(),
                                                     vm.data_ptr(), 0.02, ws.data_ptr(), exp.data_ptr(), stream) == 0
        assert L.r3dgw_gaussian_regularisers_forward_noalloc(p_, rad.data_ptr(), sky_ptr, scales.data_ptr(),
                                                             means3D.data_ptr(), vm.data_ptr(), 0.02, out.data_ptr(),
                                                             stream) == 0
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), exp.cpu().numpy(), equal_nan=True), (p_, out, exp)
    del zero_sky
