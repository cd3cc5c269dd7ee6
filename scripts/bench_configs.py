#!/usr/bin/env python
"""The BASELINE.json configs bench.py's default line does not cover (one JSON line each, CUDA-event timing):

  A   200 k Gaussians, 800x600, 1 view: U3 (render() forward, frames/s) on the native arm and on the reference arm
      (stock render() + reference extension), + the reference's torch shading (light.py / sh_utils.py) timed on the
      HOST cores with 1 and all threads (reported baseline)
  D   relit-eval grid: 1 M Gaussians x 8 environment maps x 64 views at 1080p, (view, envmap) cells sharded
      view-major over the ranks (r3dgw_b200.parallel.grid_cells_for_rank), no collective on the data path; cells/s
  E   stress: 6 M Gaussians at 3840x2160, one raster call (U1) forward and forward+backward, native and reference

usage:  python scripts/bench_configs.py A|D|E [--views 64 --envmaps 8]      (D under torchrun for N > 1)
"""
import argparse
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200"), os.path.join(ROOT, "oracle")]
import torch  # noqa: E402

import bench  # noqa: E402


def timed(fn, n, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in evs:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in evs)
    return sum(ms) / len(ms), ms[len(ms) // 2]


def config_A():
    from r3dgw_b200.synthetic import make_scene
    dev = torch.device("cuda")
    sc, cam, gt = bench.build_workload("A", 0, dev)
    out = {"config": "A: 200000 Gaussians, 800x600, 1 view, relit forward (U3)", "unit": "frames/s"}
    st = bench.make_stepper("native", "dropin", sc, cam, dev)

    def frame():
        st.raster.clear_geometry_cache()
        with torch.no_grad():
            st.forward(None, st.sky2)
    mean, med = timed(frame, 30, 5)
    out["native_fps"] = 1e3 / mean
    try:
        sr = bench.make_stepper("reference", "dropin", sc, cam, dev)

        def frame_ref():
            with torch.no_grad():
                sr.forward(None, sr.sky2)
        mean, med = timed(frame_ref, 30, 5)
        out["reference_fps"] = 1e3 / mean
        out["reference_caller"] = "stock render()" if isinstance(sr, bench.StockStepper) else "mirror"
    except Exception as e:   # noqa: BLE001
        out["reference_fps"] = None
        out["reference_error"] = str(e)[:200]
    # ---- the reference's CPU-executable piece: torch shading on host cores (north_star) ----------------
    import stock_caller as sc_
    ns = sc_.load_ref_python()
    cores = os.cpu_count() or 1
    fgm = ~sc.is_sky
    n = int(fgm.sum())
    pos, alb, kr, km = sc.means3D[fgm], sc.albedo, sc.roughness, sc.metalness
    nrm = torch.nn.functional.normalize(torch.randn(n, 3, generator=torch.Generator().manual_seed(1)), dim=1)
    campos = cam.camera_center.cpu()
    env = ns.light.EnvironmentLight.__new__(ns.light.EnvironmentLight)
    torch.nn.Module.__init__(env)
    env.sh_degree, env.sh_dim, env.base, env.NUM_CHANNELS = 4, 25, sc.env_sh.clone(), 3
    env.C1, env.C2, env.C3, env.C4, env.C5 = 0.429043, 0.511664, 0.743125, 0.886227, 0.247708
    env._FG_LUT = sc_.real_fg_lut("cpu")
    orig_cuda = torch.Tensor.cuda
    rows = {}
    try:
        torch.Tensor.cuda = lambda self, *a, **k: self        # gauss_kernel hard-codes .cuda() (sh_utils.py:173-175)
        for threads in (1, cores):
            torch.set_num_threads(threads)
            for spec in (False, True):
                def run():
                    with torch.no_grad():
                        env.shade(pos[None, None], nrm[None, None], alb[None, None], campos.repeat(n, 1)[None, None],
                                  kr[None, None], km[None, None], specular=spec)
                run()
                t0 = time.perf_counter()
                reps = 3
                for _ in range(reps):
                    run()
                dt = (time.perf_counter() - t0) / reps
                rows[f"shade_specular_{spec}_threads_{threads}"] = {"ms": dt * 1e3, "gaussians_per_s": n / dt}

            def run_sh():
                with torch.no_grad():
                    ns.sh_utils.eval_sh(4, sc.env_sh.unsqueeze(0).expand(n, -1, -1).transpose(1, 2), nrm)
            run_sh()
            t0 = time.perf_counter()
            for _ in range(3):
                run_sh()
            dt = (time.perf_counter() - t0) / 3
            rows[f"eval_sh_deg4_threads_{threads}"] = {"ms": dt * 1e3, "gaussians_per_s": n / dt}
    finally:
        torch.Tensor.cuda = orig_cuda
    out["cpu_shading_reference_torch"] = {"cores": cores, "foreground_gaussians": n, "rows": rows,
                                          "code": "scene/NVDIFFREC/light.py EnvironmentLight.shade + utils/sh_utils.py eval_sh, "
                                                  "unmodified (oracle/_ref/py), CPU tensors, no_grad"}
    # the same shading on the GPU through the fused kernel, for the ratio
    from r3dgw_b200.shading import shade_gaussians
    d = lambda t: t.to(dev)   # noqa: E731
    lut = sc_.real_fg_lut(dev)
    args = (d(pos), d(nrm), d(alb), d(kr), d(km), d(sc.env_sh), d(campos), lut, 4)
    mean, med = timed(lambda: shade_gaussians(*args, specular=True), 50, 5)
    out["gpu_shading_fused_kernel"] = {"ms": mean, "gaussians_per_s": n / (mean * 1e-3)}
    print(json.dumps(out))


def config_D(views, envmaps, mode="dropin"):
    from r3dgw_b200.parallel import grid_cells_for_rank
    from r3dgw_b200.synthetic import CONFIGS, arc_cameras, make_env_sh, make_scene
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    sc = make_scene("B")
    cfg = CONFIGS["B"]
    cams = [c.to(dev) for c in arc_cameras(cfg["W"], cfg["H"], views, centroid_z=8.0, step_deg=1.0)]
    envs = make_env_sh(torch.Generator().manual_seed(20261104 + 3), envmaps).to(dev)       # seeds +0..+7 of config 4
    st = bench.make_stepper("native", mode, sc, cams[0], dev)
    cells = grid_cells_for_rank(views, envmaps, rank, world)
    sink = torch.zeros((), device=dev)

    def grid():
        with torch.no_grad():
            for v, e in cells:
                st.env_sh.copy_(envs[e])
                out = st.forward(cams[v], st.sky2)
                sink.add_(out["render"].mean())          # PSNR-style scalar per cell stays on the device
    for _ in range(1):
        grid()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 2
    a.record()
    for _ in range(reps):
        grid()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    if dist is not None:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    if rank == 0:
        import diff_gaussian_rasterization as dgr
        print(json.dumps({"config": f"D: 1000000 Gaussians x {envmaps} envmaps x {views} views at 1920x1080, view-major "
                                    f"cells over {world} rank(s), {'render_fused()' if mode == 'fused' else 'render()'} forward, geometry reused across the envmaps of a view",
                          "metric": "relit_eval_cells_per_s", "value": views * envmaps / (ms * 1e-3), "unit": "cells/s",
                          "n_gpus": world, "ms_per_grid": ms, "cells": views * envmaps, "scaling": "strong",
                          "stats_last_grid": dgr.stats()}))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def config_E(arms=("native", "reference")):
    import diff_gaussian_rasterization as dgr
    import ref_loader
    from r3dgw_b200.synthetic import make_scene
    dev = torch.device("cuda")
    sc = make_scene("E").to(dev)
    cam = sc.camera
    H, W = cam.image_height, cam.image_width
    dL = torch.full((3, H, W), 1e-3, device=dev)
    out = {"config": f"E: {sc.P} Gaussians at {W}x{H}, one raster call (U1), C = 3", "unit": "ms"}
    for tag, mod in (("native", dgr), ("reference", ref_loader.load_ref_dgr() if ref_loader.available() else None)):
        if mod is None or tag not in arms:
            continue
        s = mod.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.world_view_transform,
                                              cam.full_proj_transform, -1, cam.camera_center, False)
        leaves = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.opacities, sc.scales, sc.rotations, sc.colors)]

        def call(bwd):
            if mod is dgr:
                dgr.clear_geometry_cache()
            m2 = torch.zeros_like(leaves[0], requires_grad=True)
            with torch.set_grad_enabled(bwd):
                img, radii = mod.GaussianRasterizer(s)(means3D=leaves[0], means2D=m2, opacities=leaves[1],
                                                       colors_precomp=leaves[4], scales=leaves[2], rotations=leaves[3])
            if bwd:
                img.backward(dL)
            return radii
        try:
            radii = call(False)
            fwd, _ = timed(lambda: call(False), 5, 2)
            fb, _ = timed(lambda: call(True), 5, 2)
            out[tag] = {"fwd_ms": fwd, "fwd_bwd_ms": fb, "P_visible": int((radii > 0).sum())}
            if mod is dgr:
                out[tag]["R"] = int(dgr.last_geometry_state().R)
        except Exception as e:   # noqa: BLE001 - the reference's own limits are part of the record
            out[tag] = {"failed": f"{type(e).__name__}: {str(e)[:200]}"}
        torch.cuda.empty_cache()
    print(json.dumps(out))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["A", "D", "E"])
    ap.add_argument("--views", type=int, default=64)
    ap.add_argument("--envmaps", type=int, default=8)
    ap.add_argument("--mode", default="dropin", choices=["dropin", "fused"], help="config D: render() or render_fused()")
    ap.add_argument("--arm", default="both", choices=["both", "native", "reference"],
                    help="config E: run one arm per process (the reference overflows its 32-bit buffer sizes at E and can "
                         "leave the CUDA context in a failed state)")
    a = ap.parse_args()
    if a.config == "E":
        config_E(("native", "reference") if a.arm == "both" else (a.arm,))
    else:
        {"A": config_A}.get(a.config, lambda: config_D(a.views, a.envmaps, a.mode))()
