#!/usr/bin/env python
"""bench.py - train iters/s of the relit Gaussian-splat hot path (BASELINE.json metric) on N B200s.

A STEP is one training iteration of the path on one view per GPU, driven through the reference's call pattern
(r3dgw_b200.render.render == gaussian_renderer/__init__.py:69-274): fused relit shading of all Gaussians, 6 raster
forwards on identical geometry (render, diffuse, specular, depth, normal, alpha), the image-space tail (normal
compositing, depth_to_normal), the training loss of train.py:78-112 (masked L1 + SSIM against a GT image, sky-BRDF,
normal-consistency, planar-scale and sky-depth regularisers), backward through 5 raster backwards + geometry
backward + shading backward (the post-15 000 regime), and - for N > 1 - an NCCL all-reduce of the Gaussian-parameter gradients (views are sharded
over ranks; weak scaling: one view per rank per step).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference] [--workload B] [--mode dropin|fused]

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for how every field is produced.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "relightable3dgaussians-w_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "train_iters_per_s_fwd+bwd_1M_gaussians_1080p"
UNIT = "iters/s"
# configs/optimizer/optimization_params.yaml:12-28 of the reference
LAMBDA = {"dssim": 0.2, "sky_brdf": 0.5, "normal": 0.05, "scale": 100.0, "sky_gauss": 0.05}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="B", help="synthetic config name (r3dgw_b200.synthetic.CONFIGS)")
    ap.add_argument("--mode", default="dropin", choices=["dropin", "fused", "stock"],
                    help="dropin: 6 raster calls through the reference API (repo's render()); fused: one 16-channel "
                         "pass; stock: the reference's unmodified render() + train.py loss on the drop-in modules")
    ap.add_argument("--views", type=int, default=0,
                    help="BASELINE.json config C: V views per step in TOTAL, sharded over the ranks (rank r renders views "
                         "r, r+N, ...; gradients accumulate, ONE all-reduce, densification statistics all-reduced): strong "
                         "scaling.  0 (default): one view per rank per step, weak scaling")
    ap.add_argument("--quick", action="store_true", help="timed step + e2e only (no U1 / U3 / fused / stock / profile extras)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    return ap.parse_args()


# ---- clocks -------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock / power / throttle-reason samples taken DURING the timed region (B200_PROFILING.md clocks line).
    A concurrent poller perturbs what it measures here: a step is ~450 kernel launches, and every NVML / nvidia-smi
    query serialises with them in the driver (measured on this box: `nvidia-smi -lms 200` +3 ms/step; an in-process
    NVML thread at 4 Hz, single stalls of 20-90 ms and +1..3 ms on the following steps).  So the samples are taken
    synchronously from the launching thread BETWEEN the per-step CUDA-event pairs, every few steps, while the GPU is
    still executing the step just enqueued: under load, inside the timed region, outside every timed interval."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.samples = []          # (sm_mhz, power_w, reasons bitmask)
        self.max_mhz = None
        self.source = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it is a list of indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = gpu_index
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                phys = int(vis.split(",")[gpu_index])
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.source = "nvml"
        except Exception:   # noqa: BLE001
            self.nv = None

    def sample(self):
        if self.nv is not None:
            nv, h = self.nv, self.h
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                try:
                    rs = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:   # noqa: BLE001
                    rs = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                self.samples.append((sm, pw, rs))
            except Exception:   # noqa: BLE001
                pass
            return
        try:   # fallback: one nvidia-smi query
            q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
                 "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
                 "clocks_event_reasons.sw_power_cap")
            out = subprocess.run(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=10).stdout.strip().splitlines()[0]
            c = [x.strip() for x in out.split(",")]
            rs = sum(b for b, v in zip([0x8, 0x40, 0x20, 0x4], c[3:7]) if v.lower().startswith("active"))
            self.samples.append((float(c[0]), float(c[2]), rs))
            self.max_mhz = float(c[1])
            self.source = "nvidia-smi"
        except Exception:   # noqa: BLE001
            pass

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["no samples (NVML and nvidia-smi unavailable)"]}
        pmax = max(p for _, p, _ in self.samples)
        busy = [s for s, p, _ in self.samples if p > 0.5 * pmax] or [s for s, _, _ in self.samples]
        seen = set()
        for _, _, rs in self.samples:
            seen |= {name for bit, name in self.REASONS.items() if rs & bit}
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": self.max_mhz, "reasons": sorted(seen),
                "samples": len(self.samples), "power_w_max": pmax, "source": self.source,
                "how": "sampled between the per-step event pairs of the timed region, GPU busy"}


# ---- workload -------------------------------------------------------------------------------------------
def build_workload(name, rank, device):
    """Seeded synthetic scene, one camera per rank on the 10-degree arc, GT image, env / sky SH."""
    import torch
    from r3dgw_b200.synthetic import CONFIGS, arc_cameras, make_scene
    sc = make_scene(name)
    cfg = CONFIGS[name]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cams = arc_cameras(cfg["W"], cfg["H"], max(world, 1), centroid_z=8.0, step_deg=2.0) if world > 1 else [sc.camera]
    cam = cams[rank].to(device)
    g = torch.Generator().manual_seed(777 + rank)
    gt = torch.rand(3, cfg["H"], cfg["W"], generator=g)
    return sc, cam, gt


def make_lut(device):
    """Split-sum FG table: the reference's bsdf_256_256.bin is a data asset we do not copy; same shape, analytic fit."""
    import torch
    res = 256
    u = (torch.arange(res, dtype=torch.float32) + 0.5) / res
    ndv, rough = u.view(1, res).expand(res, res), u.view(res, 1).expand(res, res)
    c0 = torch.tensor([-1.0, -0.0275, -0.572, 0.022])
    c1 = torch.tensor([1.0, 0.0425, 1.04, -0.04])
    r = rough.unsqueeze(-1) * c0 + c1
    a004 = torch.minimum(r[..., 0] * r[..., 0], torch.exp2(-9.28 * ndv)) * r[..., 0] + r[..., 1]
    return torch.stack([-1.04 * a004 + r[..., 2], 1.04 * a004 + r[..., 3]], dim=-1).unsqueeze(0).contiguous().to(device)


class Stepper:
    """One training iteration of the path for either arm (same loss, same call pattern)."""

    def __init__(self, impl, mode, sc, cam, device):
        import torch
        from r3dgw_b200 import render as R
        self.torch, self.R = torch, R
        self.impl, self.mode, self.cam, self.device = impl, mode, cam, device
        self.pc = R.GaussianParams.from_scene(sc.to(device), requires_grad=True)
        self.env_sh = sc.env_sh.to(device).clone().requires_grad_(True)
        self.sky_sh = sc.sky_sh.to(device).clone().requires_grad_(True)
        self.bg = torch.zeros(3, device=device)
        self.lut = make_lut(device)
        H, W = int(cam.image_height), int(cam.image_width)
        # masks all-ones except a 20 % top band of sky (sky_mask == 0 there); [H,W] / [1,H,W] like the
        # reference's viewpoint_camera.sky_mask.squeeze() and occluders_mask
        self.sky2 = torch.ones(H, W, device=device)
        self.sky2[: H // 5] = 0
        self.occl1 = torch.ones(1, H, W, device=device)
        if impl == "native":
            import diff_gaussian_rasterization as dgr
            from r3dgw_b200 import image_ops
            from r3dgw_b200.losses import training_loss
            self.train_loss = training_loss
            self.raster = dgr
            self.shade_fn = R.fused_shade_fn(self.lut, 4)
            self.ops = image_ops
        else:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import ref_loader
            import shade_oracle as so
            self.raster = ref_loader.load_ref_dgr()
            lut = self.lut

            def ref_shade(pc, normals, env_sh, sky_sh, cam_pos, specular, fix_sky):
                # the reference's torch graph (light.py:131-193 + render():119-156), eager ops on the GPU
                return so.shade_all(env_sh, 4, lut, pc.means3D, normals, pc.albedo, pc.roughness, pc.metalness,
                                    pc.is_sky, cam_pos, sky_sh, specular, fix_sky)
            self.shade_fn = ref_shade
            import image_oracle as io

            class RefOps:
                """the reference's torch graphs for the image-space ops (utils/loss_utils.py, graphics_utils.py),
                restated in oracle/image_oracle.py - including their host-side `mask.sum() == 0` checks"""
                l1_loss = staticmethod(lambda x, gt, mask=None: io.masked_l1(x, gt, mask))
                ssim = staticmethod(lambda a, b, mask=None: io.ssim(a, b, mask))
                depth_to_normal = staticmethod(lambda view, d: io.depth_to_normal(
                    view.world_view_transform, int(view.image_width), int(view.image_height), view.FoVx, view.FoVy, d))
            self.ops = RefOps
        self.params = self.pc.leaves() + [self.env_sh, self.sky_sh]

    def raster_inputs(self):
        pc = self.pc
        return pc.means3D, pc.opacity, pc.scales, pc.rotations

    def forward(self, cam=None, sky_mask=None):
        cam = cam or self.cam
        d2n = self.ops.depth_to_normal if sky_mask is not None else None
        if self.mode == "fused" and self.impl == "native":
            return self.R.render_fused(cam, self.pc, self.env_sh, self.sky_sh, self.bg, self.shade_fn,
                                       sky_mask=sky_mask, depth_to_normal=d2n)
        return self.R.render(cam, self.pc, self.env_sh, self.sky_sh, self.raster, self.bg, self.shade_fn,
                             sky_mask=sky_mask, depth_to_normal=d2n)

    def loss(self, out, gt, cam, sky2, occl1):
        """train.py:78-112 in the post-15 000 regime: photometric (masked L1 + SSIM), sky-BRDF, normal-consistency,
        planar-scale and sky-depth regularisers (the env-SH positivity term, train.py:99-102, samples 10 random
        directions on a [10,25,3] tensor - negligible and stochastic - and is left out of both arms)."""
        torch, ops, pc = self.torch, self.ops, self.pc
        if self.impl == "native":
            # the whole loss as ONE autograd node over the fused kernels (r3dgw_b200.losses): same terms, same kernels
            # as the separate drop-ins, without ~8 Function calls and ~25 scalar torch kernels per step
            return self.train_loss(out, gt, occl1, sky2, pc.scales, pc.means3D, pc.sky_index()[0],
                                   cam.world_view_transform, LAMBDA)[0]
        gt3 = gt
        occl3, sky3 = occl1.expand_as(gt3), sky2.unsqueeze(0).expand_as(gt3)
        Ll1 = ops.l1_loss(out["render"], gt3, mask=occl3)
        Ssim = 1.0 - ops.ssim(out["render"], gt3, mask=occl3)
        loss = Ll1 * (1 - LAMBDA["dssim"]) + LAMBDA["dssim"] * Ssim
        loss_sky = ops.l1_loss(out["diffuse_color"], torch.zeros_like(out["diffuse_color"]), mask=1 - sky3) + \
            ops.l1_loss(out["specular_color"], torch.zeros_like(out["specular_color"]), mask=1 - sky3)
        loss = loss + LAMBDA["sky_brdf"] * loss_sky
        rn = out["normal"] * occl3 * sky3
        rs = out["normal_ref"] * occl3 * sky3
        loss = loss + LAMBDA["normal"] * (1 - (rn * rs).sum(dim=0))[None].mean()
        # min_scale_loss (utils/loss_utils.py:210-220) and depth_loss_gaussians (:140-148)
        # the reference's caller-side torch on [P] arrays
        vis = out["radii"] > 0
        is_sky = pc.is_sky.reshape(-1)
        wvt = cam.world_view_transform
        depth = pc.means3D @ wvt[:3, 2:3] + wvt[3, 2]
        if vis.sum() > 0:
            sc = pc.scales[vis & ~is_sky]
            loss = loss + LAMBDA["scale"] * torch.sort(sc, dim=-1)[0][..., 0].mean()
        avg_sky = torch.mean(depth[is_sky & vis])
        avg_non = torch.mean(depth[(~is_sky) & vis]).detach()
        return loss + LAMBDA["sky_gauss"] * torch.exp(-0.02 * (avg_sky - avg_non))

    def step(self, gt, cam=None, sky2=None, occl1=None, on_loss=None, first=True):
        # optimizer-style in-place touch of every parameter (x *= 1): bumps the tensors' version counters exactly as
        # an optimizer step does, so nothing computed for the previous iteration (geometry cache) can be reused.
        # first=False: a further view of the same step (gradients accumulate, parameters untouched)
        if first:
            with self.torch.no_grad():
                self.torch._foreach_mul_(self.params, 1.0)
            for p in self.params:
                p.grad = None
        sky2 = self.sky2 if sky2 is None else sky2
        occl1 = self.occl1 if occl1 is None else occl1
        out = self.forward(cam, sky2)
        loss = self.loss(out, gt, cam or self.cam, sky2, occl1)
        if on_loss is not None:
            on_loss(loss)      # e2e: the loss value starts its way to the host before the backward pass is enqueued
        loss.backward()
        self.last_out = out
        return loss

    def flat_grads(self):
        return self.torch.cat([p.grad.reshape(-1) for p in self.params if p.grad is not None])


class StockStepper:
    """One training iteration through the reference's UNMODIFIED caller code (oracle/_ref/py, installed by
    oracle/build_ref.py): gaussian_renderer.render() on a reference GaussianModel + the loss of train.py:78-112.
    impl == "reference": reference extension, reference EnvironmentLight (torch graph), reference utils.loss_utils -
    the stock path, nothing of this repo in it.  impl == "native": the same render() bound to the drop-in
    diff_gaussian_rasterization, the drop-in EnvironmentLight and the drop-in loss functions (what a user gets by
    swapping the modules and touching nothing else)."""

    def __init__(self, impl, sc, cam, device):
        import types
        import torch
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import stock_caller as sc_
        self.torch, self.sc_, self.impl, self.cam, self.device = torch, sc_, impl, cam, device
        self.mode = "stock"
        self.ns = sc_.load_ref_python()
        if impl == "reference":
            import ref_loader
            self.raster = ref_loader.load_ref_dgr()
            self.renderer = sc_.load_stock_renderer(self.raster, "stock_renderer_ref")
            self.fns = self.ns.loss_utils
        else:
            import diff_gaussian_rasterization as dgr
            from r3dgw_b200 import image_ops, regularisers
            self.raster = dgr
            self.renderer = sc_.load_stock_renderer(dgr, "stock_renderer_dropin")
            self.fns = types.SimpleNamespace(l1_loss=image_ops.l1_loss, ssim=image_ops.ssim,
                                             min_scale_loss=regularisers.min_scale_loss,
                                             depth_loss_gaussians=regularisers.depth_loss_gaussians)
        self.gm = sc_.make_ref_model(sc, device)
        self.env_sh = sc.env_sh.to(device).clone().requires_grad_(True)
        self.sky_sh = sc.sky_sh.to(device).reshape(1, 4, 3).clone().requires_grad_(True)
        if impl == "reference":
            self.env = sc_.make_ref_envlight(self.env_sh, 4)
        else:
            from r3dgw_b200.shading import EnvironmentLight
            self.env = EnvironmentLight(self.env_sh, 4, fg_lut=sc_.real_fg_lut(device))
        self.bg = torch.zeros(3, device=device)
        H, W = int(cam.image_height), int(cam.image_width)
        self.sky2 = torch.ones(H, W, device=device)
        self.sky2[: H // 5] = 0
        self.occl1 = torch.ones(1, H, W, device=device)
        self.params = sc_.model_leaves(self.gm) + [self.env_sh, self.sky_sh]

    def raster_inputs(self):
        gm = self.gm
        with self.torch.no_grad():
            return gm.get_xyz.detach(), gm.get_opacity.detach(), gm.get_scaling.detach(), gm.get_rotation.detach()

    def forward(self, cam=None, sky_mask=None):
        cam = cam or self.cam
        sky2 = self.sky2 if sky_mask is None else sky_mask
        view = self.sc_.make_view(cam, sky2.unsqueeze(0))
        return self.sc_.stock_render(self.renderer, view, self.gm, self.env, self.sky_sh, self.bg, debug=False)

    def step(self, gt, cam=None, sky2=None, occl1=None, on_loss=None, first=True):
        torch = self.torch
        if first:
            with torch.no_grad():
                torch._foreach_mul_(self.params, 1.0)
            for p in self.params:
                p.grad = None
        cam = cam or self.cam
        sky2 = self.sky2 if sky2 is None else sky2
        occl1 = self.occl1 if occl1 is None else occl1
        view = self.sc_.make_view(cam, sky2.unsqueeze(0), occl1, gt)
        self.env.set_base(self.env_sh)
        pkg = self.sc_.stock_render(self.renderer, view, self.gm, self.env, self.sky_sh, self.bg, debug=False)
        loss = self.sc_.train_loss(pkg, view, self.gm, self.fns, LAMBDA)
        if on_loss is not None:
            on_loss(loss)
        loss.backward()
        self.last_out = pkg
        return loss


def make_stepper(impl, mode, sc, cam, device):
    """reference arm: the stock caller when oracle/_ref/py is installed (else the mirror + oracle restatements, noted
    in the line); native arm: `stock` = stock caller on the drop-ins, `dropin` / `fused` = the repo's own render()."""
    stock_ok = False
    try:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import stock_caller
        stock_ok = stock_caller.available()
    except Exception:   # noqa: BLE001
        pass
    if stock_ok and (impl == "reference" or mode == "stock"):
        return StockStepper(impl, sc, cam, device)
    if mode == "stock":
        raise SystemExit("bench.py --mode stock: oracle/_ref/py is absent (run oracle/build_ref.py in the build container)")
    return Stepper(impl, mode, sc, cam, device)


# ---- timing helpers ---------------------------------------------------------------------------------------
def l2_flush(torch, buf):
    buf.zero_()          # 512 MiB write > 126 MB L2, outside every timed event pair


RANK0_SAMPLES = int(os.environ.get("RANK", "0")) == 0     # clocks are sampled by rank 0 only: 8 processes querying
# NVML in the same step serialise in the driver (r01: 85-100 ms single-step stalls at N = 8)
SAMPLE_EVERY_STEP = os.environ.get("R3DGW_BENCH_SAMPLE_EVERY", "0") == "1"      # diagnostics: clocks / reasons per step
LAST_RANK_STATS = []   # per rank [mean step ms, max single step ms] of the most recent timed_steps() call (N > 1)
LAST_STEP_MS = []      # per-step device times of the most recent timed_steps() call (diagnostics on stderr)


def timed_steps(torch, fn, steps, warmup, flush_buf, dist=None, sampler=None):
    """W untimed steps, then EXACTLY K steps each bracketed by CUDA events on the current stream (L2 flushed
    between event pairs); barrier + synchronize on both sides; returns total ms (MAX over ranks)."""
    import gc
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    # a generation-2 pass of Python's cyclic collector over the interpreter's ~10^5 tracked objects is a 60-150 ms
    # host stall (seen as single-step outliers): collect now, keep the collector off inside the timed region
    gc.collect()
    gc.disable()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for k, (a, b) in enumerate(evs):
        l2_flush(torch, flush_buf)
        a.record()
        fn()
        b.record()
        if sampler is not None and (SAMPLE_EVERY_STEP or (k % 4 == 1 and k // 4 < 6)):
            if RANK0_SAMPLES:
                sampler.sample()      # GPU still busy with the step just enqueued; outside every event pair
            if dist is not None:
                # An NVML query costs rank 0's host 3 ms as a rule and 150 ms now and then.  Without this barrier the
                # other ranks enqueue their next step at once and its event pair then spans the wait for rank 0 inside
                # the all-reduce: rank 0's own step list stays flat while every other rank reports one 9-150 ms step,
                # and the MAX over ranks charges the query to the job (N = 4 on an 8-GPU box: 12.8 instead of 5.45
                # ms/step).  With it every rank absorbs the query between its event pairs, like rank 0 does.
                dist.barrier()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    gc.enable()
    per_step = [a.elapsed_time(b) for a, b in evs]
    LAST_STEP_MS[:] = per_step
    total = sum(per_step)
    LAST_RANK_STATS[:] = []
    if dist is not None:
        t = torch.tensor([total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # per-rank (mean, max single step): the reported time is the MAX over ranks, so one rank's hiccup shows here
        mine = torch.tensor([total / max(steps, 1), max(per_step)], device="cuda", dtype=torch.float64)
        allr = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
        dist.all_gather(allr, mine)
        LAST_RANK_STATS[:] = [[round(float(x[0]), 3), round(float(x[1]), 3)] for x in allr]
        total = float(t.item())
    return total


def algorithmic_bytes(kernel, P, Pv, R, N, tiles, C, n_pass):
    """SURVEY.md 8(d) per-unit figures x the units one launch processes (also in DESIGN.md)."""
    table = {
        "preprocess_scan": 44 * P + 8 * P + 52 * Pv + 8 * P,                   # K1 + fused scan (B_pre + B_scan)
        "duplicate_keys": 20 * P + 12 * R,
        # depth-first binning: the per-Gaussian depth sort (keys kernel + histogram + 4 passes over P 8-byte pairs) is
        # work the reference does not have; the R-sized sort moves 32-bit tile ids + ids (16 B per pass)
        "depth_sort": (8 + 8 + 4 + 4 * 16) * P,
        "sort_histogram": 4 * R,
        "sort_onesweep_pass": 16 * R,                                          # per pass: read 8 + write 8
        "ranges_gather": 12 * R + 8 * tiles + 24 * R + 32 * R,                 # keys+ids, gather xy/conic, write stream
        "blend_forward": (32 + 4 * C) * R + (4 * C + 8) * N,
        "blend_backward": (32 + 4 * C) * R + (4 * C + 8) * N + (4 * C + 24) * Pv,
        # deferred drop-in backward: ONE geometry-only walk over the 5 calls' concatenated channels (C = 15 here)
        "blend_backward_geometry": (32 + 4 * C) * R + (4 * C + 8) * N + 24 * Pv,
        "geometry_backward": 160 * Pv,
        "shade_forward": 80 * P,
        "shade_backward": 100 * P,
        "ssim_l1_forward": (2 * 4 * C + 4 + 3 * 4 * C) * N,                     # img, gt, mask in; 3 derivative maps out
        "ssim_l1_backward": (3 * 4 * C + 2 * 4 * C + 4 + 4 * C) * N,            # 3 maps, img, gt, mask in; grad out
        "masked_l1": (4 * C + 4) * N,
        "depth_to_normal": 16 * N,
    }
    return table.get(kernel)


def alpha_cache_rows(st, dgr):
    """Number of 128-byte alpha-cache rows a replay kernel reads for the geometry state `st`: (list entry, warp) pairs
    whose footprint-cull bit is set and whose list position lies before that warp's deepest contributor.  Counted
    with torch ops from the state buffers (B stream cull bits, tile ranges, n_contrib) - bench bookkeeping only."""
    import ctypes
    import torch
    from r3dgw_b200 import native
    L = native.lib()
    BL, IL = native.BinningLayout(), native.ImageLayout()
    L.r3dgw_get_binning_layout(st.R, st.W, st.H, ctypes.byref(BL))
    L.r3dgw_get_image_layout(st.W, st.H, ctypes.byref(IL))
    R, W, H = st.R, st.W, st.H
    gx, gy = (W + 15) // 16, (H + 15) // 16
    ranges = st.binning[BL.ranges:BL.ranges + gx * gy * 8].view(torch.int32).view(-1, 2).long()
    Bz = st.binning[BL.splats + 16 * R:BL.splats + 32 * R].view(torch.int32).view(R, 4)[:, 2]
    ncon = st.alpha_img[IL.n_contrib:IL.n_contrib + 4 * W * H].view(torch.int32).view(H, W).long()
    pad = torch.zeros(gy * 16, gx * 16, dtype=torch.long, device=ncon.device)
    pad[:H, :W] = ncon
    # warp w of a tile covers columns (w&1)*8.., rows (w>>1)*4..  ->  [tiles, 8] deepest contributor per warp
    wl = pad.view(gy, 4, 4, gx, 2, 8).amax(dim=(2, 5)).permute(0, 2, 1, 3).reshape(gy * gx, 8)
    lens = (ranges[:, 1] - ranges[:, 0]).clamp_min(0)
    tile_of = torch.repeat_interleave(torch.arange(gx * gy, device=ncon.device), lens)
    local = torch.arange(R, device=ncon.device) - ranges[tile_of, 0]
    rows = 0
    for w in range(8):
        rows += int((((Bz >> w) & 1).bool() & (local < wl[tile_of, w])).sum().item())
    return rows


def cpu_baseline(workload, budget_s=20.0):
    """The CPU port of the reference path (oracle/, OpenMP over all host cores) on a bounded sample of the workload:
    the first P/f Gaussians at full resolution, one raster fwd+bwd (C=3) + torch-CPU shading fwd+bwd, extrapolated to
    the full step (x f for the per-Gaussian / per-duplicate work; 6 raster forwards and 5 backwards per step, no
    geometry reuse - the reference recomputes everything on every call)."""
    import numpy as np
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import raster_oracle as ro
    import shade_oracle as so
    from r3dgw_b200.synthetic import make_scene
    cores = os.cpu_count() or 1
    ro.set_threads(cores)
    torch.set_num_threads(cores)
    sc = make_scene(workload)
    c = sc.camera
    f = 2 if sc.P >= 500_000 else 1
    Ps = sc.P // f
    a = lambda t: t[:Ps].numpy()   # noqa: E731
    t0 = time.perf_counter()
    fw = ro.forward(a(sc.means3D), a(sc.scales), a(sc.rotations), a(sc.opacities), a(sc.colors),
                    c.world_view_transform.numpy(), c.full_proj_transform.numpy(), c.image_width, c.image_height,
                    c.tanfovx, c.tanfovy, sc.bg.numpy())
    t_fwd = time.perf_counter() - t0
    dL = np.full((3, c.image_height, c.image_width), 1e-3, np.float32)
    t0 = time.perf_counter()
    ro.backward(fw, dL)
    t_bwd = time.perf_counter() - t0
    # shading, torch on CPU (the reference's own CPU-executable piece), on the same sample
    fg = ~sc.is_sky[:Ps]
    nfg = int(fg.sum())
    leaves = [x.clone().requires_grad_(True) for x in (sc.means3D[:Ps], sc.albedo[:nfg], sc.roughness[:nfg],
                                                       sc.metalness[:nfg], sc.env_sh, sc.sky_sh)]
    nrm = torch.nn.functional.normalize(sc.means3D[:Ps] * -1.0 + 0.1, dim=1)
    lut = so.synthetic_fg_lut()
    t0 = time.perf_counter()
    cols = so.shade_all(leaves[4], 4, lut, leaves[0], nrm, leaves[1], leaves[2], leaves[3], sc.is_sky[:Ps],
                        c.camera_center, leaves[5], True, False)
    sum(x.sum() for x in cols).backward()
    t_shade = time.perf_counter() - t0
    step_s = f * (6 * t_fwd + 5 * t_bwd + t_shade)
    ro.set_threads(1)
    return {"value": 1.0 / step_s, "unit": UNIT, "cores": cores, "kind": "port", "extrapolated": True,
            "sample": f"first {Ps} of {sc.P} Gaussians at {c.image_width}x{c.image_height}: 1 raster fwd {t_fwd:.2f}s + "
                      f"bwd {t_bwd:.2f}s (oracle/raster_oracle.c, OpenMP) + torch-CPU shading fwd+bwd {t_shade:.2f}s; "
                      f"extrapolated x{f} and to 6 fwd + 5 bwd per step",
            "measured_s": round(t_fwd + t_bwd + t_shade, 3)}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference" and rank != 0:
        return 0                     # the reference is single-GPU: rank 0 alone runs it
    import torch
    if not torch.cuda.is_available():
        if args.impl == "reference":
            cb = cpu_baseline(args.workload)
            print(json.dumps({"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT,
                              "n_gpus": 0, "steps": 1, "warmup": 0, "higher_is_better": True, "cpu_baseline": cb,
                              "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0,
                                      "d2h_bytes_per_step": 0}, "note": "no CUDA device: CPU port only"}))
            return 0
        raise SystemExit("bench.py: no CUDA device - the product path is CUDA only (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1 and args.impl == "native":
        import torch.distributed as dist_mod
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"       # NCCL prints its version banner on stdout: keep stdout to the one JSON line
        dist_mod.init_process_group("nccl", device_id=device)
        dist = dist_mod

    from r3dgw_b200 import native
    from r3dgw_b200.synthetic import CONFIGS, Camera
    L = native.lib() if args.impl == "native" else None
    ref_ok = True
    if args.impl == "reference":
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import ref_loader
        ref_ok = ref_loader.available()
        if not ref_ok:
            cb = cpu_baseline(args.workload)
            print(json.dumps({"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": 1,
                              "steps": 1, "warmup": 0, "ms_per_step": 1e3 / cb["value"], "higher_is_better": True,
                              "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                              "config": {"workload": args.workload}, "cpu_baseline": cb,
                              "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0,
                                      "d2h_bytes_per_step": 0},
                              "note": "oracle/_ref (reference CUDA extension) not built: CPU port of the path timed"}))
            return 0

    sc, cam, gt_cpu = build_workload(args.workload, rank, device)
    cfg = CONFIGS[args.workload]
    W, H, P = cfg["W"], cfg["H"], cfg["P"]
    st = make_stepper(args.impl, args.mode, sc, cam, device)
    gt = gt_cpu.to(device)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=device)

    from r3dgw_b200.parallel import GradBucket
    bucket = GradBucket(st.params) if dist is not None else None

    def step_resident():
        st.step(gt)
        if bucket is not None:
            bucket.all_reduce(average_over=world)   # gradient exchange (64 B per Gaussian) over NVLink, ONE all-reduce

    if args.views > 0:
        # config C: V views per step, view-sharded: rank r renders views r, r + world, ...
        from r3dgw_b200.parallel import sync_densification_stats, views_for_rank
        from r3dgw_b200.synthetic import arc_cameras
        V = args.views
        vcams = [c.to(device) for c in arc_cameras(W, H, V, centroid_z=8.0, step_deg=2.0)]
        mine = views_for_rank(V, rank, world)
        vgts = {v: torch.rand(3, H, W, generator=torch.Generator().manual_seed(777 + v)).to(device) for v in mine}
        if bucket is None:
            bucket = GradBucket(st.params)
        dens = {"accum": torch.zeros(P, 1, device=device), "denom": torch.zeros(P, 1, device=device),
                "max_radii": torch.zeros(P, device=device)}

        def step_resident():   # noqa: F811
            for j, v in enumerate(mine):
                st.step(vgts[v], vcams[v], first=(j == 0))
                out = st.last_out
                with torch.no_grad():
                    # add_densification_stats (scene/gaussian_model.py:627-629) + max_radii2D (train.py:130), mask-multiplied
                    vis = out["radii"] > 0
                    g2 = out["viewspace_points"].grad
                    dens["accum"].add_(g2[:, :2].norm(dim=-1, keepdim=True) * vis.unsqueeze(1))
                    dens["denom"].add_(vis.unsqueeze(1))
                    torch.maximum(dens["max_radii"], out["radii"] * vis, out=dens["max_radii"])
            bucket.all_reduce(average_over=V)
            sync_densification_stats(dens["accum"], dens["denom"], dens["max_radii"])

    sampler = ClockSampler(local_rank) if os.environ.get("R3DGW_BENCH_NO_SAMPLER", "0") != "1" else None
    # setup (not part of W): iterations so that torch's caching allocator has grown its pools (the first ones
    # otherwise pay cudaMalloc, a device-wide sync each, inside the step) and so that the timed region starts in
    # the sustained regime: on this pool the first ~100 ms of back-to-back steps run ~7 % faster than steady state
    for _ in range(25):
        step_resident()
    torch.cuda.synchronize()
    for _ in range(args.warmup):
        step_resident()
    torch.cuda.synchronize()
    launches0 = L.r3dgw_launch_count() if L else 0
    ms0 = torch.cuda.memory_stats(device)
    total_ms = timed_steps(torch, step_resident, args.steps, 0, flush, dist, sampler)
    ms1 = torch.cuda.memory_stats(device)
    allocator = {"cudaMalloc_calls_in_timed_region": int(ms1.get("num_device_alloc", 0) - ms0.get("num_device_alloc", 0)),
                 "cudaFree_calls_in_timed_region": int(ms1.get("num_device_free", 0) - ms0.get("num_device_free", 0)),
                 "alloc_retries": int(ms1.get("num_alloc_retries", 0)),
                 "reserved_gb": round(ms1.get("reserved_bytes.all.current", 0) / 2**30, 2),
                 "peak_allocated_gb": round(ms1.get("allocated_bytes.all.peak", 0) / 2**30, 2)}
    rank_stats = list(LAST_RANK_STATS)
    if rank == 0:
        print("step ms (resident):", [round(x, 2) for x in LAST_STEP_MS], file=sys.stderr)
        if rank_stats:
            print("per-rank [mean, max] step ms (resident):", rank_stats, file=sys.stderr)
        if SAMPLE_EVERY_STEP and sampler is not None:
            print("per-step (sm_mhz, power_w, reasons):", [(int(a), int(b), hex(c)) for a, b, c in sampler.samples], file=sys.stderr)
    clocks = sampler.summary() if sampler is not None else {"sm_mhz": None, "reasons": ["sampler disabled (R3DGW_BENCH_NO_SAMPLER=1)"]}
    gpu_launches = (L.r3dgw_launch_count() - launches0) if L else 0
    ms_per_step = total_ms / args.steps
    value = world * 1e3 / ms_per_step if args.impl == "native" else 1e3 / ms_per_step
    if args.views > 0:
        value = 1e3 / ms_per_step          # strong scaling: a step is V views whatever the rank count
        if rank == 0:
            print(json.dumps({
                "metric": "train_steps_per_s_8views_fwd+bwd_1M_gaussians_1080p", "value": value, "unit": "steps/s",
                "views_per_s": value * args.views, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"C: {P} Gaussians, {W}x{H}, {args.views} views/step sharded over {world} rank(s) "
                                       f"(rank r renders views r, r+N, ...), gradient accumulation, one all-reduce of the "
                                       f"parameter gradients (ReduceOp.SUM + 1/V) and of the densification statistics per step",
                           "mode": args.mode, "l2": "512 MiB flush between timed steps"},
                "clocks": clocks, "gpu_launches": int(gpu_launches), "impl": args.impl, "allocator": allocator,
                "step_ms": [round(x, 2) for x in LAST_STEP_MS]}))
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- e2e: host buffers in, loss out, through the public API ----------------------------------------
    pin = lambda t: t.contiguous().pin_memory()   # noqa: E731
    # the GT image travels as 8-bit RGB, the way image datasets are stored; it is converted to float on the device
    h_gt = pin((gt_cpu * 255.0).round().to(torch.uint8))
    h_view, h_proj = pin(cam.world_view_transform.cpu()), pin(cam.full_proj_transform.cpu())
    h_cpos, h_env, h_sky = pin(cam.camera_center.cpu()), pin(sc.env_sh), pin(sc.sky_sh)
    d_gt8 = torch.empty(gt.shape, dtype=torch.uint8, device=device)
    d_view, d_proj = torch.empty(4, 4, device=device), torch.empty(4, 4, device=device)
    d_cpos = torch.empty(3, device=device)
    h_loss = torch.zeros(1).pin_memory()
    # per-view masks travel as 8-bit images too (viewpoint_cam.sky_mask / occluders_mask, train.py:64-65)
    h_skym, h_occl = pin(st.sky2.cpu().to(torch.uint8)), pin(st.occl1.cpu().to(torch.uint8))
    d_skym8, d_occl8 = torch.empty_like(st.sky2, dtype=torch.uint8), torch.empty_like(st.occl1, dtype=torch.uint8)
    h2d = sum(t.numel() * t.element_size() for t in (h_gt, h_view, h_proj, h_cpos, h_env, h_sky, h_skym, h_occl))
    # measured pinned H2D bandwidth of this box (reported beside e2e: some hosts deliver a fraction of PCIe 5)
    probe_h = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
    probe_d = torch.empty(64 << 20, dtype=torch.uint8, device=device)
    probe_d.copy_(probe_h, non_blocking=True); torch.cuda.synchronize()
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ea.record(); probe_d.copy_(probe_h, non_blocking=True); eb.record(); torch.cuda.synchronize()
    h2d_gbps = (64 << 20) / (ea.elapsed_time(eb) * 1e-3) / 1e9
    del probe_h, probe_d

    loss_ev = torch.cuda.Event()

    def send_loss(loss):
        h_loss.copy_(loss.detach().reshape(1), non_blocking=True)
        loss_ev.record()

    # Input pipeline: two device-side input slots; the per-view inputs of step i+1 (GT image, masks, camera) travel on
    # a copy stream while step i computes, like a prefetching data loader.  Each copy is enqueued at the START of the
    # step whose event pair times it and is fenced behind an event recorded there, so every step's H2D traffic lies
    # inside a timed pair (never in the untimed L2-flush gap).
    copy_stream = torch.cuda.Stream()
    slots = [dict(gt8=torch.empty_like(d_gt8), view=torch.empty_like(d_view), proj=torch.empty_like(d_proj),
                  cpos=torch.empty_like(d_cpos), skym8=torch.empty_like(d_skym8), occl8=torch.empty_like(d_occl8))
             for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    released = [torch.cuda.Event(), torch.cuda.Event()]
    pipe = {"k": 0}

    def enqueue_inputs(slot):
        here = torch.cuda.Event()
        here.record()
        copy_stream.wait_event(here)              # not earlier than this point of the compute stream
        copy_stream.wait_event(released[slot])    # the step that last read this slot is done with it
        with torch.cuda.stream(copy_stream):
            b = slots[slot]
            b["gt8"].copy_(h_gt, non_blocking=True)
            b["view"].copy_(h_view, non_blocking=True)
            b["proj"].copy_(h_proj, non_blocking=True)
            b["cpos"].copy_(h_cpos, non_blocking=True)
            b["skym8"].copy_(h_skym, non_blocking=True)
            b["occl8"].copy_(h_occl, non_blocking=True)
            ready[slot].record(copy_stream)

    def step_e2e():
        slot = pipe["k"] & 1
        pipe["k"] += 1
        if pipe["k"] == 1:
            enqueue_inputs(slot)                  # very first call only: nothing was prefetched yet
        enqueue_inputs(slot ^ 1)                  # the next step's inputs
        cur = torch.cuda.current_stream()
        cur.wait_event(ready[slot])
        b = slots[slot]
        d_gt = b["gt8"].to(torch.float32).mul_(1.0 / 255.0)
        with torch.no_grad():                     # the lighting parameters (25x3 + 9x3 floats) ride on the compute stream
            st.env_sh.copy_(h_env, non_blocking=True)
            st.sky_sh.copy_(h_sky, non_blocking=True)
        c = Camera(W, H, cam.FoVx, cam.FoVy, b["view"], b["proj"], b["cpos"])
        st.step(d_gt, c, b["skym8"].to(torch.float32), b["occl8"].to(torch.float32), on_loss=send_loss)
        released[slot].record()
        if bucket is not None:
            bucket.all_reduce(average_over=world)
        # the user reads the loss every step: its 4-byte D2H copy was enqueued right after the forward pass, so the
        # host waits for that copy only and the backward pass keeps running while the next step is being enqueued
        loss_ev.synchronize()
        return float(h_loss[0])

    for _ in range(4):          # setup: pinned-copy path and allocator pools for the e2e variant of the step
        step_e2e()
    e2e_ms = timed_steps(torch, step_e2e, args.steps, max(3, args.warmup), flush, dist) / args.steps
    e2e_value = (world if args.impl == "native" else 1) * 1e3 / e2e_ms

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world if args.impl == "native" else 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {P} Gaussians, {W}x{H}, 1 view/GPU/step, shade + 6 raster fwd + "
                                   f"5 raster bwd + masked L1/SSIM, sky-BRDF, normal (depth_to_normal), scale and sky-depth "
                                   f"regularisers as train.py:74-116 ({args.mode})", "l2": "512 MiB flush between timed steps",
                       "host": "Python cyclic GC collected before and disabled inside every timed region",
                       "mode": args.mode, "views_per_step": world,
                       "caller": ("reference render() + train.py loss, unmodified (oracle/_ref/py)" if isinstance(st, StockStepper)
                                  else "r3dgw_b200.render (the repo's render() / render_fused(), reference call pattern)")},
            "clocks": clocks, "gpu_launches": int(gpu_launches), "allocator": allocator,
            "per_rank_mean_max_step_ms": rank_stats,
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": 4, "h2d_gbps_measured": round(h2d_gbps, 2),
                    "loss_read": "this step's loss is copied D2H right after its forward pass and read on the host "
                                 "before the next step is enqueued (no full-stream sync per step)",
                    "input_pipeline": "2 input slots; step i+1's GT image / masks / camera are copied on a side stream "
                                      "while step i computes, fenced inside step i's timed event pair"}}
    if args.impl == "reference":
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": value, "unit": UNIT, "cores": 0, "kind": "reference",
                                "sample": "the reference has no CPU rasterizer: its UNMODIFIED CUDA extension "
                                          "(oracle/_ref, built from /root/reference for sm_100a) driven by its own "
                                          "unmodified render() / GaussianModel / EnvironmentLight / loss_utils, same "
                                          "B200, same steps; CPU port figure is in the native arm's line"}
        line["e2e"]["note"] = "same host-buffer protocol as the native arm, through the reference's own code"
        if not isinstance(st, StockStepper):
            line["config"]["caller"] = ("oracle/_ref/py absent: the repo's render() mirror + oracle/ restatements of the "
                                        "reference torch graphs around the reference extension")

    if args.impl == "reference" and isinstance(st, StockStepper):
        # attribution only: the stock render() calls torch.cuda.empty_cache() on every frame
        # (gaussian_renderer/__init__.py:237), so every step hands its cached blocks back to the driver and cudaMallocs
        # them again (see `allocator`); the same loop with that one call patched out is the reference's most favourable
        # variant - the ratio quoted against it is the conservative one
        real_empty = torch.cuda.empty_cache
        try:
            torch.cuda.empty_cache = lambda: None
            ne = max(5, args.steps // 2)
            line["reference_without_empty_cache"] = {"ms_per_step": timed_steps(torch, step_resident, ne, 3, flush, None) / ne,
                                                     "note": "NOT the stock path: render()'s empty_cache() patched to a no-op"}
        finally:
            torch.cuda.empty_cache = real_empty
    if args.quick:
        if rank == 0:
            line["step_ms"] = [round(x, 2) for x in LAST_STEP_MS]
            print(json.dumps(line))
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- extras: fwd-only frames/s (U3), single raster call fwd / fwd+bwd (U1) ---------------------------
    def frame(new_view=True):
        if new_view and args.impl == "native":
            st.raster.clear_geometry_cache()     # a new view every frame: nothing carried over
        with torch.no_grad():
            st.forward(None, st.sky2)
    nf = max(5, args.steps // 2)
    scale_n = world if args.impl == "native" else 1
    line["fps_fwd"] = scale_n * 1e3 / (timed_steps(torch, frame, nf, 4, flush, dist) / nf)
    print("fps_fwd per-frame ms:", [round(x, 2) for x in LAST_STEP_MS], file=sys.stderr)
    # relit-eval grid cell (config 4): same view, new environment SH each frame -> geometry (K1..K3) is reused
    line["fps_fwd_relit_same_view"] = scale_n * 1e3 / (timed_steps(torch, lambda: frame(False), nf, 2, flush, dist) / nf)
    s = st.raster.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, st.bg, 1.0, cam.world_view_transform,
                                                cam.full_proj_transform, -1, cam.camera_center, False)
    col = sc.colors.to(device).requires_grad_(True)
    dL = torch.full((3, H, W), 1e-3, device=device)
    u_m3, u_op, u_sc, u_rot = [t.detach().clone().requires_grad_(True) for t in st.raster_inputs()]

    def u1(bwd):
        if args.impl == "native":
            st.raster.clear_geometry_cache()
        m2 = torch.zeros_like(u_m3, requires_grad=True)
        with torch.set_grad_enabled(bwd):
            img, _ = st.raster.GaussianRasterizer(s)(means3D=u_m3, means2D=m2, opacities=u_op,
                                                     colors_precomp=col, scales=u_sc, rotations=u_rot)
        if bwd:
            img.backward(dL)
    n1 = max(5, args.steps // 2)
    line["u1_raster_call"] = {"fwd_ms": timed_steps(torch, lambda: u1(False), n1, 4, flush, None) / n1}
    print("u1 fwd per-call ms:", [round(x, 2) for x in LAST_STEP_MS], file=sys.stderr)
    line["u1_raster_call"]["fwd_bwd_ms"] = timed_steps(torch, lambda: u1(True), n1, 4, flush, None) / n1

    # ---- the single-pass alternative API (render_fused: one 16-channel K4/K5), same step otherwise ------------
    if args.impl == "native" and args.mode == "dropin":
        st_f = Stepper("native", "fused", sc, cam, device)

        def step_fused():
            st_f.step(gt)
            if dist is not None:
                dist.all_reduce(st_f.flat_grads())
        nfz = max(5, args.steps // 2)
        ms_f = timed_steps(torch, step_fused, nfz, 8, flush, dist) / nfz
        print("fused step ms:", [round(x, 2) for x in LAST_STEP_MS], file=sys.stderr)
        line["fused_single_pass"] = {"value": world * 1e3 / ms_f, "unit": UNIT, "ms_per_step": ms_f,
                                     "api": "r3dgw_b200.render.render_fused (SURVEY 8f N1/N2), same loss and gradients"}
        del st_f
        # the reference's UNMODIFIED render() + train.py loss on top of the drop-in modules (module swap, nothing else)
        try:
            st_s = make_stepper("native", "stock", sc, cam, device)
            ms_s = timed_steps(torch, lambda: st_s.step(gt), nfz, 4, flush, None) / nfz
            line["stock_render_on_dropins"] = {"value": 1e3 / ms_s, "unit": UNIT, "ms_per_step": ms_s, "n_gpus": 1,
                                               "api": "reference gaussian_renderer.render() + GaussianModel getters, "
                                                      "drop-in rasterizer / EnvironmentLight / loss functions"}
            # attribution only: the same with render()'s torch.cuda.empty_cache() (gaussian_renderer/__init__.py:237)
            # patched out - every step otherwise hands all cached blocks back to the driver and re-cudaMallocs them
            real_empty = torch.cuda.empty_cache
            try:
                torch.cuda.empty_cache = lambda: None
                ms_n = timed_steps(torch, lambda: st_s.step(gt), nfz, 4, flush, None) / nfz
            finally:
                torch.cuda.empty_cache = real_empty
            line["stock_render_on_dropins"]["ms_per_step_without_empty_cache"] = ms_n
            del st_s
        except (SystemExit, Exception) as e:   # noqa: BLE001
            line["stock_render_on_dropins"] = {"unavailable": str(e)[:200]}

    # ---- per-kernel attribution + roofline of the dominant kernel (native arm) ---------------------------
    if args.impl == "native" and not args.no_profile:
        import diff_gaussian_rasterization as dgr
        st.step(gt)
        gs = dgr.last_geometry_state()
        Rn, Pv = int(gs.R), int((gs.radii > 0).sum().item())
        N, tiles = W * H, ((W + 15) // 16) * ((H + 15) // 16)
        n_pass = (int(tiles).bit_length() + 7) // 8      # passes of the R-sized tile-id sort
        C = 16 if args.mode == "fused" else 3
        torch.cuda.synchronize()
        native.profile_read(reset=True)
        native.profile_enable(True)
        nprof = max(3, min(args.steps, 10))
        for _ in range(nprof):
            l2_flush(torch, flush)
            st.step(gt)
        torch.cuda.synchronize()
        prof = native.profile_read(reset=True)
        native.profile_enable(False)
        step_kernel_ms = sum(ms for ms, _ in prof.values()) / nprof
        kern = {k: {"ms_per_step": ms / nprof, "launches_per_step": n / nprof, "avg_launch_ms": ms / max(n, 1)}
                for k, (ms, n) in prof.items() if n}
        # kernels that replay the alpha cache have no closed-form byte count (rows surviving the footprint cull are
        # data dependent): they are listed in `kernels` but the roofline object is for the top kernel with one
        gs_now = dgr.last_geometry_state()
        rows = alpha_cache_rows(gs_now, dgr) if (gs_now is not None and gs_now.alpha_cache is not None) else 0
        extra = {   # replay kernels: cache rows (128 B each) + B stream + per-pixel inputs / outputs (+ reduced gradient)
            "blend_forward_cached": 128 * rows + (16 + 4 * 3) * Rn + (4 * 3 + 8) * N,
            "blend_backward_color_cached": 128 * rows + 16 * Rn + (4 * 3 + 4) * N + 4 * 3 * Pv,
        }
        _ab = algorithmic_bytes
        algorithmic_bytes_x = lambda k, *a: extra[k] if (k in extra and rows) else _ab(k, *a)   # noqa: E731
        known = [k for k in kern if algorithmic_bytes_x(k, P, Pv, Rn, N, tiles, 3, n_pass) is not None]
        top = max(known, key=lambda k: kern[k]["ms_per_step"])
        if top == "blend_backward_geometry":
            C = 15
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:   # noqa: BLE001
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        ab = algorithmic_bytes_x(top, P, Pv, Rn, N, tiles, C, n_pass)
        ach = ab / (kern[top]["avg_launch_ms"] * 1e-3) / 1e9
        traffic = None       # dram__bytes_read+write per launch from the committed `ncu --set full` capture (C = 3, config B)
        try:
            if args.workload == "B":
                traffic = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")))[top]["dram_traffic_bytes_per_launch"]
        except Exception:   # noqa: BLE001
            pass
        # SURVEY 8(d)'s closed forms for the REFERENCE algorithm's compulsory bytes of the same stage: a replay kernel reads
        # its alphas from HBM instead of recomputing them, so the bytes it must move (above) exceed these
        survey = {"blend_forward_cached": (28 + 4 * 3) * Rn + (4 * 3 + 8) * N,
                  "blend_backward_color_cached": (28 + 4 * 3) * Rn + (4 * 3 + 8) * N + (4 * 3 + 24) * Pv}
        ab_survey = survey.get(top, _ab(top, P, Pv, Rn, N, tiles, C, n_pass))
        line["roofline"] = {"bound": "hbm", "kernel": top, "achieved": ach, "peak": peak, "unit": "GB/s",
                            "frac": ach / peak, "traffic": traffic,
                            "frac_algorithmic": ab_survey / (kern[top]["avg_launch_ms"] * 1e-3) / 1e9 / peak,
                            "survey_8d_bytes_per_launch": ab_survey,
                            "traffic_over_algorithmic": (traffic / ab_survey) if traffic else ab / ab_survey,
                            "frac_definition": "frac: bytes this kernel must move (alpha-cache rows it replays + streams) / "
                                               "time / peak; frac_algorithmic: SURVEY 8(d) compulsory bytes of the "
                                               "reference stage it replaces / time / peak",
                            "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650",
                            "algorithmic_bytes_per_launch": ab, "avg_launch_ms": kern[top]["avg_launch_ms"],
                            "share_of_step_kernel_time": kern[top]["ms_per_step"] / step_kernel_ms}
        # whole forward U1 (K1..K4, C=3) against the HBM roofline - north_star's "render fwd" figure
        # (a) SURVEY 8(d)'s closed form for the reference algorithm (one 6-pass sort of 64-bit keys): comparable
        #     across rounds and designs;  (b) the bytes THIS pipeline's forward kernels must move (depth-first binning)
        fwd_bytes = 132 * P + (200 + 4 * 3) * Rn + (4 * 3 + 8) * N + 8 * tiles
        own_bytes = sum(algorithmic_bytes(k, P, Pv, Rn, N, tiles, 3, n_pass) * (n_pass if k == "sort_onesweep_pass" else 1)
                        for k in ("preprocess_scan", "depth_sort", "duplicate_keys", "sort_histogram",
                                  "sort_onesweep_pass", "ranges_gather", "blend_forward"))
        fwd_s = line["u1_raster_call"]["fwd_ms"] * 1e-3
        fwd_gbs = fwd_bytes / fwd_s / 1e9
        line["roofline_fwd_call"] = {"achieved": fwd_gbs, "peak": peak, "frac": fwd_gbs / peak, "unit": "GB/s",
                                     "algorithmic_bytes": fwd_bytes, "bytes_definition": "SURVEY 8d B_fwd(C=3), reference algorithm",
                                     "own_pipeline_bytes": own_bytes, "own_frac": own_bytes / fwd_s / 1e9 / peak}
        line["kernels"] = {k: {a: round(b, 5) for a, b in v.items()} for k, v in kern.items()}
        line["scene"] = {"P": P, "P_visible": Pv, "R": Rn, "pixels": N, "tiles": tiles, "sort_passes": n_pass,
                         "alpha_cache_rows": rows}
        # every kernel of the step against the HBM roofline (algorithmic bytes / measured launch time / peak)
        line["roofline_all"] = {k: round(algorithmic_bytes_x(k, P, Pv, Rn, N, tiles, 15 if k == "blend_backward_geometry" else C, n_pass)
                                         / (kern[k]["avg_launch_ms"] * 1e-3) / 1e9 / peak, 4) for k in known}
        # (masked_l1 / ssim rows count each image once; a launch that re-reads L2-resident images can exceed 1)

    if args.impl == "native" and rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.workload)

    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
