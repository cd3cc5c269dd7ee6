"""Copy the round-end pass (scripts/gpu_final.sh -> gpurun_out/) into profiles/ and regenerate the ncu summaries and the
measured table of BASELINE.md.  usage: python scripts/refresh_profiles.py"""
import json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
for src, dst in (("bench_native.json", "r01_bench_native.json"), ("bench_fused.json", "r01_bench_fused.json"),
                 ("bench_ref.json", "r01_bench_reference.json"), ("launches.csv", "r01_launches_bench_dropin.csv"),
                 ("pytest_gpu.log", "r01_pytest_gpu.log")):
    shutil.copy(os.path.join(G, src), os.path.join(P, dst))
subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_summarize.py"), os.path.join(G, "prof_u2_blend.ncu-rep"),
                "/tmp/ncu_sum.txt", "/tmp/ncu_traffic.json"], check=True, stdout=subprocess.DEVNULL)
with open(os.path.join(P, "r01_ncu_blend_summary.txt"), "w") as f:
    f.write("# ncu --set full --clock-control none, scripts/prof_u2.py 3 (U2 training step, drop-in mode, config B), B200, round 1\n"
            "# source: gpurun_out/prof_u2_blend.ncu-rep, launches of the 3rd step (-k regex:blend_ -s 24 -c 12); one launch per "
            "kernel shown (scripts/ncu_summarize.py)\n" + open("/tmp/ncu_sum.txt").read())
t = json.load(open("/tmp/ncu_traffic.json"))
old = json.load(open(os.path.join(P, "r01_ncu_traffic.json")))
def key(k):
    if k.startswith("blend_forward_kernel"): return "blend_forward_record" if k.rstrip(">").endswith("1") else "blend_forward"
    if k.startswith("blend_forward_cached"): return "blend_forward_cached"
    if k.startswith("blend_backward_color_cached"): return "blend_backward_color_cached"
    if k.startswith("blend_backward_kernel"): return "blend_backward_geometry" if ", 0," in k else "blend_backward"
    return k
for k, v in t.items():
    old[key(k)] = {"dram_traffic_bytes_per_launch": v["dram_traffic_bytes_per_launch"], "duration_ms": v["duration_ms"], "kernel": k}
json.dump(old, open(os.path.join(P, "r01_ncu_traffic.json"), "w"), indent=1)

nat, fus, ref = (json.loads(open(os.path.join(P, f"r01_bench_{n}.json")).read().strip().splitlines()[-1])
                 for n in ("native", "fused", "reference"))
s = open(os.path.join(ROOT, "BASELINE.md")).read()
def row(prefix, new):
    global s
    i = s.index(prefix); j = s.index("\n", i)
    s = s[:i] + new + s[j:]
r = lambda x, n=1: f"{x:.{n}f}"
row("| U1 raster call fwd |", f"| U1 raster call fwd | B: 1 M @1080p, C=3 | {r(ref['u1_raster_call']['fwd_ms'],2)} ms | {r(nat['u1_raster_call']['fwd_ms'],2)} ms | {r(ref['u1_raster_call']['fwd_ms']/nat['u1_raster_call']['fwd_ms'])}x |")
row("| U1 raster call fwd+bwd |", f"| U1 raster call fwd+bwd | B | {r(ref['u1_raster_call']['fwd_bwd_ms'],2)} ms | {r(nat['u1_raster_call']['fwd_bwd_ms'],2)} ms | {r(ref['u1_raster_call']['fwd_bwd_ms']/nat['u1_raster_call']['fwd_bwd_ms'])}x |")
row("| **U2 train step**", f"| **U2 train step** as `train.py:74-116` (shade + 6 raster fwd + image-space tail + masked L1/SSIM, sky-BRDF, normal, scale, sky-depth losses + 5 raster bwd), drop-in API, inputs resident | B | {r(ref['ms_per_step'])} ms = {r(ref['value'])} iters/s | {r(nat['ms_per_step'],2)} ms = **{r(nat['value'])} iters/s** | **{r(nat['value']/ref['value'])}x** |")
row("| same step end to end", f"| same step end to end (GT image + masks + camera + SH from pinned host memory each step, loss read back) | B | {r(ref['e2e']['ms_per_step'])} ms = {r(ref['e2e']['value'])} iters/s | {r(nat['e2e']['ms_per_step'],2)} ms = {r(nat['e2e']['value'])} iters/s | {r(nat['e2e']['value']/ref['e2e']['value'])}x |")
row("| same step, single-pass", f"| same step, single-pass 16-channel API (`render_fused`) | B | - | {r(fus['ms_per_step'],2)} ms = {r(fus['value'])} iters/s | {r(fus['value']/ref['value'])}x |")
row("| U3 frame (", f"| U3 frame (`render()` + image-space tail under no_grad, new view each frame) | B | {r(1e3/ref['fps_fwd'])} ms = {r(ref['fps_fwd'])} fps | {r(1e3/nat['fps_fwd'],2)} ms = {r(nat['fps_fwd'],0)} fps | {r(nat['fps_fwd']/ref['fps_fwd'])}x |")
row("| U3 relit-eval cell", f"| U3 relit-eval cell (same view, new envmap; alpha cache replayed) | B | {r(ref['fps_fwd_relit_same_view'])} fps | {r(nat['fps_fwd_relit_same_view'],0)} fps ({r(fus['fps_fwd_relit_same_view'],0)} fps through `render_fused`) | {r(nat['fps_fwd_relit_same_view']/ref['fps_fwd_relit_same_view'])}x |")
open(os.path.join(ROOT, "BASELINE.md"), "w").write(s)
print("native", r(nat["value"]), "it/s", r(nat["ms_per_step"], 2), "ms | e2e", r(nat["e2e"]["value"]), "| fused", r(fus["value"]),
      "| ref", r(ref["value"]), "| roofline", nat["roofline"]["kernel"], r(nat["roofline"]["frac"], 3))
