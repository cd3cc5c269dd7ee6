"""GPU-busy vs wall for the bench step via torch.profiler (CUPTI): where the non-library time goes (GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
from torch.profiler import profile, ProfilerActivity
import bench

mode = sys.argv[1] if len(sys.argv) > 1 else "dropin"
sc, cam, gt = bench.build_workload("B", 0, torch.device("cuda"))
st = bench.Stepper("native", mode, sc, cam, torch.device("cuda"))
gt = gt.cuda()
for _ in range(8):
    st.step(gt)
torch.cuda.synchronize()
N = 5
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(N):
        st.step(gt)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0, t1 = evs[0].time_range.start, max(e.time_range.end for e in evs)
busy = 0.0; cur_end = t0
for e in evs:
    s, en = e.time_range.start, e.time_range.end
    if en > cur_end:
        busy += en - max(s, cur_end); cur_end = en
print(f"mode {mode}: span {1e-3*(t1-t0)/N:.3f} ms/step, GPU busy {1e-3*busy/N:.3f} ms/step, idle {1e-3*((t1-t0)-busy)/N:.3f} ms/step, kernels/step {len(evs)/N:.0f}")
agg = {}
for e in evs:
    k = e.name[:70]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += e.time_range.end - e.time_range.start
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print(f"{1e-3*t/N:8.3f} ms/step  x{n/N:5.1f}  {k}")
