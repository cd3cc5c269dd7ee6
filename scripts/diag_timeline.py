"""GPU-busy vs wall time of the U2 step (torch.profiler): how much of the step is launch/CPU-bound."""
import gc, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
from torch.profiler import profile, ProfilerActivity
import bench
dev = torch.device("cuda")
sc, cam, gt = bench.build_workload("B", 0, dev)
st = bench.make_stepper("native", os.environ.get("DIAG_MODE", "dropin"), sc, cam, dev)
gt = gt.cuda()
for _ in range(int(os.environ.get("DIAG_WARM", "15"))):
    st.step(gt)
torch.cuda.synchronize()
gc.collect(); gc.disable()
N = 5
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(N):
        st.step(gt)
    torch.cuda.synchronize()
ev = prof.events()
kern = [e for e in ev if e.device_type == torch.autograd.DeviceType.CUDA]
busy = sum(e.device_time for e in kern if hasattr(e, "device_time")) if kern else 0
# merge intervals for true busy time
iv = sorted((e.time_range.start, e.time_range.end) for e in kern)
merged, cur_s, cur_e = 0.0, None, None
for s, e in iv:
    if cur_e is None or s > cur_e:
        if cur_e is not None:
            merged += cur_e - cur_s
        cur_s, cur_e = s, e
    else:
        cur_e = max(cur_e, e)
if cur_e is not None:
    merged += cur_e - cur_s
span = (iv[-1][1] - iv[0][0]) if iv else 0
print(f"steps {N}: gpu span/step {span / N / 1000:.3f} ms, gpu busy/step {merged / N / 1000:.3f} ms, kernels/step {len(kern) / N:.0f}")
agg = {}
for e in kern:
    a = agg.setdefault(e.name[:70], [0.0, 0])
    a[0] += (e.time_range.end - e.time_range.start); a[1] += 1
for k, (t, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
    print(f"{t / N / 1000:8.3f} ms  {n / N:6.1f}x  {k}")
# gaps: largest idle gaps between kernels
gaps = []
prev = None
for s, e in iv:
    if prev is not None and s > prev:
        gaps.append(s - prev)
    prev = max(prev, e) if prev is not None else e
gaps.sort(reverse=True)
print("idle/step ms", sum(gaps) / N / 1000, "top gaps us", [round(g, 1) for g in gaps[:20]])

# which kernels sit on either side of the large idle gaps (one step's worth)
kk = sorted(kern, key=lambda e: e.time_range.start)
t0 = kk[0].time_range.start
rows = []
prev_end, prev_name = None, None
for e in kk:
    s_, e_ = e.time_range.start, e.time_range.end
    if prev_end is not None and s_ - prev_end > 25:
        rows.append((s_ - prev_end, (s_ - t0) / 1000, prev_name, e.name[:60]))
    if prev_end is None or e_ > prev_end:
        prev_end, prev_name = e_, e.name[:60]
per_step = span / N
print("gaps > 25 us in the first step:")
for g, t, a, b in rows:
    if t * 1000 < per_step:
        print(f"  {g:7.1f} us at {t:7.3f} ms  after [{a}]  before [{b}]")
