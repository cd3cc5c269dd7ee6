"""Which NVML query perturbs a running step loop?  (GPU box)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch, pynvml
import bench
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
sc, cam, gt = bench.build_workload("B", 0, torch.device("cuda"))
st = bench.Stepper("native", "dropin", sc, cam, torch.device("cuda"))
gt = gt.cuda()
for _ in range(10):
    st.step(gt)
torch.cuda.synchronize()
calls = {
    "none": lambda: None,
    "clock": lambda: pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
    "power": lambda: pynvml.nvmlDeviceGetPowerUsage(h),
    "reasons": lambda: pynvml.nvmlDeviceGetCurrentClocksEventReasons(h),
}
import gc
gc.collect(); gc.disable()
for name, fn in calls.items():
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(16)]
    durs = []
    for k, (a, b) in enumerate(evs):
        a.record(); st.step(gt); b.record()
        if k % 4 == 1:
            t0 = time.perf_counter(); fn(); durs.append(round(1e3 * (time.perf_counter() - t0), 2))
    torch.cuda.synchronize()
    print(name, "call ms", durs, "steps", [round(a.elapsed_time(b), 1) for a, b in evs])
