"""Per-kernel ms/step of the U2 step (drop-in mode, config B) from the library's own CUDA-event profile, plus the step
time.  usage: diag_kernels.py [steps]   (env vars select kernel variants under test)"""
import gc, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import bench
from r3dgw_b200 import native
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dev = torch.device("cuda")
sc, cam, gt = bench.build_workload("B", 0, dev)
st = bench.Stepper("native", os.environ.get("DIAG_MODE", "dropin"), sc, cam, dev)
gt = gt.cuda()
for _ in range(15):
    st.step(gt)
torch.cuda.synchronize()
gc.collect(); gc.disable()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    st.step(gt)
e1.record(); torch.cuda.synchronize()
step_ms = e0.elapsed_time(e1) / steps
native.profile_enable(True)
for _ in range(steps):
    st.step(gt)
torch.cuda.synchronize()
prof = native.profile_read(reset=True)
native.profile_enable(False)
out = {k: round(ms / steps, 4) for k, (ms, n) in prof.items() if n}
out["_step_ms"] = round(step_ms, 3)
out["_kernel_sum_ms"] = round(sum(v for k, v in out.items() if not k.startswith("_")), 3)
print(json.dumps({"variant": {k: v for k, v in os.environ.items() if k.startswith("R3DGW_")}, **out}))
