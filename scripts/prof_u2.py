"""A few U2 training steps (bench.Stepper, drop-in mode, config B) - the short command ncu wraps.  usage: prof_u2.py [steps] [mode]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import bench
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
mode = sys.argv[2] if len(sys.argv) > 2 else "dropin"
sc, cam, gt = bench.build_workload("B", 0, torch.device("cuda"))
st = bench.make_stepper("native", mode, sc, cam, torch.device("cuda"))
gt = gt.cuda()
for _ in range(steps):
    st.step(gt)
torch.cuda.synchronize()
print("ok")
