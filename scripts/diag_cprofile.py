"""cProfile of the U2 step's host side (drop-in mode): where the Python time between launches goes."""
import cProfile, gc, os, pstats, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import bench
dev = torch.device("cuda")
sc, cam, gt = bench.build_workload("B", 0, dev)
st = bench.Stepper("native", os.environ.get("DIAG_MODE", "dropin"), sc, cam, dev)
gt = gt.cuda()
for _ in range(15):
    st.step(gt)
torch.cuda.synchronize()
gc.collect(); gc.disable()
N = 20
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    st.step(gt)
torch.cuda.synchronize()
pr.disable()
ps = pstats.Stats(pr)
ps.sort_stats("cumulative").print_stats(45)
ps.sort_stats("tottime").print_stats(30)
