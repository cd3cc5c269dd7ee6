"""Per-frame times of the U3 loop (new view each frame) after a few training steps - looks for one-off stalls."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import bench
sc, cam, gt = bench.build_workload("B", 0, torch.device("cuda"))
st = bench.Stepper("native", "dropin", sc, cam, torch.device("cuda"))
gt = gt.cuda()
for _ in range(6):
    st.step(gt)
torch.cuda.synchronize()
print("mem after steps: alloc %.2f GB reserved %.2f GB" % (torch.cuda.memory_allocated() / 2**30, torch.cuda.memory_reserved() / 2**30))
def frame():
    st.raster.clear_geometry_cache()
    with torch.no_grad():
        st.forward(None, st.sky2)
for i in range(12):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    frame()
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"frame {i}: {1e3*(t1-t0):.2f} ms  reserved {torch.cuda.memory_reserved()/2**30:.2f} GB  mallocs {torch.cuda.memory_stats()['num_device_alloc']} frees {torch.cuda.memory_stats()['num_device_free']}")
