"""Which reference cycles keep GPU memory alive between steps?  Runs steps with the cyclic collector off, then collects
with DEBUG_SAVEALL and lists the garbage by type (and the CUDA bytes the tensors in it hold)."""
import collections, gc, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "relightable3dgaussians-w_b200")]
import torch
import bench
dev = torch.device("cuda")
sc, cam, gt = bench.build_workload(os.environ.get("DIAG_WORKLOAD", "small"), 0, dev)
st = bench.make_stepper("native", os.environ.get("DIAG_MODE", "dropin"), sc, cam, dev)
gt = gt.cuda()
for _ in range(3):
    st.step(gt)
torch.cuda.synchronize()
gc.collect()
gc.disable()
m0 = torch.cuda.memory_allocated()
N = 4
for _ in range(N):
    st.step(gt)
torch.cuda.synchronize()
m1 = torch.cuda.memory_allocated()
print(f"allocated growth over {N} steps with gc off: {(m1 - m0) / 2**20:.1f} MiB")
gc.set_debug(gc.DEBUG_SAVEALL)
n = gc.collect()
cnt = collections.Counter(type(o).__name__ for o in gc.garbage)
print("unreachable objects:", n, cnt.most_common(25))
tb = sum(o.numel() * o.element_size() for o in gc.garbage if isinstance(o, torch.Tensor) and o.is_cuda)
print(f"CUDA tensor bytes inside the garbage: {tb / 2**20:.1f} MiB")
for o in gc.garbage:
    if type(o).__name__ in ("GeometryState", "_GeometryFnBackward", "_BlendFnBackward", "_PreshadeFnBackward", "_RegFnBackward",
                            "_Holder", "function", "cell", "CFunctionType"):
        refs = [type(r).__name__ for r in gc.get_referents(o)][:12]
        print("  ", type(o).__name__, "->", refs)
gc.set_debug(0)
gc.garbage.clear()
m2 = torch.cuda.memory_allocated()
print(f"after collect: {(m2 - m0) / 2**20:.1f} MiB above the baseline")
