"""Developer tool: per-row against per-edge edge kernel at several row counts (c4lite world, 1M samples)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, bench
from motion_planners_b200 import ops
from motion_planners_b200.roadmap import DeviceRoadmap
world, X, k, _ = bench.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c4lite")
rm = DeviceRoadmap(world, k=k); x = torch.from_numpy(X).cuda(); rm.build(x)
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / reps
for rows in (50000, 100000, 125000, 250000, 500000, 1000000):
    nbr = rm.nbr_idx[:rows].contiguous()
    src = torch.arange(rows, device="cuda").repeat_interleave(nbr.shape[1])
    a = t(lambda: rm.obstacles.edges_check_knn(x, nbr, row_offset=0))
    b = t(lambda: rm.obstacles.edges_check(x, src, nbr.view(-1), want_counts=False))
    print("rows %8d  per-row %.3f ms  per-edge %.3f ms" % (rows, a, b))
