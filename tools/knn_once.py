"""Developer tool: one prepared self-search of the kNN (for ncu captures).  python tools/knn_once.py [N] [D] [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from motion_planners_b200 import ops
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
D = int(sys.argv[2]) if len(sys.argv) > 2 else 7
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
x = torch.from_numpy(np.random.RandomState(0).uniform(size=(N, D))).cuda()
index = ops.KnnIndex(x)
stats = torch.zeros(2, dtype=torch.int64, device="cuda")
for _ in range(reps):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e2 = torch.cuda.Event(enable_timing=True)
    index._prepared = False
    e0.record(); index.prepare(N); e1.record(); index.search(x, 11, stats=stats); e2.record(); torch.cuda.synchronize()
    print("prepare %.3f ms  search %.3f ms  fallbacks %d  tiles per 128 queries %.1f" % (e0.elapsed_time(e1), e1.elapsed_time(e2), int(stats[0]), int(stats[1]) * 128.0 / N))
