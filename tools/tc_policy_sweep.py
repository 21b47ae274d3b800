"""Developer tool: prepare + self-search times over (N, D) under the current MPB200_KNN_TC policy.
   MPB200_KNN_TC=0 python tools/tc_policy_sweep.py ; MPB200_KNN_TC=1 python tools/tc_policy_sweep.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from motion_planners_b200 import ops, _lib
dims = [int(v) for v in os.environ.get("SWEEP_DIMS", "2,3,4,5,6,7").split(",")]
sizes = [int(v) for v in os.environ.get("SWEEP_SIZES", "20000,50000,100000,200000,500000,1000000").split(",")]
k = int(os.environ.get("SWEEP_K", "11"))
for D in dims:
    for N in sizes:
        x = torch.from_numpy(np.random.RandomState(0).uniform(size=(N, D))).cuda()
        index = ops.KnnIndex(x)
        stats = torch.zeros(2, dtype=torch.int64, device="cuda")
        best = (1e9, 1e9)
        for _ in range(3):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e2 = torch.cuda.Event(enable_timing=True)
            index._prepared = False
            e0.record(); index.prepare(N); e1.record(); index.search(x, k, stats=stats); e2.record(); torch.cuda.synchronize()
            t = (e0.elapsed_time(e1), e1.elapsed_time(e2))
            if sum(t) < sum(best):
                best = t
        kind = _lib.lib().mpb_knn_filter_kind(N, D, k)
        print("D=%d N=%7d kind=%d prepare %.3f search %.3f total %.3f fallbacks %d" % (D, N, kind, best[0], best[1], sum(best), int(stats[0])), flush=True)
        del index, x
