"""Developer tool: one roadmap build of a bench workload (for ncu captures).  python tools/roadmap_once.py [c4|c4lite|c2] [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from motion_planners_b200.roadmap import DeviceRoadmap
name = sys.argv[1] if len(sys.argv) > 1 else "c4"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
world, X, k, desc = bench.make_workload(name)
rm = DeviceRoadmap(world, k=k)
x = torch.from_numpy(X).cuda()
rm.profile = True
for _ in range(reps):
    rm.build(x)
    torch.cuda.synchronize()
    print({a: round(b, 3) for a, b in rm.phase_ms().items()}, "fallbacks", int(rm.stats[0]))
