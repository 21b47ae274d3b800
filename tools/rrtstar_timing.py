"""Developer tool: where the persistent RRT* kernel spends its time (a library built with -DMPB_RRTSTAR_TIMING)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from motion_planners_b200 import _build
out = os.path.join(_build.LIB_DIR, "libmpb200_rt.so")
if "--build-only" in sys.argv or not os.path.exists(out):
    srcs = [os.path.join(_build.CSRC, s) for s in _build.SOURCES]
    subprocess.run(["nvcc"] + _build.NVCC_FLAGS + ["-DMPB_RRTSTAR_TIMING", "-shared", "-o", out] + srcs + ["-lcudart"], check=True)
    if "--build-only" in sys.argv:
        sys.exit(0)
os.environ["MPB200_LIB"] = out
import numpy as np, torch, bench
import motion_planners_b200.rrt_star as rs
orig = torch.zeros
keep = {}
def zeros(*a, **k):
    t = orig(*a, **k)
    if len(a) and isinstance(a[0], int) and k.get("dtype") == torch.uint8:
        keep["state"] = t
    return t
torch.zeros = zeros
iters = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 50000
print(bench.measure_rrt_star(iters))
st = keep["state"]
off = (-st.data_ptr()) % 256
raw = st[off:].cpu().numpy()
tail = raw[: (len(raw) // 8) * 8].view(np.int64)
struct_words = tail[12856 // 8: 12856 // 8 + 8]     # offsetof(RunState, t_phase)
names = ["A sample (every block)", "grid.sync waits", "B nearest scan", "C extension", "D radius scan + candidate edges", "E parent choice (every block)", "-", "rejected samples"]
tot = sum(int(v) for v in struct_words[:6])
for nm, v in zip(names, struct_words):
    print("%-34s %12d  %5.1f%%" % (nm, int(v), 100.0 * int(v) / tot if nm != "rejected samples" else 0))
print("ns per iteration", tot / iters)
