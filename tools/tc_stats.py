"""Developer tool: per-CTA counters of the tensor-core kNN scan (a library built with -DMPB_TC_STATS).
    python tools/tc_stats.py [N] [D]
Builds lib/libmpb200_stats.so next to the product library, runs one self-search and prints the counters."""
import os, subprocess, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from motion_planners_b200 import _build

def build_stats():
    out = os.path.join(_build.LIB_DIR, "libmpb200_stats.so")
    srcs = [os.path.join(_build.CSRC, s) for s in _build.SOURCES]
    if not os.path.exists(out) or any(os.path.getmtime(os.path.join(_build.CSRC, f)) > os.path.getmtime(out) for f in os.listdir(_build.CSRC)):
        cmd = ["nvcc"] + _build.NVCC_FLAGS + ["-DMPB_TC_STATS", "-shared", "-o", out] + srcs + ["-lcudart"]
        subprocess.run(cmd, check=True)
    return out

if __name__ == "__main__":
    if "--build-only" in sys.argv:
        print(build_stats()); sys.exit(0)
    os.environ["MPB200_LIB"] = build_stats()
    import numpy as np, torch
    from motion_planners_b200 import ops, _lib
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
    D = int(sys.argv[2]) if len(sys.argv) > 2 else 7
    x = torch.from_numpy(np.random.RandomState(0).uniform(size=(N, D))).cuda()
    index = ops.KnnIndex(x)
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    for _ in range(2):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        index._prepared = False
        index.prepare(N)
        e0.record(); index.search(x, 11, stats=stats); e1.record(); torch.cuda.synchronize()
    ws, _ = index._workspace(N)
    hdr = torch.empty(1024, dtype=torch.uint8, device="cuda")
    ctypes.cdll.LoadLibrary("libcudart.so.12").cudaMemcpy(ctypes.c_void_p(hdr.data_ptr()), ws, 1024, 3)
    raw = hdr.cpu().numpy()
    off = 16 * 8 * 2 + 16 * 8 + 8 + 8 + 16 * 4
    u = raw[off:off + 136].view(np.uint64)
    ctas = max(1, int(u[1]))
    print("search ms %.3f  fallbacks %d  tiles/CTA %.1f" % (e0.elapsed_time(e1), int(stats[0]), int(u[0]) / ctas))
    names = ["ctas", "flagged-lane events", "takers inserted", "rounds", "cycles total", "cycles plan", "cycles round0", "tiles listed", "producer wait empty", "mma wait full", "mma wait tempty", "epi wait tfull (16 warps)", "epi examine (16 warps)", "bulk chunks (thread count)", "-", "-"]
    for n, v in zip(names, u[1:]):
        print("  %-22s %14d   per CTA %10.1f" % (n, int(v), int(v) / ctas))
