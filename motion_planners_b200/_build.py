"""Compile the CUDA sources under csrc/ into lib/libmpb200.so for sm_100a (in-tree, so the
built library travels with the repository snapshot to the GPU box)."""
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
OBJ_DIR = os.path.join(LIB_DIR, "obj")
LIB = os.path.join(LIB_DIR, "libmpb200.so")
SOURCES = ["world.cu", "knn.cu", "trees.cu", "rrt.cu", "graph.cu", "peaks.cu", "circular.cu", "rrtstar.cu", "arm.cu", "rrtstar_run.cu"]

# -fmad=false: nothing is contracted behind our back; every fused operation in the kernels is an
# explicit __fma_rn / __fmaf_rn (the fp32 filter loop, the sphere test, the peak benchmarks).
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "mpb200.h"))
    srcs = list(SOURCES)
    objs = [os.path.join(OBJ_DIR, s[:-3] + ".o") for s in srcs]

    def compile_one(pair):
        src, obj = pair
        path = os.path.join(CSRC, src)
        if force or _stale(obj, [path] + headers):
            cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
            if verbose:
                print(r.stderr)
            return True
        return False

    with ThreadPoolExecutor(max_workers=len(srcs)) as ex:
        rebuilt = list(ex.map(compile_one, zip(srcs, objs)))
    if force or any(rebuilt) or _stale(LIB, objs):
        cmd = ["nvcc", "-shared", "-o", LIB] + objs + ["-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
