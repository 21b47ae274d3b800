"""Developer tool: build variants of the library with extra -D flags and time the 1M x 7-DoF self-search with each.
    python tools/ab.py build "NAME=-DMPB_TC_BULK_MIN=3" "NAME2=-DX=1 -DY=2" ...   (here, no GPU needed)
    python tools/ab.py run [N] [D]                                               (on the GPU box)"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from motion_planners_b200 import _build
VDIR = os.path.join(_build.LIB_DIR, "variants")

if sys.argv[1] == "build":
    os.makedirs(VDIR, exist_ok=True)
    for f in os.listdir(VDIR):
        os.remove(os.path.join(VDIR, f))
    srcs = [os.path.join(_build.CSRC, s) for s in _build.SOURCES if s != "knn.cu"]
    objs = [os.path.join(_build.OBJ_DIR, s[:-3] + ".o") for s in _build.SOURCES if s != "knn.cu"]
    procs = []
    for spec in sys.argv[2:]:
        name, flags = spec.split("=", 1)
        obj = os.path.join(VDIR, name + ".o")
        cmd = ["nvcc"] + _build.NVCC_FLAGS + flags.split() + ["-c", os.path.join(_build.CSRC, "knn.cu"), "-o", obj]
        procs.append((name, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for name, obj, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode != 0:
            print(name, "FAILED\n", out[-2000:]); continue
        so = os.path.join(VDIR, name + ".so")
        subprocess.run(["nvcc", "-shared", "-o", so, obj] + objs + ["-lcudart"], check=True)
        os.remove(obj)
        print("built", so)
else:
    args = sys.argv[2:]
    for f in sorted(os.listdir(VDIR)):
        if f.endswith(".so"):
            env = dict(os.environ, MPB200_LIB=os.path.join(VDIR, f))
            r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "knn_once.py")] + args, env=env, capture_output=True, text=True)
            print("%-28s %s" % (f[:-3], (r.stdout.strip().splitlines() or [r.stderr[-300:]])[-1]))
