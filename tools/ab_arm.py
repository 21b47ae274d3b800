"""Developer tool: variants of one source file (extra -D flags; AB_SRC, default arm.cu) timed on a roadmap build
(AB_WORKLOAD, default c4).
    python tools/ab_arm.py build "NAME=-DMPB_ARM_GRID=32 -DMPB_ARM_GRID_GLOBAL=1" ...   |   python tools/ab_arm.py run"""
import os, subprocess, sys
SRC = os.environ.get("AB_SRC", "arm.cu")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from motion_planners_b200 import _build
VDIR = os.path.join(_build.LIB_DIR, "variants")
if sys.argv[1] == "build":
    os.makedirs(VDIR, exist_ok=True)
    objs = [os.path.join(_build.OBJ_DIR, s[:-3] + ".o") for s in _build.SOURCES if s != SRC]
    for spec in sys.argv[2:]:
        name, flags = spec.split("=", 1)
        obj = os.path.join(VDIR, name + ".o")
        subprocess.run(["nvcc"] + _build.NVCC_FLAGS + flags.split() + ["-c", os.path.join(_build.CSRC, SRC), "-o", obj], check=True)
        subprocess.run(["nvcc", "-shared", "-o", os.path.join(VDIR, name + ".so"), obj] + objs + ["-lcudart"], check=True)
        os.remove(obj)
        print("built", name)
else:
    for f in sorted(os.listdir(VDIR)):
        if f.endswith(".so"):
            env = dict(os.environ, MPB200_LIB=os.path.join(VDIR, f))
            r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "roadmap_once.py"), os.environ.get("AB_WORKLOAD", "c4"), "2"], env=env, capture_output=True, text=True)
            print("%-24s %s" % (f[:-3], (r.stdout.strip().splitlines() or [r.stderr[-300:]])[-1]))
