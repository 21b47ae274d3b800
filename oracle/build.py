"""Build recipe for the CPU oracle (TEST INFRASTRUCTURE, never shipped or imported by the product).

Compiles oracle/mp_oracle.c into oracle/_build/libmp_oracle.so.  The reference
(caelan/motion-planners) is pure Python, so there is no reference C/C++ source to
compile into oracle/_ref; the oracle is instead pinned against outputs of the
Python reference itself (oracle/make_golden.py -> tests/golden/).
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "mp_oracle.c")
SRCS = [SRC, os.path.join(HERE, "arm_oracle.c")]      # arm_oracle.c: the collision model of configs[3]
OUT_DIR = os.path.join(HERE, "_build")
OUT = os.path.join(OUT_DIR, "libmp_oracle.so")

# -ffp-contract=off: the reference's numpy ufuncs round every operation separately.
# -mfma: fma() calls (the restated BLAS dot) become one instruction instead of a libm call.
CFLAGS = ["-O2", "-ffp-contract=off", "-mfma", "-fopenmp", "-shared", "-fPIC", "-fvisibility=hidden",
          "-Wall", "-Wextra"]


def build(force=False):
    os.makedirs(OUT_DIR, exist_ok=True)
    if (not force and os.path.exists(OUT)
            and all(os.path.getmtime(OUT) >= os.path.getmtime(f) for f in SRCS)):
        return OUT
    cmd = ["gcc"] + CFLAGS + SRCS + ["-o", OUT, "-lm"]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force=True))
