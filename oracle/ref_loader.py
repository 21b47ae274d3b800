"""Import the UNMODIFIED reference (caelan/motion-planners) headless.  TEST INFRASTRUCTURE ONLY.

Used only by oracle/make_golden.py, in the build container where the reference
checkout exists; nothing on the GPU box reads the reference at run time (the
golden vectors travel as committed fixtures under tests/golden/).

motion_planners/tkinter/viewer.py:1-7 imports Tk at module load; Tk is not
installed, so stub modules are registered first.  Nothing of Tk is ever called:
every drawing helper is skipped when viewer is None.
"""
import os
import sys
import types
import warnings

REFERENCE_ROOT = os.environ.get("MP_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "motion_planners"))


def load():
    """-> the imported `motion_planners` package of the reference."""
    if not available():
        raise RuntimeError("reference checkout not found at %s" % REFERENCE_ROOT)
    if "tkinter" not in sys.modules or not hasattr(sys.modules["tkinter"], "Tk"):
        tk = types.ModuleType("tkinter")

        class _Dummy(object):
            def __init__(self, *a, **k):
                raise RuntimeError("Tk is not available (headless)")

        tk.Tk = tk.Canvas = tk.Toplevel = _Dummy
        tk.LAST = "last"
        sys.modules["tkinter"] = tk
    if "_tkinter" not in sys.modules:
        _tk = types.ModuleType("_tkinter")
        _tk.TclError = type("TclError", (Exception,), {})
        sys.modules["_tkinter"] = _tk
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import motion_planners  # noqa: F401
        import motion_planners.tkinter.samplers  # noqa: F401
        import motion_planners.tkinter.viewer  # noqa: F401
        import motion_planners.tkinter.run  # noqa: F401
        import motion_planners.lazy_prm  # noqa: F401
        import motion_planners.prm  # noqa: F401
        import motion_planners.rrt_connect  # noqa: F401
        import motion_planners.rrt_star  # noqa: F401
        import motion_planners.nearest  # noqa: F401
        import motion_planners.primitives  # noqa: F401
        import motion_planners.meta  # noqa: F401
        import motion_planners.smoothing  # noqa: F401
    return sys.modules["motion_planners"]
