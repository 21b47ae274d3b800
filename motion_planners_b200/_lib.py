"""ctypes loader for lib/libmpb200.so (the C ABI declared in include/mpb200.h).

There is no CPU fallback: if the library has not been built, or a call fails, an exception
is raised.  Build with `python -c "import __graft_entry__ as g; g.build()"` or
`python -m motion_planners_b200._build`.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# MPB200_LIB overrides the in-tree library (A/B runs of kernel variants); there is still no other fallback
LIB_PATH = os.environ.get("MPB200_LIB") or os.path.join(HERE, "lib", "libmpb200.so")

_lib = None

c_f64p = ctypes.c_void_p
c_i64 = ctypes.c_int64
c_int = ctypes.c_int
c_dbl = ctypes.c_double
c_vp = ctypes.c_void_p
c_sz = ctypes.c_size_t
c_hostf64 = ctypes.POINTER(ctypes.c_double)

# name -> (restype, argtypes); mirrors include/mpb200.h one to one
SIGNATURES = {
    "mpb_version": (c_int, []),
    "mpb_last_error": (ctypes.c_char_p, []),
    "mpb_world_create": (c_int, [c_hostf64, c_hostf64, c_int, c_hostf64, c_hostf64, c_int, c_int, c_hostf64,
                                 c_hostf64, c_dbl, ctypes.POINTER(c_vp)]),
    "mpb_world_destroy": (c_int, [c_vp]),
    "mpb_arm_world_create": (c_int, [c_vp, c_hostf64, c_hostf64, c_int, c_hostf64, c_hostf64, c_dbl,
                                     ctypes.POINTER(c_vp)]),
    "mpb_arm_fk": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    "mpb_points_check": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    "mpb_edges_check": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_i64, c_dbl, c_int, c_int, c_vp, c_vp, c_vp]),
    "mpb_edges_check_knn": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_int, c_i64, c_dbl, c_int, c_int, c_vp, c_vp,
                                    c_vp]),
    "mpb_edges_check_knn_rows": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_i64, c_int, c_dbl, c_int, c_int, c_vp, c_vp,
                                         c_vp]),
    "mpb_edges_check_pairs": (c_int, [c_vp, c_vp, c_vp, c_i64, c_dbl, c_int, c_int, c_vp, c_vp, c_vp]),
    "mpb_edges_count": (c_int, [c_vp, c_vp, c_i64, c_int, c_dbl, c_vp, c_vp]),
    "mpb_edges_waypoints": (c_int, [c_vp, c_vp, c_i64, c_int, c_dbl, c_vp, c_vp, c_vp]),
    "mpb_edges_check_circular": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_hostf64, c_hostf64, c_dbl, c_int,
                                         c_int, c_vp, c_vp, c_vp]),
    "mpb_edges_count_circular": (c_int, [c_vp, c_vp, c_i64, c_int, c_hostf64, c_hostf64, c_dbl, c_vp, c_vp]),
    "mpb_edges_waypoints_circular": (c_int, [c_vp, c_vp, c_i64, c_int, c_hostf64, c_hostf64, c_dbl, c_vp, c_vp,
                                             c_vp]),
    "mpb_knn_f64_circular": (c_int, [c_vp, c_i64, c_vp, c_i64, c_int, c_hostf64, c_hostf64, c_hostf64, c_int, c_dbl,
                                     c_vp, c_vp, c_vp]),
    "mpb_rrt_star_workspace_bytes": (c_sz, [c_i64]),
    "mpb_rrt_star_record_bytes": (c_sz, [c_i64]),
    "mpb_rrt_star_step": (c_int, [c_vp, c_vp, c_i64, c_i64, c_hostf64, c_dbl, c_hostf64, c_int, c_dbl, c_dbl, c_vp,
                                  c_sz, c_vp, c_sz, c_vp]),
    "mpb_rrt_star_run_state_bytes": (c_sz, []),
    "mpb_rrt_star_run": (c_int, [c_vp, c_hostf64, c_hostf64, c_dbl, c_hostf64, c_int, c_dbl, c_dbl, c_dbl, c_int, c_i64,
                                 c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp, c_int, c_vp]),
    "mpb_rrt_star_run_read": (c_int, [c_vp, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_double), c_vp]),
    "mpb_knn_filter_kind": (c_int, [c_i64, c_int, c_int]),
    "mpb_knn_tc_key_error": (c_int, [c_vp, c_i64, c_vp, c_i64, c_int, c_hostf64, c_vp, c_sz, c_vp, c_vp]),
    "mpb_knn_launch_count": (c_int, []),
    "mpb_knn_part_rows": (c_i64, [c_i64, c_int, c_int, c_int]),
    "mpb_knn_search_part": (c_int, [c_vp, c_i64, c_int, c_hostf64, c_int, c_dbl, c_vp, c_sz, c_int, c_int, c_vp, c_vp,
                                    c_vp, c_vp, c_vp]),
    "mpb_knn_sample_order": (c_int, [c_vp, c_sz, c_i64, c_int, c_vp, c_vp]),
    "mpb_knn_f64": (c_int, [c_vp, c_i64, c_vp, c_i64, c_int, c_hostf64, c_int, c_dbl, c_vp, c_vp, c_vp]),
    "mpb_knn_f64_metric": (c_int, [c_vp, c_i64, c_vp, c_i64, c_int, c_hostf64, c_int, c_int, c_dbl, c_vp, c_vp, c_vp]),
    "mpb_knn_workspace_bytes": (c_sz, [c_i64, c_i64, c_int]),
    "mpb_knn_prepare": (c_int, [c_vp, c_i64, c_int, c_hostf64, c_vp, c_sz, c_vp]),
    "mpb_knn_search": (c_int, [c_vp, c_i64, c_vp, c_i64, c_int, c_hostf64, c_int, c_dbl, c_vp, c_sz, c_vp, c_vp,
                               c_vp, c_vp]),
    "mpb_knn_f32filter": (c_int, [c_vp, c_i64, c_vp, c_i64, c_int, c_hostf64, c_int, c_dbl, c_vp, c_sz, c_vp, c_vp,
                                  c_vp, c_vp]),
    "mpb_knn_f32filter_static": (c_int, [c_vp, c_i64, c_vp, c_i64, c_int, c_hostf64, c_int, c_dbl, c_vp, c_vp, c_vp,
                                         c_vp]),
    "mpb_trees_nearest": (c_int, [c_vp, c_vp, c_i64, c_i64, c_int, c_vp, c_hostf64, c_int, c_vp, c_vp, c_vp, c_vp]),
    "mpb_trees_append": (c_int, [c_vp, c_vp, c_vp, c_i64, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64,
                                 c_dbl, c_vp, c_vp, c_vp]),
    "mpb_rrt_connect_batch": (c_int, [c_vp, c_vp, c_vp, c_i64, c_hostf64, c_int, c_dbl, c_vp, c_vp, c_int, c_vp, c_vp,
                                      c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "mpb_scan_nearest_radius": (c_int, [c_vp, c_i64, c_int, c_vp, c_hostf64, c_int, c_dbl, c_vp, c_vp, c_vp, c_vp,
                                        c_vp]),
    "mpb_edge_costs": (c_int, [c_vp, c_i64, c_int, c_hostf64, c_int, c_vp, c_i64, c_int, c_i64, c_vp, c_vp]),
    "mpb_roadmap_relax": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_int, c_i64, c_vp, c_vp, c_int, c_vp]),
    "mpb_roadmap_parents": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_int, c_i64, c_vp, c_i64, c_vp, c_vp]),
    "mpb_peak_fma_f32": (c_int, [c_int, c_int, c_int, c_vp, c_vp]),
    "mpb_peak_fma_f64": (c_int, [c_int, c_int, c_int, c_vp, c_vp]),
    "mpb_peak_fma_f32x2": (c_int, [c_int, c_int, c_int, c_vp, c_vp]),
}


class MpbError(RuntimeError):
    pass


def lib():
    """The loaded library.  Raises if it is missing -- never falls back to anything else."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MpbError("CUDA library %s has not been built (run __graft_entry__.build()); "
                           "there is no CPU fallback" % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)      # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().mpb_last_error()
        raise MpbError("%s failed (%d): %s" % (what, rc, msg.decode() if msg else ""))
