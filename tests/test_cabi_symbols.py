"""CPU-side checks of the C-ABI library: it loads and exports every symbol include/mpb200.h declares.
No compute call is made here (no GPU in this tier)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "mpb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mpb_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_entry_points():
    names = _declared()
    for must in ("mpb_world_create", "mpb_points_check", "mpb_edges_check", "mpb_edges_check_pairs",
                 "mpb_knn_f64", "mpb_knn_f32filter", "mpb_trees_nearest", "mpb_scan_nearest_radius"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from motion_planners_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = _lib.lib()
    for name in _declared():
        assert hasattr(lib, name), name
    # the binding table mirrors the header one to one
    assert sorted(_lib.SIGNATURES) == _declared()
    assert lib.mpb_version() == 1


def test_bad_arguments_are_rejected_without_a_gpu():
    from motion_planners_b200 import _lib
    lib = _lib.lib()
    # sizes/dims are validated before any CUDA call
    assert lib.mpb_knn_workspace_bytes(1000, 10, 0) == 0
    assert lib.mpb_knn_workspace_bytes(1000, 10, 17) == 0
    assert lib.mpb_knn_workspace_bytes(1000, 10, 7) > 1000 * 32
    rc = lib.mpb_knn_f64(None, 10, None, 10, 99, None, 3, 1.0, None, None, None)
    assert rc == -22 and b"dim" in lib.mpb_last_error()
    rc = lib.mpb_knn_f64(None, 10, None, 10, 3, None, 33, 1.0, None, None, None)
    assert rc == -22 and b"k" in lib.mpb_last_error()
    rc = lib.mpb_points_check(None, None, 5, None, None)
    assert rc == -22


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "motion_planners_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "mp_oracle" not in src, f


def test_knn_filter_kind_policy_is_a_pure_host_query():
    """mpb_knn_filter_kind touches no device: which filter a search gets depends on size, dimension and k only."""
    import ctypes
    import os
    if os.environ.get("MPB200_KNN_TC") is not None:
        import pytest
        pytest.skip("the policy is overridden through the environment")
    from motion_planners_b200 import _lib
    lib = _lib.lib()
    kind = lambda n, d, k: lib.mpb_knn_filter_kind(ctypes.c_int64(n), d, k)
    FP32, TENSOR, EXACT = 0, 1, 2
    assert kind(1000000, 7, 11) == TENSOR and kind(100000, 6, 11) == TENSOR and kind(32768, 7, 11) == TENSOR
    assert kind(32767, 7, 11) == FP32 and kind(32767, 6, 11) == FP32
    assert kind(24576, 5, 11) == TENSOR and kind(24575, 5, 11) == FP32
    assert kind(16384, 4, 11) == TENSOR and kind(16383, 4, 11) == FP32
    assert kind(1000000, 3, 11) == FP32 and kind(1000000, 8, 11) == FP32
    assert kind(1000000, 7, 13) == FP32          # 32-entry lists: packed-fp32 kernel
    assert kind(1000000, 7, 29) == EXACT and kind(1000, 2, 32) == EXACT
    assert kind(1000, 0, 3) < 0 and kind(1000, 3, 0) < 0 and kind(1000, 3, 33) < 0


def test_knn_part_rows_partition_the_set():
    """mpb_knn_part_rows is host arithmetic: the shares of the k-d leaves that mpb_knn_search_part deals to the ranks
    cover every sample exactly once and are equal to within one per cent (leaves hold 122 or 123 samples at 1M)."""
    import ctypes
    import os
    if os.environ.get("MPB200_KNN_TC") is not None:
        import pytest
        pytest.skip("the policy is overridden through the environment")
    from motion_planners_b200 import _lib
    lib = _lib.lib()
    rows = lambda n, d, r, g: int(lib.mpb_knn_part_rows(ctypes.c_int64(n), d, r, g))
    for n, d in [(1000000, 7), (200001, 7), (50001, 6), (100000, 6), (32768, 7), (777777, 5)]:
        for g in (1, 2, 3, 4, 8):
            sizes = [rows(n, d, r, g) for r in range(g)]
            assert sum(sizes) == n and min(sizes) > 0, (n, d, g, sizes)
            assert max(sizes) - min(sizes) <= max(2 * 128, n // g // 100), (n, d, g, sizes)
    assert rows(20000, 6, 0, 2) == -1 and rows(1000000, 3, 0, 2) == -1      # sets the tensor-core scan does not serve
    assert rows(1000000, 7, 2, 2) == -1 and rows(1000000, 7, -1, 2) == -1
