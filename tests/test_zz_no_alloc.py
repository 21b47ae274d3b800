"""The kNN entry point that brings its own scratch (mpb_knn_f32filter_static).

It exists for users who cannot give the library a workspace because they cannot allocate device memory
at the moment of the call: a planner thread in a control loop that must keep clear of the driver's
allocator, a process whose device memory was handed out in full to another pool at start-up, a thin
binding that owns no device allocator.  The test drives it through the C ABI and requires the very same
bits as mpb_knn_f32filter with a caller-owned workspace, on two calls in a row (the second one smaller, so
that anything left in the shared array by the first would show), and checks that sizes beyond the built-in
array are refused.  Collected last (file name) so that a run stopped at its first failure has already run
everything else.
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _vp(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def test_static_scratch_knn_equals_workspace_knn_on_consecutive_calls():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from motion_planners_b200 import _lib
    lib = _lib.lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    nullw = ctypes.POINTER(ctypes.c_double)()
    # the claim in this file's name, asked of the driver itself: free device memory before and after the calls
    rt = ctypes.CDLL("libcudart.so.12")
    free_b, total_b = ctypes.c_size_t(), ctypes.c_size_t()

    def free_bytes():
        torch.cuda.synchronize()
        assert rt.cudaMemGetInfo(ctypes.byref(free_b), ctypes.byref(total_b)) == 0
        return free_b.value

    for N, M, D, k, seed in [(60000, 20000, 6, 11, 0), (3000, 3000, 7, 11, 1), (60000, 20000, 6, 11, 0)]:
        rs = np.random.RandomState(seed)
        xs = torch.from_numpy(rs.uniform(size=(N, D))).cuda()
        qs = torch.from_numpy(np.concatenate([xs[:M // 2].cpu().numpy(), rs.uniform(size=(M - M // 2, D))])).cuda()
        # (the caller-owned workspace of the 60000-sample set is the tensor-core scan's, larger than the built-in
        # array: the static entry point keeps to the packed-fp32 filter and sizes its scratch for that)
        need = lib.mpb_knn_workspace_bytes(N, M, D)

        # existing entry point, caller-owned workspace
        ws = torch.empty(need + 256, dtype=torch.uint8, device="cuda")
        off = (-ws.data_ptr()) % 256
        idx_a = torch.empty((M, k), dtype=torch.int64, device="cuda")
        dist_a = torch.empty((M, k), dtype=torch.float64, device="cuda")
        rc = lib.mpb_knn_f32filter(_vp(xs), N, _vp(qs), M, D, nullw, k, float("inf"),
                                   ctypes.c_void_p(ws.data_ptr() + off), need, _vp(idx_a), _vp(dist_a), None, stream)
        assert rc == 0, lib.mpb_last_error()

        # new entry point: no workspace argument at all
        idx_b = torch.full((M, k), -7, dtype=torch.int64, device="cuda")
        dist_b = torch.full((M, k), -7.0, dtype=torch.float64, device="cuda")
        stats = torch.full((2,), -1, dtype=torch.int64, device="cuda")
        # one untimed call first: the first call on a device resolves the array's address (documented set-up step)
        rc = lib.mpb_knn_f32filter_static(_vp(xs), N, _vp(qs), M, D, nullw, k, float("inf"), _vp(idx_b), _vp(dist_b),
                                          _vp(stats), stream)
        assert rc == 0, lib.mpb_last_error()
        before = free_bytes()
        rc = lib.mpb_knn_f32filter_static(_vp(xs), N, _vp(qs), M, D, nullw, k, float("inf"), _vp(idx_b), _vp(dist_b),
                                          _vp(stats), stream)
        assert rc == 0, lib.mpb_last_error()
        assert free_bytes() == before                 # the call took nothing from the device allocator
        torch.cuda.synchronize()
        assert torch.equal(idx_a, idx_b)            # bit-exact, like every other parity test here
        assert torch.equal(dist_a, dist_b)
        assert int(stats[0]) == 0
        assert (idx_b[: M // 2, 0] == torch.arange(M // 2, device="cuda")).all()


def test_static_scratch_knn_refuses_sizes_beyond_its_array():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from motion_planners_b200 import _lib
    lib = _lib.lib()
    xs = torch.zeros((8, 7), dtype=torch.float64, device="cuda")
    out = torch.zeros((8, 3), dtype=torch.int64, device="cuda")
    # claimed sizes only: the call must be refused before anything is read
    rc = lib.mpb_knn_f32filter_static(_vp(xs), 1000000, _vp(xs), 8, 7, ctypes.POINTER(ctypes.c_double)(), 3,
                                      float("inf"), _vp(out), None, None, None)
    assert rc == -12
    msg = lib.mpb_last_error().decode()
    assert "bytes of scratch" in msg and "mpb_knn_f32filter" in msg
