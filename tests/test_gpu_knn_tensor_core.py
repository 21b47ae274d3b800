"""The tensor-core form of the kNN filter (tcgen05.mma kind::tf32 on split operands, accumulator in tensor
memory) forced onto small sets with MPB200_KNN_TC=1, against the CPU oracle (nearest.py:74-85 restated in
oracle/mp_oracle.c): same neighbours, bit-identical distances.  Includes inputs built to defeat an fp32 /
TF32 filter, where the exactness certificate has to FAIL and the exact kernel has to take the query over.
The switch is read once per process, so the checks run in a child process."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r)
import oracle
from motion_planners_b200 import _lib, ops
oracle.use_all_cores()
lib = _lib.lib()
host = lambda t: t.cpu().numpy()

for N, M, D, k, seed in [(30000, 3000, 7, 11, 0), (20011, 2500, 6, 11, 1), (9000, 1500, 3, 5, 2),
                         (8192, 130, 2, 12, 3), (8193, 129, 7, 1, 4), (9000, 700, 5, 8, 5)]:
    assert lib.mpb_knn_filter_kind(N, D, k) == 1, (N, D, k)
    rs = np.random.RandomState(seed)
    X = rs.uniform(size=(N, D))
    Q = np.concatenate([X[: M // 2], rs.uniform(-0.1, 1.1, size=(M - M // 2, D))])
    w = rs.uniform(0.5, 2.0, size=D)
    x, q = torch.from_numpy(X).cuda(), torch.from_numpy(Q).cuda()
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    index = ops.KnnIndex(x, w)
    idx, dist = index.search(q, k, float("inf"), stats=stats)
    o_idx, o_dist = oracle.knn(X, Q, k, w)
    assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all(), (N, D)
    assert int(stats[0]) < M // 50, int(stats[0])        # the certificate holds for (almost) every query
    r = float(dist[:, k // 2].median())
    idx_r, dist_r = index.search(q, k, r)
    o_idx_r, o_dist_r = oracle.knn(X, Q, k, w, r)
    assert (host(idx_r) == o_idx_r).all() and (host(dist_r) == o_dist_r).all(), (N, D, "bounded")

# -- the certificate's failing branch on the tensor path
# (a) a cluster 1e-9 wide, off centre, inside a unit cloud: fp32 coordinates cannot tell its members apart
rs = np.random.RandomState(5)
N, D, k = 12000, 7, 11
X = rs.uniform(size=(N, D))
X[:4000] = 0.9 + 1e-9 * rs.uniform(size=(4000, D))
assert lib.mpb_knn_filter_kind(N, D, k) == 1
stats = torch.zeros(2, dtype=torch.int64, device="cuda")
x = torch.from_numpy(X).cuda()
idx, dist = ops.knn(x, x, k, stats=stats)
o_idx, o_dist = oracle.knn(X, X, k)
assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all()
assert int(stats[0]) >= 3000, int(stats[0])

# (b) two groups 2e3 apart: the bounding-box centre sits 1e3 away from every sample, so the centred fp32
# coordinates carry an absolute error near 6e-5 while neighbours are about 1e-4..1e-3 apart
rs = np.random.RandomState(6)
N, D, k = 16384, 6, 11
X = 0.02 * rs.uniform(size=(N, D))
X[N // 2:, 0] += 2000.0
assert lib.mpb_knn_filter_kind(N, D, k) == 1
stats = torch.zeros(2, dtype=torch.int64, device="cuda")
x = torch.from_numpy(X).cuda()
idx, dist = ops.knn(x, x, k, stats=stats)
o_idx, o_dist = oracle.knn(X, X, k)
assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all()
assert int(stats[0]) > 0, int(stats[0])

# (c) mixed magnitudes per coordinate (weights 1e3 and 1e-3) and a bounded search
rs = np.random.RandomState(7)
N, D, k = 10000, 7, 12
X = rs.uniform(size=(N, D))
w = np.array([1e3, 1.0, 1e-3, 1.0, 30.0, 1.0, 0.1])
x = torch.from_numpy(X).cuda()
stats = torch.zeros(2, dtype=torch.int64, device="cuda")
idx, dist = ops.knn(x, x[:3000], k, weights=w, stats=stats)
o_idx, o_dist = oracle.knn(X, X[:3000], k, w)
assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all()
print("TC_OK")
"""


def test_tensor_core_filter_matches_oracle():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    env = dict(os.environ, MPB200_KNN_TC="1")
    out = subprocess.run([sys.executable, "-c", CHILD % ROOT], env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "TC_OK" in out.stdout


FP32_CHILD = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r)
import oracle
from motion_planners_b200 import _lib, ops
oracle.use_all_cores()
lib = _lib.lib()
host = lambda t: t.cpu().numpy()
for N, M, D, k, seed in [(60000, 2000, 6, 11, 0), (40000, 1500, 7, 11, 1), (50000, 1000, 4, 12, 2)]:
    assert lib.mpb_knn_filter_kind(N, D, k) == 0, (N, D, k)
    rs = np.random.RandomState(seed)
    X = rs.uniform(size=(N, D))
    Q = np.concatenate([X[: M // 2], rs.uniform(-0.1, 1.1, size=(M - M // 2, D))])
    w = rs.uniform(0.5, 2.0, size=D)
    x, q = torch.from_numpy(X).cuda(), torch.from_numpy(Q).cuda()
    idx, dist = ops.KnnIndex(x, w).search(q, k, float("inf"))
    o_idx, o_dist = oracle.knn(X, Q, k, w)
    assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all(), (N, D)
print("FP32_OK")
"""


def test_packed_fp32_filter_still_serves_mid_sized_sets_when_asked():
    """With MPB200_KNN_TC=0 the FFMA2 filter takes the sets the tensor-core scan serves by default (it remains the
    path of 2-D / 3-D sets, of more than 7 dimensions and of 12 < k <= 28): same answers as the oracle."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    env = dict(os.environ, MPB200_KNN_TC="0")
    out = subprocess.run([sys.executable, "-c", FP32_CHILD % ROOT], env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "FP32_OK" in out.stdout
