"""The opt-in tensor-core form of the kNN filter (MPB200_KNN_TC=1; tcgen05.mma kind::tf32 on three-way split
operands, accumulator in tensor memory) returns the same neighbours and bit-identical distances as the exact
fp64 scan.  The switch is read once per process, so the check runs in a child process."""
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
from motion_planners_b200 import _lib, ops
# the switch took: the workspace now has room for the 96-byte-per-sample operand tiles
assert _lib.lib().mpb_knn_workspace_bytes(30000, 3000, 7) > 30000 * 96
for N, M, D, k, seed in [(30000, 3000, 7, 11, 0), (20011, 2500, 6, 11, 1), (9000, 1500, 3, 5, 2), (8192, 130, 2, 12, 3)]:
    rs = np.random.RandomState(seed)
    x = torch.from_numpy(rs.uniform(size=(N, D))).cuda()
    q = torch.cat([x[: M // 2], torch.from_numpy(rs.uniform(-0.1, 1.1, size=(M - M // 2, D))).cuda()])
    w = rs.uniform(0.5, 2.0, size=D)
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    index = ops.KnnIndex(x, w)
    idx, dist = index.search(q, k, float("inf"), stats=stats)
    ref_idx, ref_dist = ops.knn(x, q, k, weights=w, method="f64")
    assert torch.equal(idx, ref_idx) and torch.equal(dist, ref_dist), (N, D)
    assert int(stats[0]) < M // 50, int(stats[0])        # the certificate holds for (almost) every query
    r = float(dist[:, k // 2].median())
    idx_r, dist_r = index.search(q, k, r)
    ref_idx_r, ref_dist_r = ops.knn(x, q, k, weights=w, max_dist=r, method="f64")
    assert torch.equal(idx_r, ref_idx_r) and torch.equal(dist_r, ref_dist_r), (N, D, "bounded")
print("TC_OK")
"""


def test_tensor_core_filter_matches_exact_scan():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    env = dict(os.environ, MPB200_KNN_TC="1")
    out = subprocess.run([sys.executable, "-c", CHILD % ROOT], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "TC_OK" in out.stdout
