"""The tensor-core kNN filter at the sizes where it is the default -- the benchmarked roadmap build
(1M samples x 7-DoF, k + 1 = 11, 256 boxes) and the policy boundary (160k samples) -- against the CPU
oracle (oracle/mp_oracle.c, the restatement of nearest.py:74-85 / samplers.py:88-127 pinned by the
reference's own outputs): neighbour indices, fp64 distances, vertex and edge verdicts, bit for bit, on
disjoint slices of the query range (first, last = the ragged block, two interior)."""
import math

import numpy as np
import pytest

import oracle

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mp():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from motion_planners_b200 import _lib, ops, roadmap, workloads
    oracle.use_all_cores()

    class NS(object):
        pass
    ns = NS()
    ns.lib, ns.ops, ns.roadmap, ns.workloads = _lib.lib(), ops, roadmap, workloads
    return ns


def slices_of(n, width):
    """first, two interior (not aligned to the 128-query blocks), last"""
    return [slice(0, width), slice(n // 3 + 17, n // 3 + 17 + width), slice(2 * n // 3 - 5, 2 * n // 3 - 5 + width),
            slice(n - width, n)]


def check_knn_slices(X, idx, dist, k, weights=None, max_dist=math.inf, width=1000):
    idx, dist = idx.cpu().numpy(), dist.cpu().numpy()
    for sl in slices_of(X.shape[0], width):
        o_idx, o_dist = oracle.knn(X, X[sl], k, weights, max_dist)
        assert (idx[sl] == o_idx).all(), sl
        assert (dist[sl] == o_dist).all(), sl


def test_headline_roadmap_build_matches_oracle(mp):
    """bench.py's default step, as bench.py runs it, checked against the oracle."""
    world = mp.workloads.c4lite_world()
    X = mp.workloads.uniform_samples(1_000_000, 7, seed=0)
    assert mp.lib.mpb_knn_filter_kind(X.shape[0], 7, 11) == 1          # the tensor-core scan serves this build
    rm = mp.roadmap.DeviceRoadmap(world, k=10)
    rm.build(torch.from_numpy(X).cuda())
    torch.cuda.synchronize()
    assert int(rm.stats[0]) < 100                                       # certified by the filter, not re-done exactly
    check_knn_slices(X, rm.nbr_idx, rm.nbr_dist, 11)
    nbr = rm.nbr_idx.cpu().numpy()
    assert (nbr[:, 0] == np.arange(X.shape[0])).all()
    ow = oracle.World(world.dim, world.env_lo, world.env_hi, world.box_lo, world.box_hi, world.sph_c, world.sph_r,
                      world.buffer)
    vhit = rm.vertex_hit.cpu().numpy()
    ehit = rm.edge_hit.cpu().numpy()
    for sl in slices_of(X.shape[0], 1000):
        assert (vhit[sl] == ow.points_check(X[sl])).all()
        src = np.repeat(np.arange(sl.start, sl.stop), 11)
        dst = nbr[sl].reshape(-1)
        _, first, _ = ow.edges_check_indexed(X, src, dst)
        # ANY mode reports some colliding waypoint (>= 0) or -1; the verdict is what is compared
        assert ((ehit[sl].reshape(-1) >= 0) == (first >= 0)).all()
    # a second build on the same object reuses every buffer and gives the same bytes
    before = (rm.nbr_idx.clone(), rm.nbr_dist.clone(), rm.edge_hit.clone())
    rm.build(torch.from_numpy(X).cuda())
    assert torch.equal(before[0], rm.nbr_idx) and torch.equal(before[1], rm.nbr_dist)
    assert torch.equal(before[2], rm.edge_hit)


@pytest.mark.parametrize("N,D", [(32_768, 7), (32_768, 6), (24_576, 5), (16_384, 4), (160_000, 7), (500_000, 4)])
def test_policy_boundary_sizes(mp, N, D):
    """The smallest set of each dimension that the tensor-core scan takes by default (and two mid-sized ones);
    non-unit weights, a bounded search and k = 12 (the largest k of the 16-entry lists) ride along."""
    assert mp.lib.mpb_knn_filter_kind(N, D, 11) == 1
    rs = np.random.RandomState(N % 1000 + D)
    X = rs.uniform(size=(N, D))
    w = rs.uniform(0.5, 2.0, size=D)
    xs = torch.from_numpy(X).cuda()
    index = mp.ops.KnnIndex(xs, w)
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    idx, dist = index.search(xs, 11, stats=stats)
    check_knn_slices(X, idx, dist, 11, w, width=400)
    assert int(stats[0]) <= N // 1000
    r = float(dist[:, 6].median())
    idx_r, dist_r = index.search(xs, 12, r)
    check_knn_slices(X, idx_r, dist_r, 12, w, max_dist=r, width=400)
    assert int((idx_r < 0).sum()) > 0                                   # the bound did cut rows short


def test_queries_that_are_not_samples(mp):
    """Queries outside the sample set's bounding box and far from it, ragged query count."""
    N, D, M = 200_000, 7, 1531
    rs = np.random.RandomState(9)
    X = rs.uniform(size=(N, D))
    Q = np.concatenate([rs.uniform(-0.5, 1.5, size=(M - 31, D)), rs.uniform(5.0, 6.0, size=(31, D))])
    assert mp.lib.mpb_knn_filter_kind(N, D, 11) == 1
    idx, dist = mp.ops.knn(torch.from_numpy(X).cuda(), torch.from_numpy(Q).cuda(), 11)
    o_idx, o_dist = oracle.knn(X, Q, 11)
    assert (idx.cpu().numpy() == o_idx).all() and (dist.cpu().numpy() == o_dist).all()


def test_clustered_samples(mp):
    """Far from uniform: a few dense blobs on top of a thin background, duplicated points included (exact
    ties: index order decides)."""
    N, D = 220_000, 7
    rs = np.random.RandomState(21)
    centres = rs.uniform(size=(12, D))
    X = np.concatenate([centres[rs.randint(12, size=N - 20_000)] + 0.01 * rs.normal(size=(N - 20_000, D)),
                        rs.uniform(size=(20_000, D))])
    X[1000:1200] = X[5000:5200]
    X = X[rs.permutation(N)]
    xs = torch.from_numpy(X).cuda()
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    idx, dist = mp.ops.knn(xs, xs, 11, stats=stats)
    check_knn_slices(X, idx, dist, 11, width=500)


def test_tensor_core_key_error_within_certificate_budget(mp, capsys):
    """What the certificate ASSUMES about tcgen05.mma kind::tf32 (its eta budgets 32 * 2^-22 of R^2 + 2|q^|R per key
    for the TF32 split of the operands and the accumulation inside the tensor core) is measured here: every key
    the scan reads from tensor memory against the same expression in fp64, over more than 1e7 pairs per set,
    including sets that stress it (off-centre clusters, coordinates of mixed magnitude, far-away queries)."""
    rs = np.random.RandomState(3)
    N, D = 160_000, 7
    budget = 32.0 * 2.0 ** -22
    sets = {}
    sets["uniform"] = (rs.uniform(size=(N, D)), None, rs.uniform(size=(256, D)))
    X = rs.uniform(size=(N, D))
    X[: N // 2] = 0.93 + 1e-3 * rs.uniform(size=(N // 2, D))
    sets["off-centre cluster"] = (X, None, X[:256].copy())
    sets["mixed magnitudes"] = (rs.uniform(size=(N, D)), np.array([1e3, 1.0, 1e-3, 1.0, 30.0, 1.0, 0.1]),
                                rs.uniform(size=(256, D)))
    X = rs.normal(size=(N, D)) * np.array([1, 10, 100, 1, 0.1, 5, 1.0])
    sets["gaussian, far queries"] = (X, None, np.concatenate([X[:128], 50.0 + rs.uniform(size=(128, D))]))
    worst = 0.0
    for name, (X, w, Q) in sets.items():
        err, pairs = mp.ops.knn_tc_key_error(torch.from_numpy(X).cuda(), torch.from_numpy(Q).cuda(), w)
        assert pairs == N * Q.shape[0] and pairs > 10_000_000, (name, pairs)
        assert 0.0 < err <= budget, (name, err / 2.0 ** -22)
        worst = max(worst, err)
        with capsys.disabled():
            print("\n[tc key error] %-22s max = %.3f x 2^-22 over %d pairs (budget 32)" % (name, err / 2.0 ** -22, pairs))
    assert worst <= budget
