"""Parity of the CUDA path (through the C ABI) with the oracle and with the reference's golden
outputs.  Bit-exact everywhere: indices, booleans, waypoint counts and fp64 distances."""
import math

import numpy as np
import pytest

import oracle
from conftest import world_from_npz

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from motion_planners_b200 import ops as _ops
    return _ops


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def host(t):
    return t.cpu().numpy()


def gpu_world(ops, g_or_world):
    from motion_planners_b200.geometry import World
    w = g_or_world if isinstance(g_or_world, World) else world_from_npz(g_or_world, World)
    return ops.ObstacleSet(w), w


def oracle_world(w):
    return oracle.World(w.dim, w.env_lo, w.env_hi, w.box_lo, w.box_hi, w.sph_c, w.sph_r, w.buffer)


# ------------------------------------------------------------------------------- kNN
@pytest.mark.parametrize("method", ["f32filter", "f64"])
@pytest.mark.parametrize("D", [2, 3, 6, 7, 14])
def test_knn_golden(ops, golden, D, method):
    g = golden("knn_d%d" % D)
    X, w, k = g["samples"], g["weights"], int(g["k"])
    xs = dev(X)
    idx, dist = ops.knn(xs, xs, k, weights=w, method=method)
    assert (host(idx) == g["idx"]).all()
    assert (host(dist) == g["dist"]).all()
    idx, dist = ops.knn(xs, xs, k, weights=w, max_dist=float(g["radius"]), method=method)
    assert (host(idx) == g["idx_r"]).all()
    assert (host(dist) == g["dist_r"]).all()


@pytest.mark.parametrize("D,N,M,k", [(1, 3000, 500, 5), (2, 20000, 2000, 11), (4, 9000, 1500, 8), (6, 30000, 3000, 11),
                                     (7, 30000, 3000, 11), (8, 8000, 1000, 16), (9, 6000, 800, 11),
                                     (12, 6000, 800, 11), (16, 5000, 600, 32)])
def test_knn_matches_oracle(ops, D, N, M, k):
    rs = np.random.RandomState(D * 1000 + k)
    X = rs.uniform(size=(N, D))
    Q = np.concatenate([X[:M // 2], rs.uniform(-0.2, 1.2, size=(M - M // 2, D))])
    w = rs.uniform(0.5, 2.0, size=D)
    o_idx, o_dist = oracle.knn(X, Q, k, w)
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    idx, dist = ops.knn(dev(X), dev(Q), k, weights=w, stats=stats)
    assert (host(idx) == o_idx).all()
    assert (host(dist) == o_dist).all()
    if k <= 16:
        assert int(stats[0]) < M // 10          # the fp32 certificate holds for almost every query
    idx, dist = ops.knn(dev(X), dev(Q), k, weights=w, method="f64")
    assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all()


@pytest.mark.parametrize("metric", [1, 2])
@pytest.mark.parametrize("D", [2, 7, 9])
def test_knn_other_metrics_match_oracle(ops, D, metric):
    """Ranking by the value distance_fn returns (primitives / samplers procedures), stable on index."""
    rs = np.random.RandomState(77 + D)
    X = rs.uniform(size=(4000, D))
    X[100:140] = X[60:100]                                      # duplicates: exact distance ties
    Q = np.concatenate([X[:300], rs.uniform(size=(200, D))])
    w = rs.uniform(0.5, 2.0, size=D)
    for md in (np.inf, 0.35):
        o_idx, o_dist = oracle.knn_metric(X, Q, 12, w, metric, md)
        idx, dist = ops.knn(dev(X), dev(Q), 12, weights=w, metric=metric, max_dist=md)
        assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all()


def test_knn_certificate_falls_back_when_fp32_cannot_separate(ops):
    """Samples 1e-7 apart around an fp32-hostile spread: the filter cannot rank them, the
    certificate must fail and the exact kernel must still give the oracle's answer."""
    rs = np.random.RandomState(5)
    N, D, k = 4000, 3, 11
    X = rs.uniform(size=(N, D))
    X[:2000] = 0.9 + 1e-9 * rs.uniform(size=(2000, D))      # a tight off-centre cluster inside a unit-size cloud
    o_idx, o_dist = oracle.knn(X, X, k)
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    idx, dist = ops.knn(dev(X), dev(X), k, stats=stats)
    assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all()
    assert int(stats[0]) >= 1000


def test_knn_ties_broken_by_index(ops):
    """Lattice points: many exactly equal distances; the contract is (distance, index) order."""
    g = np.stack(np.meshgrid(*[np.arange(12) / 8.0] * 3, indexing="ij"), -1).reshape(-1, 3)
    X = np.concatenate([g, g[:50]])                           # plus exact duplicates
    o_idx, o_dist = oracle.knn(X, X, 9)
    for method in ("f32filter", "f64"):
        idx, dist = ops.knn(dev(X), dev(X), 9, method=method)
        assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all()


def test_knn_edge_cases(ops):
    rs = np.random.RandomState(0)
    X = rs.uniform(size=(5, 6))
    for method in ("f32filter", "f64"):
        idx, dist = ops.knn(dev(X), dev(X), 11, method=method)        # fewer samples than k
        o_idx, o_dist = oracle.knn(X, X, 11)
        assert (host(idx) == o_idx).all() and (host(dist) == o_dist).all()
        assert (host(idx)[:, 5:] == -1).all() and np.isinf(host(dist)[:, 5:]).all()
        idx, dist = ops.knn(dev(X), dev(np.zeros((0, 6))), 3, method=method)   # no queries
        assert idx.shape == (0, 3)
        idx, dist = ops.knn(dev(np.zeros((0, 6))), dev(X), 3, method=method)   # no samples
        assert (host(idx) == -1).all()
        idx, dist = ops.knn(dev(X), dev(X), 1, method=method)
        assert (host(idx)[:, 0] == np.arange(5)).all() and (host(dist) == 0).all()
        idx, dist = ops.knn(dev(X), dev(X), 3, max_dist=0.0, method=method)    # nothing is closer than 0
        assert (host(idx) == -1).all()
    with pytest.raises(Exception):
        ops.knn(dev(X), dev(X), 33)
    with pytest.raises(Exception):
        ops.knn(dev(X), dev(X[:, :5]), 3)
    with pytest.raises(TypeError):
        ops.knn(dev(X).float(), dev(X), 3)


def test_knn_full_size_properties(ops):
    """Config-2 scale (100k x 6-DoF, k = 11): properties that need no CPU answer, plus the
    exact kernel and the oracle on a slice."""
    from motion_planners_b200 import workloads
    X = workloads.uniform_samples(100000, 6, seed=0)
    xs = dev(X)
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    index = ops.KnnIndex(xs)
    idx, dist = index.search(xs, 11, stats=stats)
    idx2, dist2 = index.search(xs, 11)                                # idempotent, workspace reused
    assert torch.equal(idx, idx2) and torch.equal(dist, dist2)
    assert (idx[:, 0] == torch.arange(100000, device="cuda")).all() and (dist[:, 0] == 0).all()
    assert (dist[:, 1:] >= dist[:, :-1]).all()
    assert int(idx.min()) >= 0 and int(idx.max()) < 100000
    # distances recomputed from the returned indices are the kd-tree metric bit for bit
    sl = slice(5000, 5400)
    o_idx, o_dist = oracle.knn(X, X[sl], 11)
    assert (host(idx[sl]) == o_idx).all() and (host(dist[sl]) == o_dist).all()
    e_idx, e_dist = ops.knn(xs, xs[20000:24000], 11, method="f64")
    assert torch.equal(e_idx, idx[20000:24000]) and torch.equal(e_dist, dist[20000:24000])
    assert int(stats[0]) == 0


# ------------------------------------------------------------------------ collision_fn
@pytest.mark.parametrize("name", ["problem1", "boxes6d", "mixed3d", "boxes7d"])
def test_points_golden(ops, golden, name):
    g = golden("collide_" + name)
    obs, _ = gpu_world(ops, g)
    assert (host(obs.points_check(dev(g["points"]))) == g["hit"]).all()


def test_points_empty_and_no_obstacles(ops):
    from motion_planners_b200.geometry import World
    w = World(3, np.zeros(3), np.ones(3))
    obs = ops.ObstacleSet(w)
    assert obs.points_check(dev(np.zeros((0, 3)))).shape == (0,)
    pts = np.array([[0.5, 0.5, 0.5], [1.0, 0.0, 1.0], [1.0000001, 0.5, 0.5], [np.nan, 0.5, 0.5]])
    assert host(obs.points_check(dev(pts))).tolist() == [False, False, True, True]


# ------------------------------------------------------------------------------- edges
@pytest.mark.parametrize("name", ["problem1", "boxes6d", "mixed3d", "boxes7d"])
def test_edges_golden(ops, golden, name):
    g = golden("edges_" + name)
    obs, _ = gpu_world(ops, g)
    q1, q2 = dev(g["q1"]), dev(g["q2"])
    first, n = obs.edges_check_pairs(q1, q2, mode=ops.EDGE_FIRST)
    assert (host(n) == g["n"]).all() and (host(first) == g["first"]).all()
    last, _ = obs.edges_check_pairs(q1, q2, mode=ops.EDGE_LAST)
    assert (host(last) == g["last"]).all()
    anyhit, _ = obs.edges_check_pairs(q1, q2, mode=ops.EDGE_ANY, want_counts=False)
    assert ((host(anyhit) >= 0) == g["any_all"]).all()
    prm_any, _ = obs.edges_check_pairs(q1, q2, mode=ops.EDGE_ANY, skip_last=True)
    assert ((host(prm_any) >= 0) == g["any_prm"]).all()


@pytest.mark.parametrize("name,E", [("problem1", 40000), ("boxes6d", 60000), ("mixed3d", 40000), ("boxes7d", 40000)])
def test_edges_match_oracle(ops, golden, name, E):
    g = golden("edges_" + name)
    obs, w = gpu_world(ops, g)
    ow = oracle_world(w)
    rs = np.random.RandomState(9)
    N = 5000
    X = rs.uniform(-0.02, 1.02, size=(N, w.dim))
    src = rs.randint(0, N, size=E)
    near = rs.randint(0, N, size=E)
    short = X[src] + 0.06 * (rs.uniform(size=(E, w.dim)) - 0.5)    # roadmap-length edges ...
    q2 = np.where((np.arange(E) % 4 == 0)[:, None], X[near], short)  # ... and a quarter long ones
    n, first, last = ow.edges_check(X[src], q2)
    for skip in (False, True):
        n_s, first_s, last_s = ow.edges_check(X[src], q2, skip_last=skip)
        f, c = obs.edges_check_pairs(dev(X[src]), dev(q2), mode=ops.EDGE_FIRST, skip_last=skip)
        assert (host(c) == n_s).all() and (host(f) == first_s).all()
        l, _ = obs.edges_check_pairs(dev(X[src]), dev(q2), mode=ops.EDGE_LAST, skip_last=skip)
        assert (host(l) == last_s).all()
        a, _ = obs.edges_check_pairs(dev(X[src]), dev(q2), mode=ops.EDGE_ANY, skip_last=skip)
        assert ((host(a) >= 0) == (first_s >= 0)).all()
    # indexed form (edges as vertex pairs of a sample table)
    n_i, first_i, _ = ow.edges_check_indexed(X, src, near)
    f, c = obs.edges_check(dev(X), dev(src.astype(np.int64)), dev(near.astype(np.int64)), mode=ops.EDGE_FIRST)
    assert (host(c) == n_i).all() and (host(f) == first_i).all()
    assert 0 < (first >= 0).sum() < E


def test_edges_many_obstacles_and_long_edges(ops):
    """More than 64 obstacles of each kind (several cull groups) and edges far longer than the world."""
    from motion_planners_b200 import workloads
    w = workloads.boxes_world(3, 150, 0.03, 0.12, seed=4, n_sph=90, sph_r=(0.02, 0.06))
    obs, _ = gpu_world(ops, w)
    ow = oracle_world(w)
    rs = np.random.RandomState(2)
    E = 20000
    q1 = rs.uniform(size=(E, 3))
    q2 = rs.uniform(size=(E, 3))
    q2[:100] = q1[:100] + 40.0 * (rs.uniform(size=(100, 3)) - 0.5)
    q2[100:200] = q1[100:200]
    for skip in (False, True):
        n, first, last = ow.edges_check(q1, q2, skip_last=skip)
        f, c = obs.edges_check_pairs(dev(q1), dev(q2), mode=ops.EDGE_FIRST, skip_last=skip)
        l, _ = obs.edges_check_pairs(dev(q1), dev(q2), mode=ops.EDGE_LAST, skip_last=skip)
        assert (host(c) == n).all() and (host(f) == first).all() and (host(l) == last).all()


@pytest.mark.parametrize("name", ["problem1", "boxes6d", "mixed3d"])
def test_edges_knn_rows_equal_indexed_edges(ops, golden, name):
    """The per-row form (culling shared by a sample's edges) gives the verdicts of the per-edge form."""
    g = golden("edges_" + name)
    obs, w = gpu_world(ops, g)
    rs = np.random.RandomState(12)
    N, kk = 6000, 7
    X = rs.uniform(-0.01, 1.01, size=(N, w.dim))
    xs = dev(X)
    idx, _ = ops.knn(xs, xs, kk)
    idx[::13, 3] = -1                                            # missing neighbours
    idx[5::17, 5] = torch.randint(0, N, (idx[5::17, 5].numel(),), device="cuda")   # a few long edges
    off, rows = 1000, 4000
    nbr = idx[off:off + rows].contiguous()
    src = (torch.arange(rows, device="cuda") + off).repeat_interleave(kk)
    for mode in (ops.EDGE_FIRST, ops.EDGE_LAST):
        for skip in (False, True):
            ref, ref_n = obs.edges_check(xs, src, nbr.view(-1), mode=mode, skip_last=skip)
            hit, cnt = obs.edges_check_knn(xs, nbr, row_offset=off, mode=mode, skip_last=skip, want_counts=True)
            assert torch.equal(hit.view(-1), ref) and torch.equal(cnt.view(-1), ref_n)
    ref, _ = obs.edges_check(xs, src, nbr.view(-1), mode=ops.EDGE_ANY)
    hit, _ = obs.edges_check_knn(xs, nbr, row_offset=off, mode=ops.EDGE_ANY)
    assert torch.equal(hit.view(-1) >= 0, ref >= 0)
    assert 0 < int((ref >= 0).sum()) < ref.numel()


def test_edges_empty(ops):
    from motion_planners_b200 import workloads
    obs, _ = gpu_world(ops, workloads.c2_world())
    hit, n = obs.edges_check_pairs(dev(np.zeros((0, 6))), dev(np.zeros((0, 6))))
    assert hit.shape == (0,) and n.shape == (0,)


@pytest.mark.parametrize("D", [2, 6, 7])
def test_waypoints_golden(ops, golden, D):
    g = golden("extend_d%d" % D)
    q1, q2 = dev(g["q1"]), dev(g["q2"])
    from motion_planners_b200.geometry import World
    obs = ops.ObstacleSet(World(D, np.zeros(D), np.ones(D)))
    _, n = obs.edges_check_pairs(q1, q2)
    assert (host(n) == g["counts"]).all()
    assert (host(ops.edges_count(q1, q2)) == g["counts"]).all()
    off, way = ops.edges_waypoints(q1, q2)
    assert (host(way) == g["waypoints"]).all()              # the serial recurrence, bit for bit
    assert int(off[-1]) == g["waypoints"].shape[0]


def test_roadmap_full_size_properties(ops):
    """Config-2 roadmap: kNN edges of 100k samples, both directions checked; an edge's colliding
    status may depend on direction only through the endpoint it omits."""
    from motion_planners_b200 import workloads
    w = workloads.c2_world()
    obs, _ = gpu_world(ops, w)
    X = workloads.uniform_samples(100000, 6, seed=0)
    xs = dev(X)
    idx, _ = ops.knn(xs, xs, 11)
    src = torch.arange(100000, device="cuda").repeat_interleave(10)
    dst = idx[:, 1:].reshape(-1)
    fwd, n = obs.edges_check(xs, src, dst, mode=ops.EDGE_FIRST)
    bwd, n2 = obs.edges_check(xs, dst, src, mode=ops.EDGE_FIRST)
    assert torch.equal(n, n2)
    vert = obs.points_check(xs)
    # a free edge in one direction with both endpoints free is free in the other direction too,
    # up to the last-ulp differences of the two recurrences; count disagreements instead of asserting 0
    both_free = ~vert[src] & ~vert[dst]
    disagree = ((fwd >= 0) != (bwd >= 0)) & both_free
    assert int(disagree.sum()) < 1e-3 * src.numel()
    # slice against the oracle
    ow = oracle_world(w)
    sl = slice(300000, 320000)
    n_o, first_o, _ = ow.edges_check_indexed(X, host(src[sl]), host(dst[sl]))
    assert (host(n[sl]) == n_o).all() and (host(fwd[sl]) == first_o).all()
    assert (host(vert[:20000]) == ow.points_check(X[:20000])).all()


# ------------------------------------------------------------------------------- trees
@pytest.mark.parametrize("metric", [0, 1, 2])
def test_trees_nearest_matches_oracle(ops, metric):
    rs = np.random.RandomState(3)
    T, cap, D = 300, 257, 7
    nodes = rs.uniform(size=(T, cap, D))
    nodes[:, 10] = nodes[:, 3]                                   # duplicate nodes: first index must win
    sizes = rs.randint(1, cap + 1, size=T).astype(np.int32)
    targets = rs.uniform(size=(T, D))
    targets[:20] = nodes[:20, 3]
    w = rs.uniform(0.5, 2.0, size=D)
    idx, dist = ops.trees_nearest(dev(nodes), dev(sizes), dev(targets), weights=w, metric=metric)
    idx, dist = host(idx), host(dist)
    for t in range(T):
        i, d = oracle.nearest(nodes[t, :sizes[t]], targets[t], w, metric)
        assert idx[t] == i and dist[t] == d


@pytest.mark.parametrize("metric", [0, 1, 2])
def test_scan_nearest_radius_matches_oracle(ops, metric):
    rs = np.random.RandomState(4)
    for n in (1, 33, 1024, 5000):
        nodes = rs.uniform(size=(n, 3))
        q = rs.uniform(size=3)
        w = rs.uniform(0.5, 2.0, size=3)
        nearest, dist, count, lst = ops.scan_nearest_radius(dev(nodes), dev(q), 0.3, weights=w, metric=metric)
        i, d = oracle.nearest(nodes, q, w, metric)
        ridx, rd = oracle.radius(nodes, q, 0.3, w, metric)
        c = int(count)
        assert int(nearest) == i and host(dist)[i] == d
        assert c == len(ridx) and (host(lst)[:c] == ridx).all()
        assert (host(dist)[ridx] == rd).all()


def test_roadmap_shortest_path_matches_host_dijkstra(ops):
    """Device Bellman-Ford sweeps over the built roadmap against scipy's Dijkstra on the same usable edges:
    identical distances (the fixed point of d[v] = min fl(d[u] + w) is order independent up to the last ulp:
    compared at 1e-12) and a returned path that is a chain of usable edges with the reported cost."""
    from scipy.sparse import coo_matrix
    from scipy.sparse.csgraph import dijkstra
    from motion_planners_b200 import workloads
    from motion_planners_b200.roadmap import DeviceRoadmap
    w = workloads.boxes_world(4, 40, 0.1, 0.3, seed=2)
    N, k = 20000, 8
    X = workloads.uniform_samples(N, 4, seed=9)
    rm = DeviceRoadmap(w, k=k).build(dev(X))
    idx, dist = host(rm.nbr_idx), host(rm.nbr_dist)
    ehit, vhit = host(rm.edge_hit), host(rm.vertex_hit)
    rows = np.repeat(np.arange(N), k + 1).reshape(N, k + 1)
    ok = (idx >= 0) & (idx != rows) & (ehit < 0) & ~vhit[rows] & ~vhit[np.clip(idx, 0, N - 1)]
    g = coo_matrix((dist[ok], (rows[ok], idx[ok])), shape=(N, N)).tocsr()
    g = g.maximum(g.T)                                   # both directions
    free = np.flatnonzero(~vhit)
    src, dst = int(free[0]), int(free[-1])
    ref = dijkstra(g, directed=False, indices=src)
    d, path = rm.shortest_path(src, dst)
    d = host(d)
    both = np.isfinite(ref)
    assert (np.isfinite(d) == both).all()
    assert np.allclose(d[both], ref[both], rtol=1e-12, atol=0)
    assert path is not None and path[0] == src and path[-1] == dst
    cost = 0.0
    for a, b in zip(path[:-1], path[1:]):
        assert g[a, b] > 0                               # a usable edge of the roadmap
        cost += g[a, b]
    assert abs(cost - d[dst]) <= 1e-9 * d[dst]
    assert both.sum() > N // 2


def test_roadmap_build_from_host_returns_what_build_leaves_on_the_device(ops):
    """The host-buffer call (bench.py's e2e): pinned samples in; neighbour matrix (int32, copied on a second stream
    while the collision checks run), vertex and edge flags back in pinned buffers -- equal to the device-resident
    results of the same build, on two calls in a row (the second reuses the host buffers)."""
    import torch
    from motion_planners_b200 import workloads
    from motion_planners_b200.roadmap import DeviceRoadmap
    w = workloads.boxes_world(6, 64, 0.2, 0.6, seed=0)
    X = workloads.uniform_samples(30000, 6, seed=4)
    ref = DeviceRoadmap(w, k=10).build(dev(X))
    pinned = torch.from_numpy(X).pin_memory()
    rm = DeviceRoadmap(w, k=10)
    for _ in range(2):
        nbr, vhit, ehit = rm.build_from_host(pinned)
        assert nbr.dtype == torch.int32 and not nbr.is_cuda
        assert (nbr.numpy() == host(ref.nbr_idx)).all()
        assert (vhit.numpy() == host(ref.vertex_hit)).all()
        assert (ehit.numpy() == host(ref.edge_hit)).all()
