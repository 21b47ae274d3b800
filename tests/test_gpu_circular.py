"""Circular (wrap-around) joints on the CUDA path: waypoints, collision verdicts along them and the
three-image kNN, against goldens made by the unmodified reference and against the oracle at larger sizes;
then whole lazy_prm() runs with circular joints.  Bit-exact everywhere."""
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


def circ_dict(g):
    lo, hi = g["circ_lo"], g["circ_hi"]
    return {d: (float(lo[d]), float(hi[d])) for d in range(len(lo)) if np.isfinite(lo[d])}


@pytest.mark.parametrize("name", ["d3", "d6", "d7"])
def test_circular_waypoints_and_verdicts_golden(ops, golden, name):
    """get_extend_fn(circular)(q1, q2) (tkinter/samplers.py:113-127) and collision_fn along it."""
    from motion_planners_b200.geometry import World
    g = golden("circular_" + name)
    circ = circ_dict(g)
    q1, q2 = dev(g["q1"]), dev(g["q2"])
    counts = ops.edges_count(q1, q2, circular=circ)
    assert (host(counts) == g["counts"]).all()
    off, way = ops.edges_waypoints(q1, q2, circular=circ)
    assert host(off)[-1] == len(g["waypoints"]) and (host(way) == g["waypoints"]).all()
    if len(g["first"]):
        obs = ops.ObstacleSet(world_from_npz(g, World))
        first, n = obs.edges_check_pairs(q1, q2, mode=ops.EDGE_FIRST, circular=circ)
        last, _ = obs.edges_check_pairs(q1, q2, mode=ops.EDGE_LAST, circular=circ)
        anyhit, _ = obs.edges_check_pairs(q1, q2, mode=ops.EDGE_ANY, circular=circ)
        assert (host(n) == g["counts"]).all()
        assert (host(first) == g["first"]).all() and (host(last) == g["last"]).all()
        assert ((host(anyhit) >= 0) == (g["first"] >= 0)).all()
        # indexed form over a sample table = the pair form
        table = torch.cat([q1, q2])
        E = q1.shape[0]
        src = torch.arange(E, device="cuda")
        dst = src + E
        dst[5] = -1
        f2, n2 = obs.edges_check(table, src, dst, mode=ops.EDGE_FIRST, circular=circ)
        keep = np.arange(E) != 5
        assert (host(f2)[keep] == g["first"][keep]).all() and host(f2)[5] == -1 and host(n2)[5] == 0


@pytest.mark.parametrize("name", ["d3", "d6", "d7"])
def test_circular_knn_golden(ops, golden, name):
    """KDNeighbors(circular=...).query_neighbors (nearest.py:7-20,74-85)."""
    g = golden("circular_" + name)
    circ = circ_dict(g)
    X, w, k = g["samples"], g["weights"], int(g["k"])
    xs = dev(X)
    idx, dist = ops.knn(xs, xs, k, weights=w, circular=circ)
    assert (host(idx) == g["idx"]).all() and (host(dist) == g["dist"]).all()
    nr = g["idx_r"].shape[0]
    idx_r, dist_r = ops.knn(xs, xs[:nr], k, weights=w, circular=circ, max_dist=float(g["radius"]))
    assert (host(idx_r) == g["idx_r"]).all() and (host(dist_r) == g["dist_r"]).all()


@pytest.mark.parametrize("D,circ", [(2, {0: (0., 1.), 1: (0., 1.)}), (7, {1: (-1., 2.), 5: (0., 1.)}),
                                    (12, {0: (0., 1.), 3: (0., 1.), 11: (0.25, 0.75)})])
def test_circular_matches_oracle(ops, D, circ):
    """Larger seeded batches: the oracle spells out all 3^c images; the kernel takes per-coordinate minima."""
    from motion_planners_b200 import workloads
    rs = np.random.RandomState(50 + D)
    N, M, k = 6000, 700, 9
    X = rs.uniform(size=(N, D))
    Q = rs.uniform(-0.2, 1.2, size=(M, D))
    w = rs.uniform(0.5, 2., size=D)
    idx, dist = ops.knn(dev(X), dev(Q), k, weights=w, circular=circ)
    oi, od = oracle.knn_circular(X, Q, k, circ, w)
    assert (host(idx) == oi).all() and (host(dist) == od).all()
    if D <= 7:
        world = workloads.boxes_world(D, 24, 0.2, 0.5, seed=4)
        obs = ops.ObstacleSet(world)
        ow = oracle.World(world.dim, world.env_lo, world.env_hi, world.box_lo, world.box_hi, world.sph_c,
                          world.sph_r, world.buffer)
        E = 20000
        q1 = rs.uniform(size=(E, D))
        q2 = rs.uniform(size=(E, D))
        q2[::2] = q1[::2] + 0.1 * (rs.uniform(size=(E // 2, D)) - 0.5)
        n, first, last = ow.edges_check_circular(q1, q2, circ)
        gf, gn = obs.edges_check_pairs(dev(q1), dev(q2), mode=ops.EDGE_FIRST, circular=circ)
        gl, _ = obs.edges_check_pairs(dev(q1), dev(q2), mode=ops.EDGE_LAST, circular=circ)
        assert (host(gn) == n).all() and (host(gf) == first).all() and (host(gl) == last).all()
        sf, _ = obs.edges_check_pairs(dev(q1), dev(q2), mode=ops.EDGE_ANY, skip_last=True, circular=circ)
        _, f_skip, _ = ow.edges_check_circular(q1, q2, circ, skip_last=True)
        assert ((host(sf) >= 0) == (f_skip >= 0)).all()


def test_circular_argument_checks(ops):
    from motion_planners_b200 import _lib
    x = dev(np.random.RandomState(0).uniform(size=(10, 3)))
    with pytest.raises(ValueError):
        ops.knn(x, x, 2, circular={5: (0., 1.)})
    with pytest.raises(_lib.MpbError):
        ops.knn(x, x, 2, circular={0: (1., 0.)})
    # no circular coordinate = the plain entry points
    a, b = ops.knn(x, x, 2, circular={}, method="f64")
    c, d = ops.knn(x, x, 2, method="f64")
    assert (host(a) == host(c)).all() and (host(b) == host(d)).all()


@pytest.mark.parametrize("name", ["lazy_prm_circ_problem1_s11", "lazy_prm_circ_mixed3d_s12"])
def test_lazy_prm_with_circular_joints_matches_reference_run(golden, name):
    """lazy_prm(..., circular=...) (lazy_prm.py:253-337): same edges, same path bit for bit, same caches."""
    import motion_planners_b200.lazy_prm as lazy_prm
    import motion_planners_b200.primitives as primitives
    from motion_planners_b200.geometry import World
    g = golden(name)
    circ = circ_dict(g)
    world = world_from_npz(g, World)
    samples = [np.array(r) for r in g["samples"]]
    start, goal = samples[0], samples[1]
    w = np.ones(len(start))
    ext = primitives.get_extend_fn(circular=circ)
    col = primitives.GpuCollisionFn(world)
    cost_fn = primitives.get_distance_fn(w, p_norm=2)
    rm = lazy_prm.Roadmap(ext, weights=w, cost_fn=cost_fn, max_degree=int(g["max_degree"]), circular=circ)
    rm.add_samples(samples)

    def no_more_samples():
        raise AssertionError("the roadmap is already full; the reference draws nothing here either")

    sol = lazy_prm.lazy_prm(start, goal, no_more_samples, ext, col, roadmap=rm, num_samples=len(samples) - 2,
                            weights=w, circular=circ, lazy=bool(g["lazy"]), verbose=False)
    assert (sol.path is not None) == bool(g["solved"])
    assert (np.array(sorted(sol.edges), np.int64).reshape(-1, 2) == g["edges"]).all()
    path = np.array(sol.path)
    assert path.shape == g["path"].shape and (path == g["path"]).all()
    cv = np.array(sorted((v, int(f)) for v, f in sol.colliding_vertices.items()), np.int64).reshape(-1, 2)
    ce = np.array(sorted((a, b, int(f)) for (a, b), f in sol.colliding_edges.items()), np.int64).reshape(-1, 3)
    assert cv.shape == g["colliding_vertices"].shape and (cv == g["colliding_vertices"]).all()
    assert ce.shape == g["colliding_edges"].shape and (ce == g["colliding_edges"]).all()
