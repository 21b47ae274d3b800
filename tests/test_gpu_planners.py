"""Planner-level parity: the drop-in Roadmap / lazy_prm / prm on the CUDA kernels against runs of the
unmodified reference stored under tests/golden/ (same seeded inputs, identical observable state)."""
import numpy as np
import pytest

from conftest import world_from_npz

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mp():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import motion_planners_b200.lazy_prm as lazy_prm
    import motion_planners_b200.primitives as primitives
    import motion_planners_b200.prm as prm
    from motion_planners_b200.geometry import World

    class NS(object):
        pass
    ns = NS()
    ns.lazy_prm, ns.primitives, ns.prm, ns.World = lazy_prm, primitives, prm, World
    return ns


def _world(mp, g):
    return world_from_npz(g, mp.World)


def _edge_array(edges):
    return np.array(sorted(edges), np.int64).reshape(-1, 2)


@pytest.mark.parametrize("name", ["problem1", "boxes6d", "mixed3d"])
def test_roadmap_add_samples_and_check_roadmap(mp, golden, name):
    """lazy_prm.py:148-160 + 211-215: identical edge set and identical colliding_* dictionaries."""
    g = golden("roadmap_" + name)
    world = _world(mp, g)
    X = g["samples"]
    D = X.shape[1]
    w = np.ones(D)
    ext = mp.primitives.get_extend_fn()
    col = mp.primitives.GpuCollisionFn(world)
    rm = mp.lazy_prm.Roadmap(ext, weights=w, cost_fn=mp.primitives.get_distance_fn(w, p_norm=2),
                             max_degree=int(g["max_degree"]))
    new_edges = rm.add_samples(list(X))
    assert (_edge_array(rm.edges) == g["edges"]).all()
    assert new_edges == rm.edges
    rm.check_roadmap(col)
    cv = np.array([rm.colliding_vertices[v] for v in range(len(X))])
    assert (cv == g["colliding_vertices"]).all()
    ce = np.array(sorted((a, b, int(f)) for (a, b), f in rm.colliding_edges.items()), np.int64)
    assert ce.shape == g["colliding_edges"].shape and (ce == g["colliding_edges"]).all()
    # incremental insertion gives the same graph as the reference would build incrementally: here just
    # check that a bounded-distance roadmap drops the far neighbours exactly like scipy's upper bound
    rm2 = mp.lazy_prm.Roadmap(ext, weights=w, cost_fn=mp.primitives.get_distance_fn(w, p_norm=2),
                              max_degree=int(g["max_degree"]), max_distance=float(g["max_distance"]))
    rm2.add_samples(list(X))
    assert (_edge_array(rm2.edges) == g["edges_max_distance"]).all()


@pytest.mark.parametrize("name", ["lazy_prm_problem1_s5", "lazy_prm_problem1_s6", "lazy_prm_boxes6d_s7"])
def test_lazy_prm_matches_reference_run(mp, golden, name):
    """Full lazy_prm() (lazy_prm.py:253-337): same path bit for bit, same cost, same caches."""
    g = golden(name)
    world = _world(mp, g)
    samples = [np.array(r) for r in g["samples"]]
    start, goal = samples[0], samples[1]
    D = len(start)
    w = np.ones(D)
    ext = mp.primitives.get_extend_fn()
    col = mp.primitives.GpuCollisionFn(world)
    cost_fn = mp.primitives.get_distance_fn(w, p_norm=2)
    rm = mp.lazy_prm.Roadmap(ext, weights=w, cost_fn=cost_fn, max_degree=int(g["max_degree"]))
    rm.add_samples(samples)

    def no_more_samples():
        raise AssertionError("the roadmap is already full; the reference draws nothing here either")

    sol = mp.lazy_prm.lazy_prm(start, goal, no_more_samples, ext, col, roadmap=rm, num_samples=len(samples) - 2,
                               weights=w, lazy=bool(g["lazy"]), verbose=False)
    assert (sol.path is not None) == bool(g["solved"])
    assert (_edge_array(sol.edges) == g["edges"]).all()
    if sol.path is not None:
        path = np.array(sol.path)
        assert path.shape == g["path"].shape and (path == g["path"]).all()
        cost = sum(cost_fn(a, b) for a, b in zip(path[:-1], path[1:]))
        assert abs(cost - float(g["cost"])) <= 1e-9 * float(g["cost"])
    cv = np.array(sorted((v, int(f)) for v, f in sol.colliding_vertices.items()), np.int64).reshape(-1, 2)
    ce = np.array(sorted((a, b, int(f)) for (a, b), f in sol.colliding_edges.items()), np.int64).reshape(-1, 3)
    assert cv.shape == g["colliding_vertices"].shape and (cv == g["colliding_vertices"]).all()
    assert ce.shape == g["colliding_edges"].shape and (ce == g["colliding_edges"]).all()


def test_roadmap_scalar_check_methods(mp, golden):
    """check_vertex / check_edge / check_intermediate called one key at a time agree with the batch."""
    g = golden("roadmap_mixed3d")
    world = _world(mp, g)
    X = g["samples"][:120]
    w = np.ones(X.shape[1])
    ext = mp.primitives.get_extend_fn()
    col = mp.primitives.GpuCollisionFn(world)
    a = mp.lazy_prm.Roadmap(ext, weights=w, max_degree=4)
    b = mp.lazy_prm.Roadmap(ext, weights=w, max_degree=4)
    a.add_samples(list(X))
    b.add_samples(list(X))
    assert a.edges == b.edges
    b.check_roadmap(col)
    for v in a.vertices:
        assert a.check_vertex(v, col) == (not b.colliding_vertices[v])
    for (v1, v2) in list(a.edges)[:150]:
        assert a.check_edge(v1, v2, col) == (not b.colliding_edges[v1, v2])
    v1, v2 = next(iter(a.edges))
    path = a.get_path(v1, v2)
    flags = [a.check_intermediate(v1, v2, i, col) for i in range(len(path))]
    assert all(flags) == (not b.colliding_edges[v1, v2])
    assert len(a.get_path(v2, v1)) == len(path) and (a.get_path(v2, v1)[0] == path[-1]).all()


@pytest.mark.parametrize("name", ["prm_problem1_s11", "prm_problem1_s12", "prm_boxes6d_s13", "prm_problem1_s0"])
def test_prm_matches_reference_run(mp, golden, name):
    """prm() = DegreePRM.grow + A* (prm.py:237-282): same edges in the same order, same path.
    prm_problem1_s0 is BASELINE.json's configs[0] at full size (2-D demo fixture, 2000 samples, degree 10)."""
    g = golden(name)
    world = _world(mp, g)
    samples = [tuple(r) for r in g["samples"]]
    D = len(samples[0])
    dist = mp.primitives.get_distance_fn(np.ones(D))
    ext = mp.primitives.get_extend_fn()
    col = mp.primitives.GpuCollisionFn(world)
    rm = mp.prm.DegreePRM(dist, ext, col, samples=samples, target_degree=int(g["target_degree"]))
    edges = np.array([(e.v1.index, e.v2.index) for e in rm.edges], np.int64).reshape(-1, 2)
    assert edges.shape == g["edges"].shape and (edges == g["edges"]).all()
    path = rm(samples[0], samples[1])
    assert (path is not None) == bool(g["solved"])
    if path is not None:
        arr = np.array([np.asarray(q) for q in path])
        assert arr.shape == g["path"].shape and (arr == g["path"]).all()
        cost = sum(dist(a, b) for a, b in zip(arr[:-1], arr[1:]))
        assert abs(cost - float(g["cost"])) <= 1e-9 * max(1.0, float(g["cost"]))
    # the entry point itself, fed by a sample_fn that replays the stored draws
    it = iter(samples[2:])
    path2 = mp.prm.prm(samples[0], samples[1], dist, lambda: next(it), ext, col,
                       target_degree=int(g["target_degree"]), num_samples=len(samples) - 2)
    assert (path2 is None) == (path is None)
    if path is not None:
        assert all((np.asarray(a) == np.asarray(b)).all() for a, b in zip(path, path2))


def test_prm_small_candidate_budget_takes_the_full_ranking_path(mp, golden):
    """With only 4 candidates per vertex in the first batch, colliding samples must fall through to the
    full distance scan and still reproduce the reference roadmap."""
    g = golden("prm_problem1_s12")                     # unfiltered samples: some lie inside obstacles
    world = _world(mp, g)
    samples = [tuple(r) for r in g["samples"]]
    dist = mp.primitives.get_distance_fn(np.ones(2))

    class Small(mp.prm.DegreePRM):
        CANDIDATES = 4

    rm = Small(dist, mp.primitives.get_extend_fn(), mp.primitives.GpuCollisionFn(world), samples=samples,
               target_degree=int(g["target_degree"]))
    edges = np.array([(e.v1.index, e.v2.index) for e in rm.edges], np.int64).reshape(-1, 2)
    assert edges.shape == g["edges"].shape and (edges == g["edges"]).all()


# ----------------------------------------------------------------------------- tree planners
def _tree_primitives(mp, world):
    D = world.dim
    return (mp.primitives.get_distance_fn(np.ones(D)), mp.primitives.get_extend_fn(),
            mp.primitives.GpuCollisionFn(world))


def _paths_from(g):
    off = np.concatenate([[0], np.cumsum(g["path_len"])])
    return [g["paths"][off[i]:off[i + 1]] if g["solved"][i] else None for i in range(len(g["solved"]))]


@pytest.mark.parametrize("name", ["problem1", "boxes7d", "mixed3d"])
def test_rrt_connect_matches_reference_runs(mp, golden, name):
    """rrt_connect() under the same seeds as the reference run (rrt_connect.py:31-72): identical paths,
    both through the drop-in entry point (global generators) and through the batched form fed with the
    pre-drawn per-query streams."""
    import random
    from motion_planners_b200.rrt_connect import rrt_connect, rrt_connect_batch
    g = golden("rrt_connect_" + name)
    world = _world(mp, g)
    dist, ext, col = _tree_primitives(mp, world)
    env, _ = world.shapes()
    ref_paths = _paths_from(g)
    iters = int(g["max_iterations"])
    D = world.dim
    # drop-in entry point: seeds the global generators exactly as the fixture generator did
    for q in range(len(ref_paths)):
        seed = int(g["seeds"][q])
        random.seed(seed)
        np.random.seed(seed)
        sample_fn = mp.primitives.get_sample_fn(env, obstacles=[])
        path = rrt_connect(g["starts"][q], g["goals"][q], dist, sample_fn, ext, col, max_iterations=iters)
        assert (path is None) == (ref_paths[q] is None)
        if path is not None:
            arr = np.array(path)
            assert arr.shape == ref_paths[q].shape and (arr == ref_paths[q]).all()
    # batched form: all queries in lock step, each with its own streams
    B = len(ref_paths)
    u = np.empty((B, iters))
    tg = np.empty((B, iters, D))
    for q in range(B):
        seed = int(g["seeds"][q])
        r = random.Random(seed)
        u[q] = [r.random() for _ in range(iters)]
        tg[q] = np.random.RandomState(seed).uniform(size=(iters, D))
    for fused in (True, False):
        paths = rrt_connect_batch(g["starts"], g["goals"], dist, ext, col, u, tg, fused=fused)
        for q in range(B):
            assert (paths[q] is None) == (ref_paths[q] is None)
            if paths[q] is not None:
                arr = np.array(paths[q])
                assert arr.shape == ref_paths[q].shape and (arr == ref_paths[q]).all()


def test_rrt_connect_fused_equals_lockstep_on_a_large_batch(mp):
    """512 7-DoF queries (some with colliding endpoints): the one-launch kernel and the composition of the
    stand-alone kernels must build identical trees and report identical results."""
    import random
    from motion_planners_b200 import workloads
    from motion_planners_b200.rrt_connect import rrt_connect_batch
    world = workloads.c3_world()
    dist, ext, col = _tree_primitives(mp, world)
    rs = np.random.RandomState(8)
    B, iters, D = 512, 40, 7
    starts, goals = rs.uniform(size=(B, D)), rs.uniform(size=(B, D))
    u = rs.uniform(size=(B, iters))
    tg = rs.uniform(size=(B, iters, D))
    _, a = rrt_connect_batch(starts, goals, dist, ext, col, u, tg, return_state=True, retrace=False, fused=True)
    _, b = rrt_connect_batch(starts, goals, dist, ext, col, u, tg, return_state=True, retrace=False, fused=False)
    assert torch.equal(a.solved, b.solved)
    assert 0 < int(a.solved.sum()) < B
    sel = a.solved
    assert torch.equal(a.last_a[sel], b.last_a[sel]) and torch.equal(a.last_b[sel], b.last_b[sel])
    assert torch.equal(a.swapped[sel], b.swapped[sel])
    # trees of solved queries are identical node for node (unsolved ones too, up to the iteration budget)
    assert torch.equal(a.sizes, b.sizes)
    for q in torch.nonzero(sel).flatten()[:40].tolist():
        for t in (2 * q, 2 * q + 1):
            n = int(a.sizes[t])
            assert torch.equal(a.nodes[t, :n], b.nodes[t, :n]) and torch.equal(a.parents[t, :n], b.parents[t, :n])


@pytest.mark.parametrize("name", ["problem1", "mixed3d"])
def test_rrt_star_matches_reference_run(mp, golden, name):
    """rrt_star() as executed on Python 3 (rrt_star.py:86-146): identical path, cost within 1e-9."""
    import random
    from motion_planners_b200.rrt_star import rrt_star
    g = golden("rrt_star_" + name)
    world = _world(mp, g)
    dist, ext, col = _tree_primitives(mp, world)
    env, _ = world.shapes()
    seed = int(g["seed"])
    random.seed(seed)
    np.random.seed(seed)
    sample_fn = mp.primitives.get_sample_fn(env, obstacles=[])
    path = rrt_star(g["start"], g["goal"], dist, sample_fn, ext, col, radius=float(g["radius"]),
                    max_iterations=int(g["max_iterations"]))
    assert (path is not None) == bool(g["solved"])
    arr = np.array(path)
    assert arr.shape == g["path"].shape and (arr == g["path"]).all()
    cost = sum(dist(a, b) for a, b in zip(arr[:-1], arr[1:]))
    assert abs(cost - float(g["cost"])) <= 1e-9 * float(g["cost"])


@pytest.mark.parametrize("name", ["problem1", "mixed3d"])
def test_birrt_matches_reference_runs(mp, golden, name):
    """birrt() = restarts of rrt_connect + shortcut smoothing (meta.py:46-86, smoothing.py:51-144)."""
    import io
    import random
    from contextlib import redirect_stdout
    from motion_planners_b200.rrt_connect import birrt
    g = golden("birrt_" + name)
    world = _world(mp, g)
    dist, ext, col = _tree_primitives(mp, world)
    env, _ = world.shapes()
    ref_paths = _paths_from(g)
    for q in range(len(ref_paths)):
        seed = int(g["seeds"][q])
        random.seed(seed)
        np.random.seed(seed)
        sample_fn = mp.primitives.get_sample_fn(env, obstacles=[])
        with redirect_stdout(io.StringIO()):
            path = birrt(g["starts"][q], g["goals"][q], dist, sample_fn, ext, col,
                         max_iterations=int(g["max_iterations"]), smooth=20, restarts=2)
        assert (path is None) == (ref_paths[q] is None)
        if path is not None:
            arr = np.array(path)
            assert arr.shape == ref_paths[q].shape and (arr == ref_paths[q]).all()


def test_solve_dispatch(mp, golden):
    from motion_planners_b200.meta import solve
    g = golden("collide_problem1")
    world = _world(mp, g)
    dist, ext, col = _tree_primitives(mp, world)
    env, _ = world.shapes()
    import random
    random.seed(3)
    np.random.seed(3)
    sample_fn = mp.primitives.get_sample_fn(env, obstacles=[])
    start, goal = np.array([0.1, 0.1]), np.array([0.8, 0.8])
    for alg in ("rrt_connect", "birrt", "prm", "lazy_prm"):
        path = solve(start, goal, dist, sample_fn, ext, col, algorithm=alg, max_iterations=200, num_samples=200,
                     weights=np.ones(2) if alg == "lazy_prm" else None)
        assert path is not None
        arr = np.array([np.asarray(q) for q in path])
        assert not col.batch(arr).any()
        assert np.allclose(arr[0], start) and np.allclose(arr[-1], goal, atol=1e-9)
    with pytest.raises(NotImplementedError):
        solve(start, goal, dist, sample_fn, ext, col, algorithm="lattice")
    assert solve(np.array([0.5, 0.5]), goal, dist, sample_fn, ext, col) is False   # start inside the centre box


@pytest.mark.parametrize("name", ["problem1", "boxes7d"])
def test_smooth_paths_batch_equals_per_path_smoothing(mp, golden, name):
    """Lock-step smoothing of many paths, each with its own seeded generators, equals smooth_path run path
    by path under the same seeds (smooth_path itself is pinned to the reference through the birrt fixtures)."""
    import random
    from motion_planners_b200.smoothing import smooth_path, smooth_paths_batch
    g = golden("rrt_connect_" + name)
    world = _world(mp, g)
    dist, ext, col = _tree_primitives(mp, world)
    paths = [list(p) for p in _paths_from(g) if p is not None]
    seeds = [500 + i for i in range(len(paths))]
    expected = []
    for p, s in zip(paths, seeds):
        random.seed(s)
        np.random.seed(s)
        expected.append(smooth_path(p, ext, col, distance_fn=dist, max_iterations=30))
    got = smooth_paths_batch(paths, ext, col, dist, seeds, max_iterations=30)
    assert len(got) == len(expected)
    shorter = 0
    for a, b, p in zip(got, expected, paths):
        a, b = np.array(a), np.array(b)
        assert a.shape == b.shape and (a == b).all()
        shorter += len(a) < len(p)
        assert not col.batch(a).any()
    assert shorter > 0


def test_device_roadmap_view_equals_dropin_roadmap(mp, golden, tmp_path):
    """DeviceRoadmap -> save/load -> reference-shaped Roadmap equals the reference's own Roadmap after
    add_samples + check_roadmap (golden), without recomputing what is already on the device."""
    from motion_planners_b200.roadmap import DeviceRoadmap
    g = golden("roadmap_boxes6d")
    world = _world(mp, g)
    X = g["samples"]
    k = int(g["max_degree"])
    built = DeviceRoadmap(world, k=k).build(torch.from_numpy(X).cuda())
    f = str(tmp_path / "roadmap.pt")
    built.save(f)
    loaded = DeviceRoadmap.load(f)
    assert torch.equal(loaded.nbr_idx, built.nbr_idx) and torch.equal(loaded.edge_hit, built.edge_hit)
    ext = mp.primitives.get_extend_fn()
    col = mp.primitives.GpuCollisionFn(world)
    view = loaded.to_reference_roadmap(ext, col)
    assert (_edge_array(view.edges) == g["edges"]).all()
    cv = np.array([view.colliding_vertices[v] for v in range(len(X))])
    assert (cv == g["colliding_vertices"]).all()
    ce = np.array(sorted((a, b, int(fl)) for (a, b), fl in view.colliding_edges.items()), np.int64)
    assert ce.shape == g["colliding_edges"].shape and (ce == g["colliding_edges"]).all()
    # and it is a working Roadmap: a query over it finds a collision-free path
    d, path = loaded.shortest_path(int(np.flatnonzero(~cv)[0]), int(np.flatnonzero(~cv)[-1]))
    assert path is None or not cv[path].any()


@pytest.mark.parametrize("name", ["problem1", "boxes7d", "mixed3d"])
def test_rrt_matches_reference_runs(mp, golden, name):
    """rrt() (rrt.py:49-85) under the same seeds as the reference run: identical paths, with and without a
    goal_test, through the accelerated primitives and through plain callables."""
    import random
    from motion_planners_b200.rrt import rrt
    g = golden("rrt_" + name)
    world = _world(mp, g)
    dist, ext, col = _tree_primitives(mp, world)
    env, _ = world.shapes()
    ref_paths = _paths_from(g)
    iters = int(g["max_iterations"])
    tol = float(g["goal_tolerance"])
    for q in range(len(ref_paths)):
        seed = int(g["seeds"][q])
        goal = g["goals"][q]
        goal_test = (lambda c, gg=goal: float(np.max(np.abs(np.array(c) - gg))) < tol) if g["use_goal_test"][q] \
            else None
        for plain in (False, True):
            if plain and q % 4:
                continue                      # the callable-by-callable form on every fourth query
            random.seed(seed)
            np.random.seed(seed)
            sample_fn = mp.primitives.get_sample_fn(env, obstacles=[])
            kw = {} if goal_test is None else {"goal_test": goal_test}
            if plain:
                kw["allow_host"] = True       # plain callables run the host loop only on request
            prims = (lambda a, b: dist(a, b), sample_fn, lambda a, b: ext(a, b), lambda c: col(c)) if plain \
                else (dist, sample_fn, ext, col)
            path = rrt(g["starts"][q], goal, prims[0], prims[1], prims[2], prims[3], goal_probability=.2,
                       max_iterations=iters, **kw)
            assert (path is None) == (ref_paths[q] is None)
            if path is not None:
                arr = np.array(path)
                assert arr.shape == ref_paths[q].shape and (arr == ref_paths[q]).all()


@pytest.mark.parametrize("name", ["lazy_prm_star_problem1_s31", "lazy_prm_star_mixed3d_s32"])
def test_lazy_prm_star_matches_reference_run(mp, golden, name):
    """lazy_prm_star() (lazy_prm.py:346-389) over a finite parameter schedule: same draws consumed, same path."""
    g = golden(name)
    world = _world(mp, g)
    D = world.dim
    w = np.ones(D)
    ext = mp.primitives.get_extend_fn()
    col = mp.primitives.GpuCollisionFn(world)
    draws = iter([np.array(r) for r in g["draws"]])
    used = [0]

    def sample_fn():
        used[0] += 1
        return next(draws)

    path = mp.lazy_prm.lazy_prm_star(g["start"], g["goal"], sample_fn, ext, col, weights=w, success_cost=0.,
                                     param_sequence=[{'num_samples': int(n)} for n in g["rounds"]], verbose=False)
    assert used[0] == len(g["draws"])
    assert (path is not None) == bool(g["solved"])
    if path is not None:
        arr = np.array(path)
        assert arr.shape == g["path"].shape and (arr == g["path"]).all()
        cost_fn = mp.primitives.get_distance_fn(w, p_norm=2)
        cost = sum(cost_fn(a, b) for a, b in zip(arr[:-1], arr[1:]))
        assert abs(cost - float(g["cost"])) <= 1e-9 * float(g["cost"])


# ------------------------------------------------------------------ BruteForceNeighbors / Roadmap(weights=None)
@pytest.mark.parametrize("D", [2, 7])
def test_brute_force_neighbors_match_reference_run(golden, D):
    """nearest.py:89-109 through the drop-in class: BruteForceNeighbors.query_neighbors with both accelerated
    distance functions, and lazy_prm.Roadmap(weights=None, distance_fn=...) (lazy_prm.py:115-116), against a run
    of the unmodified reference (oracle/make_golden.py gen_brute_force)."""
    from motion_planners_b200 import primitives
    from motion_planners_b200.lazy_prm import Roadmap
    from motion_planners_b200.nearest import BruteForceNeighbors
    g = golden("brute_force_d%d" % D)
    X, Q, w, k = g["samples"], g["queries"], g["weights"], int(g["k"])
    for name, fn in (("primitives", primitives.get_distance_fn(w, p_norm=2)), ("samplers", primitives.get_distance_fn(w))):
        nn = BruteForceNeighbors(fn)
        handed = list(nn.add_data(list(X)))
        assert [i for i, _ in handed] == list(range(len(X)))
        for r in range(0, len(Q), 7):                                  # the per-query call of the reference API
            got = nn.query_neighbors(Q[r], k=k)
            assert [i for _, i, _ in got] == list(g["idx_" + name][r])
            assert [d for d, _, _ in got] == list(g["dist_" + name][r])
            assert all(x is nn.data[i] for _, i, x in got)
        idx, dist = nn.query_batch(list(Q), k)
        assert (idx == g["idx_" + name]).all() and (dist == g["dist_" + name]).all()
    rm = Roadmap(primitives.get_extend_fn(), weights=None, distance_fn=primitives.get_distance_fn(w), max_degree=k - 1)
    rm.add_samples(list(X))
    edges = np.array(sorted(rm.edges), np.int64).reshape(-1, 2)
    assert (edges == g["roadmap_edges"]).all()


def test_planners_refuse_plain_callables_unless_asked():
    """No silent CPU path: handed primitives that are not the batched ones, the planners raise; allow_host=True runs
    the reference's statement order on the host (and finds the same path as with the batched primitives)."""
    import random as pyrandom
    from motion_planners_b200 import primitives, workloads
    from motion_planners_b200.rrt_connect import rrt_connect
    world, start, goal = workloads.problem1_world()
    col, ext, dfn = primitives.GpuCollisionFn(world), primitives.get_extend_fn(), primitives.get_distance_fn(np.ones(2))
    env, obstacles = world.shapes()
    plain_col = lambda q: bool(col(q))
    with pytest.raises(TypeError):
        rrt_connect(start, goal, dfn, primitives.get_sample_fn(env), ext, plain_col, max_iterations=5)
    paths = []
    for c, kw in ((col, {}), (plain_col, {"allow_host": True})):
        pyrandom.seed(3)
        np.random.seed(3)
        paths.append(rrt_connect(start, goal, dfn, primitives.get_sample_fn(env), ext, c, max_iterations=50, **kw))
    assert paths[0] is not None and len(paths[0]) == len(paths[1])
    assert all((np.asarray(a) == np.asarray(b)).all() for a, b in zip(*paths))


def test_lazy_prm_device_cost_to_go_equals_host_dijkstra(mp):
    """lazy_prm's goal-rooted shortest-path pass (lazy_prm.py:295-296) on the device: the heuristic array is the
    host dijkstra's, bit for bit (same fixed point, same edge costs), and the planner returns the same Solution
    whichever of the two ran."""
    import random
    from motion_planners_b200 import workloads
    lp, pr = mp.lazy_prm, mp.primitives
    world = workloads.c2_world()
    env, obstacles = world.shapes()
    col, ext = pr.GpuCollisionFn(world), pr.get_extend_fn()
    D = world.dim
    w = np.ones(D)
    sols, heur = [], []
    for device in (True, False):
        random.seed(11)
        np.random.seed(11)
        cfree = pr.get_sample_fn(env, obstacles=obstacles)
        sample_fn = pr.get_sample_fn(env, obstacles=[])
        start, goal = cfree(), cfree()
        rm = lp.Roadmap(ext, weights=w, cost_fn=pr.get_distance_fn(w, p_norm=2), max_degree=10)
        rm.DEVICE_SEARCH_MIN = 2048 if device else 10 ** 9
        rm.add_samples([start, goal] + lp.sample_until(sample_fn, 6000))
        goal_v = rm.index(goal)
        h = rm.cost_to_go(goal_v, rm.cost_fn)
        assert (h is not None) == device
        if device:
            visited = lp.dijkstra(goal_v, rm.neighbors_fn, rm.get_cost)
            ref = np.array([visited[v].g if v in visited else np.inf for v in rm.vertices])
            assert (h == ref).all()
        sols.append(lp.lazy_prm(start, goal, sample_fn, ext, col, roadmap=rm, num_samples=6000, weights=w, verbose=False))
    a, b = sols
    assert (a.path is None) == (b.path is None) and a.path is not None
    assert len(a.path) == len(b.path) and all((np.asarray(x) == np.asarray(y)).all() for x, y in zip(a.path, b.path))
    assert a.colliding_vertices == b.colliding_vertices and a.colliding_edges == b.colliding_edges


# ------------------------------------------------------------------ MultiRRT / MultiBiRRT, curve collision
@pytest.mark.parametrize("name", ["problem1", "mixed3d"])
def test_multi_rrt_matches_reference_runs(mp, golden, name):
    """multi_rrt.py:71-159: a tree grown towards six goals in turn, under the reference run's seeds: the same paths
    (configurations as tuples, the form in which the reference's planner works) and the same tree sizes."""
    import random
    from motion_planners_b200.multi_rrt import MultiBiRRT, MultiRRT
    g = golden("multi_rrt_" + name)
    world = _world(mp, g)
    dist, ext, col = _tree_primitives(mp, world)
    env, obstacles = world.shapes()
    for kind, cls in (("rrt", MultiRRT), ("birrt", MultiBiRRT)):
        seed = int(g[kind + "_seed"])
        random.seed(seed)
        np.random.seed(seed)
        cfree = mp.primitives.get_sample_fn(env, obstacles=obstacles)
        sample_fn = mp.primitives.get_sample_fn(env, obstacles=[])
        start = tuple(float(x) for x in cfree())
        goals = [tuple(float(x) for x in cfree()) for _ in range(6)]
        assert start == tuple(g[kind + "_start"]) and np.array_equal(np.array(goals), g[kind + "_goals"])
        tree = cls(start, dist, sample_fn, ext, col)
        for i, goal in enumerate(goals):
            path = tree.grow((lambda gg=goal: gg), iterations=40) if kind == "rrt" else tree.grow(goal, iterations=40)
            assert (path is not None) == bool(g[kind + "_solved"][i]), (kind, i)
            if path is not None:
                arr = np.array([n.config for n in path])
                ref = g["%s_path%d" % (kind, i)]
                assert arr.shape == ref.shape and (arr == ref).all(), (kind, i)
            assert len(tree) == int(g[kind + "_sizes"][i]), (kind, i)
        # the Mapping interface of multi_rrt.py:29-38 and the two-point query of :40-59
        assert start in tree and tree[start].parent is None
        assert tree(start) == [start]


def test_curve_collision_fn_matches_reference(mp, golden):
    """trajectory/smooth.py:32-50: verdicts of the reference's curve_collision_fn on 40 spline curves, from the
    samples its distance_discretize_curve took (stored), one launch per curve and one launch for all."""
    from motion_planners_b200.trajectory import get_curve_collision_fn
    g = golden("curve_collision_mixed3d")
    world = _world(mp, g)
    col = mp.primitives.GpuCollisionFn(world)
    off = np.concatenate([[0], np.cumsum(g["counts"])])
    lists = [g["samples"][off[i]:off[i + 1]] for i in range(len(g["counts"]))]
    served = iter(lists)
    fn = get_curve_collision_fn(col, lambda curve, start_t=None, end_t=None: (None, next(served)))
    got = [fn(object(), None, None) for _ in lists]
    assert got == [bool(v) for v in g["verdicts"]]
    assert fn.batch(lists) == got
    assert fn(None) is True


def test_rrt_star_device_loop_equals_stepwise(mp):
    """The persistent-kernel loop (mpb_rrt_star_run) against the one-iteration-per-call path (mpb_rrt_star_step) on a
    longer run than the goldens: same path, same number of nodes, and the global generators left in the same state
    (the bulk draws ahead of the kernel are given back)."""
    import random
    from motion_planners_b200 import workloads
    from motion_planners_b200.rrt_star import rrt_star
    world = workloads.c5_world()
    dist, ext, col = _tree_primitives(mp, world)
    env, obstacles = world.shapes()
    random.seed(0)
    np.random.seed(0)
    cfree = mp.primitives.get_sample_fn(env, obstacles=obstacles)
    while True:
        s, g = cfree(), cfree()
        if np.linalg.norm(s - g) > 0.8:
            break
    got = []
    for device_loop in (True, False):
        random.seed(5)
        np.random.seed(5)
        sample_fn = mp.primitives.get_sample_fn(env, obstacles=[])
        path = rrt_star(s, g, dist, sample_fn, ext, col, radius=0.1, max_iterations=3000, device_loop=device_loop)
        if device_loop:
            assert rrt_star.last_launches < 50          # a handful of kernel launches, not four per iteration
        got.append((path, random.random(), float(np.random.uniform())))
    (pa, ra, na), (pb, rb, nb) = got
    assert pa is not None and pb is not None and len(pa) == len(pb)
    assert all((np.asarray(x) == np.asarray(y)).all() for x, y in zip(pa, pb))
    assert ra == rb and na == nb
