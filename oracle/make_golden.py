"""Generate tests/golden/*.npz by RUNNING THE UNMODIFIED REFERENCE.  TEST INFRASTRUCTURE ONLY.

Run in the build container (the reference checkout is at /root/reference there):

    python -m oracle.make_golden            # all fixtures
    python -m oracle.make_golden knn edges  # a subset

Every array under tests/golden/ is an output of the reference's own functions
(KDNeighbors.query_neighbors, get_extend_fn, get_collision_fn, lazy_prm.Roadmap,
prm, rrt_connect, rrt_star, ...) on seeded inputs that are stored beside it.  The
reference has no tests or fixtures of its own, so these files
are what pins the oracle, and through it the CUDA path, to the reference.
"""
import io
import os
import random
import sys
from contextlib import redirect_stdout

import numpy as np

from . import ref_loader

INF_ = float("inf")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)

from motion_planners_b200 import workloads  # noqa: E402  (seeded generators only; no GPU code)


def _save(name, **arrays):
    os.makedirs(GOLDEN, exist_ok=True)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **arrays)
    print("wrote %s (%.1f KB)" % (path, os.path.getsize(path) / 1024.))


def _ref_world(world):
    """World -> the reference's (environment Box, [Box|Circle]) objects."""
    from motion_planners.tkinter.viewer import Box, Circle
    env = Box(world.env_lo.copy(), world.env_hi.copy())
    obs = [Box(lo.copy(), hi.copy()) for lo, hi in zip(world.box_lo, world.box_hi)]
    obs += [Circle(c.copy(), float(r)) for c, r in zip(world.sph_c, world.sph_r)]
    return env, obs


def _world_arrays(world):
    return dict(dim=np.int64(world.dim), env_lo=world.env_lo, env_hi=world.env_hi, box_lo=world.box_lo,
                box_hi=world.box_hi, sph_c=world.sph_c, sph_r=world.sph_r, buffer=np.float64(world.buffer))


WORLDS = {
    "problem1": lambda: workloads.problem1_world()[0],
    "boxes6d": lambda: workloads.boxes_world(6, 64, 0.2, 0.6, seed=0),
    "mixed3d": lambda: workloads.boxes_world(3, 12, 0.1, 0.3, seed=3, n_sph=6),
    "boxes7d": lambda: workloads.boxes_world(7, 64, 0.3, 0.7, seed=0),
}


# --------------------------------------------------------------------------- kNN
def gen_knn():
    """KDNeighbors.query_neighbors called once per stored sample, as Roadmap.add_samples does
    (lazy_prm.py:148-160 -> nearest.py:74-85)."""
    from motion_planners.nearest import KDNeighbors
    from motion_planners.primitives import get_embed_fn
    for D, N, weighted in [(2, 500, False), (3, 500, True), (6, 600, False), (7, 600, True), (14, 400, True)]:
        rs = np.random.RandomState(100 + D)
        X = rs.uniform(size=(N, D))
        w = rs.uniform(0.5, 2.0, size=D) if weighted else np.ones(D)
        k = 11
        nn = KDNeighbors(embed_fn=get_embed_fn(w))
        nn.add_data(list(X))
        idx = np.empty((N, k), np.int64)
        dist = np.empty((N, k))
        for v, x in enumerate(X):
            res = nn.query_neighbors(x, k=k, eps=0., p=2, distance_upper_bound=float("inf"))
            idx[v] = [i for _, i, _ in res]
            dist[v] = [d for d, _, _ in res]
        # bounded-radius variant: missing neighbours are dropped (nearest.py:82)
        rad = float(np.median(dist[:, 5]))
        idx_r = np.full((N, k), -1, np.int64)
        dist_r = np.full((N, k), np.inf)
        for v, x in enumerate(X):
            res = nn.query_neighbors(x, k=k, eps=0., p=2, distance_upper_bound=rad)
            idx_r[v, :len(res)] = [i for _, i, _ in res]
            dist_r[v, :len(res)] = [d for d, _, _ in res]
        _save("knn_d%d" % D, samples=X, weights=w, k=np.int64(k), idx=idx, dist=dist, radius=np.float64(rad),
              idx_r=idx_r, dist_r=dist_r)


# ------------------------------------------------------------------------ extend
def gen_extend():
    """get_extend_fn(circular={})(q1, q2) waypoint lists (tkinter/samplers.py:113-127)."""
    from motion_planners.tkinter.samplers import get_extend_fn
    ext = get_extend_fn()
    for D in (2, 6, 7):
        rs = np.random.RandomState(200 + D)
        E = 90
        q1 = rs.uniform(size=(E, D))
        scale = np.array([0.05, 0.3, 1.0])[np.arange(E) % 3]
        q2 = q1 + scale[:, None] * (rs.uniform(size=(E, D)) - q1)
        q2[0] = q1[0]                      # degenerate edge: one waypoint
        q2[1] = q1[1] + 0.01               # exactly on a step boundary
        counts, flat = [], []
        for a, b in zip(q1, q2):
            pts = np.array(list(ext(a, b)))
            counts.append(len(pts))
            flat.append(pts)
        _save("extend_d%d" % D, q1=q1, q2=q2, counts=np.array(counts, np.int64), waypoints=np.concatenate(flat))


# --------------------------------------------------------------------- collision
def gen_collide():
    """get_collision_fn(environment, obstacles)(q) (tkinter/samplers.py:88-98)."""
    from motion_planners.tkinter.samplers import get_collision_fn
    for name, mk in WORLDS.items():
        world = mk()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        rs = np.random.RandomState(300)
        n = 1500
        pts = rs.uniform(-0.05, 1.05, size=(n, world.dim))
        # points exactly on inflated / raw faces exercise the closed-interval compares
        if world.n_box:
            pts[0] = world.box_lo[0] - world.buffer
            pts[1] = world.box_hi[0] + world.buffer
            pts[2] = world.box_lo[0]
            pts[3] = np.nextafter(world.box_hi[0] + world.buffer, np.inf)
        pts[4] = world.env_lo
        pts[5] = world.env_hi
        pts[6] = np.nextafter(world.env_hi, np.inf)
        if world.n_sph:
            d = np.zeros(world.dim)
            d[0] = 1.0
            pts[7] = world.sph_c[0] + (world.sph_r[0] + world.buffer) * d
            pts[8] = world.sph_c[0]
        hit = np.array([bool(col(q)) for q in pts])
        _save("collide_%s" % name, points=pts, hit=hit, **_world_arrays(world))


# ------------------------------------------------------------------------- edges
def gen_edges():
    """Per directed edge: number of waypoints, first and last colliding waypoint index of
    [collision_fn(q) for q in extend_fn(q1, q2)] -- everything check_edge (lazy_prm.py:189-195),
    DegreePRM.grow (prm.py:250-251), extend_towards (primitives.py:19-29) and safe_path
    (rrt_star.py:76-82) derive their answers from."""
    from motion_planners.tkinter.samplers import get_collision_fn, get_extend_fn
    ext = get_extend_fn()
    for name, mk in WORLDS.items():
        world = mk()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        rs = np.random.RandomState(400)
        E = 240
        q1 = rs.uniform(size=(E, world.dim))
        scale = np.array([0.08, 0.25, 1.0])[np.arange(E) % 3]
        q2 = q1 + scale[:, None] * (rs.uniform(size=(E, world.dim)) - q1)
        q2[0] = q1[0]
        q2[3] = q1[3] + 1.5 * (rs.uniform(size=world.dim) - 0.5)     # leaves the environment
        n, first, last, any_all, any_prm = [], [], [], [], []
        for a, b in zip(q1, q2):
            flags = [bool(col(q)) for q in ext(a, b)]
            n.append(len(flags))
            hits = [i for i, f in enumerate(flags) if f]
            first.append(hits[0] if hits else -1)
            last.append(hits[-1] if hits else -1)
            any_all.append(any(flags))
            any_prm.append(any(flags[:-1]))
        _save("edges_%s" % name, q1=q1, q2=q2, n=np.array(n, np.int32), first=np.array(first, np.int32),
              last=np.array(last, np.int32), any_all=np.array(any_all), any_prm=np.array(any_prm),
              **_world_arrays(world))


# --------------------------------------------------------------------- distances
def gen_dist():
    """The three distance procedures: primitives.get_distance_fn,
    tkinter.samplers.get_distance_fn, and the kd-tree metric via KDNeighbors."""
    from motion_planners.primitives import get_distance_fn as prim
    from motion_planners.tkinter.samplers import get_distance_fn as samp
    out = {}
    for D in (2, 3, 7, 14):
        rs = np.random.RandomState(500 + D)
        n = 400
        a = rs.uniform(size=(n, D))
        b = rs.uniform(size=(n, D))
        w = rs.uniform(0.5, 2.0, size=D)
        f1, f2 = prim(w), samp(w)
        out["a%d" % D] = a
        out["b%d" % D] = b
        out["w%d" % D] = w
        out["prim%d" % D] = np.array([f1(x, y) for x, y in zip(a, b)])
        out["samp%d" % D] = np.array([f2(x, y) for x, y in zip(a, b)])
    _save("dist", **out)


# ---------------------------------------------------------------------- lazy_prm
def _edge_array(edges):
    return np.array(sorted(edges), np.int64).reshape(-1, 2)


def gen_lazy_prm():
    """lazy_prm.Roadmap.add_samples + check_roadmap, and full lazy_prm() runs
    (lazy_prm.py:102-237, 253-337) under fixed seeds."""
    from motion_planners.lazy_prm import Roadmap, lazy_prm
    from motion_planners.primitives import get_distance_fn as prim
    from motion_planners.tkinter.samplers import get_collision_fn, get_extend_fn, get_sample_fn
    ext = get_extend_fn()
    cases = [("problem1", 400, 10, 1), ("boxes6d", 250, 10, 2), ("mixed3d", 300, 5, 3)]
    for name, N, deg, seed in cases:
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        D = world.dim
        X = np.random.RandomState(seed).uniform(size=(N, D))
        w = np.ones(D)
        rm = Roadmap(ext, weights=w, cost_fn=prim(w), max_degree=deg)
        rm.add_samples(list(X))
        edges = _edge_array(rm.edges)
        rm.check_roadmap(col)
        cv = np.array([rm.colliding_vertices[v] for v in range(N)])
        ce = np.array(sorted((a, b, int(f)) for (a, b), f in rm.colliding_edges.items()), np.int64)
        # bounded-distance roadmap
        rm2 = Roadmap(ext, weights=w, cost_fn=prim(w), max_degree=deg,
                      max_distance=float(np.sqrt(D) * 0.08))
        rm2.add_samples(list(X))
        edges_md = _edge_array(rm2.edges)
        _save("roadmap_%s" % name, samples=X, max_degree=np.int64(deg), edges=edges, colliding_vertices=cv,
              colliding_edges=ce, max_distance=np.float64(np.sqrt(D) * 0.08), edges_max_distance=edges_md,
              **_world_arrays(world))

    # full planner runs: start/goal + num_samples draws of the reference sample_fn
    runs = [("problem1", 200, 10, 5, True), ("problem1", 200, 10, 6, False), ("boxes6d", 150, 10, 7, True)]
    for name, num, deg, seed, lazy in runs:
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        D = world.dim
        random.seed(seed)
        np.random.seed(seed)
        sample_fn = get_sample_fn(env, obstacles=[], use_halton=False)
        cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
        start, goal = cfree(), cfree()
        w = np.ones(D)
        rm = Roadmap(ext, weights=w, cost_fn=prim(w), max_degree=deg)
        state = np.random.get_state()
        with redirect_stdout(io.StringIO()):
            from motion_planners.lazy_prm import sample_until
            rm.add_samples([start, goal] + sample_until(sample_fn, num))
            sol = lazy_prm(start, goal, sample_fn, ext, col, roadmap=rm, num_samples=num, weights=w,
                           lazy=lazy, verbose=False)
        np.random.set_state(state)
        drawn = np.random.uniform(size=(num, D))      # the same stream sample_fn consumed
        assert all((a == b).all() for a, b in zip(drawn, sol.samples[2:]))
        path = np.array(sol.path) if sol.path is not None else np.zeros((0, D))
        cv = np.array(sorted((v, int(f)) for v, f in sol.colliding_vertices.items()), np.int64).reshape(-1, 2)
        ce = np.array(sorted((a, b, int(f)) for (a, b), f in sol.colliding_edges.items()), np.int64).reshape(-1, 3)
        cost = sum(prim(w)(a, b) for a, b in zip(path[:-1], path[1:])) if len(path) else np.inf
        _save("lazy_prm_%s_s%d" % (name, seed), start=start, goal=goal, samples=np.array(sol.samples),
              max_degree=np.int64(deg), lazy=np.bool_(lazy), solved=np.bool_(sol.path is not None), path=path,
              cost=np.float64(cost), edges=_edge_array(sol.edges), colliding_vertices=cv, colliding_edges=ce,
              **_world_arrays(world))


# --------------------------------------------------------------------------- prm
def gen_prm():
    """prm() = DegreePRM.grow + A* (prm.py:237-282) on the demo fixture and a 6-D box world."""
    from motion_planners.prm import DegreePRM
    from motion_planners.tkinter.samplers import get_collision_fn, get_distance_fn, get_extend_fn, get_sample_fn
    ext = get_extend_fn()
    for name, num, deg, seed, cfree_samples in [("problem1", 250, 10, 11, True), ("problem1", 120, 4, 12, False),
                                                ("boxes6d", 120, 6, 13, True),
                                                ("problem1", 2000, 10, 0, True)]:     # configs[0] at full size
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        D = world.dim
        dist = get_distance_fn(np.ones(D))
        random.seed(seed)
        np.random.seed(seed)
        cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
        sample_fn = cfree if cfree_samples else get_sample_fn(env, obstacles=[], use_halton=False)
        start, goal = tuple(cfree()), tuple(cfree())
        samples = [start, goal] + [tuple(sample_fn()) for _ in range(num)]
        rm = DegreePRM(dist, ext, col, samples=samples, target_degree=deg)
        index = {q: i for i, q in enumerate(samples)}
        edges = np.array([(index[e.v1.q], index[e.v2.q]) for e in rm.edges], np.int64).reshape(-1, 2)
        path = rm(start, goal)
        path_arr = np.array(path) if path is not None else np.zeros((0, D))
        cost = sum(dist(a, b) for a, b in zip(path_arr[:-1], path_arr[1:])) if len(path_arr) else np.inf
        _save("prm_%s_s%d" % (name, seed), samples=np.array(samples), target_degree=np.int64(deg), edges=edges,
              solved=np.bool_(path is not None), path=path_arr, cost=np.float64(cost), **_world_arrays(world))


# ------------------------------------------------------------------- rrt_connect
def gen_rrt_connect():
    """rrt_connect() (rrt_connect.py:31-72) under per-query seeds; stores the draws it consumed
    so the batched port can replay the same streams."""
    from motion_planners.rrt_connect import rrt_connect
    from motion_planners.tkinter.samplers import get_collision_fn, get_distance_fn, get_extend_fn, get_sample_fn
    ext = get_extend_fn()
    for name, nq, iters, seed0 in [("problem1", 12, 60, 20), ("boxes7d", 10, 120, 40), ("mixed3d", 10, 80, 60)]:
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        D = world.dim
        dist = get_distance_fn(np.ones(D))
        starts, goals, solved, lens, flat, seeds, sizes = [], [], [], [], [], [], []
        for q in range(nq):
            seed = seed0 + q
            random.seed(seed)
            np.random.seed(seed)
            cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
            start, goal = cfree(), cfree()
            random.seed(seed)
            np.random.seed(seed)
            sample_fn = get_sample_fn(env, obstacles=[], use_halton=False)
            path = rrt_connect(start, goal, dist, sample_fn, ext, col, max_iterations=iters)
            starts.append(start)
            goals.append(goal)
            seeds.append(seed)
            solved.append(path is not None)
            p = np.array(path) if path is not None else np.zeros((0, D))
            lens.append(len(p))
            flat.append(p)
        _save("rrt_connect_%s" % name, starts=np.array(starts), goals=np.array(goals), seeds=np.array(seeds, np.int64),
              max_iterations=np.int64(iters), solved=np.array(solved), path_len=np.array(lens, np.int64),
              paths=np.concatenate(flat), **_world_arrays(world))


# --------------------------------------------------------------------------- rrt
def gen_rrt():
    """rrt() (rrt.py:49-85), goal-biased single tree, under per-query seeds; one case passes a goal_test."""
    from motion_planners.rrt import rrt
    from motion_planners.tkinter.samplers import get_collision_fn, get_distance_fn, get_extend_fn, get_sample_fn
    ext = get_extend_fn()
    for name, nq, iters, seed0 in [("problem1", 10, 150, 120), ("boxes7d", 8, 200, 140), ("mixed3d", 8, 150, 160)]:
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        D = world.dim
        dist = get_distance_fn(np.ones(D))
        starts, goals, solved, lens, flat, seeds, use_test = [], [], [], [], [], [], []
        for q in range(nq):
            seed = seed0 + q
            random.seed(seed)
            np.random.seed(seed)
            cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
            start, goal = cfree(), cfree()
            random.seed(seed)
            np.random.seed(seed)
            sample_fn = get_sample_fn(env, obstacles=[], use_halton=False)
            with_test = (q % 3 == 2)
            goal_test = (lambda c, g=goal: float(np.max(np.abs(np.array(c) - g))) < 0.15) if with_test \
                else (lambda c: False)
            path = rrt(start, goal, dist, sample_fn, ext, col, goal_test=goal_test, goal_probability=.2,
                       max_iterations=iters)
            starts.append(start)
            goals.append(goal)
            seeds.append(seed)
            use_test.append(with_test)
            solved.append(path is not None)
            p = np.array(path) if path is not None else np.zeros((0, D))
            lens.append(len(p))
            flat.append(p)
        print("  rrt %s: solved %d of %d" % (name, sum(solved), nq))
        _save("rrt_%s" % name, starts=np.array(starts), goals=np.array(goals), seeds=np.array(seeds, np.int64),
              use_goal_test=np.array(use_test), goal_tolerance=np.float64(0.15), max_iterations=np.int64(iters),
              solved=np.array(solved), path_len=np.array(lens, np.int64), paths=np.concatenate(flat),
              **_world_arrays(world))


# ---------------------------------------------------------------------- rrt_star
def gen_rrt_star():
    """rrt_star() as executed on Python 3 (rrt_star.py:86-146; the second rewire loop never
    runs because `filter` is a one-shot iterator)."""
    from motion_planners.rrt_star import rrt_star
    from motion_planners.tkinter.samplers import get_collision_fn, get_distance_fn, get_extend_fn, get_sample_fn
    ext = get_extend_fn()
    for name, iters, radius, seed in [("problem1", 300, 0.2, 70), ("mixed3d", 250, 0.3, 71)]:
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        D = world.dim
        dist = get_distance_fn(np.ones(D))
        random.seed(seed)
        np.random.seed(seed)
        cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
        # A start/goal pair joined by a free straight line makes the reference spin forever:
        # the goal is reached at iteration 0 with the straight-line cost, after which the
        # informed test (rrt_star.py:108) rejects every sample without counting the iteration.
        while True:
            start, goal = cfree(), cfree()
            if any(col(q) for q in ext(start, goal)):
                break
        random.seed(seed)
        np.random.seed(seed)
        sample_fn = get_sample_fn(env, obstacles=[], use_halton=False)
        with redirect_stdout(io.StringIO()):
            path = rrt_star(start, goal, dist, sample_fn, ext, col, radius=radius, max_iterations=iters)
        p = np.array(path) if path is not None else np.zeros((0, D))
        cost = sum(dist(a, b) for a, b in zip(p[:-1], p[1:])) if len(p) else np.inf
        _save("rrt_star_%s" % name, start=start, goal=goal, seed=np.int64(seed), max_iterations=np.int64(iters),
              radius=np.float64(radius), solved=np.bool_(path is not None), path=p, cost=np.float64(cost),
              **_world_arrays(world))


# ------------------------------------------------------------------------- birrt
def gen_birrt():
    """birrt() = random_restarts(rrt_connect) + smooth_path (rrt_connect.py:76-93, meta.py:46-86,
    smoothing.py:51-144) under fixed seeds; both global generators are consumed."""
    from motion_planners.rrt_connect import birrt
    from motion_planners.tkinter.samplers import get_collision_fn, get_distance_fn, get_extend_fn, get_sample_fn
    ext = get_extend_fn()
    for name, nq, iters, seed0 in [("problem1", 6, 80, 120), ("mixed3d", 5, 100, 140)]:
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        D = world.dim
        dist = get_distance_fn(np.ones(D))
        starts, goals, solved, lens, flat, seeds = [], [], [], [], [], []
        for q in range(nq):
            seed = seed0 + q
            random.seed(seed)
            np.random.seed(seed)
            cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
            start, goal = cfree(), cfree()
            random.seed(seed)
            np.random.seed(seed)
            sample_fn = get_sample_fn(env, obstacles=[], use_halton=False)
            with redirect_stdout(io.StringIO()):
                path = birrt(start, goal, dist, sample_fn, ext, col, max_iterations=iters, smooth=20, restarts=2)
            starts.append(start)
            goals.append(goal)
            seeds.append(seed)
            solved.append(path is not None)
            p = np.array(path) if path is not None else np.zeros((0, D))
            lens.append(len(p))
            flat.append(p)
        _save("birrt_%s" % name, starts=np.array(starts), goals=np.array(goals), seeds=np.array(seeds, np.int64),
              max_iterations=np.int64(iters), solved=np.array(solved), path_len=np.array(lens, np.int64),
              paths=np.concatenate(flat), **_world_arrays(world))


# ---------------------------------------------------------------------- circular
CIRCULAR_CASES = {
    # name: (D, {coordinate: (lower, upper)}, world name or None)
    "d3": (3, {0: (0., 1.), 2: (0., 1.)}, "mixed3d"),
    "d6": (6, {1: (0., 1.), 4: (-0.25, 1.25)}, "boxes6d"),
    "d7": (7, {0: (0., 1.), 3: (0., 1.), 6: (-np.pi, np.pi)}, None),
}


def gen_circular():
    """Circular joints: get_extend_fn(circular=...) waypoints (tkinter/samplers.py:113-127 with
    primitives.py:74-78 and utils.py:364-384), collision flags along them, and
    KDNeighbors(circular=...).query_neighbors (nearest.py:7-20,74-85)."""
    from motion_planners.nearest import KDNeighbors
    from motion_planners.primitives import get_embed_fn
    from motion_planners.tkinter.samplers import get_collision_fn, get_extend_fn
    for name, (D, circular, wname) in CIRCULAR_CASES.items():
        rs = np.random.RandomState(900 + D)
        lo = np.full(D, -np.inf)
        hi = np.full(D, np.inf)
        for d, (a, b) in circular.items():
            lo[d], hi[d] = a, b
        span_lo = np.where(np.isfinite(lo), lo, 0.)
        span_hi = np.where(np.isfinite(hi), hi, 1.)
        # -- extend (+ collision flags where the case has a world)
        ext = get_extend_fn(circular=circular)
        E = 120
        q1 = span_lo + (span_hi - span_lo) * rs.uniform(size=(E, D))
        q2 = span_lo + (span_hi - span_lo) * rs.uniform(size=(E, D))
        short = np.arange(E) % 3 == 0
        q2[short] = q1[short] + 0.08 * (rs.uniform(size=(short.sum(), D)) - 0.5)   # may leave the interval
        q2[0] = q1[0]
        col = None
        world_arrays = {}
        if wname is not None:
            world = WORLDS[wname]()
            env, obs = _ref_world(world)
            col = get_collision_fn(env, obs)
            world_arrays = _world_arrays(world)
        counts, flat, first, last = [], [], [], []
        for a, b in zip(q1, q2):
            pts = np.array(list(ext(a, b)))
            counts.append(len(pts))
            flat.append(pts)
            if col is not None:
                hits = [i for i, q in enumerate(pts) if col(q)]
                first.append(hits[0] if hits else -1)
                last.append(hits[-1] if hits else -1)
        # -- kNN
        N, k = 700, 11
        X = span_lo + (span_hi - span_lo) * rs.uniform(size=(N, D))
        w = rs.uniform(0.5, 2.0, size=D)
        nn = KDNeighbors(circular=circular, embed_fn=get_embed_fn(w))
        nn.add_data(list(X))
        idx = np.full((N, k), -1, np.int64)
        dist = np.full((N, k), np.inf)
        for v, x in enumerate(X):
            res = nn.query_neighbors(x, k=k, eps=0., p=2, distance_upper_bound=float("inf"))
            idx[v, :len(res)] = [i for _, i, _ in res]
            dist[v, :len(res)] = [d for d, _, _ in res]
        rad = float(np.median(dist[:, 5]))
        idx_r = np.full((N, k), -1, np.int64)
        dist_r = np.full((N, k), np.inf)
        for v, x in enumerate(X[:200]):
            res = nn.query_neighbors(x, k=k, eps=0., p=2, distance_upper_bound=rad)
            idx_r[v, :len(res)] = [i for _, i, _ in res]
            dist_r[v, :len(res)] = [d for d, _, _ in res]
        _save("circular_%s" % name, circ_lo=lo, circ_hi=hi, q1=q1, q2=q2, counts=np.array(counts, np.int64),
              waypoints=np.concatenate(flat), first=np.array(first, np.int32), last=np.array(last, np.int32),
              samples=X, weights=w, k=np.int64(k), idx=idx, dist=dist, radius=np.float64(rad), idx_r=idx_r[:200],
              dist_r=dist_r[:200], **world_arrays)


def gen_lazy_prm_circular():
    """lazy_prm() with circular joints (lazy_prm.py:253-337 with circular=..., as run.py:254 sketches):
    wrapped kNN when the roadmap is built, wrapped waypoints when edges are checked."""
    from motion_planners.lazy_prm import Roadmap, lazy_prm, sample_until
    from motion_planners.primitives import get_distance_fn as prim
    from motion_planners.tkinter.samplers import get_collision_fn, get_extend_fn, get_sample_fn
    runs = [("problem1", {0: (0., 1.), 1: (0., 1.)}, 200, 10, 11, True),
            ("mixed3d", {0: (0., 1.), 2: (0., 1.)}, 200, 8, 12, False)]
    for name, circular, num, deg, seed, lazy in runs:
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        ext = get_extend_fn(circular=circular)
        D = world.dim
        random.seed(seed)
        np.random.seed(seed)
        sample_fn = get_sample_fn(env, obstacles=[], use_halton=False)
        cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
        start, goal = cfree(), cfree()
        w = np.ones(D)
        rm = Roadmap(ext, weights=w, cost_fn=prim(w), max_degree=deg, circular=circular)
        with redirect_stdout(io.StringIO()):
            rm.add_samples([start, goal] + sample_until(sample_fn, num))
            sol = lazy_prm(start, goal, sample_fn, ext, col, roadmap=rm, num_samples=num, weights=w,
                           circular=circular, lazy=lazy, verbose=False)
        lo = np.full(D, -np.inf)
        hi = np.full(D, np.inf)
        for d, (a, b) in circular.items():
            lo[d], hi[d] = a, b
        path = np.array(sol.path) if sol.path is not None else np.zeros((0, D))
        cv = np.array(sorted((v, int(f)) for v, f in sol.colliding_vertices.items()), np.int64).reshape(-1, 2)
        ce = np.array(sorted((a, b, int(f)) for (a, b), f in sol.colliding_edges.items()), np.int64).reshape(-1, 3)
        cost = sum(prim(w)(a, b) for a, b in zip(path[:-1], path[1:])) if len(path) else np.inf
        print("  circular %s: solved=%s, %d path points, %d edges" % (name, sol.path is not None, len(path),
                                                                        len(sol.edges)))
        _save("lazy_prm_circ_%s_s%d" % (name, seed), start=start, goal=goal, samples=np.array(sol.samples),
              circ_lo=lo, circ_hi=hi, max_degree=np.int64(deg), lazy=np.bool_(lazy),
              solved=np.bool_(sol.path is not None), path=path, cost=np.float64(cost),
              edges=_edge_array(sol.edges), colliding_vertices=cv, colliding_edges=ce, **_world_arrays(world))


def gen_lazy_prm_star():
    """lazy_prm_star() (lazy_prm.py:346-389): rounds of lazy_prm on a growing roadmap, here along a finite
    param_sequence (so the run ends without a wall-clock limit) and with success_cost = 0 (every round runs).
    Stores every sample the run drew, in order, so the port can replay the stream."""
    from motion_planners.lazy_prm import lazy_prm_star
    from motion_planners.primitives import get_distance_fn as prim
    from motion_planners.tkinter.samplers import get_collision_fn, get_extend_fn, get_sample_fn
    ext = get_extend_fn()
    for name, rounds, seed in [("problem1", [6, 40, 120], 31), ("mixed3d", [10, 60, 150, 260], 32)]:
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        col = get_collision_fn(env, obs)
        D = world.dim
        random.seed(seed)
        np.random.seed(seed)
        cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
        start, goal = cfree(), cfree()
        inner = get_sample_fn(env, obstacles=[], use_halton=False)
        drawn = []

        def sample_fn():
            q = inner()
            drawn.append(np.array(q))
            return q
        w = np.ones(D)
        with redirect_stdout(io.StringIO()):
            path = lazy_prm_star(start, goal, sample_fn, ext, col, weights=w, success_cost=0.,
                                 param_sequence=[{'num_samples': n} for n in rounds], verbose=False)
        p = np.array(path) if path is not None else np.zeros((0, D))
        cost = sum(prim(w)(a, b) for a, b in zip(p[:-1], p[1:])) if len(p) else np.inf
        print("  lazy_prm_star %s: %d draws, solved=%s, cost %.4f" % (name, len(drawn), path is not None, cost))
        _save("lazy_prm_star_%s_s%d" % (name, seed), start=start, goal=goal, draws=np.array(drawn),
              rounds=np.array(rounds, np.int64), solved=np.bool_(path is not None), path=p, cost=np.float64(cost),
              **_world_arrays(world))


# --------------------------------------------------------------------------- BruteForceNeighbors
def gen_brute_force():
    """nearest.py:89-109 BruteForceNeighbors.query_neighbors (full distance_fn scan, stable sort, first k) with the
    reference's two distance functions, and lazy_prm.Roadmap(weights=None, distance_fn=...) .add_samples
    (lazy_prm.py:115-116,148-160), which is the caller that reaches it."""
    from motion_planners.lazy_prm import Roadmap
    from motion_planners.nearest import BruteForceNeighbors
    from motion_planners.primitives import get_distance_fn as prim
    from motion_planners.tkinter.samplers import get_distance_fn as samp, get_extend_fn
    for D, N, M, k, seed in [(2, 400, 60, 6, 41), (7, 500, 50, 11, 42)]:
        rs = np.random.RandomState(seed)
        X = rs.uniform(size=(N, D))
        X[40:60] = X[10:30]                                   # exact duplicates: ties go to the lower index
        Q = np.concatenate([X[: M // 2], rs.uniform(size=(M - M // 2, D))])
        w = rs.uniform(0.5, 2.0, size=D)
        out = dict(samples=X, queries=Q, weights=w, k=np.int64(k))
        for name, fn in (("primitives", prim(w)), ("samplers", samp(w))):
            nn = BruteForceNeighbors(fn)
            list(nn.add_data(list(X)))
            idx = np.full((M, k), -1, np.int64)
            dist = np.full((M, k), np.inf)
            for r, q in enumerate(Q):
                for j, (d, i, _) in enumerate(nn.query_neighbors(q, k=k)):
                    idx[r, j], dist[r, j] = i, d
            out["idx_" + name], out["dist_" + name] = idx, dist
        rm = Roadmap(get_extend_fn(), weights=None, distance_fn=samp(w), max_degree=k - 1)
        rm.add_samples(list(X))
        out["roadmap_edges"] = _edge_array(rm.edges)
        _save("brute_force_d%d" % D, **out)


# --------------------------------------------------------------------------- MultiRRT / MultiBiRRT
def gen_multi_rrt():
    """multi_rrt.py:71-159 under fixed seeds.  The reference keys its node table by configuration, so its own
    numpy-yielding extend_fn cannot be used (unhashable); the runs go through tuple-yielding wrappers of the
    reference's primitives, which is the form in which the unmodified planner works."""
    from motion_planners.multi_rrt import MultiBiRRT, MultiRRT
    from motion_planners.tkinter.samplers import get_collision_fn, get_distance_fn, get_extend_fn, get_sample_fn
    for name in ("problem1", "mixed3d"):
        world = WORLDS[name]()
        env, obs = _ref_world(world)
        D = world.dim
        col_np = get_collision_fn(env, obs)
        ext_np = get_extend_fn()
        dist = get_distance_fn(np.ones(D))
        col = lambda q: col_np(np.array(q))
        ext = lambda a, b: (tuple(float(x) for x in q) for q in ext_np(np.array(a), np.array(b)))
        dfn = lambda a, b: dist(np.array(a), np.array(b))
        out = {}
        for kind, cls in (("rrt", MultiRRT), ("birrt", MultiBiRRT)):
            seed = 50 + (0 if kind == "rrt" else 1)
            random.seed(seed)
            np.random.seed(seed)
            cfree = get_sample_fn(env, obstacles=obs, use_halton=False)
            sample_np = get_sample_fn(env, obstacles=[], use_halton=False)
            sample = lambda: tuple(float(x) for x in sample_np())
            start = tuple(float(x) for x in cfree())
            goals = [tuple(float(x) for x in cfree()) for _ in range(6)]
            tree = cls(start, dfn, sample, ext, col)
            paths, sizes = [], []
            for g in goals:
                if kind == "rrt":
                    path = tree.grow((lambda gg=g: gg), iterations=40)
                else:
                    path = tree.grow(g, iterations=40)
                paths.append(None if path is None else np.array([n.config for n in path]))
                sizes.append(len(tree))
            out[kind + "_seed"] = np.int64(seed)
            out[kind + "_start"] = np.array(start)
            out[kind + "_goals"] = np.array(goals)
            out[kind + "_sizes"] = np.array(sizes, np.int64)
            out[kind + "_solved"] = np.array([p is not None for p in paths])
            for i, pth in enumerate(paths):
                out["%s_path%d" % (kind, i)] = pth if pth is not None else np.zeros((0, D))
        _save("multi_rrt_%s" % name, **out, **_world_arrays(world))


# --------------------------------------------------------------------------- curve collision
def gen_curve_collision():
    """trajectory/smooth.py:32-50 get_curve_collision_fn: distance_discretize_curve samples of spline curves and
    the verdict any(map(collision_fn, default_selector(samples)))."""
    from scipy.interpolate import CubicSpline
    from motion_planners.tkinter.samplers import get_collision_fn
    from motion_planners.trajectory.discretize import distance_discretize_curve
    from motion_planners.trajectory.smooth import get_curve_collision_fn
    world = WORLDS["mixed3d"]()
    env, obs = _ref_world(world)
    col = get_collision_fn(env, obs)
    fn = get_curve_collision_fn(col)
    rs = np.random.RandomState(60)
    knots, verdicts, samples, counts, windows = [], [], [], [], []
    for c in range(40):
        pts = rs.uniform(0.05, 0.95, size=(4, world.dim))
        curve = CubicSpline(np.arange(4.0), pts)
        t0, t1 = (None, None) if c % 3 else (0.5, 2.5)
        _, smp = distance_discretize_curve(curve, start_t=t0, end_t=t1)
        smp = np.array(smp)
        knots.append(pts)
        windows.append([np.nan if t0 is None else t0, np.nan if t1 is None else t1])
        verdicts.append(bool(fn(curve, t0, t1)))
        samples.append(smp)
        counts.append(len(smp))
    _save("curve_collision_mixed3d", knots=np.array(knots), windows=np.array(windows), verdicts=np.array(verdicts),
          samples=np.concatenate(samples), counts=np.array(counts, np.int64), **_world_arrays(world))


GENERATORS = {"multi_rrt": gen_multi_rrt, "curve_collision": gen_curve_collision, "brute_force": gen_brute_force, "knn": gen_knn, "extend": gen_extend, "collide": gen_collide, "edges": gen_edges, "dist": gen_dist,
              "lazy_prm": gen_lazy_prm, "prm": gen_prm, "rrt_connect": gen_rrt_connect, "rrt_star": gen_rrt_star,
              "birrt": gen_birrt, "circular": gen_circular,
              "lazy_prm_circular": gen_lazy_prm_circular, "rrt": gen_rrt,
              "lazy_prm_star": gen_lazy_prm_star}


def main(argv):
    ref_loader.load()
    names = argv or list(GENERATORS)
    for n in names:
        GENERATORS[n]()


if __name__ == "__main__":
    main(sys.argv[1:])
