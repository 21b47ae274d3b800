"""CPU baselines timed by bench.py beside the GPU numbers.  TEST / MEASUREMENT INFRASTRUCTURE ONLY.

Two ports of the reference's path (the reference checkout itself does not travel to the GPU box):

native_port  -- the fastest faithful CPU rendering: scipy's cKDTree (the same C++ the reference calls,
                nearest.py:59-61,81) queried in one batch on all cores, and the C oracle's edge /
                vertex checks under OpenMP.  This is what `--impl reference` and `cpu_baseline` report.
python_port  -- the reference's own per-call structure (one KDTree.query per sample as
                lazy_prm.py:155-157 does; a Python generator per edge and numpy compares per obstacle
                per waypoint as tkinter/samplers.py:113-127 and tkinter/viewer.py:72-92 do), single
                threaded like the reference.  Reported as an extra, on a much smaller sample.
"""
import os
import time

import numpy as np
from scipy.spatial import KDTree

import oracle


def _oracle_world(world):
    if hasattr(world, "model"):                      # configs[3]: the sphere-approximated arm (oracle/arm_oracle.c)
        return oracle.ArmWorld(world)
    return oracle.World(world.dim, world.env_lo, world.env_hi, world.box_lo, world.box_hi, world.sph_c,
                        world.sph_r, world.buffer)


def native_port_step(world, X, q_lo, q_hi, k, tree=None):
    """One pass of the hot path for queries [q_lo, q_hi): kNN (k+1 incl. self), vertex checks, directed
    edge checks.  -> dict(seconds, knn_s, edges_s, n_queries, n_edges)."""
    ow = _oracle_world(world)
    t0 = time.perf_counter()
    if tree is None:
        tree = KDTree(X)
    _, idx = tree.query(X[q_lo:q_hi], k=k + 1, workers=-1)
    t1 = time.perf_counter()
    src = np.repeat(np.arange(q_lo, q_hi), k)
    dst = idx[:, 1:].reshape(-1)
    ow.points_check(X[q_lo:q_hi])
    _, first, _ = ow.edges_check_indexed(X, src, dst)
    t2 = time.perf_counter()
    return dict(seconds=t2 - t0, knn_s=t1 - t0, edges_s=t2 - t1, n_queries=q_hi - q_lo, n_edges=len(src),
                n_colliding=int((first >= 0).sum()))


def _py_contains_box(point, lower, upper, buffer):
    lo = lower - buffer * np.ones(len(lower))
    hi = upper + buffer * np.ones(len(upper))
    return np.less_equal(lo, point).all() and np.less_equal(point, hi).all()


def _py_collision_arm(world, q):
    """The arm's collision_fn written the way a user of the reference would: numpy per call, one forward
    kinematics and a Python loop over the boxes per configuration."""
    from oracle import np_oracle
    if not _py_contains_box(q, world.env_lo, world.env_hi, 0.):
        return True
    cen = np_oracle.arm_centres(world.model, np.asarray(q, dtype=np.float64).reshape(1, -1))[0]
    rb = np.asarray(world.model.sph_r) + world.buffer
    for lo, hi in zip(world.box_lo, world.box_hi):
        d = np.minimum(np.maximum(cen, lo), hi) - cen
        if (np.sqrt((d * d).sum(1)) <= rb).any():
            return True
    return False


def _py_collision(world, q):
    if hasattr(world, "model"):
        return _py_collision_arm(world, q)
    if not _py_contains_box(q, world.env_lo, world.env_hi, 0.):
        return True
    for lo, hi in zip(world.box_lo, world.box_hi):
        if _py_contains_box(q, lo, hi, world.buffer):
            return True
    for c, r in zip(world.sph_c, world.sph_r):
        if np.linalg.norm(np.array(q) - np.array(c)) <= (r + world.buffer):
            return True
    return False


def _py_extend(q1, q2, step=1e-2):
    steps = int(np.linalg.norm((q2 - q1) / step, ord=np.inf))
    n = steps + 1
    q = q1
    for i in range(n):
        q = (1. / (n - i)) * (q2 - q) + q
        yield q


def python_port_step(world, X, q_lo, q_hi, k):
    t0 = time.perf_counter()
    tree = KDTree(X)
    nbrs = []
    for v in range(q_lo, q_hi):
        d, i = tree.query(X[v], k=k + 1)
        nbrs.append([j for j in i if j != v])
    t1 = time.perf_counter()
    n_edges = 0
    for v, row in zip(range(q_lo, q_hi), nbrs):
        _py_collision(world, X[v])
        for j in row:
            any(_py_collision(world, q) for q in _py_extend(X[v], X[j]))
            n_edges += 1
    t2 = time.perf_counter()
    return dict(seconds=t2 - t0, knn_s=t1 - t0, edges_s=t2 - t1, n_queries=q_hi - q_lo, n_edges=n_edges)


def reference_step(world, X, q_lo, q_hi, k):
    """The same slice of work through the UNMODIFIED reference's own functions, imported from the reference
    checkout (oracle/ref_loader.py): KDNeighbors.add_data + query_neighbors per sample (nearest.py:36-85, as
    lazy_prm.py:155-157 calls it), get_extend_fn (tkinter/samplers.py:113-127) and, for configuration-space
    worlds, get_collision_fn (tkinter/samplers.py:88-98); for the arm world the reference has no collision_fn and
    takes the user's callable (_py_collision_arm).  Single-threaded, like the reference.  Returns None when the
    checkout is not reachable (the GPU box): callers then report the python_port figure and say so."""
    from . import ref_loader
    if not ref_loader.available():
        return None
    ref_loader.load()
    from motion_planners.nearest import KDNeighbors
    from motion_planners.primitives import get_embed_fn
    from motion_planners.tkinter.samplers import get_collision_fn, get_extend_fn
    from motion_planners.tkinter.viewer import Box, Circle
    from motion_planners.utils import default_selector
    t0 = time.perf_counter()
    nearest = KDNeighbors(embed_fn=get_embed_fn(np.ones(X.shape[1])))
    list(nearest.add_data(list(X)))
    nbrs = []
    for v in range(q_lo, q_hi):
        nbrs.append([i for _, i, _ in nearest.query_neighbors(X[v], k=k + 1) if i != v])
    t1 = time.perf_counter()
    if hasattr(world, "model"):
        col = lambda q: _py_collision_arm(world, q)
    else:
        env = Box(world.env_lo.copy(), world.env_hi.copy())
        obs = [Box(lo.copy(), hi.copy()) for lo, hi in zip(world.box_lo, world.box_hi)]
        obs += [Circle(c.copy(), float(r)) for c, r in zip(world.sph_c, world.sph_r)]
        col = get_collision_fn(env, obs)
    ext = get_extend_fn()
    n_edges = 0
    for v, row in zip(range(q_lo, q_hi), nbrs):
        col(X[v])
        for j in row:
            any(col(q) for q in default_selector(list(ext(X[v], X[j]))))    # lazy_prm.py:189-195 check_edge
            n_edges += 1
    t2 = time.perf_counter()
    return dict(seconds=t2 - t0, knn_s=t1 - t0, edges_s=t2 - t1, n_queries=q_hi - q_lo, n_edges=n_edges)


def cores():
    return oracle.use_all_cores()
