"""The oracle (oracle/mp_oracle.c and oracle/np_oracle.py) against outputs of the unmodified
reference committed under tests/golden/ (made by oracle/make_golden.py).  CPU only."""
import numpy as np
import pytest

import oracle
from oracle import np_oracle
from conftest import world_from_npz


@pytest.mark.parametrize("D", [2, 3, 6, 7, 14])
def test_knn_matches_reference_kdtree(golden, D):
    g = golden("knn_d%d" % D)
    X, w, k = g["samples"], g["weights"], int(g["k"])
    idx, dist = oracle.knn(X, X, k, w)
    assert (idx == g["idx"]).all()
    assert (dist == g["dist"]).all()          # bit-exact distances
    ni, nd = np_oracle.knn(X, X[:64], k, w)
    assert (ni == g["idx"][:64]).all() and (nd == g["dist"][:64]).all()
    idx_r, dist_r = oracle.knn(X, X, k, w, max_dist=float(g["radius"]))
    assert (idx_r == g["idx_r"]).all()
    assert (dist_r == g["dist_r"]).all()


@pytest.mark.parametrize("D", [2, 6, 7])
def test_extend_matches_reference(golden, D):
    g = golden("extend_d%d" % D)
    q1, q2, counts, way = g["q1"], g["q2"], g["counts"], g["waypoints"]
    off = np.concatenate([[0], np.cumsum(counts)])
    n_np, way_np = np_oracle.extend_batch(q1, q2)
    assert (n_np == counts).all()
    for e in range(len(q1)):
        got = oracle.extend(q1[e], q2[e])
        ref = way[off[e]:off[e + 1]]
        assert got.shape == ref.shape and (got == ref).all()
        assert (way_np[e, :counts[e]] == ref).all()
    assert counts[0] == 1                      # q1 == q2 yields a single waypoint


@pytest.mark.parametrize("name", ["problem1", "boxes6d", "mixed3d", "boxes7d"])
def test_collision_matches_reference(golden, name):
    g = golden("collide_" + name)
    w = world_from_npz(g, oracle.World)
    assert (w.points_check(g["points"]) == g["hit"]).all()
    hit_np = np_oracle.points_check(g["points"], w.env_lo, w.env_hi, w.box_lo, w.box_hi, w.sph_c, w.sph_r, w.buffer)
    assert (hit_np == g["hit"]).all()
    assert 0 < g["hit"].sum() < len(g["hit"])


@pytest.mark.parametrize("name", ["problem1", "boxes6d", "mixed3d", "boxes7d"])
def test_edges_match_reference(golden, name):
    g = golden("edges_" + name)
    w = world_from_npz(g, oracle.World)
    n, first, last = w.edges_check(g["q1"], g["q2"])
    assert (n == g["n"]).all() and (first == g["first"]).all() and (last == g["last"]).all()
    assert ((first >= 0) == g["any_all"]).all()
    n2, first2, _ = w.edges_check(g["q1"], g["q2"], skip_last=True)
    assert (n2 == g["n"]).all() and ((first2 >= 0) == g["any_prm"]).all()
    assert 0 < g["any_all"].sum() < len(g["any_all"])


def test_distance_procedures_match_reference(golden):
    g = golden("dist")
    for D in (2, 3, 7, 14):
        a, b, w = g["a%d" % D], g["b%d" % D], g["w%d" % D]
        p = np.array([oracle.dist_primitives(x, y, w) for x, y in zip(a, b)])
        s = np.array([oracle.dist_samplers(x, y, w) for x, y in zip(a, b)])
        assert (p == g["prim%d" % D]).all()
        assert (s == g["samp%d" % D]).all()


def test_nearest_and_radius_tiebreak():
    nodes = np.array([[0.5, 0.5], [0.2, 0.2], [0.8, 0.8], [0.2, 0.2]])
    i, d = oracle.nearest(nodes, np.array([0.0, 0.0]))
    assert i == 1                               # first minimum wins (utils.py:61-65)
    idx, dd = oracle.radius(nodes, np.array([0.2, 0.2]), 0.5)
    assert idx.tolist() == [0, 1, 3]


def _circ_dict(g):
    lo, hi = g["circ_lo"], g["circ_hi"]
    return {d: (float(lo[d]), float(hi[d])) for d in range(len(lo)) if np.isfinite(lo[d])}


@pytest.mark.parametrize("name", ["d3", "d6", "d7"])
def test_circular_extend_matches_reference(golden, name):
    g = golden("circular_%s" % name)
    circ = _circ_dict(g)
    q1, q2, counts, way = g["q1"], g["q2"], g["counts"], g["waypoints"]
    off = np.concatenate([[0], np.cumsum(counts)])
    wrapped = 0
    for e in range(len(q1)):
        got = oracle.extend_circular(q1[e], q2[e], circ)
        ref = way[off[e]:off[e + 1]]
        assert got.shape == ref.shape and (got == ref).all()
        plain = oracle.extend(q1[e], q2[e])
        wrapped += int(plain.shape != ref.shape or (plain != ref).any())
    assert wrapped > len(q1) // 4              # the fixtures do exercise the wrap-around
    if len(g["first"]):
        world = world_from_npz(g, oracle.World)
        n, first, last = world.edges_check_circular(q1, q2, circ)
        assert (n == counts).all() and (first == g["first"]).all() and (last == g["last"]).all()
        assert (first >= 0).any() and (first < 0).any()


@pytest.mark.parametrize("name", ["d3", "d6", "d7"])
def test_circular_knn_matches_reference(golden, name):
    g = golden("circular_%s" % name)
    circ = _circ_dict(g)
    X, w, k = g["samples"], g["weights"], int(g["k"])
    idx, dist = oracle.knn_circular(X, X, k, circ, w)
    assert (dist == g["dist"]).all()           # bit-exact distances (min over the 3^c images)
    assert (idx == g["idx"]).all()
    plain, _ = oracle.knn(X, X, k, w)
    assert (plain != idx).any()                # the images change the answer on this data
    nr = g["idx_r"].shape[0]
    idx_r, dist_r = oracle.knn_circular(X, X[:nr], k, circ, w, max_dist=float(g["radius"]))
    assert (idx_r == g["idx_r"]).all() and (dist_r == g["dist_r"]).all()


@pytest.mark.parametrize("D", [2, 7])
def test_brute_force_neighbors_golden(golden, D):
    """nearest.py:89-109 BruteForceNeighbors with both reference distance functions: the oracle's metric kNN
    (stable on the index) returns the reference's neighbours and distances, duplicates included."""
    g = golden("brute_force_d%d" % D)
    X, Q, w, k = g["samples"], g["queries"], g["weights"], int(g["k"])
    for name, mode in (("primitives", oracle.METRIC_PRIMITIVES), ("samplers", oracle.METRIC_SAMPLERS)):
        idx, dist = oracle.knn_metric(X, Q, k, w, mode)
        assert (idx == g["idx_" + name]).all()
        assert (dist == g["dist_" + name]).all()
