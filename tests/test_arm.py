"""configs[3]'s collision model: a 7-DoF arm approximated by spheres among workspace boxes.

The reference has no robot model (its collision_fn is user-supplied), so there is no reference output to pin
this model to: oracle/arm_oracle.c DEFINES it in the reference's terms, oracle/np_oracle.py restates it in numpy,
and csrc/arm.cu must reproduce both bit for bit.  CPU tests: the two oracle statements against each other and
against a plain-libm forward kinematics; GPU tests: the kernels through the C ABI against the C oracle."""
import numpy as np
import pytest

import oracle
from oracle import np_oracle


def world(n_box=256, seed=0):
    from motion_planners_b200 import workloads
    return workloads.c4_world(seed=seed, n_box=n_box)


def test_arm_oracle_statements_agree():
    w = world()
    o = oracle.ArmWorld(w)
    q = np.random.RandomState(3).uniform(size=(4000, 7))
    cen = o.centres(q)
    assert np.array_equal(cen, np_oracle.arm_centres(w.model, q))                 # C == numpy, bit for bit
    assert np.array_equal(o.points_check(q), np_oracle.arm_points_check(w, q))
    # and the fixed-polynomial sincos is sin / cos: centres within 1e-14 of a libm forward kinematics
    th = q * w.model.theta_span + w.model.theta_lo
    R = np.tile(np.eye(3), (q.shape[0], 1, 1))
    p = np.tile(w.model.base, (q.shape[0], 1))
    ref = np.empty_like(cen)
    for i in range(7):
        c, s = np.cos(th[:, i]), np.sin(th[:, i])
        T = np.zeros((q.shape[0], 3, 3))
        if w.model.axis[i] == 2:
            T[:, 0, 0], T[:, 0, 1], T[:, 1, 0], T[:, 1, 1], T[:, 2, 2] = c, -s, s, c, 1
        else:
            T[:, 0, 0], T[:, 0, 2], T[:, 2, 0], T[:, 2, 2], T[:, 1, 1] = c, s, -s, c, 1
        R = R @ T
        for j in range(w.model.n_sph):
            if w.model.sph_link[j] == i:
                ref[:, j] = p + w.model.sph_u[j] * R[:, :, 0] + w.model.sph_w[j] * R[:, :, 2]
        p = p + w.model.a[i] * R[:, :, 0] + w.model.d[i] * R[:, :, 2]
    assert np.abs(cen - ref).max() < 1e-14


def test_arm_oracle_edges_are_the_reference_recurrence():
    """The arm's edge check walks the waypoints of the reference's extend_fn (oracle.extend is pinned to
    samplers.py:113-127 by tests/golden/extend_*.npz) and ORs the point verdicts."""
    w = world(64)
    o = oracle.ArmWorld(w)
    rs = np.random.RandomState(4)
    X = rs.uniform(size=(300, 7))
    src = rs.randint(300, size=200)
    dst = (src + 1 + rs.randint(299, size=200)) % 300
    X[dst[:100]] = X[src[:100]] + rs.uniform(-0.08, 0.08, size=(100, 7))          # short edges too
    n, first, last = o.edges_check_indexed(X, src, dst)
    for e in range(200):
        wp = oracle.extend(X[src[e]], X[dst[e]])
        hits = np.nonzero(o.points_check(wp))[0]
        assert n[e] == len(wp)
        assert first[e] == (hits[0] if len(hits) else -1) and last[e] == (hits[-1] if len(hits) else -1)


# ------------------------------------------------------------------------------------------------ GPU
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from motion_planners_b200 import ops
    oracle.use_all_cores()
    return ops


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
def test_arm_fk_and_points_match_oracle(gpu):
    w = world()
    o = oracle.ArmWorld(w)
    obs = gpu.ObstacleSet(w)
    rs = np.random.RandomState(11)
    q = np.concatenate([rs.uniform(size=(200_000, 7)), rs.uniform(-0.05, 1.05, size=(2000, 7)),
                        np.zeros((1, 7)), np.ones((1, 7))])
    assert np.array_equal(obs.arm_centres(dev(q)).cpu().numpy(), o.centres(q))     # every sphere centre, bit for bit
    hit = obs.points_check(dev(q)).cpu().numpy()
    ref = o.points_check(q)
    assert np.array_equal(hit, ref)
    assert 0.3 < 1 - ref[:200_000].mean() < 0.7                                     # a workload, not a wall
    assert ref[200_000:202_000].sum() > (~ref[200_000:202_000]).sum()               # outside the joint limits collides


@pytest.mark.gpu
@pytest.mark.parametrize("n_box", [0, 1, 256])
def test_arm_edges_match_oracle(gpu, n_box):
    """>= 1e5 directed edges (kNN-length and long ones), all three modes, skip_last, the three edge forms."""
    w = world(n_box)
    o = oracle.ArmWorld(w)
    obs = gpu.ObstacleSet(w)
    rs = np.random.RandomState(12 + n_box)
    N, E = 20_000, 120_000
    X = rs.uniform(size=(N, 7))
    src = rs.randint(N, size=E)
    dst = rs.randint(N, size=E)
    near = X[src[: E // 2]] + rs.uniform(-0.12, 0.12, size=(E // 2, 7))            # roadmap-length edges
    X = np.concatenate([X, near])
    dst[: E // 2] = N + np.arange(E // 2)
    dst[5::1000] = -1                                                                # "no neighbour" slots
    xs = dev(X)
    for skip_last in (False, True):
        n, first, last = o.edges_check_indexed(X, src, dst, skip_last=skip_last)
        for mode, want in ((gpu.EDGE_FIRST, first), (gpu.EDGE_LAST, last), (gpu.EDGE_ANY, None)):
            hit, cnt = obs.edges_check(xs, dev(src), dev(dst), mode=mode, skip_last=skip_last)
            hit, cnt = hit.cpu().numpy(), cnt.cpu().numpy()
            assert np.array_equal(cnt, n)
            if want is None:
                assert np.array_equal(hit >= 0, first >= 0)
                ok = hit >= 0
                assert (hit[ok] >= first[ok]).all() and (hit[ok] <= last[ok]).all()
            else:
                assert np.array_equal(hit, want)
    n, first, last = o.edges_check_indexed(X, src, dst)
    keep = dst >= 0
    hit, cnt = obs.edges_check_pairs(dev(X[src[keep]]), dev(X[dst[keep]]), mode=gpu.EDGE_FIRST)
    assert np.array_equal(hit.cpu().numpy(), first[keep]) and np.array_equal(cnt.cpu().numpy(), n[keep])
    # neighbour-matrix form: rows of 6 neighbours of the first 5000 samples
    nbr = dst[: 5000 * 6].reshape(5000, 6).copy()
    hit, cnt = obs.edges_check_knn(xs, dev(nbr), row_offset=0, mode=gpu.EDGE_FIRST, want_counts=True)
    rsrc = np.repeat(np.arange(5000), 6)
    n2, first2, _ = o.edges_check_indexed(X, rsrc, nbr.reshape(-1))
    assert np.array_equal(hit.cpu().numpy().reshape(-1), first2) and np.array_equal(cnt.cpu().numpy().reshape(-1), n2)


@pytest.mark.gpu
def test_arm_roadmap_build_and_planner(gpu):
    """The arm world behind the drop-in API: DeviceRoadmap.build on it, and lazy_prm with a GpuCollisionFn."""
    from motion_planners_b200 import primitives, roadmap
    w = world()
    o = oracle.ArmWorld(w)
    X = np.random.RandomState(5).uniform(size=(30_000, 7))
    rm = roadmap.DeviceRoadmap(w, k=10)
    rm.build(dev(X))
    nbr = rm.nbr_idx.cpu().numpy()
    o_idx, _ = oracle.knn(X, X[:500], 11)
    assert np.array_equal(nbr[:500], o_idx)
    assert np.array_equal(rm.vertex_hit.cpu().numpy(), o.points_check(X))
    src = np.repeat(np.arange(30_000), 11)
    _, first, _ = o.edges_check_indexed(X, src, nbr.reshape(-1))
    assert np.array_equal(rm.edge_hit.cpu().numpy().reshape(-1) >= 0, first >= 0)
    col = primitives.GpuCollisionFn(w)
    q = X[:64]
    assert [bool(col(x)) for x in q] == list(o.points_check(q))
    assert np.array_equal(col.batch(q).cpu().numpy(), o.points_check(q))
