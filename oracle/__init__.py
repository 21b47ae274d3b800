"""CPU oracle for the kNN + edge-collision hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; the product (motion_planners_b200) never does.

Thin ctypes front-end over oracle/mp_oracle.c (a C restatement of the reference
arithmetic, each function citing the reference file:line it follows) plus, in
oracle/np_oracle.py, an independent vectorised numpy restatement used to
cross-check the C one.  Pinned against outputs of the reference itself: see
oracle/make_golden.py and tests/golden/.
"""
import ctypes
import os

import numpy as np

from . import build as _build

_lib = None

_f64p = ctypes.POINTER(ctypes.c_double)
_i64p = ctypes.POINTER(ctypes.c_int64)
_i32p = ctypes.POINTER(ctypes.c_int32)
_u8p = ctypes.POINTER(ctypes.c_uint8)


def lib():
    global _lib
    if _lib is None:
        path = _build.OUT
        if not os.path.exists(path) or any(os.path.getmtime(path) < os.path.getmtime(f) for f in _build.SRCS):
            path = _build.build()
        _lib = ctypes.CDLL(path)
        _lib.orc_sqdist_scipy.restype = ctypes.c_double
        _lib.orc_dist_primitives.restype = ctypes.c_double
        _lib.orc_dist_samplers.restype = ctypes.c_double
        _lib.orc_extend_count.restype = ctypes.c_int64
        _lib.orc_extend.restype = ctypes.c_int64
        _lib.orc_nearest.restype = ctypes.c_int64
        _lib.orc_radius.restype = ctypes.c_int64
        _lib.orc_extend_circ_count.restype = ctypes.c_int64
        _lib.orc_extend_circ.restype = ctypes.c_int64
    return _lib


METRIC_SCIPY, METRIC_PRIMITIVES, METRIC_SAMPLERS = 0, 1, 2


def _p(a, t):
    return a.ctypes.data_as(t)


def _c(a, dtype=np.float64):
    return np.ascontiguousarray(a, dtype=dtype)


class World:
    """Environment box + box and sphere obstacles (tkinter/viewer.py:16-17,69-92)."""

    def __init__(self, dim, env_lo, env_hi, box_lo=None, box_hi=None, sph_c=None, sph_r=None,
                 buffer=1e-3):
        self.dim = int(dim)
        self.env_lo = _c(env_lo).reshape(self.dim)
        self.env_hi = _c(env_hi).reshape(self.dim)
        self.box_lo = _c(box_lo if box_lo is not None else np.zeros((0, self.dim))).reshape(-1, self.dim)
        self.box_hi = _c(box_hi if box_hi is not None else np.zeros((0, self.dim))).reshape(-1, self.dim)
        self.sph_c = _c(sph_c if sph_c is not None else np.zeros((0, self.dim))).reshape(-1, self.dim)
        self.sph_r = _c(sph_r if sph_r is not None else np.zeros((0,))).reshape(-1)
        assert self.box_lo.shape == self.box_hi.shape
        assert self.sph_c.shape[0] == self.sph_r.shape[0]
        self.buffer = float(buffer)

    def _args(self):
        return (ctypes.c_int(self.dim), ctypes.c_int(self.box_lo.shape[0]), _p(self.box_lo, _f64p),
                _p(self.box_hi, _f64p), ctypes.c_int(self.sph_c.shape[0]), _p(self.sph_c, _f64p),
                _p(self.sph_r, _f64p), _p(self.env_lo, _f64p), _p(self.env_hi, _f64p),
                ctypes.c_double(self.buffer))

    def points_check(self, q):
        q = _c(q).reshape(-1, self.dim)
        out = np.empty(q.shape[0], dtype=np.uint8)
        lib().orc_points_check(*self._args(), _p(q, _f64p), ctypes.c_int64(q.shape[0]), _p(out, _u8p))
        return out.astype(bool)

    def edges_check(self, q1, q2, step=1e-2, skip_last=False):
        """-> (n_waypoints, first_hit, last_hit) int32 arrays; -1 = no colliding waypoint."""
        q1 = _c(q1).reshape(-1, self.dim)
        q2 = _c(q2).reshape(-1, self.dim)
        E = q1.shape[0]
        n = np.empty(E, np.int32)
        first = np.empty(E, np.int32)
        last = np.empty(E, np.int32)
        rc = lib().orc_edges_check(*self._args(), _p(q1, _f64p), _p(q2, _f64p), ctypes.c_int64(E),
                                   ctypes.c_double(step), ctypes.c_int(int(skip_last)), _p(n, _i32p),
                                   _p(first, _i32p), _p(last, _i32p))
        assert rc == 0
        return n, first, last

    def edges_check_circular(self, q1, q2, circular, step=1e-2, skip_last=False):
        """edges_check with circular coordinates ({coordinate: (lower, upper)})."""
        q1 = _c(q1).reshape(-1, self.dim)
        q2 = _c(q2).reshape(-1, self.dim)
        E = q1.shape[0]
        lo, hi = _circ_arrays(circular, self.dim)
        n = np.empty(E, np.int32)
        first = np.empty(E, np.int32)
        last = np.empty(E, np.int32)
        rc = lib().orc_edges_check_circ(*self._args(), _p(q1, _f64p), _p(q2, _f64p), ctypes.c_int64(E),
                                        _p(lo, _f64p), _p(hi, _f64p), ctypes.c_double(step),
                                        ctypes.c_int(int(skip_last)), _p(n, _i32p), _p(first, _i32p),
                                        _p(last, _i32p))
        assert rc == 0
        return n, first, last

    def edges_check_indexed(self, samples, src, dst, step=1e-2, skip_last=False):
        samples = _c(samples).reshape(-1, self.dim)
        src = _c(src, np.int64)
        dst = _c(dst, np.int64)
        E = src.shape[0]
        n = np.empty(E, np.int32)
        first = np.empty(E, np.int32)
        last = np.empty(E, np.int32)
        rc = lib().orc_edges_check_indexed(*self._args(), _p(samples, _f64p), _p(src, _i64p),
                                           _p(dst, _i64p), ctypes.c_int64(E), ctypes.c_double(step),
                                           ctypes.c_int(int(skip_last)), _p(n, _i32p), _p(first, _i32p),
                                           _p(last, _i32p))
        assert rc == 0
        return n, first, last


def knn(samples, queries, k, weights=None, max_dist=np.inf):
    """k smallest (distance, index) per query under the scipy-cKDTree arithmetic
    (nearest.py:74-85 with p=2, eps=0).  -> (idx[M,k] int64, dist[M,k] float64)."""
    samples = _c(samples)
    queries = _c(queries)
    N, D = samples.shape
    M = queries.shape[0]
    w = np.ones(D) if weights is None else _c(weights)
    emb = np.ascontiguousarray(w * samples)      # primitives.py:55-57
    qemb = np.ascontiguousarray(w * queries)
    idx = np.empty((M, k), np.int64)
    dist = np.empty((M, k), np.float64)
    rc = lib().orc_knn(_p(emb, _f64p), ctypes.c_int64(N), _p(qemb, _f64p), ctypes.c_int64(M),
                       ctypes.c_int(D), ctypes.c_int(k), ctypes.c_double(max_dist), _p(idx, _i64p),
                       _p(dist, _f64p))
    assert rc == 0
    return idx, dist


def knn_metric(samples, queries, k, weights=None, mode=2, max_dist=np.inf):
    """k smallest (distance_fn value, index) per query for the primitives / samplers metrics
    (nearest.py:102-109, prm.py:244-246); mode 0 falls back to the kd-tree metric's knn()."""
    if mode == METRIC_SCIPY:
        return knn(samples, queries, k, weights, max_dist)
    samples = _c(samples)
    queries = _c(queries)
    N, D = samples.shape
    M = queries.shape[0]
    w = np.ones(D) if weights is None else _c(weights)
    idx = np.empty((M, k), np.int64)
    dist = np.empty((M, k), np.float64)
    rc = lib().orc_knn_metric(_p(samples, _f64p), ctypes.c_int64(N), _p(queries, _f64p), ctypes.c_int64(M),
                              ctypes.c_int(D), _p(w, _f64p), ctypes.c_int(mode), ctypes.c_int(k),
                              ctypes.c_double(max_dist), _p(idx, _i64p), _p(dist, _f64p))
    assert rc == 0
    return idx, dist


def extend(q1, q2, step=1e-2):
    """Waypoints of get_extend_fn(circular={})(q1, q2) (tkinter/samplers.py:113-127)."""
    q1 = _c(q1)
    q2 = _c(q2)
    D = q1.shape[0]
    n = lib().orc_extend_count(_p(q1, _f64p), _p(q2, _f64p), ctypes.c_int(D), ctypes.c_double(step))
    out = np.empty((n, D))
    got = lib().orc_extend(_p(q1, _f64p), _p(q2, _f64p), ctypes.c_int(D), ctypes.c_double(step),
                           _p(out, _f64p), ctypes.c_int64(n))
    assert got == n
    return out


def _circ_arrays(circular, D):
    """{coordinate: (lower, upper)} -> (lo[D], hi[D]) with (-inf, +inf) where not circular."""
    lo = np.full(D, -np.inf)
    hi = np.full(D, np.inf)
    for d, interval in dict(circular or {}).items():
        lo[int(d)], hi[int(d)] = float(interval[0]), float(interval[1])
    return lo, hi


def extend_circular(q1, q2, circular, step=1e-2):
    """Waypoints of get_extend_fn(circular)(q1, q2) (tkinter/samplers.py:113-127 with primitives.py:74-78,
    utils.py:364-384)."""
    q1 = _c(q1)
    q2 = _c(q2)
    D = q1.shape[0]
    lo, hi = _circ_arrays(circular, D)
    n = lib().orc_extend_circ_count(_p(q1, _f64p), _p(q2, _f64p), ctypes.c_int(D), _p(lo, _f64p), _p(hi, _f64p),
                                    ctypes.c_double(step))
    out = np.empty((n, D))
    got = lib().orc_extend_circ(_p(q1, _f64p), _p(q2, _f64p), ctypes.c_int(D), _p(lo, _f64p), _p(hi, _f64p),
                                ctypes.c_double(step), _p(out, _f64p), ctypes.c_int64(n))
    assert got == n
    return out


def knn_circular(samples, queries, k, circular, weights=None, max_dist=np.inf):
    """KDNeighbors(circular=...).query_neighbors for a batch (nearest.py:7-20,74-85), 3^c images spelled out."""
    samples = _c(samples)
    queries = _c(queries)
    N, D = samples.shape
    M = queries.shape[0]
    w = np.ones(D) if weights is None else _c(weights)
    lo, hi = _circ_arrays(circular, D)
    idx = np.empty((M, k), np.int64)
    dist = np.empty((M, k), np.float64)
    rc = lib().orc_knn_circ(_p(samples, _f64p), ctypes.c_int64(N), _p(queries, _f64p), ctypes.c_int64(M),
                            ctypes.c_int(D), _p(w, _f64p), _p(lo, _f64p), _p(hi, _f64p), ctypes.c_int(k),
                            ctypes.c_double(max_dist), _p(idx, _i64p), _p(dist, _f64p))
    assert rc == 0
    return idx, dist


def dist_primitives(q1, q2, w):
    q1, q2, w = _c(q1), _c(q2), _c(w)
    return lib().orc_dist_primitives(_p(q1, _f64p), _p(q2, _f64p), _p(w, _f64p), ctypes.c_int(q1.shape[0]))


def dist_samplers(q1, q2, w):
    q1, q2, w = _c(q1), _c(q2), _c(w)
    return lib().orc_dist_samplers(_p(q1, _f64p), _p(q2, _f64p), _p(w, _f64p), ctypes.c_int(q1.shape[0]))


def dist_scipy(q1, q2, w):
    q1, q2, w = _c(q1), _c(q2), _c(w)
    e1 = np.ascontiguousarray(w * q1)
    e2 = np.ascontiguousarray(w * q2)
    return float(np.sqrt(lib().orc_sqdist_scipy(_p(e1, _f64p), _p(e2, _f64p), ctypes.c_int(q1.shape[0]))))


def nearest(nodes, target, w=None, mode=METRIC_SAMPLERS):
    nodes = _c(nodes)
    target = _c(target)
    n, D = nodes.shape
    w = np.ones(D) if w is None else _c(w)
    d = ctypes.c_double()
    i = lib().orc_nearest(_p(nodes, _f64p), ctypes.c_int64(n), _p(target, _f64p), _p(w, _f64p),
                          ctypes.c_int(D), ctypes.c_int(mode), ctypes.byref(d))
    return int(i), d.value


def radius(nodes, q, r, w=None, mode=METRIC_SAMPLERS):
    nodes = _c(nodes)
    q = _c(q)
    n, D = nodes.shape
    w = np.ones(D) if w is None else _c(w)
    idx = np.empty(n, np.int64)
    dd = np.empty(n, np.float64)
    c = lib().orc_radius(_p(nodes, _f64p), ctypes.c_int64(n), _p(q, _f64p), _p(w, _f64p),
                         ctypes.c_int(D), ctypes.c_int(mode), ctypes.c_double(r), _p(idx, _i64p),
                         _p(dd, _f64p))
    return idx[:c].copy(), dd[:c].copy()


def num_threads():
    return int(lib().orc_num_threads())


def use_all_cores():
    """Let the OpenMP loops of the C oracle use every host core (launchers such as torchrun set
    OMP_NUM_THREADS=1 for their workers)."""
    lib().orc_set_threads(int(os.cpu_count() or 1))
    return num_threads()


# ------------------------------------------------------------------ configs[3]: the sphere-approximated arm
class _ArmModelC(ctypes.Structure):
    _fields_ = [("n_joints", ctypes.c_int), ("axis", ctypes.c_int * 8), ("theta_lo", ctypes.c_double * 8),
                ("theta_span", ctypes.c_double * 8), ("a", ctypes.c_double * 8), ("d", ctypes.c_double * 8),
                ("base", ctypes.c_double * 3), ("n_sph", ctypes.c_int), ("sph_link", ctypes.c_int * 32),
                ("sph_u", ctypes.c_double * 32), ("sph_w", ctypes.c_double * 32), ("sph_r", ctypes.c_double * 32)]


def _arm_model_c(model):
    """model: any object with the fields of motion_planners_b200.arm.ArmModel (plain data)."""
    m = _ArmModelC()
    m.n_joints = int(model.n_joints)
    for i in range(m.n_joints):
        m.axis[i] = int(model.axis[i])
        m.theta_lo[i] = float(model.theta_lo[i])
        m.theta_span[i] = float(model.theta_span[i])
        m.a[i] = float(model.a[i])
        m.d[i] = float(model.d[i])
    for k in range(3):
        m.base[k] = float(model.base[k])
    m.n_sph = int(model.n_sph)
    for j in range(m.n_sph):
        m.sph_link[j] = int(model.sph_link[j])
        m.sph_u[j] = float(model.sph_u[j])
        m.sph_w[j] = float(model.sph_w[j])
        m.sph_r[j] = float(model.sph_r[j])
    return m


class ArmWorld:
    """The collision model of configs[3] as oracle/arm_oracle.c defines it: a serial arm approximated by
    spheres against workspace boxes.  `world`: anything with .model, .box_lo/.box_hi [n_box, 3], .env_lo/.env_hi
    [n_joints], .buffer (e.g. motion_planners_b200.arm.ArmWorld; only its plain data is read)."""

    def __init__(self, world):
        self.m = _arm_model_c(world.model)
        self.dim = int(world.model.n_joints)
        self.n_sph = int(world.model.n_sph)
        self.box_lo = _c(world.box_lo).reshape(-1, 3)
        self.box_hi = _c(world.box_hi).reshape(-1, 3)
        self.env_lo = _c(world.env_lo).reshape(self.dim)
        self.env_hi = _c(world.env_hi).reshape(self.dim)
        self.buffer = float(world.buffer)

    def _args(self):
        return (ctypes.byref(self.m), ctypes.c_int(self.box_lo.shape[0]), _p(self.box_lo, _f64p), _p(self.box_hi, _f64p),
                _p(self.env_lo, _f64p), _p(self.env_hi, _f64p), ctypes.c_double(self.buffer))

    def centres(self, q):
        q = _c(q).reshape(-1, self.dim)
        out = np.empty((q.shape[0], self.n_sph, 3))
        lib().orc_arm_centres(ctypes.byref(self.m), _p(q, _f64p), ctypes.c_int64(q.shape[0]), _p(out, _f64p))
        return out

    def points_check(self, q):
        q = _c(q).reshape(-1, self.dim)
        out = np.empty(q.shape[0], dtype=np.uint8)
        lib().orc_arm_points_check(*self._args(), _p(q, _f64p), ctypes.c_int64(q.shape[0]), _p(out, _u8p))
        return out.astype(bool)

    def edges_check_indexed(self, samples, src, dst, step=1e-2, skip_last=False):
        samples = _c(samples).reshape(-1, self.dim)
        src = _c(src, np.int64)
        dst = _c(dst, np.int64)
        E = src.shape[0]
        n = np.empty(E, np.int32)
        first = np.empty(E, np.int32)
        last = np.empty(E, np.int32)
        rc = lib().orc_arm_edges_check(*self._args(), _p(samples, _f64p), _p(src, _i64p), _p(dst, _i64p),
                                       ctypes.c_int64(E), ctypes.c_double(step), ctypes.c_int(int(skip_last)),
                                       _p(n, _i32p), _p(first, _i32p), _p(last, _i32p))
        assert rc == 0
        return n, first, last
