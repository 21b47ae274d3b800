"""Batched primitives on torch CUDA tensors, dispatched through the C ABI (include/mpb200.h).

torch is plumbing here: it owns device memory and streams; all arithmetic happens in the
hand-written kernels of csrc/.  Tensors are float64, contiguous, row-major [n, D].
Every function raises if the CUDA library is missing or a call fails -- there is no CPU
path in this package.
"""
import ctypes
import math

import numpy as np
import torch

from . import _lib
from .arm import ArmWorld
from .geometry import STEP_SIZE, World

EDGE_ANY, EDGE_FIRST, EDGE_LAST = 0, 1, 2
METRIC_KDTREE, METRIC_PRIMITIVES, METRIC_SAMPLERS = 0, 1, 2
MAX_K = 32
MAX_DIM = 16


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _f64(t, name, dim=None):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise TypeError("%s must be a CUDA tensor" % name)
    if t.dtype != torch.float64:
        raise TypeError("%s must be float64 (the reference computes in fp64)" % name)
    if t.dim() != 2 or (dim is not None and t.shape[1] != dim):
        raise ValueError("%s must have shape [n, %s]" % (name, "D" if dim is None else dim))
    return t.contiguous()


def _i64(t, name):
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.int64 and t.dim() == 1):
        raise TypeError("%s must be a 1-D int64 CUDA tensor" % name)
    return t.contiguous()


def _host_weights(weights, dim):
    if weights is None:
        return None, ctypes.POINTER(ctypes.c_double)()
    w = np.ascontiguousarray(np.asarray(weights, dtype=np.float64).reshape(dim))
    return w, w.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _host_circular(circular, dim):
    """circular: {coordinate: (lower, upper)} as the reference passes it (nearest.py:43, samplers.py:113);
    -> (keep-alive arrays, lo pointer, hi pointer) or None when no coordinate is circular."""
    if not circular:
        return None
    lo = np.full(dim, -np.inf)
    hi = np.full(dim, np.inf)
    for d, interval in dict(circular).items():
        d = int(d)
        if not 0 <= d < dim:
            raise ValueError("circular coordinate %d outside [0, %d)" % (d, dim))
        lo[d], hi[d] = float(interval[0]), float(interval[1])
    if not (np.isfinite(lo) | np.isfinite(hi)).any():
        return None
    hp = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
    return (lo, hi), hp(lo), hp(hi)


class _ArmModelC(ctypes.Structure):
    """mpb_arm_model of include/mpb200.h"""
    _fields_ = [("n_joints", ctypes.c_int), ("axis", ctypes.c_int * 8), ("theta_lo", ctypes.c_double * 8),
                ("theta_span", ctypes.c_double * 8), ("a", ctypes.c_double * 8), ("d", ctypes.c_double * 8),
                ("base", ctypes.c_double * 3), ("n_sph", ctypes.c_int), ("sph_link", ctypes.c_int * 32),
                ("sph_u", ctypes.c_double * 32), ("sph_w", ctypes.c_double * 32), ("sph_r", ctypes.c_double * 32)]


def _arm_model_struct(model):
    m = _ArmModelC()
    m.n_joints = model.n_joints
    for i in range(model.n_joints):
        m.axis[i] = model.axis[i]
        m.theta_lo[i], m.theta_span[i] = float(model.theta_lo[i]), float(model.theta_span[i])
        m.a[i], m.d[i] = float(model.a[i]), float(model.d[i])
    for k in range(3):
        m.base[k] = float(model.base[k])
    m.n_sph = model.n_sph
    for j in range(model.n_sph):
        m.sph_link[j] = model.sph_link[j]
        m.sph_u[j], m.sph_w[j], m.sph_r[j] = float(model.sph_u[j]), float(model.sph_w[j]), float(model.sph_r[j])
    return m


class ObstacleSet(object):
    """Device-resident obstacle set (mpb_world) for one World on the current CUDA device."""

    def __init__(self, world):
        if not isinstance(world, (World, ArmWorld)):
            raise TypeError("world must be a geometry.World or an arm.ArmWorld")
        if not torch.cuda.is_available():
            raise _lib.MpbError("no CUDA device: the collision kernels have no CPU fallback")
        self.world = world
        self.dim = world.dim
        self.device = torch.cuda.current_device()
        hp = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        handle = ctypes.c_void_p()
        if isinstance(world, ArmWorld):
            # the sphere-approximated arm among workspace boxes (mpb_arm_world_create)
            model = _arm_model_struct(world.model)
            rc = _lib.lib().mpb_arm_world_create(ctypes.byref(model), hp(world.box_lo), hp(world.box_hi), world.n_box,
                                                 hp(world.env_lo), hp(world.env_hi), world.buffer, ctypes.byref(handle))
            _lib.check(rc, "mpb_arm_world_create")
            self._h = handle
            return
        rc = _lib.lib().mpb_world_create(hp(world.box_lo), hp(world.box_hi), world.n_box, hp(world.sph_c),
                                         hp(world.sph_r), world.n_sph, world.dim, hp(world.env_lo),
                                         hp(world.env_hi), world.buffer, ctypes.byref(handle))
        _lib.check(rc, "mpb_world_create")
        self._h = handle

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.lib().mpb_world_destroy(h)
            except Exception:
                pass
            self._h = None

    # -- sphere centres of an arm world's configurations: [n, n_sph, 3]
    def arm_centres(self, q):
        if not isinstance(self.world, ArmWorld):
            raise TypeError("not an arm world")
        q = _f64(q, "q", self.dim)
        out = torch.empty((q.shape[0], self.world.n_sph, 3), dtype=torch.float64, device=q.device)
        rc = _lib.lib().mpb_arm_fk(self._h, _ptr(q), q.shape[0], _ptr(out), _stream())
        _lib.check(rc, "mpb_arm_fk")
        return out

    # -- collision_fn over a batch (tkinter/samplers.py:88-98)
    def points_check(self, q):
        q = _f64(q, "q", self.dim)
        out = torch.empty(q.shape[0], dtype=torch.uint8, device=q.device)
        rc = _lib.lib().mpb_points_check(self._h, _ptr(q), q.shape[0], _ptr(out), _stream())
        _lib.check(rc, "mpb_points_check")
        return out.bool()

    # -- extend_fn + collision_fn over directed edges given as index pairs
    def edges_check(self, samples, src, dst, mode=EDGE_ANY, skip_last=False, step=STEP_SIZE, want_counts=True,
                    circular=None):
        samples = _f64(samples, "samples", self.dim)
        src = _i64(src, "src")
        dst = _i64(dst, "dst")
        if src.shape != dst.shape:
            raise ValueError("src and dst differ in length")
        E = src.shape[0]
        hit = torch.empty(E, dtype=torch.int32, device=samples.device)
        cnt = torch.empty(E, dtype=torch.int32, device=samples.device) if want_counts else None
        circ = _host_circular(circular, self.dim)
        if circ is not None:
            rc = _lib.lib().mpb_edges_check_circular(self._h, _ptr(samples), None, _ptr(src), _ptr(dst), E, circ[1],
                                                     circ[2], float(step), int(mode), int(bool(skip_last)),
                                                     _ptr(hit), _ptr(cnt), _stream())
            _lib.check(rc, "mpb_edges_check_circular")
            return hit, cnt
        rc = _lib.lib().mpb_edges_check(self._h, _ptr(samples), samples.shape[0], _ptr(src), _ptr(dst), E,
                                        float(step), int(mode), int(bool(skip_last)), _ptr(hit), _ptr(cnt),
                                        _stream())
        _lib.check(rc, "mpb_edges_check")
        return hit, cnt

    # -- the same for the rows of a neighbour matrix (culling once per row)
    def edges_check_knn(self, samples, nbr, row_offset=0, mode=EDGE_ANY, skip_last=False, step=STEP_SIZE,
                        want_counts=False, row_ids=None):
        samples = _f64(samples, "samples", self.dim)
        if not (nbr.is_cuda and nbr.dtype == torch.int64 and nbr.dim() == 2 and nbr.is_contiguous()):
            raise TypeError("nbr must be a contiguous 2-D int64 CUDA tensor")
        M, kk = nbr.shape
        hit = torch.empty((M, kk), dtype=torch.int32, device=samples.device)
        cnt = torch.empty((M, kk), dtype=torch.int32, device=samples.device) if want_counts else None
        if row_ids is not None:
            # row r leaves sample row_ids[r] (rows of a spatially sharded roadmap)
            row_ids = _i64(row_ids, "row_ids")
            if row_ids.shape[0] != M:
                raise ValueError("row_ids and nbr differ in length")
            rc = _lib.lib().mpb_edges_check_knn_rows(self._h, _ptr(samples), samples.shape[0], _ptr(row_ids), _ptr(nbr), M,
                                                     kk, float(step), int(mode), int(bool(skip_last)), _ptr(hit),
                                                     _ptr(cnt), _stream())
            _lib.check(rc, "mpb_edges_check_knn_rows")
            return hit, cnt
        rc = _lib.lib().mpb_edges_check_knn(self._h, _ptr(samples), samples.shape[0], _ptr(nbr), M, kk,
                                            int(row_offset), float(step), int(mode), int(bool(skip_last)),
                                            _ptr(hit), _ptr(cnt), _stream())
        _lib.check(rc, "mpb_edges_check_knn")
        return hit, cnt

    # -- the same for explicit endpoint arrays
    def edges_check_pairs(self, q1, q2, mode=EDGE_ANY, skip_last=False, step=STEP_SIZE, want_counts=True,
                          circular=None):
        q1 = _f64(q1, "q1", self.dim)
        q2 = _f64(q2, "q2", self.dim)
        if q1.shape != q2.shape:
            raise ValueError("q1 and q2 differ in shape")
        E = q1.shape[0]
        hit = torch.empty(E, dtype=torch.int32, device=q1.device)
        cnt = torch.empty(E, dtype=torch.int32, device=q1.device) if want_counts else None
        circ = _host_circular(circular, self.dim)
        if circ is not None:
            rc = _lib.lib().mpb_edges_check_circular(self._h, _ptr(q1), _ptr(q2), None, None, E, circ[1], circ[2],
                                                     float(step), int(mode), int(bool(skip_last)), _ptr(hit),
                                                     _ptr(cnt), _stream())
            _lib.check(rc, "mpb_edges_check_circular")
            return hit, cnt
        rc = _lib.lib().mpb_edges_check_pairs(self._h, _ptr(q1), _ptr(q2), E, float(step), int(mode),
                                              int(bool(skip_last)), _ptr(hit), _ptr(cnt), _stream())
        _lib.check(rc, "mpb_edges_check_pairs")
        return hit, cnt


def edges_count(q1, q2, step=STEP_SIZE, circular=None):
    """Number of waypoints extend_fn(q1[e], q2[e]) yields (tkinter/samplers.py:119-120) -> int32 [E]."""
    q1 = _f64(q1, "q1")
    q2 = _f64(q2, "q2", q1.shape[1])
    E, D = q1.shape
    out = torch.empty(E, dtype=torch.int32, device=q1.device)
    circ = _host_circular(circular, D)
    if circ is not None:
        rc = _lib.lib().mpb_edges_count_circular(_ptr(q1), _ptr(q2), E, D, circ[1], circ[2], float(step), _ptr(out),
                                                 _stream())
        _lib.check(rc, "mpb_edges_count_circular")
        return out
    rc = _lib.lib().mpb_edges_count(_ptr(q1), _ptr(q2), E, D, float(step), _ptr(out), _stream())
    _lib.check(rc, "mpb_edges_count")
    return out


def edges_waypoints(q1, q2, counts=None, step=STEP_SIZE, circular=None):
    """extend_fn(q1[e], q2[e]) for a batch -> (offsets[E+1] int64, waypoints[sum(counts), D])."""
    q1 = _f64(q1, "q1")
    q2 = _f64(q2, "q2", q1.shape[1])
    E, D = q1.shape
    if counts is None:
        counts = edges_count(q1, q2, step, circular=circular)
    offsets = torch.zeros(E + 1, dtype=torch.int64, device=q1.device)
    if E:
        offsets[1:] = torch.cumsum(counts.to(torch.int64), 0)
    total = int(offsets[-1].item()) if E else 0
    out = torch.empty((total, D), dtype=torch.float64, device=q1.device)
    circ = _host_circular(circular, D)
    if circ is not None:
        rc = _lib.lib().mpb_edges_waypoints_circular(_ptr(q1), _ptr(q2), E, D, circ[1], circ[2], float(step),
                                                     _ptr(offsets), _ptr(out), _stream())
        _lib.check(rc, "mpb_edges_waypoints_circular")
        return offsets, out
    rc = _lib.lib().mpb_edges_waypoints(_ptr(q1), _ptr(q2), E, D, float(step), _ptr(offsets), _ptr(out), _stream())
    _lib.check(rc, "mpb_edges_waypoints")
    return offsets, out


class KnnIndex(object):
    """Packed fp32 tiles of one sample set, reusable across query batches
    (the device-side counterpart of nearest.py:36-71 KDNeighbors.add_data + update)."""

    def __init__(self, samples, weights=None):
        self.samples = _f64(samples, "samples")
        self.n, self.dim = self.samples.shape
        if not (1 <= self.dim <= MAX_DIM):
            raise ValueError("dim %d outside [1, %d]" % (self.dim, MAX_DIM))
        self.weights = None if weights is None else np.asarray(weights, dtype=np.float64).reshape(self.dim).copy()
        self._ws = None
        self._ws_queries = -1
        self._prepared = False

    def _workspace(self, n_queries):
        if self._ws is None or n_queries > self._ws_queries:
            nbytes = _lib.lib().mpb_knn_workspace_bytes(self.n, n_queries, self.dim)
            self._ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=self.samples.device)
            self._ws_queries = n_queries
            self._prepared = False
        off = (-self._ws.data_ptr()) % 256
        return ctypes.c_void_p(self._ws.data_ptr() + off), self._ws.numel() - off

    def prepare(self, n_queries=0):
        ws, nbytes = self._workspace(n_queries)
        if not self._prepared:
            _, wp = _host_weights(self.weights, self.dim)
            rc = _lib.lib().mpb_knn_prepare(_ptr(self.samples), self.n, self.dim, wp, ws, nbytes, _stream())
            _lib.check(rc, "mpb_knn_prepare")
            self._prepared = True
        return ws, nbytes

    def spatial_order(self):
        """Sample indices in the k-d leaf order of the prepared set (int64 [N], neighbours in the list are
        neighbours in space), or None when the set is too small for the scan that builds it."""
        if _lib.lib().mpb_knn_filter_kind(self.n, self.dim, 11) != 1:
            return None
        ws, nbytes = self.prepare(self._ws_queries if self._ws_queries > 0 else 0)
        out = torch.empty(self.n, dtype=torch.int32, device=self.samples.device)
        rc = _lib.lib().mpb_knn_sample_order(ws, nbytes, self.n, self.dim, _ptr(out), _stream())
        _lib.check(rc, "mpb_knn_sample_order")
        return out.long()

    def part_rows(self, part, parts):
        """Rows of share `part` of `parts` of the set queried against itself (-1: the set has no spatial order)."""
        return int(_lib.lib().mpb_knn_part_rows(self.n, self.dim, int(part), int(parts)))

    def search_part(self, part, parts, k, max_dist=math.inf, stats=None):
        """Share `part` of `parts` of the set's own kNN graph (mpb_knn_search_part): -> (row_ids [rows] int64 =
        sample index per row, idx [rows, k], dist [rows, k]); the parts are shares of the k-d leaves, dealt in rotation."""
        rows = self.part_rows(part, parts)
        if rows < 0:
            raise _lib.MpbError("this sample set has no spatial order (the tensor-core scan does not serve it)")
        ws, nbytes = self.prepare(self.n)
        dev = self.samples.device
        row_ids = torch.empty(rows, dtype=torch.int64, device=dev)
        idx = torch.empty((rows, k), dtype=torch.int64, device=dev)
        dist = torch.empty((rows, k), dtype=torch.float64, device=dev)
        _, wp = _host_weights(self.weights, self.dim)
        rc = _lib.lib().mpb_knn_search_part(_ptr(self.samples), self.n, self.dim, wp, int(k), float(max_dist), ws, nbytes,
                                            int(part), int(parts), _ptr(row_ids), _ptr(idx), _ptr(dist), _ptr(stats),
                                            _stream())
        _lib.check(rc, "mpb_knn_search_part")
        return row_ids, idx, dist

    def search(self, queries, k, max_dist=math.inf, want_dist=True, stats=None, out_idx=None, out_dist=None):
        queries = _f64(queries, "queries", self.dim)
        M = queries.shape[0]
        ws, nbytes = self.prepare(M)
        idx = out_idx if out_idx is not None else torch.empty((M, k), dtype=torch.int64, device=queries.device)
        dist = out_dist if out_dist is not None else (
            torch.empty((M, k), dtype=torch.float64, device=queries.device) if want_dist else None)
        _, wp = _host_weights(self.weights, self.dim)
        rc = _lib.lib().mpb_knn_search(_ptr(self.samples), self.n, _ptr(queries), M, self.dim, wp, int(k),
                                       float(max_dist), ws, nbytes, _ptr(idx), _ptr(dist), _ptr(stats), _stream())
        _lib.check(rc, "mpb_knn_search")
        return idx, dist


def knn(samples, queries, k, weights=None, max_dist=math.inf, method="f32filter", stats=None, metric=METRIC_KDTREE,
        circular=None):
    """k nearest stored samples per query, bit-exact to nearest.py:74-85 (see include/mpb200.h).
    method: "f32filter" (tiled fp32 filter + fp64 re-rank) or "f64" (exact brute force).
    metric: METRIC_KDTREE ranks like scipy's cKDTree; METRIC_PRIMITIVES / METRIC_SAMPLERS rank by the
    value the corresponding distance_fn returns (exact kernel only).
    circular: {coordinate: (lower, upper)}; the three-image metric of nearest.py:7-20 (kd-tree metric, exact
    kernel)."""
    samples = _f64(samples, "samples")
    queries = _f64(queries, "queries", samples.shape[1])
    circ = _host_circular(circular, samples.shape[1])
    if circ is not None:
        if metric != METRIC_KDTREE:
            raise ValueError("circular coordinates are defined for the kd-tree metric only (nearest.py:74-85)")
        N, D = samples.shape
        M = queries.shape[0]
        idx = torch.empty((M, k), dtype=torch.int64, device=queries.device)
        dist = torch.empty((M, k), dtype=torch.float64, device=queries.device)
        _, wp = _host_weights(weights, D)
        rc = _lib.lib().mpb_knn_f64_circular(_ptr(samples), N, _ptr(queries), M, D, wp, circ[1], circ[2], int(k),
                                             float(max_dist), _ptr(idx), _ptr(dist), _stream())
        _lib.check(rc, "mpb_knn_f64_circular")
        return idx, dist
    if metric != METRIC_KDTREE:
        N, D = samples.shape
        M = queries.shape[0]
        idx = torch.empty((M, k), dtype=torch.int64, device=queries.device)
        dist = torch.empty((M, k), dtype=torch.float64, device=queries.device)
        _, wp = _host_weights(weights, D)
        rc = _lib.lib().mpb_knn_f64_metric(_ptr(samples), N, _ptr(queries), M, D, wp, int(metric), int(k),
                                           float(max_dist), _ptr(idx), _ptr(dist), _stream())
        _lib.check(rc, "mpb_knn_f64_metric")
        return idx, dist
    if method == "f32filter":
        return KnnIndex(samples, weights).search(queries, k, max_dist, stats=stats)
    if method != "f64":
        raise ValueError("unknown method %r" % (method,))
    N, D = samples.shape
    M = queries.shape[0]
    idx = torch.empty((M, k), dtype=torch.int64, device=queries.device)
    dist = torch.empty((M, k), dtype=torch.float64, device=queries.device)
    _, wp = _host_weights(weights, D)
    rc = _lib.lib().mpb_knn_f64(_ptr(samples), N, _ptr(queries), M, D, wp, int(k), float(max_dist), _ptr(idx),
                                _ptr(dist), _stream())
    _lib.check(rc, "mpb_knn_f64")
    return idx, dist


def knn_tc_key_error(samples, queries, weights=None):
    """Measured rounding of the tensor-core filter's keys (mpb_knn_tc_key_error): -> (max over pairs of
    |key_tc - key_fp64| / (R^2 + 2 |q^| R), pairs measured).  The certificate budgets 32 * 2^-22 for it."""
    samples = _f64(samples, "samples")
    queries = _f64(queries, "queries", samples.shape[1])
    N, D = samples.shape
    M = queries.shape[0]
    nbytes = _lib.lib().mpb_knn_workspace_bytes(N, M, D)
    ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=samples.device)
    off = (-ws.data_ptr()) % 256
    out = torch.zeros(2, dtype=torch.int64, device=samples.device)
    _, wp = _host_weights(weights, D)
    rc = _lib.lib().mpb_knn_tc_key_error(_ptr(samples), N, _ptr(queries), M, D, wp, ctypes.c_void_p(ws.data_ptr() + off),
                                         ws.numel() - off, _ptr(out), _stream())
    _lib.check(rc, "mpb_knn_tc_key_error")
    host = out.cpu().numpy()
    return float(host[:1].view(np.float64)[0]), int(host[1])


def knn_static(samples, queries, k, weights=None, max_dist=math.inf, stats=None):
    """knn() on the library's built-in scratch array, for callers that must not allocate device memory
    (see mpb_knn_f32filter_static in include/mpb200.h for who those are and for the size limit).
    The output tensors are still allocated here by torch; pass through the C ABI directly to avoid that too."""
    samples = _f64(samples, "samples")
    queries = _f64(queries, "queries", samples.shape[1])
    N, D = samples.shape
    M = queries.shape[0]
    idx = torch.empty((M, k), dtype=torch.int64, device=queries.device)
    dist = torch.empty((M, k), dtype=torch.float64, device=queries.device)
    _, wp = _host_weights(weights, D)
    rc = _lib.lib().mpb_knn_f32filter_static(_ptr(samples), N, _ptr(queries), M, D, wp, int(k), float(max_dist),
                                             _ptr(idx), _ptr(dist), _ptr(stats), _stream())
    _lib.check(rc, "mpb_knn_f32filter_static")
    return idx, dist


def trees_nearest(nodes, sizes, targets, weights=None, metric=METRIC_SAMPLERS, active=None):
    """First arg-min nearest node per tree (primitives.py:21).  nodes [T, cap, D], sizes [T] int32,
    targets [T, D] -> (idx [T] int32, dist [T] float64)."""
    if nodes.dim() != 3 or nodes.dtype != torch.float64 or not nodes.is_cuda:
        raise TypeError("nodes must be a [T, cap, D] float64 CUDA tensor")
    nodes = nodes.contiguous()
    T, cap, D = nodes.shape
    targets = _f64(targets, "targets", D)
    sizes = sizes.contiguous()
    if sizes.dtype != torch.int32:
        raise TypeError("sizes must be int32")
    idx = torch.empty(T, dtype=torch.int32, device=nodes.device)
    dist = torch.empty(T, dtype=torch.float64, device=nodes.device)
    _, wp = _host_weights(weights, D)
    rc = _lib.lib().mpb_trees_nearest(_ptr(nodes), _ptr(sizes), T, cap, D, _ptr(targets), wp, int(metric),
                                      _ptr(active), _ptr(idx), _ptr(dist), _stream())
    _lib.check(rc, "mpb_trees_nearest")
    return idx, dist


def trees_append(nodes, parents, sizes, tree_ids, frm, to, lo, hi, reverse, parent0, overflow, step=STEP_SIZE):
    """Store the kept waypoints of one extension per entry at the end of resident trees (see
    mpb_trees_append).  -> out_last [B] int32; `overflow` (int32 [1]) counts entries that did not fit."""
    T, cap, D = nodes.shape
    B = tree_ids.shape[0]
    out_last = torch.empty(B, dtype=torch.int32, device=nodes.device)
    rc = _lib.lib().mpb_trees_append(_ptr(nodes), _ptr(parents), _ptr(sizes), T, cap, D, _ptr(tree_ids), _ptr(frm),
                                     _ptr(to), _ptr(lo), _ptr(hi), _ptr(reverse), _ptr(parent0), B, float(step),
                                     _ptr(out_last), _ptr(overflow), _stream())
    _lib.check(rc, "mpb_trees_append")
    return out_last


def scan_nearest_radius(nodes, q, radius, weights=None, metric=METRIC_SAMPLERS):
    """Nearest node and every node closer than `radius` (rrt_star.py:117,128).
    -> (nearest [1] int32, dist [n] float64, count [1] int32, list [n] int32)."""
    nodes = _f64(nodes, "nodes")
    n, D = nodes.shape
    q = q.contiguous()
    nearest = torch.empty(1, dtype=torch.int32, device=nodes.device)
    count = torch.empty(1, dtype=torch.int32, device=nodes.device)
    dist = torch.empty(n, dtype=torch.float64, device=nodes.device)
    lst = torch.empty(n, dtype=torch.int32, device=nodes.device)
    _, wp = _host_weights(weights, D)
    rc = _lib.lib().mpb_scan_nearest_radius(_ptr(nodes), n, D, _ptr(q), wp, int(metric), float(radius),
                                            _ptr(nearest), _ptr(dist), _ptr(count), _ptr(lst), _stream())
    _lib.check(rc, "mpb_scan_nearest_radius")
    return nearest, dist, count, lst


def peak_fma(dtype=torch.float32, iters=4096, repeats=5, packed=False):
    """Measured FMA issue peak of this GPU in FLOP/s (2 flops per FMA) -- bench.py's roofline denominator.
    packed=True measures the fp32 pipe through FFMA2 (two fused multiply-adds per instruction)."""
    props = torch.cuda.get_device_properties(torch.cuda.current_device())
    blocks, threads = props.multi_processor_count * 8, 256
    sink = torch.empty(blocks * threads, dtype=dtype, device="cuda")
    fn = _lib.lib().mpb_peak_fma_f32 if dtype == torch.float32 else _lib.lib().mpb_peak_fma_f64
    if packed:
        fn = _lib.lib().mpb_peak_fma_f32x2
    best = 0.0
    for _ in range(repeats + 1):
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(fn(blocks, threads, iters, _ptr(sink), _stream()), "mpb_peak_fma")
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        best = max(best, (2.0 if packed else 1.0) * 2.0 * 8 * iters * blocks * threads / (ms * 1e-3))
    return best
