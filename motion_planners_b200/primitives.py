"""The four primitives of the planner API, backed by the CUDA kernels.

Drop-in counterparts of the reference's factories
    motion_planners/tkinter/samplers.py:11-16   get_distance_fn(weights)
    motion_planners/tkinter/samplers.py:32-47   get_sample_fn(region, obstacles, only_cfree)
    motion_planners/tkinter/samplers.py:88-98   get_collision_fn(environment, obstacles)
    motion_planners/tkinter/samplers.py:113-127 get_extend_fn(circular, step_size, norm)
    motion_planners/primitives.py:55-62         get_embed_fn / get_distance_fn(weights, p_norm)
Each returned object is still callable one configuration at a time exactly like the reference's
closure (so unmodified reference planners accept it), and additionally carries the batched form the
drop-in planners of this package detect and use: `.batch(...)` methods taking [n, D] arrays/tensors.
A per-configuration call is one tiny kernel launch plus a device->host read: correct, not fast; the
speed is in the batched forms.
"""
import numpy as np
import torch

from . import ops
from .geometry import STEP_SIZE, World
from .utils import INF


def _to_dev(a, dim=None):
    if isinstance(a, torch.Tensor):
        t = a if a.is_cuda else a.cuda()
        return t.to(torch.float64).contiguous()
    arr = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    if arr.ndim == 1:
        arr = arr.reshape(1, -1)
    return torch.from_numpy(arr).cuda()


class GpuCollisionFn(object):
    """collision_fn(q) -> bool with a device-resident obstacle set (samplers.py:88-98)."""

    def __init__(self, world):
        self.world = world
        self._obstacles = None

    @property
    def obstacles(self):
        if self._obstacles is None:
            self._obstacles = ops.ObstacleSet(self.world)
        return self._obstacles

    def __call__(self, q):
        return bool(self.batch(np.asarray(q, dtype=np.float64).reshape(1, -1))[0].item())

    def batch(self, configs):
        """[n, D] configurations -> bool CUDA tensor [n]."""
        return self.obstacles.points_check(_to_dev(configs))


class GpuExtendFn(object):
    """extend_fn(q1, q2) -> waypoints (q1 excluded, last one ~ q2), samplers.py:113-127 with norm = INF:
    n = int(max|diff/step|) + 1 points of the serial recurrence.  `circular` = {coordinate: (lower, upper)}
    switches on the wrapped differences and positions of primitives.py:74-78 / utils.py:364-384."""

    def __init__(self, step_size=STEP_SIZE, circular=None):
        self.step_size = float(step_size)
        self.circular = dict(circular) if circular else {}

    def __call__(self, q1, q2):
        _, way = self.batch(np.asarray(q1, dtype=np.float64).reshape(1, -1),
                            np.asarray(q2, dtype=np.float64).reshape(1, -1))
        for row in way.cpu().numpy():
            yield row

    def batch(self, q1, q2):
        """-> (offsets [E+1] int64, waypoints [total, D]) CUDA tensors."""
        return ops.edges_waypoints(_to_dev(q1), _to_dev(q2), step=self.step_size, circular=self.circular)

    def counts(self, q1, q2):
        return ops.edges_count(_to_dev(q1), _to_dev(q2), step=self.step_size, circular=self.circular)


def require_plain_extend(extend_fn, who):
    """Planners whose batched edge checks are not wired for circular joints say so instead of ignoring them."""
    if getattr(extend_fn, 'circular', None):
        raise NotImplementedError('%s: circular joints are supported by lazy_prm and the primitives only' % who)


_HOST_DOT_CHECKED = False


def check_host_rounding():
    """The scalar distance_fn below evaluates the reference's own numpy expressions on the host, while the ranking
    kernels restate them operation by operation: numpy's `dot` / `linalg.norm` go through BLAS, and on the host
    the fixtures were made on that is a sequential chain of fused multiply-adds, s = fma(a_k, b_k, s).  That is a
    property of the host's BLAS build.  Here it is checked once, exactly, on a fixed vector set, so that a host
    whose `dot` rounds differently is reported instead of silently making the scalar and the batched form of the
    same distance_fn disagree in the last bit.  Raises RuntimeError."""
    global _HOST_DOT_CHECKED
    if _HOST_DOT_CHECKED:
        return
    from fractions import Fraction
    rs = np.random.RandomState(20261107)
    for D in (2, 3, 7, 14):
        for _ in range(8):
            a = rs.uniform(0.5, 2.0, size=D)
            b = rs.uniform(-1.0, 1.0, size=D)
            sq = b * b
            s = 0.0
            for k in range(D):                                   # exactly rounded fused multiply-add chain
                s = float(Fraction(a[k]) * Fraction(sq[k]) + Fraction(s))
            t = 0.0
            for k in range(D):
                t = float(Fraction(b[k]) * Fraction(b[k]) + Fraction(t))
            if float(np.dot(a, sq)) != s or float(np.linalg.norm(b, ord=2)) != float(np.sqrt(t)):
                raise RuntimeError(
                    "this host's numpy dot / linalg.norm do not round like a sequential fused multiply-add chain "
                    "(BLAS build differs from the one the reference fixtures were generated with): the scalar "
                    "distance_fn would disagree with the batched kernels in the last bit.  Use the batched forms "
                    "(distance_fn.nearest, ops.knn) or rebuild the fixtures with oracle/make_golden.py on this host.")
    _HOST_DOT_CHECKED = True


class GpuDistanceFn(object):
    """distance_fn(q1, q2) -> float for the weighted Euclidean family, tagged with the reference
    rounding procedure it follows (`metric`) so that batched ranking kernels reproduce its values bit
    for bit.  The scalar call evaluates the reference's own numpy expression on the host (it is the
    user-facing scalar primitive, e.g. for path costs); ranking and arg-min scans never go through it."""

    def __init__(self, weights, metric):
        self.weights = np.array(weights, dtype=np.float64)
        self.metric = int(metric)

    def __call__(self, q1, q2):
        if self.metric != ops.METRIC_KDTREE:
            check_host_rounding()                                   # once per process
        if self.metric == ops.METRIC_SAMPLERS:                      # samplers.py:13-15
            diff = np.array(q2) - np.array(q1)
            return np.sqrt(np.dot(self.weights, diff * diff))
        if self.metric == ops.METRIC_PRIMITIVES:                    # primitives.py:60-62
            return np.linalg.norm(self.weights * q2 - self.weights * q1, ord=2)
        e1, e2 = self.weights * q1, self.weights * q2               # the kd-tree metric (nearest.py:81)
        d = np.asarray(e1, dtype=np.float64) - np.asarray(e2, dtype=np.float64)
        sq = d * d
        n4 = (len(sq) // 4) * 4
        acc = [0.0, 0.0, 0.0, 0.0]
        for i in range(0, n4, 4):
            for j in range(4):
                acc[j] = acc[j] + sq[i + j]
        s = ((acc[0] + acc[1]) + acc[2]) + acc[3]
        for i in range(n4, len(sq)):
            s = s + sq[i]
        return np.sqrt(s)

    def nearest(self, nodes, targets, sizes=None, active=None):
        """Batched first arg-min (utils.py:61-65) over resident trees; see ops.trees_nearest."""
        return ops.trees_nearest(nodes, sizes, targets, weights=self.weights, metric=self.metric, active=active)


class EmbedFn(object):
    """get_embed_fn(weights) (primitives.py:55-57): q -> weights * q, with the weights exposed."""

    def __init__(self, weights):
        self.weights = np.array(weights, dtype=np.float64)

    def __call__(self, q):
        return self.weights * q


def get_embed_fn(weights):
    return EmbedFn(weights)


def get_collision_fn(environment, obstacles, buffer=None):
    """environment: a Box (or anything with .lower/.upper); obstacles: list of Box / Circle."""
    world = World.from_shapes(environment, obstacles) if buffer is None else \
        World.from_shapes(environment, obstacles, buffer)
    return GpuCollisionFn(world)


def get_extend_fn(circular={}, step_size=STEP_SIZE, norm=INF):
    if norm != INF:
        raise NotImplementedError('only the reference default norm=INF is accelerated')
    return GpuExtendFn(step_size, circular=circular)


def get_distance_fn(weights, p_norm=None):
    """With one argument: tkinter.samplers.get_distance_fn(weights).  With p_norm=2:
    primitives.get_distance_fn(weights, p_norm=2)."""
    if p_norm is None:
        return GpuDistanceFn(weights, ops.METRIC_SAMPLERS)
    if p_norm != 2:
        raise NotImplementedError('only p_norm=2 is accelerated')
    return GpuDistanceFn(weights, ops.METRIC_PRIMITIVES)


def get_sample_fn(region, obstacles=[], only_cfree=True, use_halton=False):
    """sample_fn() -> configuration drawn from numpy's GLOBAL generator exactly as the reference does
    (utils.py:246-249 uniform_generator + :242-243 convex_combination, samplers.py:32-47), so seeded
    runs consume the same stream; rejection against obstacles uses the GPU collision_fn."""
    if use_halton:
        raise NotImplementedError('ghalton is not available; the reference itself degrades to uniform sampling')
    lower = np.asarray(region.lower, dtype=np.float64)
    upper = np.asarray(region.upper, dtype=np.float64)
    collision_fn = get_collision_fn(region, obstacles) if (only_cfree and len(obstacles)) else None
    # without obstacles the reference's rejection test is the closed environment test alone (samplers.py:92):
    # two comparisons per coordinate, done here on the host instead of one kernel launch per drawn sample
    env_only = (lambda q: not (np.less_equal(lower, q).all() and np.less_equal(q, upper).all())) \
        if (only_cfree and not len(obstacles)) else None
    degenerate = bool(np.equal(lower, upper).all())
    state = {'used': False}

    def sample_fn():
        while True:
            if degenerate:                                    # utils.py:284-285: a one-shot iterator
                if state['used']:
                    return None
                state['used'] = True
                q = lower
            else:
                w = np.random.uniform(size=len(lower))
                q = (1 - w) * np.array(lower) + w * np.array(upper)
            test = collision_fn or env_only
            if test is not None and test(q):
                continue
            return q

    # what a caller needs to draw the same stream in bulk (rrt_star's device loop): n calls of a plain box sampler
    # = np.random.uniform(size=(n, D)) mapped by the same convex combination
    sample_fn.lower, sample_fn.upper = lower, upper
    sample_fn.plain_box = (collision_fn is None) and not degenerate

    def bulk(n):
        w = np.random.uniform(size=(int(n), len(lower)))
        return (1 - w) * np.array(lower) + w * np.array(upper)
    sample_fn.bulk = bulk
    return sample_fn
