"""Small host-side helpers shared by the planner mirrors.

Behavioural counterparts of a few functions in the reference's motion_planners/utils.py; only what
the hot-path callers need.  Cited lines give the behaviour each one must keep."""
import time
from collections import deque

import numpy as np

INF = float('inf')

# defaults of the tree planners (utils.py:17-19)
RRT_ITERATIONS = 20
RRT_RESTARTS = 2
RRT_SMOOTHING = 20


def elapsed_time(start_time):
    return time.time() - start_time


def first_argmin(scores):
    """Index of the first minimum (utils.py:61-65: scores.index(min(scores)))."""
    best, where = None, -1
    for i, s in enumerate(scores):
        if best is None or s < best:
            best, where = s, i
    return where


def get_pairs(sequence):
    seq = list(sequence)
    return [(seq[i], seq[i + 1]) for i in range(len(seq) - 1)]


def bisect_order(n):
    """Index visiting order of the reference's default collision-check selector for a sequence of
    length n (utils.py:99-114, 303-307): breadth-first midpoints, index = int((lo + hi) / 2.)."""
    order = []
    queue = deque([(0, n - 1)])
    while queue:
        lo, hi = queue.popleft()
        if lo > hi:
            continue
        mid = int((lo + hi) / 2.)
        order.append(mid)
        queue.append((lo, mid - 1))
        queue.append((mid + 1, hi))
    return order


def default_selector(sequence):
    seq = list(sequence)
    return [seq[i] for i in bisect_order(len(seq))]


def compute_path_cost(path, cost_fn):
    """Sum of cost_fn over consecutive pairs; INF for a missing path (utils.py:191-196)."""
    if path is None:
        return INF
    return sum(cost_fn(a, b) for a, b in get_pairs(path))


def flatten(list_of_lists):
    return [item for sub in list_of_lists for item in sub]


def as_config_array(configs, dim=None):
    """Stack a list of configurations (tuples, lists or arrays) into a contiguous float64 [n, D] array."""
    if len(configs) == 0:
        return np.zeros((0, dim or 0))
    arr = np.ascontiguousarray(np.array([np.asarray(c, dtype=np.float64) for c in configs], dtype=np.float64))
    return arr.reshape(len(configs), -1)


ALLOW_HOST_PLANNERS = False


def require_device_primitives(who, allow_host=False):
    """The planners of this package run their nearest-neighbour and collision work on the GPU; that needs the
    batched primitives of motion_planners_b200.primitives (GpuDistanceFn / GpuExtendFn / GpuCollisionFn).  Handed
    anything else they refuse -- there is no silent CPU path -- unless the caller opts into the plain host loop
    (the reference's own statement order, useful to cross-check a custom callable) with allow_host=True or by
    setting motion_planners_b200.utils.ALLOW_HOST_PLANNERS = True."""
    if allow_host or ALLOW_HOST_PLANNERS:
        return
    raise TypeError("%s: distance_fn / extend_fn / collision_fn must be the batched primitives of "
                    "motion_planners_b200.primitives (there is no CPU fallback); pass allow_host=True to run the "
                    "host loop on arbitrary callables" % who)
