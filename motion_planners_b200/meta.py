"""Planner dispatch and restarts (the reference's motion_planners/meta.py), on the accelerated planners.

direct_path / check_direct (meta.py:15-42), random_restarts (meta.py:46-86) and the string dispatcher
solve (meta.py:94-146) keep their signatures and return conventions: None = no path, False from
check_direct = an endpoint collides."""
import time

import numpy as np

from . import ops
from .primitives import GpuCollisionFn, GpuExtendFn
from .smoothing import smooth_path
from .utils import INF, RRT_RESTARTS, RRT_SMOOTHING, bisect_order, compute_path_cost, elapsed_time


def direct_path(start, goal, extend_fn, collision_fn):
    """The straight extension start -> goal if it is collision free, else None (meta.py:15-31)."""
    if collision_fn(start) or collision_fn(goal):
        return None
    path = [start] + list(extend_fn(start, goal))
    if isinstance(collision_fn, GpuCollisionFn):
        pts = np.ascontiguousarray(np.array([np.asarray(q, dtype=np.float64) for q in path]))
        if bool(collision_fn.batch(pts).any().item()):
            return None
        return path
    if any(collision_fn(path[i]) for i in bisect_order(len(path))):
        return None
    return path


def check_direct(start, goal, extend_fn, collision_fn):
    if any(collision_fn(q) for q in [start, goal]):
        return False
    return direct_path(start, goal, extend_fn, collision_fn)


def random_restarts(solve_fn, start, goal, distance_fn, sample_fn, extend_fn, collision_fn, restarts=RRT_RESTARTS,
                    smooth=RRT_SMOOTHING, success_cost=0., max_time=INF, max_solutions=1, verbose=False, **kwargs):
    """:return: list of paths sorted by cost, or None when an endpoint collides (meta.py:46-86)."""
    start_time = time.time()
    solutions = []
    path = check_direct(start, goal, extend_fn, collision_fn)
    if path is False:
        return None
    if path is not None:
        solutions.append(path)
    attempt = 0
    while attempt < restarts + 1:
        if (len(solutions) >= max_solutions) or (elapsed_time(start_time) >= max_time):
            break
        attempt_time = max_time - elapsed_time(start_time)
        path = solve_fn(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, max_time=attempt_time,
                        verbose=verbose, **kwargs)
        attempt += 1
        if path is None:
            continue
        if smooth:
            path = smooth_path(path, extend_fn, collision_fn, distance_fn=distance_fn, max_iterations=smooth,
                               max_time=max_time - elapsed_time(start_time), verbose=verbose)
        solutions.append(path)
        if compute_path_cost(path, distance_fn) < success_cost:
            break
    solutions = sorted(solutions, key=lambda p: compute_path_cost(p, distance_fn))
    if verbose:
        print('Attempts: {} | Solutions: {} | Time: {:.3f}'.format(attempt, len(solutions), elapsed_time(start_time)))
    return solutions


def solve(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, algorithm='birrt', max_time=INF,
          max_iterations=INF, num_samples=100, smooth=None, smooth_time=INF, weights=None, circular={}, cost_fn=None,
          success_cost=INF, verbose=False, **kwargs):
    """String dispatch over the planners (meta.py:94-146)."""
    from .lazy_prm import lazy_prm
    from .prm import prm
    from .rrt_connect import birrt, rrt_connect
    from .rrt_star import rrt_star
    start_time = time.time()
    path = check_direct(start, goal, extend_fn, collision_fn)
    if (path is not None) or (algorithm == 'direct'):
        return path
    remaining = max_time - elapsed_time(start_time)
    if algorithm == 'prm':
        path = prm(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, num_samples=num_samples)
    elif algorithm == 'lazy_prm':
        if weights is not None:
            distance_fn = None
        path = lazy_prm(start, goal, sample_fn, extend_fn, collision_fn, num_samples=num_samples, max_time=remaining,
                        weights=weights, circular=circular, distance_fn=distance_fn, cost_fn=cost_fn,
                        verbose=verbose)[0]
    elif algorithm == 'lazy_prm_star':
        from .lazy_prm import create_param_sequence, lazy_prm_star
        if weights is not None:
            distance_fn = None
        path = lazy_prm_star(start, goal, sample_fn, extend_fn, collision_fn,
                             param_sequence=create_param_sequence(initial_samples=num_samples), max_time=remaining,
                             weights=weights, circular=circular, distance_fn=distance_fn, cost_fn=cost_fn,
                             success_cost=success_cost, verbose=verbose)
    elif algorithm == 'rrt':
        from .rrt import rrt
        path = rrt(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, max_iterations=max_iterations,
                   max_time=remaining)
    elif algorithm == 'rrt_connect':
        path = rrt_connect(start, goal, distance_fn, sample_fn, extend_fn, collision_fn,
                           max_iterations=max_iterations, max_time=remaining)
    elif algorithm == 'birrt':
        path = birrt(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, max_iterations=max_iterations,
                     max_time=remaining, smooth=None, **kwargs)
    elif algorithm == 'rrt_star':
        path = rrt_star(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, radius=1,
                        max_iterations=max_iterations, max_time=remaining)
    else:
        raise NotImplementedError(algorithm)     # meta.py:139 (lattice: out of scope here)
    if not smooth:
        return path
    remaining = min(smooth_time, max_time - elapsed_time(start_time))
    return smooth_path(path, extend_fn, collision_fn, distance_fn=distance_fn, cost_fn=cost_fn, max_iterations=smooth,
                       max_time=remaining, verbose=verbose)
