"""Shortcut smoothing of a path (the step right after every planner).

Same contract as the reference's motion_planners/smoothing.py:51-144 smooth_path: per iteration two
segments are drawn with probability proportional to their length (np.random.choice), one point on each
(random.choice over the segment's waypoints), and the bridge between the two points replaces the part of
the path between them when it is shorter and collision free.  The random draws are made by the same
library calls in the same order as the reference, so a seeded run consumes the same streams.
The two point checks and the bridge's edge check of an iteration go to the device in one launch each
when the accelerated primitives are in use.
"""
import time
from random import choice

import numpy as np

from . import ops
from .primitives import GpuCollisionFn, GpuExtendFn, require_plain_extend
from .utils import INF, bisect_order, compute_path_cost, elapsed_time, get_pairs


def _unit(vec):
    norm = np.linalg.norm(vec)
    return vec if norm == 0 else np.array(vec) / norm


def remove_redundant(path, tolerance=1e-3):
    """Drop configurations closer than `tolerance` (per coordinate) to the last kept one (utils.py:205-212)."""
    assert path
    kept = [path[0]]
    for conf in path[1:]:
        diff = np.array(kept[-1]) - np.array(conf)
        if not np.allclose(np.zeros(len(diff)), diff, atol=tolerance, rtol=0):
            kept.append(conf)
    return kept


def waypoints_from_path(path, tolerance=1e-3):
    """Corner points of a densely sampled path: a configuration is kept where the unit direction from the
    last kept corner changes by more than `tolerance` (utils.py:215-232)."""
    path = remove_redundant(path, tolerance=tolerance)
    if len(path) < 2:
        return path
    corners = [path[0]]
    last_conf = path[1]
    last_dir = _unit(np.array(last_conf) - np.array(corners[-1]))
    for conf in path[2:]:
        direction = _unit(np.array(conf) - np.array(corners[-1]))
        if not np.allclose(last_dir, direction, atol=tolerance, rtol=0):
            corners.append(last_conf)
            direction = _unit(np.array(conf) - np.array(corners[-1]))
        last_conf = conf
        last_dir = direction
    corners.append(last_conf)
    return corners


def refine_waypoints(waypoints, extend_fn):
    """Concatenated extend_fn output over consecutive corners (utils.py:235-238)."""
    out = []
    for q1, q2 in get_pairs(waypoints):
        out.extend(extend_fn(q1, q2))
    return out


def _bridge_collides(shortcut, extend_fn, collision_fn):
    """any(map(collision_fn, shortcut)) or any collision along refine_waypoints(shortcut)[:-1]."""
    if isinstance(collision_fn, GpuCollisionFn) and isinstance(extend_fn, GpuExtendFn):
        pts = np.ascontiguousarray(np.array([np.asarray(p, dtype=np.float64) for p in shortcut]))
        if bool(collision_fn.batch(pts).any().item()):
            return True
        import torch
        hit, _ = collision_fn.obstacles.edges_check_pairs(torch.from_numpy(pts[:1]).cuda(),
                                                          torch.from_numpy(pts[1:2]).cuda(), mode=ops.EDGE_ANY,
                                                          skip_last=True, step=extend_fn.step_size, want_counts=False)
        return bool(hit[0].item() >= 0)
    if any(collision_fn(q) for q in shortcut):
        return True
    dense = refine_waypoints(shortcut, extend_fn)[:-1]
    return any(collision_fn(dense[i]) for i in bisect_order(len(dense)))


def smooth_path(path, extend_fn, collision_fn, distance_fn=None, cost_fn=None, sample_fn=None, max_iterations=100,
                max_time=INF, converge_fraction=0.0, converge_time=INF, print_frequency=1., verbose=False):
    """:return: smoothed, re-discretised path (smoothing.py:51-144)."""
    require_plain_extend(extend_fn, 'smooth_path')
    start_time = last_time = time.time()
    if not path or not max_iterations:
        return path
    assert (max_iterations < INF) or (max_time < INF) or (converge_time < INF)
    if distance_fn is None:
        distance_fn = lambda q1, q2: len(list(extend_fn(q1, q2)))      # primitives.py:65-70
    if cost_fn is None:
        cost_fn = distance_fn
    waypoints = waypoints_from_path(path)
    cost = compute_path_cost(waypoints, cost_fn=cost_fn)
    iteration = 0
    while iteration < max_iterations:
        iteration += 1
        if (max_time - elapsed_time(start_time) <= 0.) or (converge_time - elapsed_time(last_time) <= 0.) or \
                (len(waypoints) <= 2):
            break
        segments = get_pairs(range(len(waypoints)))
        weights = [distance_fn(waypoints[i], waypoints[j]) for i, j in segments]
        probabilities = np.array(weights) / sum(weights)
        seg_indices = list(range(len(segments)))
        seg_idx1, seg_idx2 = np.random.choice(seg_indices, size=2, replace=True, p=probabilities)
        if seg_idx1 == seg_idx2:
            continue
        if seg_idx2 < seg_idx1:
            seg_idx1, seg_idx2 = seg_idx2, seg_idx1
        segment1, segment2 = segments[seg_idx1], segments[seg_idx2]
        paths = [list(extend_fn(waypoints[i], waypoints[j]))[:-1] for i, j in [segment1, segment2]]
        if not all(paths):
            continue
        point1, point2 = [choice(p) for p in paths]
        i, _ = segment1
        _, j = segment2
        shortcut = [point1, point2]
        new_waypoints = waypoints[:i + 1] + shortcut + waypoints[j:]
        new_cost = compute_path_cost(new_waypoints, cost_fn=cost_fn)
        improve_fraction = (cost - new_cost) / cost
        if (new_cost >= cost) or (improve_fraction <= converge_fraction) or \
                _bridge_collides(shortcut, extend_fn, collision_fn):
            continue
        waypoints = new_waypoints
        cost = new_cost
        last_time = time.time()
    return [waypoints[0]] + refine_waypoints(waypoints, extend_fn)


def smooth_paths_batch(paths, extend_fn, collision_fn, distance_fn, seeds, max_iterations=100, cost_fn=None):
    """smooth_path for MANY paths in lock step (configs[4]: "linear shortcut smoothing of 10k paths").

    Path i consumes its own generators random.Random(seeds[i]) / numpy RandomState(seeds[i]) with the same
    calls in the same order as smooth_path does with the global ones, so the result for path i equals

        random.seed(seeds[i]); np.random.seed(seeds[i]); smooth_path(paths[i], ..., max_iterations)

    Per iteration the device sees three launches for the whole batch instead of three per path: the waypoints
    of the two drawn segments of every path (extend_fn), the collision verdicts of all chosen points, and the
    verdicts of all bridges.  The per-path bookkeeping (segment draw, cost test, splice) stays on the host.
    Requires the accelerated primitives."""
    require_plain_extend(extend_fn, 'smooth_paths_batch')
    import random as pyrandom
    import torch
    if not (isinstance(collision_fn, GpuCollisionFn) and isinstance(extend_fn, GpuExtendFn)):
        raise TypeError('smooth_paths_batch needs the accelerated primitives of motion_planners_b200.primitives')
    if cost_fn is None:
        cost_fn = distance_fn
    n = len(paths)
    way = [waypoints_from_path(p) if p else p for p in paths]
    same_fn = cost_fn is distance_fn
    # per path: cost_fn of every segment, in order (compute_path_cost is their sequential sum from the left, which
    # the candidate test below re-does with three new terms in the middle instead of calling cost_fn on every
    # segment again), and distance_fn of every segment (the draw's weights)
    segc = [[cost_fn(w[a], w[a + 1]) for a in range(len(w) - 1)] if w else None for w in way]
    cost = [sum(c) if c is not None else INF for c in segc]
    rpy = [pyrandom.Random(int(s)) for s in seeds]
    rnp = [np.random.RandomState(int(s)) for s in seeds]
    weights = [None] * n                     # cached segment lengths, recomputed when a path changes
    cdf = [None] * n                         # ... and the cumulative distribution RandomState.choice builds from them
    obs = collision_fn.obstacles
    for _ in range(int(max_iterations)):
        # 1. per path: draw the two segments (host; the generator calls of smooth_path: choice(p=...) of the legacy
        # RandomState is cdf = p.cumsum(); cdf /= cdf[-1]; cdf.searchsorted(random_sample(size), side='right'))
        drawn = []
        for i in range(n):
            w = way[i]
            if not w or len(w) <= 2:
                continue
            if cdf[i] is None:
                if weights[i] is None:
                    weights[i] = segc[i] if same_fn else [distance_fn(w[a], w[a + 1]) for a in range(len(w) - 1)]
                prob = np.array(weights[i]) / sum(weights[i])
                c = prob.cumsum()
                c /= c[-1]
                cdf[i] = c
            s1, s2 = cdf[i].searchsorted(rnp[i].random_sample(2), side='right')
            if s1 == s2:
                continue
            if s2 < s1:
                s1, s2 = s2, s1
            drawn.append((i, int(s1), int(s2)))
        if not drawn:
            if all((not w) or len(w) <= 2 for w in way):
                break
            continue
        # 2. waypoints of both segments of every drawn path in one launch
        q1 = np.array([way[i][s] for i, s1, s2 in drawn for s in (s1, s2)], dtype=np.float64)
        q2 = np.array([way[i][s + 1] for i, s1, s2 in drawn for s in (s1, s2)], dtype=np.float64)
        off, pts = extend_fn.batch(q1, q2)
        off, pts = off.cpu().numpy(), pts.cpu().numpy()
        cand = []
        for j, (i, s1, s2) in enumerate(drawn):
            a0, a1 = int(off[2 * j]), int(off[2 * j + 1]) - 1       # list(extend_fn(...))[:-1]
            b0, b1 = int(off[2 * j + 1]), int(off[2 * j + 2]) - 1
            if a1 <= a0 or b1 <= b0:
                continue
            p1 = pts[a0 + rpy[i].choice(range(a1 - a0))]              # random.choice(seq) = seq[_randbelow(len(seq))]
            p2 = pts[b0 + rpy[i].choice(range(b1 - b0))]
            w, sc = way[i], segc[i]
            three = [cost_fn(w[s1], p1), cost_fn(p1, p2), cost_fn(p2, w[s2 + 1])]
            new_cost = 0
            for a in range(s1):
                new_cost = new_cost + sc[a]
            for c in three:
                new_cost = new_cost + c
            for a in range(s2 + 1, len(sc)):
                new_cost = new_cost + sc[a]
            if new_cost >= cost[i] or (cost[i] - new_cost) / cost[i] <= 0.0:
                continue
            cand.append((i, p1, p2, s1, s2, three, new_cost))
        if not cand:
            continue
        # 3. verdicts of all bridges: both end points, then the edge between them without its last waypoint
        a = torch.from_numpy(np.ascontiguousarray(np.array([c[1] for c in cand]))).cuda()
        b = torch.from_numpy(np.ascontiguousarray(np.array([c[2] for c in cand]))).cuda()
        bad = obs.points_check(a) | obs.points_check(b)
        hit, _ = obs.edges_check_pairs(a, b, mode=ops.EDGE_ANY, skip_last=True, step=extend_fn.step_size,
                                       want_counts=False)
        bad = (bad | (hit >= 0)).cpu().numpy()
        for ok, (i, p1, p2, s1, s2, three, new_cost) in zip(~bad, cand):
            if ok:
                w = way[i]
                way[i] = w[:s1 + 1] + [p1, p2] + w[s2 + 1:]
                segc[i] = segc[i][:s1] + three + segc[i][s2 + 1:]
                cost[i], weights[i], cdf[i] = new_cost, None, None
    # 4. final re-discretisation of every path, one launch
    out = [None] * n
    idx = [i for i in range(n) if way[i]]
    pairs = [(i, a) for i in idx for a in range(len(way[i]) - 1)]
    if pairs:
        q1 = np.array([way[i][a] for i, a in pairs], dtype=np.float64)
        q2 = np.array([way[i][a + 1] for i, a in pairs], dtype=np.float64)
        off, pts = extend_fn.batch(q1, q2)
        off, pts = off.cpu().numpy(), pts.cpu().numpy()
        for j, (i, a) in enumerate(pairs):
            if out[i] is None:
                out[i] = [way[i][0]]
            out[i].extend(list(pts[off[j]:off[j + 1]]))
    for i in range(n):
        if out[i] is None:
            out[i] = [way[i][0]] if way[i] else paths[i]
    return out
