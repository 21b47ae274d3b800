"""RRT* with the node scan and the edge checks on the device.

Same entry point and behaviour as the reference's motion_planners/rrt_star.py:86-146 AS EXECUTED ON
PYTHON 3: per iteration the nearest node (first arg-min of distance_fn), the collision-free prefix of the
extension towards the sample, one new node, and the choose-parent pass over the nodes closer than
`radius` (strict <, the new node included).  The reference's second, "rewire the neighbours" loop iterates
an already exhausted `filter` object and therefore never runs (rrt_star.py:138-143); it is not
reproduced.

Device work per iteration: one fused distance scan over all nodes (mpb_scan_nearest_radius: arg-min and
the ordered radius list share the distances), the first-hit edge check of the extension, and one batched
first-hit check of every candidate parent -> new edge.  The tree bookkeeping (costs, children, solution
flags: OptimalNode, rrt_star.py:11-73) is sequential and stays on the host; node configurations live in a
device array that grows geometrically.
"""
import time
from random import random

import numpy as np
import torch

from . import ops
from .primitives import GpuCollisionFn, GpuDistanceFn, GpuExtendFn, require_plain_extend
from .utils import INF, elapsed_time, first_argmin

EPSILON = 1e-6          # rrt_star.py:8
PRINT_FREQUENCY = 100


class OptimalNode(object):
    """Tree node with cost-to-come (rrt_star.py:11-73).  `path` holds the waypoints strictly between the
    parent's configuration and this one; it is kept as a recipe (from, to, count) and expanded on retrace."""

    def __init__(self, config, parent=None, d=0, path=None, iteration=None):
        self.config = config
        self.parent = parent
        self.children = set()
        self.d = d
        self.path = path
        if parent is not None:
            self.cost = parent.cost + d
            parent.children.add(self)
        else:
            self.cost = d
        self.solution = False
        self.creation = iteration
        self.last_rewire = iteration

    def set_solution(self, solution):
        node = self
        while node is not None and node.solution is not solution:
            node.solution = solution
            node = node.parent

    def rewire(self, parent, d, path, iteration=None):
        if self.solution:
            self.parent.set_solution(False)
        self.parent.children.remove(self)
        self.parent = parent
        parent.children.add(self)
        if self.solution:
            parent.set_solution(True)
        self.d = d
        self.path = path
        self.update()
        self.last_rewire = iteration

    def update(self):
        stack = [self]
        while stack:
            n = stack.pop()
            n.cost = n.parent.cost + n.d
            stack.extend(n.children)

    def retrace(self, expand):
        chain = []
        node = self
        while node is not None:
            chain.append(node)
            node = node.parent
        out = []
        for node in reversed(chain):
            out.extend(expand(node.path))
            out.append(node.config)
        return out


def _safe_path_reference(sequence, collision):
    path = []
    for q in sequence:
        if collision(q):
            break
        path.append(q)
    return path


def _rrt_star_reference_order(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, radius, max_time,
                              max_iterations, goal_probability, informed):
    """Arbitrary callables: the same algorithm call by call."""
    if collision_fn(start) or collision_fn(goal):
        return None
    nodes = [OptimalNode(start)]
    goal_n = None
    start_time = time.time()
    iteration = 0
    while (elapsed_time(start_time) < max_time) and (iteration < max_iterations):
        do_goal = goal_n is None and (iteration == 0 or random() < goal_probability)
        s = goal if do_goal else sample_fn()
        if informed and (goal_n is not None) and (distance_fn(start, s) + distance_fn(s, goal) >= goal_n.cost):
            continue
        iteration += 1
        nearest = nodes[first_argmin([distance_fn(n.config, s) for n in nodes])]
        path = _safe_path_reference(extend_fn(nearest.config, s), collision_fn)
        if len(path) == 0:
            continue
        new = OptimalNode(path[-1], parent=nearest, d=distance_fn(nearest.config, path[-1]), path=path[:-1],
                          iteration=iteration)
        if do_goal and (distance_fn(new.config, goal) < EPSILON):
            goal_n = new
            goal_n.set_solution(True)
        nodes.append(new)
        for n in [m for m in nodes if distance_fn(m.config, new.config) < radius]:
            d = distance_fn(n.config, new.config)
            if (n.cost + d) < new.cost:
                path = _safe_path_reference(extend_fn(n.config, new.config), collision_fn)
                if (len(path) != 0) and (distance_fn(new.config, path[-1]) < EPSILON):
                    new.rewire(n, d, path[:-1], iteration=iteration)
    if goal_n is None:
        return None
    return goal_n.retrace(lambda p: p or [])


def rrt_star(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, radius, max_time=INF, max_iterations=INF,
             goal_probability=.2, informed=True, verbose=False):
    """:return: path [q', ..., q"] or None (rrt_star.py:86-146)."""
    require_plain_extend(extend_fn, 'rrt_star')
    if not (isinstance(distance_fn, GpuDistanceFn) and isinstance(extend_fn, GpuExtendFn)
            and isinstance(collision_fn, GpuCollisionFn)):
        return _rrt_star_reference_order(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, radius,
                                         max_time, max_iterations, goal_probability, informed)
    start = np.asarray(start, dtype=np.float64)
    goal = np.asarray(goal, dtype=np.float64)
    D = len(start)
    if bool(collision_fn.batch(np.stack([start, goal])).any().item()):
        return None
    obs = collision_fn.obstacles
    step = extend_fn.step_size
    dev = torch.device('cuda')
    cap = 4096
    table = torch.zeros((cap, D), dtype=torch.float64, device=dev)
    table[0] = torch.from_numpy(start)
    nodes = [OptimalNode(start)]
    goal_n = None
    launches = 0

    def waypoint(frm, to, index):
        """Waypoint `index` of extend_fn(frm, to) (host copy)."""
        off, way = extend_fn.batch(frm.reshape(1, -1), to.reshape(1, -1))
        return way[index].cpu().numpy()

    def expand(recipe):
        if not recipe:
            return []
        frm, to, count = recipe
        if count <= 0:
            return []
        _, way = extend_fn.batch(frm.reshape(1, -1), to.reshape(1, -1))
        return [w for w in way[:count].cpu().numpy()]

    start_time = time.time()
    iteration = 0
    while (elapsed_time(start_time) < max_time) and (iteration < max_iterations):
        do_goal = goal_n is None and (iteration == 0 or random() < goal_probability)
        s = goal if do_goal else np.asarray(sample_fn(), dtype=np.float64)
        if informed and (goal_n is not None) and (distance_fn(start, s) + distance_fn(s, goal) >= goal_n.cost):
            continue
        if verbose and iteration % PRINT_FREQUENCY == 0:
            print('Iteration: {} | Time: {:.3f} | Success: {} | Cost: {:.3f}'.format(
                iteration, elapsed_time(start_time), goal_n is not None, goal_n.cost if goal_n else INF))
        iteration += 1

        n_nodes = len(nodes)
        s_dev = torch.from_numpy(s).to(dev)
        nearest_i, _, _, _ = ops.scan_nearest_radius(table[:n_nodes], s_dev, 0.0, weights=distance_fn.weights,
                                                     metric=distance_fn.metric)
        nearest = nodes[int(nearest_i.item())]
        near_dev = torch.from_numpy(np.asarray(nearest.config, dtype=np.float64)).to(dev).reshape(1, -1)
        first, cnt = obs.edges_check_pairs(near_dev, s_dev.reshape(1, -1), mode=ops.EDGE_FIRST, step=step)
        launches += 3
        first, cnt = int(first.item()), int(cnt.item())
        safe = cnt if first < 0 else first               # len(safe_path(...)), rrt_star.py:118
        if safe == 0:
            continue
        new_config = waypoint(np.asarray(nearest.config, dtype=np.float64), s, safe - 1)
        new = OptimalNode(new_config, parent=nearest, d=distance_fn(nearest.config, new_config),
                          path=(np.asarray(nearest.config, dtype=np.float64), s, safe - 1), iteration=iteration)
        if do_goal and (distance_fn(new.config, goal) < EPSILON):
            goal_n = new
            goal_n.set_solution(True)
        if n_nodes == cap:                                # grow the device table
            cap *= 2
            bigger = torch.zeros((cap, D), dtype=torch.float64, device=dev)
            bigger[:n_nodes] = table[:n_nodes]
            table = bigger
        new_dev = torch.from_numpy(new_config).to(dev)
        table[n_nodes] = new_dev
        nodes.append(new)

        # choose-parent pass over the nodes strictly closer than `radius` (rrt_star.py:128-137)
        _, dist, count, lst = ops.scan_nearest_radius(table[:n_nodes + 1], new_dev, radius,
                                                      weights=distance_fn.weights, metric=distance_fn.metric)
        c = int(count.item())
        launches += 2
        if c == 0:
            continue
        cand = lst[:c].long()
        d_all = dist[cand].cpu().numpy()
        cand_host = cand.cpu().numpy()
        q1 = table[cand]
        q2 = new_dev.reshape(1, -1).expand(c, D).contiguous()
        first_c, cnt_c = obs.edges_check_pairs(q1, q2, mode=ops.EDGE_FIRST, step=step)
        first_c, cnt_c = first_c.cpu().numpy(), cnt_c.cpu().numpy()
        launches += 1
        for j in range(c):
            n = nodes[int(cand_host[j])]
            d = float(d_all[j])
            if (n.cost + d) < new.cost:
                safe_j = int(cnt_c[j]) if first_c[j] < 0 else int(first_c[j])
                if safe_j == 0:
                    continue
                n_cfg = np.asarray(n.config, dtype=np.float64)
                end = new_config if (first_c[j] < 0 and _ends_on_target(n_cfg, new_config, step)) else \
                    waypoint(n_cfg, new_config, safe_j - 1)
                if distance_fn(new.config, end) < EPSILON:
                    new.rewire(n, d, (n_cfg, new_config, safe_j - 1), iteration=iteration)
    rrt_star.last_launches = launches
    if goal_n is None:
        return None
    return goal_n.retrace(expand)


def _ends_on_target(q1, q2, step):
    """False: always fetch the true last waypoint (the recurrence's final value may differ from q2 by an
    ulp; the EPSILON test is insensitive to it, but the stored configuration must be the real one)."""
    return False
