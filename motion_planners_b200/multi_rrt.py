"""Multi-goal trees: MultiTree / MultiRRT / MultiBiRRT (motion_planners/multi_rrt.py:16-159) on the device
primitives.

The reference keeps every node in a dictionary keyed by its configuration (multi_rrt.py:25-27), so configurations
must be hashable: with the reference's own numpy-yielding extend_fn (tkinter/samplers.py:113-127) `add` raises
`TypeError: unhashable type`.  Here every configuration that enters the tree is stored as a tuple of Python
floats -- the form in which the reference's planner works unchanged -- and results are compared with runs of the
unmodified reference driven through tuple-yielding wrappers of its own primitives (tests/golden/multi_rrt_*.npz).

Statement order, random-number order (`random()` per iteration, `shuffle` of the node list per grow call, the
caller's sample_fn) and return conventions are the reference's.  Per iteration the device does what the Python
loops of the reference do: the first arg-min of distance_fn over `nodes + new_nodes` (mpb_trees_nearest on a
resident table in exactly that order) and extend_fn + collision_fn up to the first colliding waypoint
(mpb_edges_waypoints + mpb_edges_check_pairs, mode FIRST).

One deviation: MultiRRT.grow given a configuration instead of a callable samples that configuration (what
multi_rrt.py:74-75 means to do; its `lambda: goal_sample` refers to itself and returns the lambda)."""
from collections.abc import Mapping
from random import random, shuffle

import numpy as np
import torch

from . import ops, utils
from .primitives import GpuCollisionFn, GpuDistanceFn, GpuExtendFn, require_plain_extend
from .rrt import TreeNode, configs
from .utils import get_pairs


class _Enum(object):
    ALL, SUCCESS, PATH, NONE = range(4)
    names = ['ALL', 'SUCCESS', 'PATH', 'NONE']


ts = _Enum


def _tup(q):
    return tuple(float(x) for x in np.asarray(q, dtype=np.float64).reshape(-1))


class _Table(object):
    """Device copy of a node list in list order (the order argmin scans): grows by doubling."""

    def __init__(self, dim, nodes):
        self.dim = dim
        self.cap = max(256, 2 * len(nodes))
        self.buf = torch.zeros((1, self.cap, dim), dtype=torch.float64, device='cuda')
        self.size = torch.zeros(1, dtype=torch.int32, device='cuda')
        self.n = 0
        self.extend([n.config for n in nodes])

    def extend(self, cfgs):
        if not cfgs:
            return
        need = self.n + len(cfgs)
        if need > self.cap:
            while self.cap < need:
                self.cap *= 2
            buf = torch.zeros((1, self.cap, self.dim), dtype=torch.float64, device='cuda')
            buf[0, :self.n] = self.buf[0, :self.n]
            self.buf = buf
        self.buf[0, self.n:need] = torch.from_numpy(np.asarray(cfgs, dtype=np.float64).reshape(len(cfgs), self.dim)).cuda()
        self.n = need
        self.size.fill_(need)


class MultiTree(Mapping, object):

    def __init__(self, start, distance_fn, sample_fn, extend_fn, collision_fn, allow_host=False):
        require_plain_extend(extend_fn, 'MultiTree')
        if not (isinstance(distance_fn, GpuDistanceFn) and isinstance(extend_fn, GpuExtendFn)
                and isinstance(collision_fn, GpuCollisionFn)):
            utils.require_device_primitives('MultiTree', allow_host)
        self.nodes = {}
        self.distance_fn = distance_fn
        self.sample_fn = sample_fn
        self.extend_fn = extend_fn
        self.collision_fn = collision_fn
        self.add(TreeNode(_tup(start)))

    def add(self, *nodes):
        for n in nodes:
            self.nodes[n.config] = n

    def __getitem__(self, q):
        return self.nodes[_tup(q)]

    def __contains__(self, q):
        return _tup(q) in self.nodes

    def __len__(self):
        return len(self.nodes)

    def __iter__(self):
        for n in self.nodes.values():
            yield n

    def __call__(self, q1, q2=None, iterations=50):
        if q1 in self:
            path1 = self[q1].retrace()
        else:
            path1 = self.grow(q1, iterations=iterations)
        if q2 is None:
            return configs(path1)
        if q2 in self:
            path2 = self[q2].retrace()
        else:
            path2 = self.grow(q2, iterations=iterations)
        if path1 is None or path2 is None:
            return None
        for i in range(min(len(path1), len(path2))):
            if path1[i] != path2[i]:
                break
        else:
            i += 1
        return configs(path1[:i - 1:-1] + path2[i - 1:])

    def clear(self):
        pass

    # -- one extension from `last` towards `s`: the nodes of the collision-free prefix, and whether all of it was free
    def _extend(self, last, s):
        q1 = np.asarray(last.config, dtype=np.float64).reshape(1, -1)
        q2 = np.asarray(s, dtype=np.float64).reshape(1, -1)
        if not isinstance(self.extend_fn, GpuExtendFn) or not isinstance(self.collision_fn, GpuCollisionFn):
            new, free = [], True                                  # allow_host: the reference's loop
            for q in self.extend_fn(last.config, s):
                if self.collision_fn(q):
                    free = False
                    break
                last = TreeNode(_tup(q), parent=last)
                new.append(last)
            return last, new, free
        _, way = self.extend_fn.batch(q1, q2)
        hit, _ = self.collision_fn.obstacles.edges_check_pairs(torch.from_numpy(q1).cuda(), torch.from_numpy(q2).cuda(),
                                                               mode=ops.EDGE_FIRST, step=self.extend_fn.step_size,
                                                               want_counts=False)
        first = int(hit.item())
        way = way.cpu().numpy()
        keep = way if first < 0 else way[:first]
        new = []
        for q in keep:
            last = TreeNode(_tup(q), parent=last)
            new.append(last)
        return last, new, first < 0

    def _nearest(self, table, pool, s):
        if table is None:                                         # allow_host
            scores = [self.distance_fn(n.config, s) for n in pool]
            return pool[scores.index(min(scores))]
        target = torch.from_numpy(np.asarray(s, dtype=np.float64).reshape(1, -1)).cuda()
        idx, _ = ops.trees_nearest(table.buf, table.size, target, weights=self.distance_fn.weights,
                                   metric=self.distance_fn.metric)
        return pool[int(idx.item())]

    def _table(self, nodes):
        if isinstance(self.distance_fn, GpuDistanceFn):
            return _Table(len(next(iter(self.nodes))), nodes)
        return None


class MultiRRT(MultiTree):

    def grow(self, goal_sample, iterations=50, goal_probability=.2, store=ts.PATH, max_tree_size=500):
        if not callable(goal_sample):
            fixed = goal_sample
            goal_sample = lambda: fixed
        pool = list(self.nodes.values())
        shuffle(pool)                                              # utils.randomize
        nodes, new_nodes = pool[:max_tree_size], []
        table = self._table(nodes)
        for i in range(iterations):
            goal = random() < goal_probability or i == 0
            s = goal_sample() if goal else self.sample_fn()
            last = self._nearest(table, nodes + new_nodes, s)
            last, grown, free = self._extend(last, s)
            new_nodes.extend(grown)
            if table is not None:
                table.extend([n.config for n in grown])
            if free and goal:
                path = last.retrace()
                if store in [ts.ALL, ts.SUCCESS]:
                    self.add(*new_nodes)
                elif store == ts.PATH:
                    new_nodes_set = set(new_nodes)
                    self.add(*[n for n in path if n in new_nodes_set])
                return path
        if store == ts.ALL:
            self.add(*new_nodes)
        return None


class MultiBiRRT(MultiTree):

    def grow(self, goal, iterations=50, store=ts.PATH, max_tree_size=500):
        if goal in self:
            return self[goal].retrace()
        if self.collision_fn(np.asarray(goal, dtype=np.float64)):
            return None
        pool = list(self.nodes.values())
        shuffle(pool)
        nodes1, new_nodes1 = pool[:max_tree_size], []
        nodes2, new_nodes2 = [], [TreeNode(_tup(goal))]
        table1, table2 = self._table(nodes1), self._table([])
        if table2 is not None:
            table2.extend([new_nodes2[0].config])
        for _ in range(iterations):
            if len(nodes1) + len(new_nodes1) > len(nodes2) + len(new_nodes2):
                nodes1, nodes2 = nodes2, nodes1
                new_nodes1, new_nodes2 = new_nodes2, new_nodes1
                table1, table2 = table2, table1
            s = self.sample_fn()
            last1 = self._nearest(table1, nodes1 + new_nodes1, s)
            last1, grown, _ = self._extend(last1, s)
            new_nodes1.extend(grown)
            if table1 is not None:
                table1.extend([n.config for n in grown])

            last2 = self._nearest(table2, nodes2 + new_nodes2, last1.config)
            last2, grown, free = self._extend(last2, last1.config)
            new_nodes2.extend(grown)
            if table2 is not None:
                table2.extend([n.config for n in grown])
            if free:
                if len(nodes1) == 0:
                    nodes1, nodes2 = nodes2, nodes1
                    new_nodes1, new_nodes2 = new_nodes2, new_nodes1
                    last1, last2 = last2, last1
                path1, path2 = last1.retrace(), last2.retrace()[:-1][::-1]
                for p, n in get_pairs(path2):
                    n.parent = p
                if len(path2) == 0:
                    for n in new_nodes2:
                        if n.parent == last2:
                            n.parent = path1[-1]
                else:
                    path2[0].parent = path1[-1]
                path = path1 + path2
                if store in [ts.ALL, ts.SUCCESS]:
                    self.add(*(new_nodes1 + new_nodes2[:-1]))
                elif store == ts.PATH:
                    new_nodes_set = set(new_nodes1 + new_nodes2[:-1])
                    self.add(*[n for n in path if n in new_nodes_set])
                return path
        if store == ts.ALL:
            self.add(*new_nodes1)
        return None
