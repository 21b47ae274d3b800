"""Goal-biased single-tree RRT on the CUDA kernels (reference: motion_planners/rrt.py:49-85).

Per iteration the reference does an arg-min over every tree node in Python, generates the waypoints towards the
sample one at a time and collision-checks each.  Here the tree lives on the device and an iteration is three
launches: first arg-min nearest node (`mpb_trees_nearest`), first colliding waypoint of the extension
(`mpb_edges_check_pairs`, mode FIRST) and the append of the kept prefix (`mpb_trees_append`).  Random numbers
are drawn from the global `random` generator and from `sample_fn` / `goal_sample` in the reference's order, so
a seeded run returns the same path bit for bit.
"""
import random
import time

import numpy as np

from . import utils
import torch

from . import ops
from .primitives import GpuCollisionFn, GpuDistanceFn, GpuExtendFn, require_plain_extend
from .utils import INF, RRT_ITERATIONS, elapsed_time, first_argmin


class TreeNode(object):
    """rrt.py:7-24 (drawing helpers left out)."""

    def __init__(self, config, parent=None):
        self.config = config
        self.parent = parent

    def retrace(self):
        sequence = []
        node = self
        while node is not None:
            sequence.append(node)
            node = node.parent
        return sequence[::-1]

    def __str__(self):
        return 'TreeNode(' + str(self.config) + ')'
    __repr__ = __str__


def configs(nodes):
    if nodes is None:
        return None
    return list(map(lambda n: n.config, nodes))


def _never(q):
    return False


def _rrt_reference_order(start, goal_sample, distance_fn, sample_fn, extend_fn, collision_fn, goal_test,
                         goal_probability, max_iterations, max_time):
    """The planner over arbitrary user callables, statement for statement in the reference's order."""
    start_time = time.time()
    if collision_fn(start):
        return None
    nodes = [TreeNode(start)]
    iteration = 0
    while iteration < max_iterations:
        if elapsed_time(start_time) >= max_time:
            break
        goal = random.random() < goal_probability or iteration == 0
        s = goal_sample() if goal else sample_fn()
        last = nodes[first_argmin([distance_fn(n.config, s) for n in nodes])]
        collided = False
        for q in extend_fn(last.config, s):
            if collision_fn(q):
                collided = True
                break
            last = TreeNode(q, parent=last)
            nodes.append(last)
            if goal_test(last.config):
                return configs(last.retrace())
        if not collided and goal:
            return configs(last.retrace())
        iteration += 1
    return None


class _DeviceTree(object):
    """One tree resident on the device: nodes [1, cap, D], parents [1, cap], size."""

    def __init__(self, root, max_new):
        self.D = root.shape[0]
        self.cap = 1 + 16 * int(max_new)
        dev = torch.device('cuda')
        self.nodes = torch.zeros((1, self.cap, self.D), dtype=torch.float64, device=dev)
        self.parents = torch.full((1, self.cap), -1, dtype=torch.int32, device=dev)
        self.sizes = torch.ones(1, dtype=torch.int32, device=dev)
        self.nodes[0, 0] = torch.from_numpy(root).cuda()
        self.size = 1
        self.overflow = torch.zeros(1, dtype=torch.int32, device=dev)
        self.tid = torch.zeros(1, dtype=torch.int32, device=dev)

    def reserve(self, extra):
        if self.size + extra <= self.cap:
            return
        cap = self.cap
        while self.size + extra > cap:
            cap *= 2
        nodes = torch.zeros((1, cap, self.D), dtype=torch.float64, device=self.nodes.device)
        parents = torch.full((1, cap), -1, dtype=torch.int32, device=self.nodes.device)
        nodes[:, :self.cap] = self.nodes
        parents[:, :self.cap] = self.parents
        self.nodes, self.parents, self.cap = nodes, parents, cap

    def retrace(self, last):
        parents = self.parents[0, :self.size].cpu().numpy()
        nodes = self.nodes[0, :self.size].cpu().numpy()
        chain = []
        v = int(last)
        while v >= 0:
            chain.append(v)
            if len(chain) > self.size:
                raise RuntimeError('parent chain does not terminate')
            v = int(parents[v])
        return [nodes[i] for i in chain[::-1]]


def rrt(start, goal_sample, distance_fn, sample_fn, extend_fn, collision_fn, goal_test=_never,
        goal_probability=.2, max_iterations=RRT_ITERATIONS, max_time=INF, allow_host=False):
    """:return: path [q', ..., q"] or None (rrt.py:49-85)."""
    require_plain_extend(extend_fn, 'rrt')
    if not callable(goal_sample):
        g = goal_sample
        goal_sample = lambda: g
    if not (isinstance(distance_fn, GpuDistanceFn) and isinstance(extend_fn, GpuExtendFn)
            and isinstance(collision_fn, GpuCollisionFn)):
        utils.require_device_primitives('rrt', allow_host)
        return _rrt_reference_order(start, goal_sample, distance_fn, sample_fn, extend_fn, collision_fn, goal_test,
                                    goal_probability, max_iterations, max_time)
    start_time = time.time()
    root = np.asarray(start, dtype=np.float64).reshape(-1)
    if bool(collision_fn.batch(root.reshape(1, -1)).any().item()):
        return None
    world = collision_fn.world
    obs = collision_fn.obstacles
    step = extend_fn.step_size
    tree = None
    zero = torch.zeros(1, dtype=torch.int32, device='cuda')
    fwd = torch.zeros(1, dtype=torch.uint8, device='cuda')
    iteration = 0
    while iteration < max_iterations:
        if elapsed_time(start_time) >= max_time:
            break
        goal = random.random() < goal_probability or iteration == 0
        s = np.asarray(goal_sample() if goal else sample_fn(), dtype=np.float64).reshape(1, -1)
        if tree is None:
            tree = _DeviceTree(root, int(float(np.max(world.env_hi - world.env_lo)) / step) + 2)
        target = torch.from_numpy(s).cuda()
        idx, _ = ops.trees_nearest(tree.nodes, tree.sizes, target, weights=distance_fn.weights,
                                   metric=distance_fn.metric)
        near = idx.clamp(min=0)
        frm = tree.nodes[0, near.long()].contiguous()
        first, n = obs.edges_check_pairs(frm, target, mode=ops.EDGE_FIRST, step=step)
        hi = torch.where(first >= 0, first, n).to(torch.int32).contiguous()
        tree.reserve(int(n.item()))           # an extension yields at most n = int(max|diff| / step) + 1 nodes
        out_last = ops.trees_append(tree.nodes, tree.parents, tree.sizes, tree.tid, frm, target, zero, hi, fwd,
                                    near.to(torch.int32).contiguous(), tree.overflow, step=step)
        old_size = tree.size
        kept = int(hi.item())
        tree.size = old_size + kept
        if int(tree.overflow.item()):
            raise RuntimeError('rrt: tree capacity exceeded')
        if goal_test is not _never and kept:
            new_cfgs = tree.nodes[0, old_size:tree.size].cpu().numpy()
            for j, q in enumerate(new_cfgs):
                if goal_test(q):
                    return tree.retrace(old_size + j)
        if goal and int(first.item()) < 0:
            return tree.retrace(int(out_last.item()) if kept else int(near.item()))
        iteration += 1
    return None
