"""RRT-Connect / BiRRT with device-resident trees, batched over independent queries.

Entry points keep the reference's signatures (motion_planners/rrt_connect.py:31-93):
    rrt_connect(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, max_iterations, max_time)
    birrt(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, **kwargs)
plus the batched form the hot path is built for:
    rrt_connect_batch(starts, goals, distance_fn, extend_fn, collision_fn, swap_draws, targets)
which advances B independent queries in lock step, every query consuming its own pre-drawn random
streams (one uniform for random_swap, one target per iteration: rrt_connect.py:26-29,61).

One iteration of a query (rrt_connect.py:56-63 -> primitives.py:19-29 extend_towards), all on the device:
  nearest node of the tree being extended   mpb_trees_nearest (first arg-min of distance_fn's value)
  waypoints + collision verdict             mpb_edges_check_pairs: FIRST hit for a forward extension,
                                            LAST hit for a backward one (goal tree: the waypoints are
                                            generated from the target towards the tree and consumed
                                            reversed, primitives.py:13-16, so the safe prefix of the
                                            reversed list is the suffix after the last colliding one)
  one TreeNode per safe waypoint            mpb_trees_append
Trees live in [2B, cap, D] arrays (tree 2q = start tree, 2q+1 = goal tree) with parent indices; paths
are retraced on the host once a query has connected.
"""
import random
import time

import numpy as np

from . import utils
import torch

from . import ops
from .primitives import GpuCollisionFn, GpuDistanceFn, GpuExtendFn, require_plain_extend
from .utils import INF, RRT_ITERATIONS, elapsed_time, first_argmin


def _accelerated(distance_fn, extend_fn, collision_fn):
    return isinstance(distance_fn, GpuDistanceFn) and isinstance(extend_fn, GpuExtendFn) and \
        isinstance(collision_fn, GpuCollisionFn)


class TreeBatch(object):
    """Device state of B bidirectional queries."""

    def __init__(self, starts, goals, distance_fn, extend_fn, collision_fn, extensions_per_tree,
                 memory_budget=32 << 30, zero_fill=True):
        self.dist, self.ext, self.col = distance_fn, extend_fn, collision_fn
        world = collision_fn.world
        self.B, self.D = starts.shape
        dev = torch.device('cuda')
        # an extension between two configurations inside the environment yields at most this many nodes
        span = float(np.max(world.env_hi - world.env_lo))
        self.max_new = int(span / extend_fn.step_size) + 2
        cap = 1 + int(extensions_per_tree) * self.max_new
        per_tree = self.D * 8 + 4
        if 2 * self.B * cap * per_tree > memory_budget:
            cap = max(self.max_new + 1, int(memory_budget // (2 * self.B * per_tree)))
        self.cap = cap
        if zero_fill:
            self.nodes = torch.zeros((2 * self.B, cap, self.D), dtype=torch.float64, device=dev)
            self.parents = torch.full((2 * self.B, cap), -1, dtype=torch.int32, device=dev)
        else:   # the one-launch kernel writes the roots itself and never reads past a tree's size
            self.nodes = torch.empty((2 * self.B, cap, self.D), dtype=torch.float64, device=dev)
            self.parents = torch.empty((2 * self.B, cap), dtype=torch.int32, device=dev)
        self.sizes = torch.ones(2 * self.B, dtype=torch.int32, device=dev)
        self.nodes[0::2, 0] = torch.as_tensor(starts, dtype=torch.float64, device=dev)
        self.nodes[1::2, 0] = torch.as_tensor(goals, dtype=torch.float64, device=dev)
        self.active = torch.ones(self.B, dtype=torch.bool, device=dev)
        self.solved = torch.zeros(self.B, dtype=torch.bool, device=dev)
        self.last_a = torch.zeros(self.B, dtype=torch.int64, device=dev)     # last node on the first-extended tree
        self.last_b = torch.zeros(self.B, dtype=torch.int64, device=dev)
        self.swapped = torch.zeros(self.B, dtype=torch.bool, device=dev)
        self.overflow = torch.zeros(1, dtype=torch.int32, device=dev)
        self.qid = torch.arange(self.B, device=dev)
        self.launches = 0
        self.edge_checks = 0

    def _extend(self, tree_ids, targets, backward, active):
        """extend_towards for one tree per query.  -> (index of the new last node, success flag)."""
        B = self.B
        tgt_all = torch.zeros((2 * B, self.D), dtype=torch.float64, device=targets.device)
        tgt_all[tree_ids] = targets
        act = torch.zeros(2 * B, dtype=torch.uint8, device=targets.device)
        act[tree_ids] = active.to(torch.uint8)
        idx, _ = ops.trees_nearest(self.nodes, self.sizes, tgt_all, weights=self.dist.weights,
                                   metric=self.dist.metric, active=act)
        near = idx[tree_ids].clamp_(min=0)
        near_cfg = self.nodes[tree_ids, near.long()]
        bw = backward.unsqueeze(1)
        frm = torch.where(bw, targets, near_cfg).contiguous()
        to = torch.where(bw, near_cfg, targets).contiguous()
        obs = self.col.obstacles
        first, n = obs.edges_check_pairs(frm, to, mode=ops.EDGE_FIRST, step=self.ext.step_size)
        last, _ = obs.edges_check_pairs(frm, to, mode=ops.EDGE_LAST, step=self.ext.step_size, want_counts=False)
        lo = torch.where(backward, last + 1, torch.zeros_like(first))
        hi = torch.where(backward, n, torch.where(first >= 0, first, n))
        tid = torch.where(active, tree_ids, torch.full_like(tree_ids, -1)).to(torch.int32)
        out_last = ops.trees_append(self.nodes, self.parents, self.sizes, tid, frm, to, lo.to(torch.int32).contiguous(),
                                    hi.to(torch.int32).contiguous(), backward.to(torch.uint8).contiguous(),
                                    near.to(torch.int32).contiguous(), self.overflow, step=self.ext.step_size)
        self.launches += 4
        self.edge_checks += 2 * int(B)
        return torch.where(active, out_last.long(), near.long()), (first < 0)

    def step(self, u, targets):
        """One rrt_connect iteration for every active query.  u: [B] uniforms, targets: [B, D]."""
        s1 = self.sizes[0::2].to(torch.float64)
        s2 = self.sizes[1::2].to(torch.float64)
        swap = u < s1 / (s1 + s2)                                   # random_swap, rrt_connect.py:26-29
        ta = 2 * self.qid + swap.long()
        tb = 2 * self.qid + (~swap).long()
        active = self.active
        last1, _ = self._extend(ta, targets, swap, active)
        cfg1 = self.nodes[ta, last1]
        last2, success = self._extend(tb, cfg1, ~swap, active)
        newly = active & success
        self.last_a = torch.where(newly, last1, self.last_a)
        self.last_b = torch.where(newly, last2, self.last_b)
        self.swapped = torch.where(newly, swap, self.swapped)
        self.solved |= newly
        self.active = active & ~success

    def run_fused(self, u, targets):
        """All iterations of all queries in one launch (mpb_rrt_connect_batch): one thread block per query.
        u: [B, iters], targets: [B, iters, D] CUDA float64."""
        import ctypes
        from . import _lib
        B, iters = u.shape
        dev = u.device
        solved = torch.empty(B, dtype=torch.uint8, device=dev)
        swapped = torch.empty(B, dtype=torch.uint8, device=dev)
        last_a = torch.empty(B, dtype=torch.int32, device=dev)
        last_b = torch.empty(B, dtype=torch.int32, device=dev)
        its = torch.empty(B, dtype=torch.int32, device=dev)
        starts = self.nodes[0::2, 0].contiguous()
        goals = self.nodes[1::2, 0].contiguous()
        w = np.ascontiguousarray(self.dist.weights, dtype=np.float64)
        p = lambda t: ctypes.c_void_p(t.data_ptr())
        rc = _lib.lib().mpb_rrt_connect_batch(
            self.col.obstacles._h, p(starts), p(goals), B, w.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
            int(self.dist.metric), float(self.ext.step_size), p(u), p(targets), int(iters), p(self.nodes),
            p(self.parents), p(self.sizes), int(self.cap), p(solved), p(last_a), p(last_b), p(swapped), p(its),
            p(self.overflow), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.check(rc, "mpb_rrt_connect_batch")
        self.solved = solved.bool()
        self.swapped = swapped.bool()
        self.last_a = last_a.long()
        self.last_b = last_b.long()
        self.iters_used = its
        self.active = ~self.solved
        self.launches += 1

    def check_overflow(self):
        if int(self.overflow.item()):
            raise MemoryError('tree capacity %d exceeded; raise extensions_per_tree / memory_budget' % self.cap)

    def path(self, q):
        """Host retrace of a connected query: configs(path1[:-1] + path2[::-1]) (rrt_connect.py:65-71)."""
        if not bool(self.solved[q].item()):
            return None
        swap = bool(self.swapped[q].item())
        ta, tb = 2 * q + int(swap), 2 * q + int(not swap)

        def retrace(tree, leaf):
            size = int(self.sizes[tree].item())
            nodes = self.nodes[tree, :size].cpu().numpy()
            par = self.parents[tree, :size].cpu().numpy()
            chain = []
            i = int(leaf)
            while i >= 0:
                if len(chain) > size:
                    raise RuntimeError('corrupt parent chain in tree %d' % tree)
                chain.append(nodes[i])
                i = int(par[i])
            return chain[::-1]

        path1 = retrace(ta, self.last_a[q].item())
        path2 = retrace(tb, self.last_b[q].item())
        if swap:
            path1, path2 = path2, path1
        return path1[:-1] + path2[::-1]


def rrt_connect_batch(starts, goals, distance_fn, extend_fn, collision_fn, swap_draws, targets, return_state=False,
                      retrace=True, fused=True):
    """B independent rrt_connect runs.  starts, goals: [B, D]; swap_draws: [B, iters] uniforms in [0, 1);
    targets: [B, iters, D] sampled configurations (query q consumes column i at its iteration i).
    -> list of B paths (lists of numpy configs) or None.  fused=True runs everything in one launch
    (mpb_rrt_connect_batch); fused=False composes the stand-alone kernels iteration by iteration."""
    require_plain_extend(extend_fn, 'rrt_connect_batch')
    if not _accelerated(distance_fn, extend_fn, collision_fn):
        raise TypeError('rrt_connect_batch needs the accelerated primitives of motion_planners_b200.primitives')
    starts = np.ascontiguousarray(np.asarray(starts, dtype=np.float64))
    goals = np.ascontiguousarray(np.asarray(goals, dtype=np.float64))
    B, D = starts.shape
    u = torch.as_tensor(np.asarray(swap_draws, dtype=np.float64)).cuda()
    tg = torch.as_tensor(np.asarray(targets, dtype=np.float64)).cuda()
    iters = u.shape[1]
    if fused:
        state = TreeBatch(starts, goals, distance_fn, extend_fn, collision_fn, extensions_per_tree=iters,
                          zero_fill=False)
        state.run_fused(u.contiguous(), tg.contiguous())
        state.check_overflow()
        state.iterations = int(state.iters_used.max().item()) if B else 0
        paths = [state.path(q) for q in range(B)] if retrace else None
        return (paths, state) if return_state else paths
    # lock-step composition of the stand-alone kernels (kept as the cross-check of the fused launch)
    state = TreeBatch(starts, goals, distance_fn, extend_fn, collision_fn, extensions_per_tree=iters)
    # rrt_connect.py:47: a colliding start or goal fails the query before any iteration
    bad = collision_fn.batch(np.concatenate([starts, goals])).view(2, B).any(0)
    state.active &= ~bad
    for i in range(iters):
        state.step(u[:, i].contiguous(), tg[:, i].contiguous())
        if (i & 15) == 15 and not bool(state.active.any().item()):
            break
    state.check_overflow()
    state.iterations = i + 1 if iters else 0
    paths = [state.path(q) for q in range(B)] if retrace else None
    return (paths, state) if return_state else paths


def _rrt_connect_reference_order(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, max_iterations,
                                 max_time, tree_frequency=1):
    """Arbitrary user callables: nothing to batch; the same steps call by call (rrt_connect.py:31-72)."""
    start_time = time.time()
    if collision_fn(start) or collision_fn(goal):
        return None
    nodes1, nodes2 = [(start, -1)], [(goal, -1)]

    def extend_towards(tree, target, backward):
        last = first_argmin([distance_fn(c, target) for c, _ in tree])
        seq = list(extend_fn(target, tree[last][0]))[::-1] if backward else list(extend_fn(tree[last][0], target))
        safe = []
        for q in seq:
            if collision_fn(q):
                break
            safe.append(q)
        for i, q in enumerate(safe):
            if (i % tree_frequency == 0) or (i == len(safe) - 1):
                tree.append((q, last))
                last = len(tree) - 1
        return last, len(seq) == len(safe)

    def retrace(tree, i):
        out = []
        while i >= 0:
            out.append(tree[i][0])
            i = tree[i][1]
        return out[::-1]

    iteration = 0
    while iteration < max_iterations:
        if elapsed_time(start_time) >= max_time:
            break
        swap = random.random() < float(len(nodes1)) / (len(nodes1) + len(nodes2))
        tree1, tree2 = (nodes2, nodes1) if swap else (nodes1, nodes2)
        target = sample_fn()
        last1, _ = extend_towards(tree1, target, swap)
        last2, success = extend_towards(tree2, tree1[last1][0], not swap)
        if success:
            path1, path2 = retrace(tree1, last1), retrace(tree2, last2)
            if swap:
                path1, path2 = path2, path1
            return path1[:-1] + path2[::-1]
        iteration += 1
    return None


def rrt_connect(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, max_iterations=RRT_ITERATIONS,
                max_time=INF, **kwargs):
    """:return: path [q', ..., q"] or None (rrt_connect.py:31-72).  Random numbers are drawn from the global
    `random` / numpy generators in the reference's order (one random.random(), then one sample_fn(), per
    iteration), so a seeded run consumes the same streams."""
    require_plain_extend(extend_fn, 'rrt_connect')
    tree_frequency = kwargs.get('tree_frequency', 1)
    if not _accelerated(distance_fn, extend_fn, collision_fn) or tree_frequency != 1:
        utils.require_device_primitives('rrt_connect' if tree_frequency == 1 else 'rrt_connect(tree_frequency != 1)',
                                        kwargs.get('allow_host', False))
        return _rrt_connect_reference_order(start, goal, distance_fn, sample_fn, extend_fn, collision_fn,
                                            max_iterations, max_time, tree_frequency)
    start_time = time.time()
    s = np.asarray(start, dtype=np.float64).reshape(1, -1)
    g = np.asarray(goal, dtype=np.float64).reshape(1, -1)
    if bool(collision_fn.batch(np.concatenate([s, g])).any().item()):
        return None
    state = None
    budget = 64                       # extensions the current trees have room for
    iteration = 0
    while iteration < max_iterations:
        if elapsed_time(start_time) >= max_time:
            break
        if state is None or iteration >= budget:
            budget = budget * 2 if state is not None else budget
            state = _regrow(state, s, g, distance_fn, extend_fn, collision_fn, budget)
        u = random.random()
        target = np.asarray(sample_fn(), dtype=np.float64).reshape(1, -1)
        state.step(torch.tensor([u], dtype=torch.float64, device='cuda'), torch.from_numpy(target).cuda())
        if bool(state.solved[0].item()):
            state.check_overflow()
            return state.path(0)
        iteration += 1
    return None


def _regrow(old, s, g, distance_fn, extend_fn, collision_fn, budget):
    """(Re)allocate the single-query trees with room for `budget` extensions each, keeping their contents."""
    new = TreeBatch(s, g, distance_fn, extend_fn, collision_fn, extensions_per_tree=budget)
    if old is not None:
        n = old.cap
        new.nodes[:, :n] = old.nodes
        new.parents[:, :n] = old.parents
        new.sizes.copy_(old.sizes)
    return new


def birrt(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, **kwargs):
    """:return: path or None (rrt_connect.py:76-93): random restarts of rrt_connect with shortcut smoothing."""
    require_plain_extend(extend_fn, 'birrt')
    from .meta import random_restarts
    solutions = random_restarts(rrt_connect, start, goal, distance_fn, sample_fn, extend_fn, collision_fn,
                                max_solutions=1, **kwargs)
    if not solutions:
        return None
    return solutions[0]
