"""RRT* with the node scan and the edge checks on the device.

Same entry point and behaviour as the reference's motion_planners/rrt_star.py:86-146 AS EXECUTED ON
PYTHON 3: per iteration the nearest node (first arg-min of distance_fn), the collision-free prefix of the
extension towards the sample, one new node, and the choose-parent pass over the nodes closer than
`radius` (strict <, the new node included).  The reference's second, "rewire the neighbours" loop iterates
an already exhausted `filter` object and therefore never runs (rrt_star.py:138-143); it is not
reproduced.

Device work per iteration: ONE library call (mpb_rrt_star_step: nearest-node scan, the extension's safe
prefix, the new node, the radius scan and the safe prefix of every candidate parent -> new edge; four launches,
one small pinned host record back).  The tree bookkeeping (costs, children, solution flags: OptimalNode,
rrt_star.py:11-73) is sequential and stays on the host; node configurations live in a device array that grows
geometrically.
"""
import time
from random import random

import numpy as np

from . import utils
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


class _StepState(object):
    """Device node table, scratch and the pinned host record of mpb_rrt_star_step; grows geometrically."""

    HEAD = np.dtype([('nearest', '<i4'), ('first', '<i4'), ('cnt', '<i4'), ('safe', '<i4'), ('n_cand', '<i4'),
                     ('pad', '<i4'), ('d_near_new', '<f8'), ('new_cfg', '<f8', (ops.MAX_DIM,))])
    CAND = np.dtype([('idx', '<i4'), ('safe', '<i4'), ('ok', '<i4'), ('pad', '<i4'), ('d', '<f8')])

    def __init__(self, obstacles, dim, cap):
        self.obs, self.D, self.cap = obstacles, dim, 0
        self.table = None
        self._alloc(cap)

    def _alloc(self, cap):
        from . import _lib
        lib = _lib.lib()
        dev = torch.device('cuda')
        table = torch.zeros((cap, self.D), dtype=torch.float64, device=dev)
        if self.table is not None:
            table[:self.cap] = self.table
        self.table, self.cap = table, cap
        self.ws_bytes = int(lib.mpb_rrt_star_workspace_bytes(cap))
        self.rec_bytes = int(lib.mpb_rrt_star_record_bytes(cap))
        self.ws = torch.empty(self.ws_bytes + 256, dtype=torch.uint8, device=dev)
        self.ws_ptr = (self.ws.data_ptr() + 255) & ~255
        self.rec = torch.empty(self.rec_bytes, dtype=torch.uint8).pin_memory()
        self.rec_np = self.rec.numpy()
        self.head = self.rec_np[:self.HEAD.itemsize].view(self.HEAD)[0]
        self.cands = self.rec_np[self.HEAD.itemsize:].view(self.CAND)

    def reserve(self, n):
        if n >= self.cap:
            self._alloc(2 * self.cap)

    def bind(self, radius, weights, metric, step, eps):
        """Fix the per-run arguments once: the per-iteration call then only passes the node count."""
        import ctypes
        from . import _lib
        dp = ctypes.POINTER(ctypes.c_double)
        self._fn = _lib.lib().mpb_rrt_star_step
        self._check = _lib.check
        self._sample = np.zeros(self.D, dtype=np.float64)
        self._weights = np.ascontiguousarray(weights, dtype=np.float64)
        self._fixed = (self._sample.ctypes.data_as(dp), float(radius), self._weights.ctypes.data_as(dp), int(metric),
                       float(step), float(eps))
        self._stream = ops._stream()

    def step(self, n_nodes, sample):
        self._sample[:] = sample
        f = self._fixed
        rc = self._fn(self.obs._h, self.table.data_ptr(), n_nodes, self.cap, f[0], f[1], f[2], f[3], f[4], f[5],
                      self.ws_ptr, self.ws_bytes, self.rec.data_ptr(), self.rec_bytes, self._stream)
        if rc:
            self._check(rc, "mpb_rrt_star_step")
        head = self.head
        n_cand = int(head['n_cand']) if head['safe'] > 0 else 0
        return head, self.cands[:n_cand]


def _rrt_star_device_loop(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, radius, max_time, max_iterations,
                          goal_probability, informed):
    """The whole loop of rrt_star.py:100-137 inside one cooperative kernel (mpb_rrt_star_run): the host only
    pre-draws the two random streams in the reference's order (random() and the box sampler, in bulk), grows the
    node arrays when they fill up, and retraces the path at the end.  The global generators are left exactly where
    the reference's loop would have left them (the unconsumed part of the last bulk draw is given back)."""
    import ctypes
    import random as pyrandom
    from . import _lib
    lib = _lib.lib()
    D = len(start)
    dev = 'cuda'
    hp = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    cap = 1 << 14
    if max_iterations < INF:
        cap = max(cap, int(max_iterations) + 2)

    def alloc(c):
        return (torch.empty((c, D), dtype=torch.float64, device=dev), torch.empty(c, dtype=torch.float64, device=dev),
                torch.empty(c, dtype=torch.int32, device=dev), torch.empty((c, D), dtype=torch.float64, device=dev),
                torch.empty(c, dtype=torch.int32, device=dev))
    table, cost, parent, path_to, path_cnt = alloc(cap)
    nbytes = lib.mpb_rrt_star_run_state_bytes()
    state_buf = torch.zeros(nbytes + 256, dtype=torch.uint8, device=dev)
    state = ctypes.c_void_p(state_buf.data_ptr() + (-state_buf.data_ptr()) % 256)
    w = np.ascontiguousarray(np.asarray(distance_fn.weights, dtype=np.float64))
    s_host, g_host = np.ascontiguousarray(start), np.ascontiguousarray(goal)
    saved_py, saved_np = pyrandom.getstate(), np.random.get_state()
    used_u = used_s = 0
    left_u, left_s = [], np.zeros((0, D))
    chunk, done, reset, launches = 4096, 0, 1, 0
    out6 = (ctypes.c_int64 * 6)()
    goal_cost = ctypes.c_double()
    start_time = time.time()
    n_nodes, goal_idx = 1, -1
    while done < max_iterations and elapsed_time(start_time) < max_time:
        u = left_u + [pyrandom.random() for _ in range(max(0, chunk - len(left_u)))]
        smp = np.concatenate([left_s, sample_fn.bulk(max(0, chunk - len(left_s)))]) if len(left_s) < chunk else left_s
        u_dev = torch.tensor(u, dtype=torch.float64, device=dev)
        s_dev = torch.from_numpy(np.ascontiguousarray(smp)).to(dev)
        limit = int(min(max_iterations, done + 8 * chunk))
        rc = lib.mpb_rrt_star_run(collision_fn.obstacles._h, hp(s_host), hp(g_host), float(radius), hp(w),
                                  int(distance_fn.metric), float(extend_fn.step_size), EPSILON, float(goal_probability),
                                  int(bool(informed)), limit, p(table), p(cost), p(parent), p(path_to), p(path_cnt), cap,
                                  p(u_dev), len(u), p(s_dev), len(smp), state, reset, st)
        _lib.check(rc, 'mpb_rrt_star_run')
        _lib.check(lib.mpb_rrt_star_run_read(state, out6, ctypes.byref(goal_cost), st), 'mpb_rrt_star_run_read')
        launches += 1
        reset = 0
        n_nodes, done, u_pos, s_pos, goal_idx, status = [int(v) for v in out6]
        used_u += u_pos
        used_s += s_pos
        left_u, left_s = u[u_pos:], smp[s_pos:]
        if status == 4:
            raise _lib.MpbError('rrt_star: more candidate parents in one iteration than the device loop holds')
        if status == 3:                                   # node arrays full: double them
            new = alloc(2 * cap)
            for old, big in zip((table, cost, parent, path_to, path_cnt), new):
                big[:cap] = old
            table, cost, parent, path_to, path_cnt = new
            cap *= 2
        elif status == 2:
            chunk = min(2 * chunk, 1 << 17)
    # give back what was drawn ahead: the generators end where the reference's loop would leave them
    pyrandom.setstate(saved_py)
    for _ in range(used_u):
        pyrandom.random()
    np.random.set_state(saved_np)
    if used_s:
        sample_fn.bulk(used_s)
    rrt_star.last_launches = launches
    rrt_star.last_iterations = done
    rrt_star.last_nodes = n_nodes
    if goal_idx < 0:
        return None
    par = parent[:n_nodes].cpu().numpy()
    chain = [goal_idx]
    while par[chain[-1]] >= 0:
        chain.append(int(par[chain[-1]]))
    chain = chain[::-1]
    sel = torch.tensor(chain, dtype=torch.int64, device=dev)
    cfg = table[sel].cpu().numpy()
    to = path_to[sel].cpu().numpy()
    cnt = path_cnt[sel].cpu().numpy()
    out = [cfg[0]]
    if len(chain) > 1:
        offsets, way = extend_fn.batch(cfg[:-1], to[1:])  # extend_fn(parent config, recipe target), all nodes at once
        offsets, way = offsets.cpu().numpy(), way.cpu().numpy()
        for i in range(1, len(chain)):
            out.extend(list(way[offsets[i - 1]:offsets[i - 1] + int(cnt[i])]))
            out.append(cfg[i])
    return out


def rrt_star(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, radius, max_time=INF, max_iterations=INF,
             goal_probability=.2, informed=True, verbose=False, allow_host=False, device_loop=True):
    """:return: path [q', ..., q"] or None (rrt_star.py:86-146)."""
    require_plain_extend(extend_fn, 'rrt_star')
    if not (isinstance(distance_fn, GpuDistanceFn) and isinstance(extend_fn, GpuExtendFn)
            and isinstance(collision_fn, GpuCollisionFn)):
        utils.require_device_primitives('rrt_star', allow_host)
        return _rrt_star_reference_order(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, radius,
                                         max_time, max_iterations, goal_probability, informed)
    start = np.asarray(start, dtype=np.float64)
    goal = np.asarray(goal, dtype=np.float64)
    D = len(start)
    if bool(collision_fn.batch(np.stack([start, goal])).any().item()):
        return None
    if device_loop and not verbose and getattr(sample_fn, 'plain_box', False) and (max_iterations < INF or max_time < INF):
        return _rrt_star_device_loop(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, radius, max_time,
                                     max_iterations, goal_probability, informed)
    obs = collision_fn.obstacles
    step = extend_fn.step_size
    state = _StepState(obs, D, 4096)
    state.table[0] = torch.from_numpy(start)
    state.bind(radius, distance_fn.weights, distance_fn.metric, step, EPSILON)
    nodes = [OptimalNode(start)]
    costs = np.zeros(4096)            # costs[i] = nodes[i].cost
    goal_n = None
    launches = 0

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
        state.reserve(n_nodes + 1)
        head, cands = state.step(n_nodes, s)
        launches += 4
        safe = int(head['safe'])                          # len(safe_path(...)), rrt_star.py:118
        if safe == 0:
            continue
        nearest = nodes[int(head['nearest'])]
        new_config = np.array(head['new_cfg'][:D])
        new = OptimalNode(new_config, parent=nearest, d=float(head['d_near_new']),
                          path=(np.asarray(nearest.config, dtype=np.float64), s, safe - 1), iteration=iteration)
        if do_goal and (distance_fn(new.config, goal) < EPSILON):
            goal_n = new
            goal_n.set_solution(True)
        nodes.append(new)

        # choose-parent pass over the nodes strictly closer than `radius` (rrt_star.py:128-137), index order.
        # A node's cost is final once its own pass is over (the reference's second loop, which would rewire
        # the neighbours, never runs), so `costs` is append-only and the candidates that cannot beat the new
        # node's first cost -- almost all of them -- are dropped in one vector operation.
        if n_nodes >= len(costs):
            costs = np.concatenate([costs, np.empty(len(costs))])
        costs[n_nodes] = new.cost          # the new node is its own candidate (distance 0): never a parent
        if len(cands) > 1:
            idx = cands['idx']
            keep = np.flatnonzero((cands['ok'] != 0) & (costs[idx] + cands['d'] < new.cost))
            if len(keep):
                keep = keep[np.argsort(idx[keep], kind='stable')]
                for j in keep:
                    c = cands[j]
                    n = nodes[int(c['idx'])]
                    d = float(c['d'])
                    if (n.cost + d) < new.cost:
                        new.rewire(n, d, (np.asarray(n.config, dtype=np.float64), new_config, int(c['safe']) - 1),
                                   iteration=iteration)
        costs[n_nodes] = new.cost
    rrt_star.last_launches = launches
    if goal_n is None:
        return None
    return goal_n.retrace(expand)
