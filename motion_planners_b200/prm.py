"""Eager PRM (degree-limited) on the CUDA kernels.

Same entry point and result as the reference's motion_planners/prm.py:260-282
`prm(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, target_degree, connect_distance,
num_samples)`, whose roadmap is grown by `DegreePRM.grow` (prm.py:237-256): vertices are processed in
insertion order; each walks its other vertices by increasing distance_fn (stable sort, so equal
distances keep insertion order) and, until `target_degree` neighbours are counted, connects every
not-yet-connected one whose interior waypoints `list(extend_fn(v1.q, v2.q))[:-1]` are collision free;
already-connected ones count too.

The sequential part (the degree walk) stays on the host; everything it consumes is computed in batch:
for all new vertices at once the `K` nearest candidates ranked on distance_fn's rounded value
(mpb_knn_f64_metric) and the directed edge verdicts v1 -> candidate with the last waypoint skipped
(mpb_edges_check, skip_last).  A vertex that exhausts its K candidates before reaching its degree (a
sample inside an obstacle tries every other vertex) gets the full ranking from
one distance scan (mpb_scan_nearest_radius) and its remaining candidates checked in blocks.
"""
from heapq import heappop, heappush

import numpy as np

from . import utils
import torch

from . import ops
from .primitives import GpuCollisionFn, GpuDistanceFn, GpuExtendFn, require_plain_extend
from .utils import INF, as_config_array, bisect_order


class Vertex(object):
    def __init__(self, q, index):
        self.q = q
        self.index = index
        self.edges = {}          # neighbour Vertex -> Edge (prm.py:19)

    def __repr__(self):
        return 'Vertex(' + str(self.q) + ')'


class Edge(object):
    """Undirected edge; the interior waypoints are produced on demand (prm.py:36-58 keeps them eagerly)."""

    def __init__(self, v1, v2, roadmap):
        self.v1, self.v2 = v1, v2
        v1.edges[v2], v2.edges[v1] = self, self
        self._roadmap = roadmap
        self._path = None

    def interior(self):
        if self._path is None:
            self._path = list(self._roadmap.extend_fn(self.v1.q, self.v2.q))[:-1]
        return self._path

    def end(self, start):
        if self.v1 is start:
            return self.v2
        assert self.v2 is start
        return self.v1

    def path(self, start):
        """Waypoints from `start` to the other end, end included (prm.py:51-58)."""
        if self.v1 is start:
            return self.interior() + [self.v2.q]
        assert self.v2 is start
        return self.interior()[::-1] + [self.v1.q]


class DegreePRM(object):
    CANDIDATES = ops.MAX_K        # candidates fetched per vertex in the first batch

    def __init__(self, distance_fn, extend_fn, collision_fn, samples=[], target_degree=4, connect_distance=INF):
        self.distance_fn = distance_fn
        self.extend_fn = extend_fn
        self.collision_fn = collision_fn
        self.target_degree = target_degree
        self.connect_distance = connect_distance
        self.vertices = {}            # q -> Vertex, insertion ordered (prm.py:86)
        self.order = []               # Vertex list in insertion order
        self.edges = []
        self.edge_checks = 0          # directed edge verdicts consumed by the walk (for the metric)
        self.grow(samples)

    def __contains__(self, q):
        return q in self.vertices

    def __getitem__(self, q):
        return self.vertices[q]

    def __len__(self):
        return len(self.vertices)

    def add(self, samples):
        new = []
        for q in samples:
            if q not in self.vertices:
                v = Vertex(q, len(self.order))
                self.vertices[q] = v
                self.order.append(v)
                new.append(v)
        return new

    def connect(self, v1, v2):
        if v1 not in v2.edges:
            edge = Edge(v1, v2, self)
            self.edges.append(edge)
            return edge
        return None

    # -- batched ingredients
    def _accelerated(self):
        return isinstance(self.distance_fn, GpuDistanceFn) and isinstance(self.extend_fn, GpuExtendFn) and \
            isinstance(self.collision_fn, GpuCollisionFn)

    def _edge_free(self, dev, src, dst):
        s = torch.tensor(src, dtype=torch.int64, device='cuda')
        d = torch.tensor(dst, dtype=torch.int64, device='cuda')
        hit, _ = self.collision_fn.obstacles.edges_check(dev, s, d, mode=ops.EDGE_ANY, skip_last=True,
                                                         step=self.extend_fn.step_size, want_counts=False)
        return (hit < 0).cpu().numpy()

    def grow(self, samples):
        new = self.add(samples)
        if self.target_degree == 0 or not new:
            return new
        if not self._accelerated():
            utils.require_device_primitives('prm', getattr(self, 'allow_host', False))
            return self._grow_reference_order(new)
        fn = self.distance_fn
        table = as_config_array([v.q for v in self.order])
        dev = torch.from_numpy(table).cuda()
        n = len(self.order)
        K = min(self.CANDIDATES, n)
        new_idx = np.array([v.index for v in new], dtype=np.int64)
        idx, dist = ops.knn(dev, dev[torch.from_numpy(new_idx).cuda()], K, weights=fn.weights, metric=fn.metric,
                            max_dist=float(self.connect_distance))
        idx = idx.cpu().numpy()
        # verdicts for every (new vertex -> candidate) pair, self and empty slots skipped by the kernel
        src = np.repeat(new_idx, K)
        dst = idx.reshape(-1).copy()
        dst[dst == src] = -1
        free = self._edge_free(dev, src.tolist(), dst.tolist()).reshape(len(new), K)

        for row, v1 in enumerate(new):
            degree = 0
            satisfied = complete = False
            for col in range(K):
                j = idx[row, col]
                if j < 0:
                    complete = True                        # every vertex within connect_distance was listed
                    break
                if j == v1.index:
                    continue
                if self.target_degree <= degree:
                    satisfied = True
                    break
                v2 = self.order[j]
                if v2 not in v1.edges:
                    self.edge_checks += 1
                    if free[row, col]:
                        self.connect(v1, v2)
                        degree += 1
                else:
                    degree += 1
            if not satisfied and not complete and K < n and degree < self.target_degree:
                self._finish_vertex(v1, degree, dev, set(int(j) for j in idx[row]))
        return new

    def _finish_vertex(self, v1, degree, dev, already):
        """Full ranking for one vertex whose first K candidates did not satisfy its degree."""
        fn = self.distance_fn
        q = torch.from_numpy(np.asarray(v1.q, dtype=np.float64)).cuda()
        _, dist, _, _ = ops.scan_nearest_radius(dev, q, 0.0, weights=fn.weights, metric=fn.metric)
        dist = dist.cpu().numpy()
        order = np.argsort(dist, kind='stable')            # stable: equal distances keep insertion order
        rest = [int(j) for j in order if int(j) not in already and int(j) != v1.index
                and dist[j] <= self.connect_distance]
        block = 256
        for b0 in range(0, len(rest), block):
            if self.target_degree <= degree:
                return
            chunk = rest[b0:b0 + block]
            free = self._edge_free(dev, [v1.index] * len(chunk), chunk)
            for j, ok in zip(chunk, free):
                if self.target_degree <= degree:
                    return
                v2 = self.order[j]
                if v2 not in v1.edges:
                    self.edge_checks += 1
                    if ok:
                        self.connect(v1, v2)
                        degree += 1
                else:
                    degree += 1

    def _grow_reference_order(self, new):
        """Arbitrary user callables: nothing to batch, follow prm.py:242-255 call by call."""
        for v1 in new:
            degree = 0
            scored = [(self.distance_fn(v1.q, v.q), v) for v in self.order]
            scored = [p for p in scored if p[1] is not v1 and p[0] <= self.connect_distance]
            for _, v2 in sorted(scored, key=lambda p: p[0]):
                if self.target_degree <= degree:
                    break
                if v2 not in v1.edges:
                    path = list(self.extend_fn(v1.q, v2.q))[:-1]
                    self.edge_checks += 1
                    if not any(self.collision_fn(path[i]) for i in bisect_order(len(path))):
                        self.connect(v1, v2)
                        degree += 1
                else:
                    degree += 1
        return new

    # -- query: A* over the roadmap with distance_fn costs and heuristic (prm.py:179-206)
    def __call__(self, q1, q2):
        self.grow(samples=[q1, q2])
        if q1 not in self or q2 not in self:
            return None
        start, goal = self[q1], self[q2]
        heuristic = lambda v: self.distance_fn(v.q, goal.q)
        counter = 0                                       # heap tie-break: Vertex objects are not orderable
        queue = [(heuristic(start), counter, start)]
        nodes, processed = {start: (0, None)}, set()
        while queue:
            _, _, cv = heappop(queue)
            if cv in processed:
                continue
            processed.add(cv)
            if cv is goal:
                return self._retrace(nodes, cv)
            for nv in cv.edges:
                cost = nodes[cv][0] + self.distance_fn(cv.q, nv.q)
                if nv not in nodes or cost < nodes[nv][0]:
                    nodes[nv] = (cost, cv)
                    counter += 1
                    heappush(queue, (cost + heuristic(nv), counter, nv))
        return None

    def _retrace(self, nodes, v):
        chain = []
        while v is not None:
            chain.append(v)
            v = nodes[v][1]
        chain.reverse()
        # waypoints of all edges on the chain in one launch
        if isinstance(self.extend_fn, GpuExtendFn):
            todo = [(a, b) for a, b in zip(chain[:-1], chain[1:]) if b.edges[a]._path is None]
            if todo:
                q1 = as_config_array([b.edges[a].v1.q for a, b in todo])
                q2 = as_config_array([b.edges[a].v2.q for a, b in todo])
                off, way = self.extend_fn.batch(q1, q2)
                off, way = off.cpu().numpy(), way.cpu().numpy()
                for e, (a, b) in enumerate(todo):
                    b.edges[a]._path = [way[i] for i in range(off[e], off[e + 1] - 1)]
        path = [chain[0].q]
        for a, b in zip(chain[:-1], chain[1:]):
            path += b.edges[a].path(a)
        return path


def prm(start, goal, distance_fn, sample_fn, extend_fn, collision_fn, target_degree=4, connect_distance=INF,
        num_samples=100):
    """:return: path [q', ..., q"] or None (prm.py:260-282)."""
    require_plain_extend(extend_fn, 'prm')
    start = tuple(start)
    goal = tuple(goal)
    samples = [start, goal] + [tuple(sample_fn()) for _ in range(num_samples)]
    if target_degree is None:
        raise NotImplementedError('DistancePRM (target_degree=None) crashes in the reference on Python 3 '
                                  '(prm.py:218: list + dict_keys); not reproduced')
    roadmap = DegreePRM(distance_fn, extend_fn, collision_fn, samples=samples, target_degree=target_degree,
                        connect_distance=connect_distance)
    return roadmap(start, goal)
