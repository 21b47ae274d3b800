"""Lazy PRM on the CUDA kNN and edge-collision kernels.

Keeps the entry points and observable state of the reference's motion_planners/lazy_prm.py:
`Roadmap(extend_fn, weights=..., max_degree=..., ...)` with `add_samples`, `check_vertex`,
`check_intermediate`, `check_edge`, `check_path`, `check_roadmap`, `get_cost`, `get_path`, `augment`
(lazy_prm.py:102-237) and `lazy_prm(...) -> Solution(path, samples, edges, colliding_vertices,
colliding_edges)` (lazy_prm.py:253-337).

How the batching keeps parity.  Every boolean the reference caches is a pure function of its key
(a vertex, a directed edge, a waypoint of a directed edge); the order in which the reference visits
keys (bisect order, set-iteration order, early returns) only decides WHICH keys end up in the caches.
So each check_* method first evaluates, in one kernel launch, the keys it may need, then walks the keys
in the reference's order, filling the caches from those results and stopping where the reference stops.
`edges` is a Python set filled by the same inserts in the same order as the reference, so its iteration
order -- which picks the direction check_roadmap evaluates first -- is the same.
Graph search (dijkstra / weighted A*) is host-side glue, as in the reference.
"""
import time
from collections import defaultdict, namedtuple
from heapq import heappop, heappush

import numpy as np
import torch

from . import ops
from .nearest import BruteForceNeighbors, GpuNeighbors
from .primitives import GpuCollisionFn, GpuExtendFn, get_distance_fn, get_embed_fn
from .utils import INF, bisect_order, compute_path_cost, elapsed_time, flatten, get_pairs

Solution = namedtuple('PRMSolution', ['path', 'samples', 'edges', 'colliding_vertices', 'colliding_edges'])
SearchNode = namedtuple('Node', ['g', 'parent'])

ORDINAL = 1e3          # lazy_prm.py:19
REVERSIBLE = True      # lazy_prm.py:20
ROADMAPS = []


def sample_until(sample_fn, num_samples, max_time=INF):
    start_time = time.time()
    samples = []
    while len(samples) < num_samples and elapsed_time(start_time) < max_time:
        samples.append(sample_fn())
    return samples


# ------------------------------------------------------------------------------------------ search
def retrace_path(visited, vertex):
    path = []
    while vertex is not None:
        path.append(vertex)
        vertex = visited[vertex].parent
    return path[::-1]


def dijkstra(start_v, neighbors_fn, cost_fn=lambda v1, v2: 1.):
    """Single-source shortest paths; heap entries (g, vertex) and the same relaxation rule as
    lazy_prm.py:41-58 so that equal-cost ties resolve identically."""
    visited = {start_v: SearchNode(0., None)}
    queue = [(0., start_v)]
    while queue:
        g, v = heappop(queue)
        if visited[v].g < g:
            continue
        for nv in neighbors_fn(v):
            ng = g + cost_fn(v, nv)
            if nv not in visited or ng < visited[nv].g:
                visited[nv] = SearchNode(ng, v)
                heappush(queue, (ng, nv))
    return visited


def _priority(w):
    assert 0. <= w <= INF
    if w == 0.:
        return lambda g, h: g
    if w == INF:
        return lambda g, h: (h, g)
    return lambda g, h: g + w * h


def wastar_search(start_v, end_v, neighbors_fn, cost_fn=lambda v1, v2: 1., heuristic_fn=lambda v: 0,
                  max_cost=INF, max_time=INF, w=1.):
    """Weighted A* with heap entries (priority, g, vertex) (lazy_prm.py:70-98)."""
    priority = _priority(w)
    start_time = time.time()
    visited = {start_v: SearchNode(0., None)}
    queue = [(priority(0., heuristic_fn(start_v)), 0., start_v)]
    while queue and elapsed_time(start_time) < max_time:
        _, g, v = heappop(queue)
        if visited[v].g < g:
            continue
        if v == end_v:
            return retrace_path(visited, v)
        for nv in neighbors_fn(v):
            ng = g + cost_fn(v, nv)
            if nv not in visited or ng < visited[nv].g:
                visited[nv] = SearchNode(ng, v)
                nh = heuristic_fn(nv)
                if ng + nh < max_cost:
                    heappush(queue, (priority(ng, nh), ng, nv))
    return None


# ----------------------------------------------------------------------------------------- roadmap
class Roadmap(object):
    def __init__(self, extend_fn, weights=None, distance_fn=None, cost_fn=None, p_norm=2, max_degree=5,
                 max_distance=INF, approximate_eps=0., **kwargs):
        assert (weights is not None) or (distance_fn is not None)
        if p_norm != 2 or approximate_eps != 0.:
            raise NotImplementedError('only p_norm=2, approximate_eps=0 (the reference defaults) are accelerated')
        self.distance_fn = distance_fn
        self.extend_fn = extend_fn
        self.weights = weights
        self.cost_fn = cost_fn
        self.p_norm = p_norm
        self.max_degree = max_degree
        self.max_distance = max_distance
        self.approximate_eps = approximate_eps
        if self.weights is None:
            self.nearest = BruteForceNeighbors(self.distance_fn)
        else:
            self.nearest = GpuNeighbors(embed_fn=get_embed_fn(self.weights), **kwargs)
        self.edges = set()
        self.outgoing_from_vertex = defaultdict(set)
        self.edge_costs = {}
        self.edge_paths = {}
        self.colliding_vertices = {}
        self.colliding_edges = {}
        self.colliding_intermediates = {}
        # direction in which the reference would have generated an edge's waypoints first (its get_path
        # mirrors the list for the opposite direction, and the mirrored list is NOT the opposite extension:
        # it starts at ~v1 and stops one step short of v2), recorded when a verdict is computed in batch
        self._path_dir = {}
        # the neighbour rows every add_samples batch produced: [(first new vertex, idx int64 [n_new, k])] -- every
        # edge of the roadmap is listed in the row of the vertex that asked for it, which is all the device search
        # needs (it relaxes both directions of every listed edge)
        self._knn_blocks = []

    # -- views (lazy_prm.py:127-140)
    @property
    def samples(self):
        return self.nearest.data

    @property
    def vertices(self):
        return list(range(len(self.samples)))

    def index(self, x):
        for i, q in enumerate(self.samples):
            if x is q:
                return i
        return ValueError(x)          # returned, not raised -- as the reference does (lazy_prm.py:138)

    def __iter__(self):
        return iter([self.samples, self.vertices, self.edges])

    # -- construction
    def add_edge(self, v1, v2):
        assert REVERSIBLE
        pair = {(v1, v2), (v2, v1)}
        self.edges.update(pair)
        self.outgoing_from_vertex[v1].add(v2)
        self.outgoing_from_vertex[v2].add(v1)
        return pair

    def add_samples(self, samples):
        """All new samples are stored first, then each one is queried against the full set
        (lazy_prm.py:148-160) -- here in a single batched kNN."""
        samples = list(samples)
        new = list(self.nearest.add_data(samples))
        new_edges = set()
        if not new:
            return new_edges
        k = min(self.max_degree + 1, len(self.samples))
        queries = [x for _, x in new]
        if isinstance(self.nearest, GpuNeighbors):
            idx, _ = self.nearest.query_batch(queries, k, distance_upper_bound=self.max_distance)
            rows = [[int(j) for j in row if j >= 0] for row in idx]
            first = new[0][0]
            if [v for v, _ in new] == list(range(first, first + len(new))):
                self._knn_blocks.append((first, np.ascontiguousarray(idx, dtype=np.int64)))
            else:
                self._knn_blocks = None
        elif k <= ops.MAX_K and getattr(self.nearest.distance_fn, 'metric', None) is not None:
            idx, _ = self.nearest.query_batch(queries, k)      # the brute-force index ignores max_distance
            rows = [[int(j) for j in row if j >= 0] for row in idx]
        else:
            rows = [[i for _, i, _ in self.nearest.query_neighbors(x, k=k)] for x in queries]
        edges, out = self.edges, self.outgoing_from_vertex       # add_edge (lazy_prm.py:141-147), inlined
        for (v1, _), row in zip(new, rows):
            for v2 in row:
                if v1 != v2:
                    pair = {(v1, v2), (v2, v1)}
                    edges.update(pair)
                    out[v1].add(v2)
                    out[v2].add(v1)
                    new_edges.update(pair)
        return new_edges

    # -- the goal-rooted shortest-path pass of lazy_prm (lazy_prm.py:295-296) on the device
    DEVICE_SEARCH_MIN = 2048         # vertices from which cost_to_go runs on the GPU

    def cost_to_go(self, goal_vertex, cost_fn):
        """dijkstra(goal, neighbors_fn, get_cost) as an array: h[v] = cost of the cheapest path v -> goal over the
        edges not known to collide (inf: unreachable), or None when this roadmap cannot be searched on the device
        (then the caller runs the host search).  The values are the host search's bit for bit: both compute the
        fixed point d[v] = min_u fl(d[u] + c(u, v)) from d[goal] = 0 with the same edge costs (mpb_edge_costs
        evaluates cost_fn's own rounding procedure), and fl(a + c) is monotone in a."""
        n = len(self.samples)
        if (n < self.DEVICE_SEARCH_MIN or not self._knn_blocks or not isinstance(self.nearest, GpuNeighbors)
                or getattr(cost_fn, 'metric', None) is None or self.colliding_vertices or self.colliding_edges
                or getattr(self.nearest, 'circular', None)):
            return None
        if sum(idx.shape[0] for _, idx in self._knn_blocks) != n:
            return None
        import ctypes
        from . import _lib
        lib = _lib.lib()
        dev = self.nearest.device_samples
        D = dev.shape[1]
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        p = lambda t: ctypes.c_void_p(t.data_ptr())
        w = np.ascontiguousarray(np.asarray(cost_fn.weights, dtype=np.float64))
        wp = w.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        blocks = []
        for first, idx in self._knn_blocks:
            nbr = torch.from_numpy(idx).cuda()
            cost = torch.empty(nbr.shape, dtype=torch.float64, device='cuda')
            _lib.check(lib.mpb_edge_costs(p(dev), n, D, wp, int(cost_fn.metric), p(nbr), nbr.shape[0], nbr.shape[1], first,
                                          p(cost), st), 'mpb_edge_costs')
            free = torch.full(nbr.shape, -1, dtype=torch.int32, device='cuda')     # nothing is known to collide yet
            blocks.append((first, nbr, cost, free))
        dist = torch.full((n,), float('inf'), dtype=torch.float64, device='cuda')
        dist[int(goal_vertex)] = 0.0
        changed = torch.zeros(1, dtype=torch.int32, device='cuda')
        for _ in range(4096):
            changed.zero_()
            for first, nbr, cost, free in blocks:
                _lib.check(lib.mpb_roadmap_relax(p(nbr), p(cost), p(free), None, nbr.shape[0], nbr.shape[1], first, p(dist),
                                                 p(changed), 8, st), 'mpb_roadmap_relax')
            if int(changed.item()) == 0:
                return dist.cpu().numpy()
        raise RuntimeError('cost_to_go: distances still changing after 32768 sweeps')

    # -- queries on the caches (lazy_prm.py:161-172)
    def is_colliding(self, v1, v2):
        return self.colliding_vertices.get(v1, False) or self.colliding_vertices.get(v2, False) or \
            self.colliding_edges.get((v1, v2), False)

    def is_safe(self, v1, v2):
        return not self.colliding_edges.get((v1, v2), True)

    def neighbors_fn(self, v1):
        for v2 in self.outgoing_from_vertex[v1]:
            if not self.is_colliding(v1, v2):
                yield v2

    # -- batching helpers
    def _batched(self, collision_fn):
        return isinstance(collision_fn, GpuCollisionFn) and isinstance(self.extend_fn, GpuExtendFn)

    def _sample_rows(self, vs):
        return np.ascontiguousarray(np.array([np.asarray(self.samples[v], dtype=np.float64) for v in vs]))

    def _prefetch_vertices(self, vs, collision_fn):
        todo = [v for v in dict.fromkeys(vs) if v not in self.colliding_vertices]
        if not todo:
            return {}
        flags = collision_fn.batch(self._sample_rows(todo)).cpu().numpy()
        return {v: bool(f) for v, f in zip(todo, flags)}

    def _prefetch_paths(self, pairs):
        """Waypoint lists for directed edges not cached yet (one launch); fills edge_paths the way
        get_path does (lazy_prm.py:223-230: the reverse edge shares the list, reversed)."""
        todo = []
        for v1, v2 in pairs:
            base = self._path_dir.get((v1, v2), (v1, v2))
            if (v1, v2) not in self.edge_paths and base not in todo and base[::-1] not in todo:
                todo.append(base)
        if not todo:
            return
        q1 = self._sample_rows([a for a, _ in todo])
        q2 = self._sample_rows([b for _, b in todo])
        offsets, way = self.extend_fn.batch(q1, q2)
        offsets = offsets.cpu().numpy()
        way = way.cpu().numpy()
        for e, (v1, v2) in enumerate(todo):
            path = [way[i] for i in range(offsets[e], offsets[e + 1])]
            self.edge_paths[v1, v2] = path
            if REVERSIBLE:
                self.edge_paths[v2, v1] = path[::-1]

    # -- the reference's check_* methods
    def check_vertex(self, v, collision_fn, _pre=None):
        if v not in self.colliding_vertices:
            if _pre is not None and v in _pre:
                self.colliding_vertices[v] = _pre[v]
            else:
                self.colliding_vertices[v] = collision_fn(self.samples[v])
        return not self.colliding_vertices[v]

    def check_intermediate(self, v1, v2, index, collision_fn, _pre=None):
        key = (v1, v2, index)
        if key not in self.colliding_intermediates:
            if _pre is not None and key in _pre:
                hit = _pre[key]
            else:
                hit = collision_fn(self.get_path(v1, v2)[index])
            self.colliding_intermediates[key] = hit
            if hit:
                self.colliding_edges[v1, v2] = hit
                if REVERSIBLE:
                    self.colliding_edges[v2, v1] = hit
        return not self.colliding_intermediates[key]

    def _note_direction(self, v1, v2):
        if (v1, v2) not in self.edge_paths and (v1, v2) not in self._path_dir:
            self._path_dir[v1, v2] = (v1, v2)
            if REVERSIBLE:
                self._path_dir[v2, v1] = (v1, v2)

    def check_edge(self, v1, v2, collision_fn, _pre=None):
        if (v1, v2) not in self.colliding_edges:
            if _pre is not None and (v1, v2) in _pre:
                hit = _pre[v1, v2]
                self._note_direction(v1, v2)
            elif self._batched(collision_fn):
                self._note_direction(v1, v2)
                rows1, rows2 = self._sample_rows([v1]), self._sample_rows([v2])
                h, _ = collision_fn.obstacles.edges_check_pairs(torch.from_numpy(rows1).cuda(),
                                                                torch.from_numpy(rows2).cuda(), mode=ops.EDGE_ANY,
                                                                step=self.extend_fn.step_size, want_counts=False,
                                                            circular=getattr(self.extend_fn, 'circular', None))
                hit = bool(h[0].item() >= 0)
            else:
                path = self.get_path(v1, v2)
                hit = any(collision_fn(path[i]) for i in bisect_order(len(path)))
            self.colliding_edges[v1, v2] = hit
            if REVERSIBLE:
                self.colliding_edges[v2, v1] = hit
        return not self.colliding_edges[v1, v2]

    def check_path(self, path, collision_fn):
        """Vertices in bisect order, then every waypoint of every edge of the path in bisect order,
        stopping at the first collision (lazy_prm.py:196-210)."""
        path = list(path)
        batched = self._batched(collision_fn)
        v_order = [path[i] for i in bisect_order(len(path))]
        pre_v = self._prefetch_vertices(v_order, collision_fn) if batched else None
        for v in v_order:
            if not self.check_vertex(v, collision_fn, pre_v):
                return False
        pairs = get_pairs(path)
        if batched:
            self._prefetch_paths(pairs)
        inter = []
        for v1, v2 in pairs:
            inter.extend((v1, v2, i) for i in range(len(self.get_path(v1, v2))))
        pre_i = None
        if batched:
            todo = [key for key in inter if key not in self.colliding_intermediates]
            if todo:
                pts = np.ascontiguousarray(np.array([self.edge_paths[a, b][i] for a, b, i in todo]))
                flags = collision_fn.batch(pts).cpu().numpy()
                pre_i = {key: bool(f) for key, f in zip(todo, flags)}
        for j in bisect_order(len(inter)):
            v1, v2, i = inter[j]
            if not self.check_intermediate(v1, v2, i, collision_fn, pre_i):
                return False
        return True

    def check_roadmap(self, collision_fn):
        """Every vertex, then every directed pair in `edges` iteration order; the first-met direction of
        an undirected edge is the one evaluated, the other is mirrored (lazy_prm.py:211-215, 189-195)."""
        batched = self._batched(collision_fn)
        vs = self.vertices
        pre_v = self._prefetch_vertices(vs, collision_fn) if batched else None
        for v in vs:
            self.check_vertex(v, collision_fn, pre_v)
        order = list(self.edges)
        pre_e = None
        if batched:
            seen = set(self.colliding_edges)
            todo = []
            for v1, v2 in order:
                if (v1, v2) not in seen:
                    todo.append((v1, v2))
                    seen.add((v1, v2))
                    seen.add((v2, v1))
            if todo:
                dev = self.nearest.device_samples if isinstance(self.nearest, GpuNeighbors) else \
                    torch.from_numpy(self._sample_rows(vs)).cuda()
                src = torch.tensor([a for a, _ in todo], dtype=torch.int64, device='cuda')
                dst = torch.tensor([b for _, b in todo], dtype=torch.int64, device='cuda')
                hit, _ = collision_fn.obstacles.edges_check(dev, src, dst, mode=ops.EDGE_ANY,
                                                            step=self.extend_fn.step_size, want_counts=False,
                                                            circular=getattr(self.extend_fn, 'circular', None))
                pre_e = {e: bool(h >= 0) for e, h in zip(todo, hit.cpu().numpy())}
        for v1, v2 in order:
            self.check_edge(v1, v2, collision_fn, pre_e)

    # -- costs and paths (lazy_prm.py:216-230)
    def get_cost(self, v1, v2):
        edge = (v1, v2)
        if edge not in self.edge_costs:
            self.edge_costs[edge] = self.cost_fn(self.samples[v1], self.samples[v2])
            if REVERSIBLE:
                self.edge_costs[edge[::-1]] = self.edge_costs[edge]
        return self.edge_costs[edge]

    def get_path(self, v1, v2):
        edge = (v1, v2)
        if edge not in self.edge_paths:
            a, b = self._path_dir.get(edge, edge)
            path = list(self.extend_fn(self.samples[a], self.samples[b]))
            self.edge_paths[a, b] = path
            if REVERSIBLE:
                self.edge_paths[b, a] = path[::-1]
        return self.edge_paths[edge]

    def augment(self, sample_fn, num_samples=100):
        n = len(self.samples)
        if n >= num_samples:
            return self
        self.add_samples(sample_until(sample_fn, num_samples=num_samples - n))
        return self


# ------------------------------------------------------------------------------------------ planner
def get_metrics(conf, weights=None, p_norm=2, distance_fn=None, cost_fn=None):
    """lazy_prm.py:241-251: unit weights when nothing is given; the cost defaults to the weighted norm."""
    if weights is None and distance_fn is None:
        weights = np.ones(len(conf))
    if cost_fn is None:
        cost_fn = get_distance_fn(weights, p_norm=p_norm) if distance_fn is None else distance_fn
    return weights, distance_fn, cost_fn


def lazy_prm(start, goal, sample_fn, extend_fn, collision_fn, distance_fn=None, cost_fn=None, roadmap=None,
             num_samples=100, weights=None, circular={}, p_norm=2, lazy=True, max_cost=INF, max_time=INF, w=1.,
             meta=False, verbose=True, **kwargs):
    """Same contract as the reference's lazy_prm (lazy_prm.py:253-337): returns a Solution whose `path` is
    None when no path exists.  As in the reference, a roadmap built here uses max_degree=5 (lazy_prm.py:272
    does not forward it); pass `roadmap=Roadmap(..., max_degree=k)` for another degree."""
    start_time = time.time()
    weights, distance_fn, cost_fn = get_metrics(start, weights=weights, p_norm=p_norm, distance_fn=distance_fn,
                                                cost_fn=cost_fn)
    if roadmap is None:
        roadmap = Roadmap(extend_fn, weights=weights, distance_fn=distance_fn, cost_fn=cost_fn, circular=circular)
        roadmap.add_samples([start, goal] + sample_until(sample_fn, num_samples))
    roadmap = roadmap.augment(sample_fn, num_samples=num_samples)

    samples, vertices, edges = roadmap
    start_vertex = roadmap.index(start)
    end_vertex = roadmap.index(goal)
    degree = np.average(list(map(len, roadmap.outgoing_from_vertex.values())))

    if not lazy:
        roadmap.check_roadmap(collision_fn)

    weight_fn = roadmap.get_cost
    if meta:
        weight_fn = lambda v1, v2: ORDINAL * int(not roadmap.is_safe(v1, v2)) + roadmap.get_cost(v1, v2)

    to_go = None if meta else roadmap.cost_to_go(end_vertex, cost_fn)      # on the device where it can be
    if to_go is not None:
        heuristic_fn = lambda v: float(to_go[v])
    else:
        visited = dijkstra(end_vertex, roadmap.neighbors_fn, weight_fn)
        heuristic_fn = lambda v: visited[v].g if v in visited else INF
    path = None
    while elapsed_time(start_time) < max_time and path is None:
        lazy_path = wastar_search(start_vertex, end_vertex, neighbors_fn=roadmap.neighbors_fn, cost_fn=weight_fn,
                                  heuristic_fn=heuristic_fn, max_cost=max_cost,
                                  max_time=max_time - elapsed_time(start_time), w=w)
        if lazy_path is None:
            break
        if verbose:
            print('Candidate | Length: {} | Cost: {:.3f} | Vertices: {} | Samples: {} | Degree: {:.3f} | Time: {:.3f}'
                  .format(len(lazy_path), compute_path_cost(lazy_path, cost_fn=weight_fn),
                          len(roadmap.colliding_vertices), len(roadmap.colliding_intermediates), degree,
                          elapsed_time(start_time)))
        if roadmap.check_path(lazy_path, collision_fn):
            path = lazy_path

    if path is None:
        if verbose:
            forward = set(dijkstra(start_vertex, roadmap.neighbors_fn))
            backward = set(dijkstra(end_vertex, roadmap.neighbors_fn))
            print('Failure | Forward: {} | Backward: {} | Vertices: {} | Samples: {} | Degree: {:.3f} | Time: {:.3f}'
                  .format(len(forward), len(backward), len(roadmap.colliding_vertices),
                          len(roadmap.colliding_intermediates), degree, elapsed_time(start_time)))
        return Solution(path, samples, edges, roadmap.colliding_vertices, roadmap.colliding_edges)

    if verbose:
        print('Solution | Length: {} | Cost: {:.3f} | Vertices: {} | Samples: {} | Degree: {:.3f} | Time: {:.3f}'
              .format(len(path), compute_path_cost(path, cost_fn=weight_fn), len(roadmap.colliding_vertices),
                      len(roadmap.colliding_intermediates), degree, elapsed_time(start_time)))
    roadmap._prefetch_paths(get_pairs(path)) if isinstance(extend_fn, GpuExtendFn) else None
    solution = [start] + flatten(roadmap.get_path(v1, v2) for v1, v2 in get_pairs(path))
    return Solution(solution, samples, edges, roadmap.colliding_vertices, roadmap.colliding_edges)


def create_param_sequence(initial_samples=100, step_samples=100, **kwargs):
    """lazy_prm.py:340-344: the endless schedule num_samples = initial, initial + step, ..."""
    num_samples = initial_samples
    while True:
        params = dict(kwargs)
        params['num_samples'] = num_samples
        yield params
        num_samples += step_samples


def lazy_prm_star(start, goal, sample_fn, extend_fn, collision_fn, distance_fn=None, cost_fn=None, max_cost=INF,
                  success_cost=INF, param_sequence=None, resuse=True, weights=None, circular={}, p_norm=2,
                  max_time=INF, verbose=False, **kwargs):
    """:return: the cheapest path found, or None (lazy_prm.py:346-389): lazy_prm on a roadmap that is augmented
    round after round along `param_sequence`, each round bounded by the best cost so far.  (The keyword really
    is spelled `resuse` in the reference.)"""
    start_time = time.time()
    weights, distance_fn, cost_fn = get_metrics(start, weights=weights, p_norm=p_norm, distance_fn=distance_fn,
                                                cost_fn=cost_fn)
    if param_sequence is None:
        param_sequence = create_param_sequence()
    roadmap = None
    best_path = None
    best_cost = max_cost
    for i, params in enumerate(param_sequence):
        remaining_time = max_time - elapsed_time(start_time)
        if remaining_time <= 0.:
            break
        if verbose:
            print('\nIteration: {} | Cost: {:.3f} | Elapsed: {:.3f} | Remaining: {:.3f} | Params: {}'.format(
                i, best_cost, elapsed_time(start_time), remaining_time, params))
        if (roadmap is None) or not resuse:
            roadmap = Roadmap(extend_fn, weights=weights, distance_fn=distance_fn, cost_fn=cost_fn, circular=circular)
            roadmap.add_samples([start, goal] + sample_until(sample_fn, params['num_samples']))
        call = dict(kwargs)
        call.update(params)
        new_path = lazy_prm(start, goal, sample_fn, extend_fn, collision_fn, roadmap=roadmap, cost_fn=cost_fn,
                            weights=weights, circular=circular, p_norm=p_norm, max_time=remaining_time,
                            max_cost=best_cost, verbose=verbose, **call)[0]
        new_cost = compute_path_cost(new_path, cost_fn=cost_fn)
        if verbose:
            print('Path: {} | Cost: {:.3f} | Length: {}'.format(new_path is not None, new_cost,
                                                                INF if new_path is None else len(new_path)))
        if new_cost < best_cost:
            best_path = new_path
            best_cost = new_cost
        if best_cost < success_cost:
            break
    if roadmap is not None:
        ROADMAPS.append(roadmap)
    return best_path
