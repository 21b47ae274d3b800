"""Device-resident kNN roadmap: the batched form of lazy_prm.Roadmap.add_samples + check_roadmap.

What the reference builds one sample and one edge at a time in Python
(motion_planners/lazy_prm.py:148-160 add_samples -> nearest.py:74-85; lazy_prm.py:211-215
check_roadmap -> :173-178 check_vertex, :189-195 check_edge) is built here in three kernel
passes over tensors that stay in HBM:

    nbr_idx  [n_local, k+1] int64   neighbour matrix (global indices, self included, -1 = none)
    nbr_dist [n_local, k+1] float64
    vertex_hit [n_local] bool       collision_fn(sample)
    edge_hit [n_local, k+1] int32   >= 0: a colliding waypoint index of sample -> neighbour,
                                    -1: free (or no edge: self / missing neighbour)

Samples are sharded across ranks; one all-gather (NCCL over NVLink) builds the global sample
set, every rank answers its share of the queries against it and checks the directed edges that
leave those samples.  Which rows a rank owns (`row_ids`): its own index shard for small sets;
for sets the tensor-core scan serves, a share of the set's k-d leaves (whole leaves, dealt to the
ranks in rotation), so that the 128-query blocks of every rank are as compact in space as a single
GPU's, the scan prunes as well at 8 GPUs as at 1, and every rank gets leaves from all over the
sampled region.  The union of the ranks' rows is identical for every
world size (`gather_global` assembles it in sample order).
"""
import math

import numpy as np
import torch

from . import _lib, ops
from .geometry import STEP_SIZE


def shard_bounds(n, world_size, rank):
    """Contiguous, balanced shard [lo, hi) of n items for `rank`."""
    base, rem = divmod(n, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def dist_info(group=None):
    """(torch.distributed or None, world size, rank) of the running job."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist, dist.get_world_size(group), dist.get_rank(group)
    return None, 1, 0


def gather_samples(local, group=None):
    """All-gather the per-rank sample shards [n_r, D] into the global [N, D] table, rank order.
    -> (table, offset of this rank's first sample).  Equal shards take one all_gather_into_tensor;
    ragged shards are padded to the longest one."""
    dist, ws, rank = dist_info(group)
    if ws == 1:
        return local, 0
    counts = torch.tensor([local.shape[0]], device=local.device, dtype=torch.int64)
    all_counts = [torch.zeros_like(counts) for _ in range(ws)]
    dist.all_gather(all_counts, counts, group=group)
    sizes = [int(c.item()) for c in all_counts]
    D = local.shape[1]
    if len(set(sizes)) == 1:
        out = torch.empty((sum(sizes), D), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
    else:
        m = max(sizes)
        pad = torch.zeros((m, D), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
        parts = [torch.empty_like(pad) for _ in range(ws)]
        dist.all_gather(parts, pad, group=group)
        out = torch.cat([p[:s] for p, s in zip(parts, sizes)], 0)
    return out, sum(sizes[:rank])


def gather_rows(local, group=None):
    """Concatenate per-rank result rows (neighbour matrices, flags) in rank order on every rank."""
    dist, ws, rank = dist_info(group)
    if ws == 1:
        return local
    out, _ = gather_samples(local.reshape(local.shape[0], -1), group)
    return out.reshape((-1,) + tuple(local.shape[1:]))


class DeviceRoadmap(object):
    ROWS_KERNEL_MIN = 50000      # rows from which the per-row edge kernel beats the per-edge one (measured: 0.47 vs 0.54 ms at 50k rows, 0.88 vs 1.26 at 125k, 5.5 vs 9.5 at 1M; tools/edges_ab.py)

    def __init__(self, world, k=10, weights=None, max_distance=math.inf, step=STEP_SIZE, group=None):
        self.world = world
        self.k = int(k)
        self.kk = self.k + 1                      # the reference asks for max_degree + 1 and drops self
        self.weights = weights
        self.max_distance = float(max_distance)
        self.step = float(step)
        self.group = group
        self.obstacles = ops.ObstacleSet(world)
        self.sharding = "auto"                     # "index": every rank keeps the rows of its own sample shard
        self.launches = 0                          # kernels of this library launched by the last build()
        self.profile = False                       # record CUDA events around each phase of build()
        self.events = {}
        self._index = None
        self._src = None

    # -- distributed plumbing (torch.distributed only; no collective inside the library)
    def _dist(self):
        return dist_info(self.group)

    def gather_samples(self, local):
        return gather_samples(local, self.group)

    def gather_samples_equal(self, local, out):
        """Fast path used by the bench: equal shards, preallocated output, one collective."""
        dist, ws, rank = self._dist()
        if ws == 1:
            return local, 0
        dist.all_gather_into_tensor(out, local, group=self.group)
        return out, rank * local.shape[0]

    # -- the build
    def build(self, local_samples, check=True, gathered=None):
        """local_samples: this rank's shard, CUDA float64 [n_local, D].  Returns self."""
        if gathered is None:
            samples, offset = self.gather_samples(local_samples)
        else:
            samples, offset = gathered
        n_local = local_samples.shape[0]
        self.samples, self.offset, self.n_local = samples, offset, n_local
        launches = 0
        mark = self._mark
        mark("start")

        if self._index is None or self._index.samples.data_ptr() != samples.data_ptr() \
                or self._index.n != samples.shape[0]:
            self._index = ops.KnnIndex(samples, self.weights)
        count0 = _lib.lib().mpb_knn_launch_count()  # the library counts the kernels its kNN entry points launch
        self._index._prepared = False
        _, ws_size, rank = self._dist()
        N = samples.shape[0]
        spatial = ws_size > 1 and n_local < N and self.sharding != "index" and self.kk <= 12 \
            and self._index.part_rows(rank, ws_size) >= 0
        if getattr(self, "stats", None) is None:
            self.stats = torch.zeros(2, dtype=torch.int64, device=samples.device)
        if spatial:
            # this rank's rows: a share of the leaves of the set's k-d order (every rank computes the same order), i.e.
            # a share of the very query blocks a single-GPU build forms
            self._index.prepare(N)
            mark("packed")
            self.row_ids, self.nbr_idx, self.nbr_dist = self._index.search_part(rank, ws_size, self.kk, self.max_distance,
                                                                                stats=self.stats)
            n_local = self.n_local = self.row_ids.shape[0]
            queries = samples.index_select(0, self.row_ids)
            self.offset = None
        else:
            self._index.prepare(n_local)
            mark("packed")
            self.row_ids = torch.arange(offset, offset + n_local, device=samples.device, dtype=torch.int64)
            queries = samples[offset:offset + n_local]
            if getattr(self, "nbr_idx", None) is None or self.nbr_idx.shape != (n_local, self.kk):
                self.nbr_idx = torch.empty((n_local, self.kk), dtype=torch.int64, device=samples.device)
                self.nbr_dist = torch.empty((n_local, self.kk), dtype=torch.float64, device=samples.device)
            self._index.search(queries, self.kk, self.max_distance, stats=self.stats, out_idx=self.nbr_idx,
                               out_dist=self.nbr_dist)
        launches += _lib.lib().mpb_knn_launch_count() - count0
        mark("knn")
        if getattr(self, "_after_knn", None) is not None:
            self._after_knn()

        if check:
            self.vertex_hit = self.obstacles.points_check(queries)
            launches += 1
            mark("vertices")
            # self-edges (one waypoint) and missing neighbours (-1, skipped) ride along, so the neighbour
            # matrix is consumed in place with no compaction pass
            if self.offset is None:
                # rows in spatial order: explicit source indices
                self.edge_hit, _ = self.obstacles.edges_check_knn(samples, self.nbr_idx, row_ids=self.row_ids,
                                                                  mode=ops.EDGE_ANY, step=self.step)
            elif n_local >= self.ROWS_KERNEL_MIN or hasattr(self.world, "model"):
                # one thread per row: obstacles culled once per sample (1.7x faster at 1M rows, measured)
                self.edge_hit, _ = self.obstacles.edges_check_knn(samples, self.nbr_idx, row_offset=offset,
                                                                  mode=ops.EDGE_ANY, step=self.step)
            else:
                # one thread per edge: 11x more threads, which is what fills the GPU at 100k rows
                if self._src is None or self._src.numel() != n_local * self.kk or self._src_offset != offset:
                    self._src = (torch.arange(n_local, device=samples.device, dtype=torch.int64) + offset) \
                        .repeat_interleave(self.kk)
                    self._src_offset = offset
                hit, _ = self.obstacles.edges_check(samples, self._src, self.nbr_idx.view(-1), mode=ops.EDGE_ANY,
                                                    step=self.step, want_counts=False)
                self.edge_hit = hit.view(n_local, self.kk)
            launches += 1
            mark("edges")
        self.launches = launches
        return self

    def _mark(self, name):
        if self.profile:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self.events[name] = ev

    def phase_ms(self):
        """Milliseconds per phase of the last profiled build() (call after a synchronize)."""
        order = ["start", "packed", "knn", "vertices", "edges"]
        have = [n for n in order if n in self.events]
        return {b: self.events[a].elapsed_time(self.events[b]) for a, b in zip(have[:-1], have[1:])}

    def build_from_host(self, pinned_samples, gather=None):
        """End-to-end call with HOST buffers: a pinned float64 [N, D] CPU tensor in; neighbour matrix (int32),
        vertex flags and edge flags back on the host.  Host->device and device->host copies are part of
        the call; this is what bench.py times as `e2e`.  `gather` (multi-GPU): called with this rank's device shard,
        returns the all-gathered sample set for build()."""
        dev = pinned_samples.cuda(non_blocking=True)
        side = getattr(self, "_side_stream", None)
        if side is None:
            side = self._side_stream = torch.cuda.Stream()
        sent = {}

        def after_knn():
            # the neighbour matrix goes home on a second stream while the collision checks run.  Sample indices are
            # below 2^31 (the C ABI refuses larger sets): it crosses PCIe as int32
            host = getattr(self, "_host_nbr", None)
            if host is None or host.shape != self.nbr_idx.shape:
                host = self._host_nbr = torch.empty(self.nbr_idx.shape, dtype=torch.int32).pin_memory()
            done = torch.cuda.Event()
            done.record()
            side.wait_event(done)
            with torch.cuda.stream(side):
                sent["nbr32"] = self.nbr_idx.to(torch.int32)
                host.copy_(sent["nbr32"], non_blocking=True)

        self._after_knn = after_knn
        try:
            if gather is not None:
                self.build(dev, gathered=gather(dev))
            else:
                self.build(dev)
        finally:
            self._after_knn = None
        outs = (self.vertex_hit, self.edge_hit)
        host = getattr(self, "_host_out", None)
        if host is None or any(h.shape != o.shape for h, o in zip(host, outs)):
            host = self._host_out = tuple(torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in outs)
        for h, o in zip(host, outs):
            h.copy_(o, non_blocking=True)          # pinned destinations
        torch.cuda.current_stream().synchronize()
        side.synchronize()
        return (self._host_nbr,) + host

    # -- shortest path over the built roadmap (lazy_prm.py:41-98: dijkstra / wastar_search, on the device)
    def shortest_path(self, source, target=None, max_rounds=4096):
        """Shortest-path distances from sample `source` over the usable edges (free verdict, free end points,
        both directions), cost = the kNN distance.  -> (dist [N] float64 CUDA tensor, path) where path is the list
        of sample indices source..target (None when `target` is None or unreachable).  Single-rank roadmaps
        only: gather nbr_idx / nbr_dist / edge_hit / vertex_hit with roadmap.gather_rows first otherwise."""
        import ctypes
        from . import _lib
        N = self.samples.shape[0]
        if self.offset != 0 or self.n_local != N:
            raise NotImplementedError('shortest_path needs the whole roadmap on this rank (gather the shards first)')
        dev = self.samples.device
        dist = torch.full((N,), float('inf'), dtype=torch.float64, device=dev)
        dist[int(source)] = 0.0
        changed = torch.zeros(1, dtype=torch.int32, device=dev)
        vhit = self.vertex_hit.to(torch.uint8).contiguous()
        p = lambda t: ctypes.c_void_p(t.data_ptr())
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        lib = _lib.lib()
        rounds = 0
        while rounds < max_rounds:
            changed.zero_()
            _lib.check(lib.mpb_roadmap_relax(p(self.nbr_idx), p(self.nbr_dist), p(self.edge_hit), p(vhit), self.n_local,
                                             self.kk, 0, p(dist), p(changed), 8, st), "mpb_roadmap_relax")
            rounds += 8
            if int(changed.item()) == 0:
                break
        else:
            raise RuntimeError('shortest_path: distances still changing after %d sweeps (max_rounds)' % max_rounds)
        self.sssp_sweeps = rounds
        if target is None or not bool(torch.isfinite(dist[int(target)]).item()):
            return dist, None
        parent = torch.full((N,), torch.iinfo(torch.int64).max, dtype=torch.int64, device=dev)
        _lib.check(lib.mpb_roadmap_parents(p(self.nbr_idx), p(self.nbr_dist), p(self.edge_hit), p(vhit), self.n_local,
                                           self.kk, 0, p(dist), int(source), p(parent), st), "mpb_roadmap_parents")
        par = parent.cpu().numpy()
        path, v = [int(target)], int(target)
        while v != int(source):
            v = int(par[v])
            if v == np.iinfo(np.int64).max or len(path) > N:
                raise RuntimeError('broken parent chain (zero-length edges between distinct samples?)')
            path.append(v)
        return dist, path[::-1]

    # -- persistence and the reference-shaped view
    def save(self, path):
        """Store the built roadmap (tensors + obstacle set) with torch.save; the reference has no on-disk
        form (lazy_prm.py:267 lists it as a TODO)."""
        w = self.world
        torch.save({'k': self.k, 'weights': self.weights, 'max_distance': self.max_distance, 'step': self.step,
                    'offset': self.offset, 'samples': self.samples.cpu(), 'nbr_idx': self.nbr_idx.cpu(),
                    'nbr_dist': self.nbr_dist.cpu(), 'vertex_hit': self.vertex_hit.cpu(),
                    'edge_hit': self.edge_hit.cpu(),
                    'world': {'dim': w.dim, 'env_lo': w.env_lo, 'env_hi': w.env_hi, 'box_lo': w.box_lo,
                              'box_hi': w.box_hi, 'sph_c': w.sph_c, 'sph_r': w.sph_r, 'buffer': w.buffer}}, path)

    @staticmethod
    def load(path, group=None):
        from .geometry import World
        d = torch.load(path, weights_only=False)
        rm = DeviceRoadmap(World(**d['world']), k=d['k'], weights=d['weights'], max_distance=d['max_distance'],
                           step=d['step'], group=group)
        rm.samples = d['samples'].cuda()
        rm.offset = int(d['offset'])
        rm.nbr_idx, rm.nbr_dist = d['nbr_idx'].cuda(), d['nbr_dist'].cuda()
        rm.vertex_hit, rm.edge_hit = d['vertex_hit'].cuda(), d['edge_hit'].cuda()
        rm.n_local = rm.nbr_idx.shape[0]
        return rm

    def to_reference_roadmap(self, extend_fn, collision_fn, cost_fn=None, samples=None):
        """Materialise the reference-shaped lazy_prm.Roadmap (edge set, adjacency, colliding_* dictionaries
        as check_roadmap would leave them) from the device tensors, without recomputing any neighbour or any
        verdict that is already on the device.  `samples`: the configuration objects to store (default: rows of
        the sample table).  Single-rank roadmaps only."""
        from .lazy_prm import Roadmap
        N = self.samples.shape[0]
        if self.offset != 0 or self.n_local != N:
            raise NotImplementedError('gather the shards first')
        host = self.samples.cpu().numpy()
        objs = list(samples) if samples is not None else [row for row in host]
        weights = self.weights if self.weights is not None else np.ones(host.shape[1])
        rm = Roadmap(extend_fn, weights=weights, cost_fn=cost_fn, max_degree=self.k, max_distance=self.max_distance)
        list(rm.nearest.add_data(objs))
        idx = self.nbr_idx.cpu().numpy()
        for v1 in range(N):                       # the insert order of lazy_prm.py:148-160
            for v2 in idx[v1]:
                if v2 >= 0 and v2 != v1:
                    rm.add_edge(v1, int(v2))
        vhit = self.vertex_hit.cpu().numpy()
        for v in range(N):
            rm.colliding_vertices[v] = bool(vhit[v])
        # check_roadmap evaluates the first-met direction of every undirected edge (set iteration order)
        ehit = self.edge_hit.cpu().numpy()
        known = {}
        for v1 in range(N):
            for j, v2 in enumerate(idx[v1]):
                if v2 >= 0 and v2 != v1:
                    known[v1, int(v2)] = bool(ehit[v1, j] >= 0)
        first, seen = [], set()
        for e in rm.edges:
            if e not in seen:
                first.append(e)
                seen.add(e)
                seen.add(e[::-1])
        missing = [e for e in first if e not in known]
        if missing:
            src = torch.tensor([a for a, _ in missing], dtype=torch.int64, device='cuda')
            dst = torch.tensor([b for _, b in missing], dtype=torch.int64, device='cuda')
            hit, _ = self.obstacles.edges_check(self.samples, src, dst, mode=ops.EDGE_ANY, step=self.step,
                                                want_counts=False)
            for e, h in zip(missing, hit.cpu().numpy()):
                known[e] = bool(h >= 0)
        for a, b in first:
            rm.colliding_edges[a, b] = known[a, b]
            rm.colliding_edges[b, a] = known[a, b]
            rm._note_direction(a, b)
        return rm

    # -- counts for the metric
    def n_edges_checked(self):
        """Directed sample->neighbour edges checked (self and missing slots excluded)."""
        own = self.row_ids.unsqueeze(1)
        return int(((self.nbr_idx >= 0) & (self.nbr_idx != own)).sum().item())

    def gather_global(self):
        """Every rank's rows assembled in sample order on every rank:
        -> (nbr_idx [N, k+1], nbr_dist, vertex_hit [N] bool, edge_hit [N, k+1])."""
        ids = gather_rows(self.row_ids.unsqueeze(1))[:, 0]
        out = []
        for t in (self.nbr_idx, self.nbr_dist, self.vertex_hit.to(torch.uint8).unsqueeze(1), self.edge_hit):
            rows = gather_rows(t)
            full = torch.empty_like(rows)
            full[ids] = rows
            out.append(full)
        return out[0], out[1], out[2][:, 0].bool(), out[3]

    # -- host views
    def undirected_edges(self):
        """Sorted unique (v1 < v2) pairs of this rank's directed edges, as a numpy int64 [E, 2]."""
        idx = self.nbr_idx.cpu().numpy()
        own = self.row_ids.cpu().numpy()[:, None]
        mask = (idx >= 0) & (idx != own)
        a = np.broadcast_to(own, idx.shape)[mask]
        b = idx[mask]
        lo, hi = np.minimum(a, b), np.maximum(a, b)
        return np.unique(np.stack([lo, hi], 1), axis=0)
