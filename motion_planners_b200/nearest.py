"""Nearest-neighbour indices over the CUDA kNN kernels.

Mirrors the reference's motion_planners/nearest.py interface: `add_data(new_data) -> zip(indices,
data)` and `query_neighbors(x, k, **kwargs) -> [(distance, index, data[index]), ...]` ascending.
`GpuNeighbors` stands where `KDNeighbors` (nearest.py:36-85, scipy cKDTree) stood and returns the same
indices and bit-identical distances; `BruteForceNeighbors` (nearest.py:89-109) ranks by an arbitrary
distance_fn and uses the exact kernel when that function is one of the accelerated metrics.
The stored configurations stay Python objects in `.data` (lazy_prm looks start/goal up by identity).
"""
import numpy as np

from . import utils
import torch

from . import ops
from .primitives import GpuDistanceFn
from .utils import INF, as_config_array


class NearestNeighbors(object):
    def add_data(self, new_data):
        raise NotImplementedError()

    def query_neighbors(self, x, k=1, **kwargs):
        raise NotImplementedError()


class GpuNeighbors(NearestNeighbors):
    def __init__(self, data=[], circular={}, embed_fn=None, weights=None, **kwargs):
        if weights is None and embed_fn is not None:
            weights = getattr(embed_fn, 'weights', None)
            if weights is None:
                raise NotImplementedError('only weighted embeddings (get_embed_fn(weights)) are accelerated')
        self.weights = None if weights is None else np.array(weights, dtype=np.float64)
        self.circular = dict(circular) if circular else {}   # {coordinate: (lower, upper)}, nearest.py:7-20
        self.data = []
        self._rows = []
        self._dev = None
        self._index = None
        self.stale = True
        self.add_data(data)

    # the reference rebuilds its kd-tree on every add_data (nearest.py:54-71); here the device copy and
    # the packed tiles are rebuilt lazily at the next query
    def add_data(self, new_data):
        new_data = list(new_data)
        first = len(self.data)
        for x in new_data:
            self.data.append(x)
            self._rows.append(np.asarray(x, dtype=np.float64))
        self.stale |= bool(new_data)
        return zip(range(first, first + len(new_data)), new_data)

    def update(self):
        if not self.stale:
            return
        self.stale = False
        if self._rows:
            self._dev = torch.from_numpy(as_config_array(self._rows)).cuda()
            self._index = ops.KnnIndex(self._dev, self.weights)

    @property
    def device_samples(self):
        self.update()
        return self._dev

    def query_batch(self, queries, k, distance_upper_bound=INF, eps=0., p=2):
        """queries: [M, D] array/tensor or None for 'every stored sample'.  -> (idx [M, k] int64, dist [M, k]) on the
        host; missing neighbours are idx = -1, dist = inf."""
        if eps != 0. or p != 2:
            raise NotImplementedError('only eps=0, p=2 (the reference defaults) are accelerated')
        self.update()
        if self._index is None:
            return np.zeros((0, k), np.int64), np.zeros((0, k))
        if queries is None:
            q = self._dev
        elif isinstance(queries, torch.Tensor):
            q = queries.cuda().to(torch.float64).contiguous()
        else:
            q = torch.from_numpy(as_config_array(list(queries))).cuda()
        kk = min(int(k), ops.MAX_K)
        if int(k) > ops.MAX_K:
            raise NotImplementedError('k > %d' % ops.MAX_K)
        if self.circular:                      # three images per circular coordinate: exact kernel
            idx, dist = ops.knn(self._dev, q, kk, weights=self.weights, max_dist=float(distance_upper_bound),
                                method='f64', circular=self.circular)
        elif q.shape[0] < 64:                  # a handful of queries: one warp each, no tile pass
            idx, dist = ops.knn(self._dev, q, kk, weights=self.weights, max_dist=float(distance_upper_bound),
                                method='f64')
        else:
            idx, dist = self._index.search(q, kk, float(distance_upper_bound))
        return idx.cpu().numpy(), dist.cpu().numpy()

    def query_neighbors(self, x, k=1, eps=0., p=2, distance_upper_bound=INF, **kwargs):
        if not self.data:
            return []
        idx, dist = self.query_batch([x], k, distance_upper_bound, eps, p)
        return [(dist[0, j], int(idx[0, j]), self.data[int(idx[0, j])]) for j in range(idx.shape[1]) if idx[0, j] >= 0]


KDNeighbors = GpuNeighbors   # the name planner code written against the reference expects


class BruteForceNeighbors(NearestNeighbors):
    def __init__(self, distance_fn, data=[], **kwargs):
        self.distance_fn = distance_fn
        self.data = []
        self._rows = []
        self._dev = None
        self.add_data(data)

    def add_data(self, new_data):
        new_data = list(new_data)
        first = len(self.data)
        for x in new_data:
            self.data.append(x)
            self._rows.append(np.asarray(x, dtype=np.float64))
        self._dev = None
        return zip(range(first, first + len(new_data)), new_data)

    def query_batch(self, queries, k, distance_upper_bound=INF, **kwargs):
        fn = self.distance_fn
        if not isinstance(fn, GpuDistanceFn):
            raise NotImplementedError('batched queries need one of the accelerated distance functions')
        if self._dev is None:
            self._dev = torch.from_numpy(as_config_array(self._rows)).cuda()
        q = self._dev if queries is None else torch.from_numpy(as_config_array(list(queries))).cuda()
        idx, dist = ops.knn(self._dev, q, min(int(k), ops.MAX_K), weights=fn.weights, metric=fn.metric,
                            max_dist=float(distance_upper_bound))
        return idx.cpu().numpy(), dist.cpu().numpy()

    def query_neighbors(self, x, k=1, **kwargs):
        if isinstance(self.distance_fn, GpuDistanceFn) and self.data and k <= ops.MAX_K:
            idx, dist = self.query_batch([x], k)
            return [(dist[0, j], int(idx[0, j]), self.data[int(idx[0, j])]) for j in range(idx.shape[1])
                    if idx[0, j] >= 0]
        # arbitrary user metric: nothing to accelerate; only on request is it scanned on the host the way the
        # reference does (nearest.py:102-109)
        utils.require_device_primitives('BruteForceNeighbors.query_neighbors', getattr(self, 'allow_host', False))
        scored = [(self.distance_fn(x, x2), i, x2) for i, x2 in enumerate(self.data)]
        return sorted(scored, key=lambda t: t[0])[:k]
