"""world_size-2 gloo tests (CPU) of the multi-rank host logic: sharding, the sample all-gather (equal and
ragged shards), and that the union of the per-rank results equals the single-rank result.  The compute
stand-in on CPU is the oracle (allowed in tests); the CUDA path itself is covered by the -m gpu tests."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.multiprocessing as mp   # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world_size, port, n, dim, k, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        import oracle
        from motion_planners_b200 import workloads
        from motion_planners_b200.roadmap import gather_rows, gather_samples, shard_bounds
        X = workloads.uniform_samples(n, dim, seed=3)
        lo, hi = shard_bounds(n, world_size, rank)
        local = torch.from_numpy(X[lo:hi].copy())
        table, offset = gather_samples(local)
        assert offset == lo
        assert table.shape == (n, dim) and (table.numpy() == X).all()
        # this rank answers the queries of its own shard against the gathered table (global indices)
        idx, dist_ = oracle.knn(table.numpy(), table.numpy()[offset:offset + (hi - lo)], k)
        world = workloads.boxes_world(dim, 16, 0.2, 0.5, seed=1)
        ow = oracle.World(world.dim, world.env_lo, world.env_hi, world.box_lo, world.box_hi, world.sph_c,
                          world.sph_r, world.buffer)
        src = np.repeat(np.arange(lo, hi), k)
        _, first, _ = ow.edges_check_indexed(table.numpy(), src, idx.reshape(-1))
        all_idx = gather_rows(torch.from_numpy(idx))
        all_first = gather_rows(torch.from_numpy(first.reshape(hi - lo, k)))
        if rank == 0:
            np.savez(os.path.join(out_dir, "gathered.npz"), idx=all_idx.numpy(), first=all_first.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [400, 401])          # equal and ragged shards
def test_two_rank_roadmap_equals_single_rank(tmp_path, n):
    import oracle
    from motion_planners_b200 import workloads
    dim, k = 4, 6
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n, dim, k, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "gathered.npz")
    X = workloads.uniform_samples(n, dim, seed=3)
    idx, _ = oracle.knn(X, X, k)
    world = workloads.boxes_world(dim, 16, 0.2, 0.5, seed=1)
    ow = oracle.World(world.dim, world.env_lo, world.env_hi, world.box_lo, world.box_hi, world.sph_c, world.sph_r,
                      world.buffer)
    _, first, _ = ow.edges_check_indexed(X, np.repeat(np.arange(n), k), idx.reshape(-1))
    assert (got["idx"] == idx).all()
    assert (got["first"] == first.reshape(n, k)).all()


def test_shard_bounds_cover_everything():
    from motion_planners_b200.roadmap import shard_bounds
    for n in (0, 1, 7, 100, 1000003):
        for ws in (1, 2, 3, 8):
            spans = [shard_bounds(n, ws, r) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans[:-1], spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_spatial_slices_partition_the_samples():
    """Rows owned by slices of a spatial order (any permutation of the samples) cover every sample exactly once for
    every world size, and scattering per-rank rows back by their ids rebuilds the global table: the host logic of
    DeviceRoadmap's spatial sharding and gather_global."""
    from motion_planners_b200.roadmap import shard_bounds
    rs = np.random.RandomState(0)
    for n in (1, 5, 1000, 4097):
        order = torch.from_numpy(rs.permutation(n))
        table = torch.from_numpy(rs.uniform(size=(n, 3)))
        for ws in (1, 2, 3, 8):
            ids = [order[slice(*shard_bounds(n, ws, r))] for r in range(ws)]
            allids = torch.cat(ids)
            assert sorted(allids.tolist()) == list(range(n))
            rows = torch.cat([table[i] for i in ids])          # what each rank computes, concatenated in rank order
            full = torch.empty_like(rows)
            full[allids] = rows
            assert torch.equal(full, table)
