"""Multi-GPU parity check, run under torchrun on N GPUs of one box (not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/mgpu_check.py

Every rank builds its shard of the roadmap through the CUDA path; the gathered neighbour matrix, vertex
flags and edge flags must equal the single-GPU build of the whole set (rank 0 computes that too)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from motion_planners_b200 import workloads  # noqa: E402
from motion_planners_b200.roadmap import DeviceRoadmap, gather_rows, shard_bounds  # noqa: E402


def main():
    rank = int(os.environ["RANK"])
    ws = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    ok = True
    # small sets (rows = the rank's own index shard), sets the tensor-core scan serves (rows = a share of the leaves of the
    # spatial order), ragged shard sizes, and the arm world of configs[3]
    cases = [(20000, 6, workloads.c2_world()), (20003, 7, workloads.c3_world()), (50001, 6, workloads.c2_world()),
             (200001, 7, workloads.c4lite_world()),
             (170000, 7, workloads.c4_world())]
    if os.environ.get("MPB_MGPU_FULL"):
        cases.append((1000000, 7, workloads.c4lite_world()))
    for n, dim, world in cases:
        X = workloads.uniform_samples(n, dim, seed=5)
        lo, hi = shard_bounds(n, ws, rank)
        rm = DeviceRoadmap(world, k=10).build(torch.from_numpy(X[lo:hi]).cuda())
        idx, _, vhit_b, ehit = rm.gather_global()
        vhit = vhit_b.to(torch.uint8).unsqueeze(1)
        if rank == 0:
            single = DeviceRoadmap(world, k=10)
            single.group = dist.new_group([0]) if False else None
            # single-GPU build of the whole set, bypassing the collective
            full = torch.from_numpy(X).cuda()
            single.build(full, gathered=(full, 0))
            # ANY-mode edge flags name SOME colliding waypoint: the two edge kernels may name different ones
            same = bool(torch.equal(idx, single.nbr_idx) and torch.equal(vhit[:, 0].bool(), single.vertex_hit)
                        and torch.equal(ehit >= 0, single.edge_hit >= 0))
            print("n=%d dim=%d world_size=%d identical=%s edges_colliding=%d" %
                  (n, dim, ws, same, int((single.edge_hit >= 0).sum())))
            ok = ok and same
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
