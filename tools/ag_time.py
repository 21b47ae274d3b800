"""Developer tool: what the all-gather of the sample shards costs (torchrun, N ranks)."""
import os, sys, time
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
n, D = 1000000, 7
local = torch.rand((n // world, D), dtype=torch.float64, device="cuda")
out = torch.empty((n, D), dtype=torch.float64, device="cuda")
for _ in range(5):
    dist.all_gather_into_tensor(out, local)
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    dist.all_gather_into_tensor(out, local)
e1.record(); e1.synchronize()
back = e0.elapsed_time(e1) / 50
tot = 0.0
for _ in range(20):
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); dist.all_gather_into_tensor(out, local); b.record(); b.synchronize()
    tot += a.elapsed_time(b)
if rank == 0:
    print("all-gather of %d MB over %d ranks: %.3f ms back to back, %.3f ms alone after a barrier" % (n * D * 8 // 1000000, world, back, tot / 20))
dist.destroy_process_group()
