#!/usr/bin/env python
"""bench.py -- one JSON line for the kNN + edge-collision hot path (BASELINE.json's metric).

    python bench.py --gpus 1 --steps K --warmup W                       (single GPU)
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...                                (CPU arm: the oracle port)

Default workload: `c4` = configs[3] as written, the samples-sharded + all-gathered PRM that BASELINE.json's
"at 1/2/4/8 B200" metric and its target are quoted on: 1M samples, 7-DoF, k=10, a sphere-approximated arm
(16 spheres) among 256 workspace boxes (the collision model is defined in oracle/arm_oracle.c: the reference has
none); it fits one GPU.  `extra_c4lite` is the same roadmap with the reference's own primitives (256 boxes in
configuration space), `extra_c2` is configs[1] (100k x 6-DoF, 64 boxes).

A step = one pass of the hot path over one synthetic roadmap: [all-gather of the sample shards when
N > 1] -> kNN (k+1 = 11) of every sample of this rank's shard against the global set -> collision_fn
on the shard's vertices -> extend_fn + collision_fn on every directed sample->neighbour edge.
`value` = directed edges collision-checked per second over the whole step, all ranks, inputs resident
in HBM; `knn_queries_per_s` is the same step's query rate.  `e2e` times the same pass through the
host-buffer API (pinned host samples in, neighbour matrix and flags back on the host).
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the kernel, from the committed `ncu --set full`
# captures under profiles/ (1 GPU, the workload's full size); None = not captured
NCU = {
    "knn_tc_kernel": {"traffic": 569.560064e6 + 68.432384e6, "tensor_pipe_active_pct": 23.03,
                      "source": "profiles/r2_ncu_full_knn_tc_raw.csv (c4, 1M x 7-DoF, 1 GPU)"},
    "arm_edges_refill_kernel": {
        "traffic": 212.128512e6 + 40.869376e6, "fp64_pipe_active_pct": 40.3,
        "source": "profiles/r2_ncu_full_arm_edges_refill_raw.csv (c4, 11M directed edges, 1 GPU; 7.75 ms, 19.1 of 32 "
                  "lanes per instruction, issue slots 55.7 % busy)"},
    "edges_check_rows_kernel": {"traffic": 219.491840e6 + 40.684288e6, "fp64_pipe_active_pct": 23.99,
                                "source": "profiles/r2_ncu_full_edges_rows_raw.csv (c4lite, 11M directed edges, 1 GPU)"},
}

METRIC = "PRM edges collision-checked/sec (kNN k=10 + vertex + edge checks per roadmap build step)"
UNIT = "edges/s"

WORKLOADS = {
    # name -> (world factory name, dim, n_samples, k, description)
    "c2": ("c2_world", 6, 100000, 10, "configs[1]: lazy-PRM roadmap, 100k samples, 6-DoF unit cube, 64 random boxes, k=10"),
    "c4": ("c4_world", 7, 1000000, 10, "configs[3]: the sharded, all-gathered PRM the 1/2/4/8-GPU metric and the north-star "
           "target are quoted on: 1M samples, 7-DoF arm approximated by 16 spheres, 256 workspace boxes, k=10"),
    "c4lite": ("c4lite_world", 7, 1000000, 10, "configs[3]'s roadmap with the reference's own primitives: 1M samples, 7-DoF, "
               "256 boxes in configuration space, k=10"),
    "c1": ("c1", 2, 2000, 10, "configs[0]: 2-D demo fixture, 2k samples, k=10"),
}


def collision_model(world):
    if hasattr(world, "model"):
        return "%d-joint arm, %d spheres vs %d workspace boxes" % (world.model.n_joints, world.n_sph, world.n_box)
    return "configuration-space obstacles: %d boxes, %d spheres" % (world.n_box, world.n_sph)


def make_workload(name):
    from motion_planners_b200 import workloads
    fac, dim, n, k, desc = WORKLOADS[name]
    world = workloads.problem1_world()[0] if name == "c1" else getattr(workloads, fac)()
    X = workloads.uniform_samples(n, dim, seed=0)
    return world, X, k, desc


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler(object):
    """Samples SM clock and throttle reasons of one GPU every 100 ms while the timed region runs."""

    def __init__(self, index):
        self.index = index
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------- CPU arms
def cpu_sample_rate(world, X, k, budget_s=12.0, first=8000):
    """The oracle port on a bounded sample: the first `first` queries (grown until ~budget_s of CPU work)."""
    from oracle import cpu_baseline
    from scipy.spatial import KDTree
    cpu_baseline.cores()
    n = X.shape[0]
    t0 = time.perf_counter()
    tree = KDTree(X)
    build_s = time.perf_counter() - t0
    hi = min(first if not hasattr(world, "model") else first // 4, n)
    r = cpu_baseline.native_port_step(world, X, 0, hi, k, tree=tree)
    while r["seconds"] < budget_s / 4 and hi < n:
        hi = min(n, hi * 4)
        r = cpu_baseline.native_port_step(world, X, 0, hi, k, tree=tree)
    r["tree_build_s"] = build_s
    r["hi"] = hi
    return r


def run_reference_arm(args, rank):
    """`--impl reference`: the reference's CPU path (oracle port: scipy cKDTree on all cores + the C oracle
    under OpenMP) on the same workload; each step is a bounded sample of it."""
    if rank != 0:
        return 0
    from oracle import cpu_baseline
    from scipy.spatial import KDTree
    cpu_baseline.cores()                                  # all host cores, whatever the launcher exported
    world, X, k, desc = make_workload(args.workload)
    n = X.shape[0]
    tree = KDTree(X)
    # the arm's collision_fn costs the CPU ~40x more per waypoint than a box test: fewer queries per step
    sample = min(n, args.ref_sample or (10000 if hasattr(world, "model") else 100000))
    for _ in range(args.warmup):
        cpu_baseline.native_port_step(world, X, 0, min(sample, 2000), k, tree=tree)
    secs, edges, queries = 0.0, 0, 0
    for s in range(args.steps):
        lo = (s * sample) % max(1, n - sample + 1)
        r = cpu_baseline.native_port_step(world, X, lo, lo + sample, k, tree=tree)
        secs += r["seconds"]
        edges += r["n_edges"]
        queries += r["n_queries"]
    value = edges / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "knn_queries_per_s": queries / secs,
        "config": {"workload": args.workload, "description": desc, "n_samples": n, "dim": X.shape[1], "k": k,
                   "n_box": world.n_box, "n_sph": world.n_sph, "collision_model": collision_model(world)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu_baseline.cores(), "kind": "port",
                         "sample": "%d of %d queries and their %d directed edges per step; kd-tree built once, "
                                   "untimed" % (sample, n, sample * k)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------- batched RRT extra
def measure_rrt_batch(world_size, rank, dist, total_queries=65536, iters=200):
    """configs[2]: a batch of independent 7-DoF RRT-Connect queries split over the ranks with no
    communication; every query consumes its own pre-drawn random streams."""
    import random as pyrandom
    import torch
    from motion_planners_b200 import primitives, workloads
    from motion_planners_b200.roadmap import shard_bounds
    from motion_planners_b200.rrt_connect import rrt_connect_batch
    world = workloads.c3_world()
    D = world.dim
    col = primitives.GpuCollisionFn(world)
    ext = primitives.get_extend_fn()
    dfn = primitives.get_distance_fn(np.ones(D))
    lo, hi = shard_bounds(total_queries, world_size, rank)
    B = hi - lo
    rs = np.random.RandomState(12345)
    pool = rs.uniform(size=(8 * total_queries, D))
    free = pool[~col.batch(pool).cpu().numpy()]
    starts, goals = free[lo:hi], free[total_queries + lo:total_queries + hi]
    u = np.empty((B, iters))
    tg = np.empty((B, iters, D))
    for q in range(B):
        r = pyrandom.Random(1000 + lo + q)
        u[q] = [r.random() for _ in range(iters)]
        tg[q] = np.random.RandomState(1000 + lo + q).uniform(size=(iters, D))
    rrt_connect_batch(starts[:64], goals[:64], dfn, ext, col, u[:64, :8], tg[:64, :8], retrace=False)   # warm-up
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    _, state = rrt_connect_batch(starts, goals, dfn, ext, col, u, tg, return_state=True, retrace=False)
    e1.record()
    e1.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    solved = torch.tensor([int(state.solved.sum().item())], device="cuda", dtype=torch.int64)
    ext_count = torch.tensor([2 * B * state.iterations], device="cuda", dtype=torch.int64)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(solved)
        dist.all_reduce(ext_count)
    sec = float(ms.item()) * 1e-3
    return {"description": "configs[2]: %d independent 7-DoF rrt_connect queries, 64 boxes, %d iterations max, "
                           "split over ranks, no communication" % (total_queries, iters),
            "seconds": sec, "queries_per_s": total_queries / sec, "solved": int(solved.item()),
            "tree_extensions_per_s": int(ext_count.item()) / sec, "iterations_run": int(state.iterations),
            "kernel_launches": int(state.launches)}


def measure_rrt_star(iters=200000):
    """configs[4] as a rate: RRT* in the 3-D world (32 boxes + 8 spheres, radius 0.1); every rank runs its
    own replica (the tree is sequential), so this is a per-GPU figure."""
    import random as pyrandom
    from motion_planners_b200 import primitives, workloads
    from motion_planners_b200.rrt_star import rrt_star
    world = workloads.c5_world()
    env, obstacles = world.shapes()
    col = primitives.GpuCollisionFn(world)
    ext = primitives.get_extend_fn()
    dfn = primitives.get_distance_fn(np.ones(world.dim))
    pyrandom.seed(0)
    np.random.seed(0)
    cfree = primitives.get_sample_fn(env, obstacles=obstacles)
    while True:
        s, g = cfree(), cfree()
        if np.linalg.norm(s - g) > 0.8:
            break
    sample_fn = primitives.get_sample_fn(env, obstacles=[])
    pyrandom.seed(1)
    np.random.seed(1)
    t0 = time.perf_counter()
    path = rrt_star(s, g, dfn, sample_fn, ext, col, radius=0.1, max_iterations=iters)
    sec = time.perf_counter() - t0
    return {"description": "configs[4]: rrt_star, 3-D, 32 boxes + 8 spheres, radius 0.1, %d iterations "
                           "(one tree per GPU: replicas only)" % iters,
            "seconds": sec, "iterations_per_s": iters / sec, "solved": path is not None,
            "nodes": int(getattr(rrt_star, "last_nodes", 0)),
            "kernel_launches": int(getattr(rrt_star, "last_launches", 0))}


def measure_lazy_prm_c2(num=100000, degree=10):
    """configs[1] through the drop-in signature: lazy_prm(start, goal, ..., roadmap=Roadmap(max_degree=10),
    num_samples=100k) on the 6-DoF world, wall clock, with the split between the roadmap's construction (batched
    kNN on the device + the reference-shaped Python edge sets), the goal-rooted shortest-path pass (device) and
    the lazy search / path checks."""
    import random as pyrandom
    import torch
    from motion_planners_b200 import lazy_prm as lp, primitives, workloads
    world = workloads.c2_world()
    env, obstacles = world.shapes()
    col, ext = primitives.GpuCollisionFn(world), primitives.get_extend_fn()
    w = np.ones(world.dim)
    pyrandom.seed(0)
    np.random.seed(0)
    cfree = primitives.get_sample_fn(env, obstacles=obstacles)
    sample_fn = primitives.get_sample_fn(env, obstacles=[])
    start, goal = cfree(), cfree()
    col.batch(np.zeros((1, world.dim)))                      # CUDA context and library up before the clock starts
    torch.cuda.synchronize()
    ts = time.perf_counter()
    drawn = lp.sample_until(sample_fn, num)                  # the reference's per-call sample_fn, on the host
    t0 = time.perf_counter()
    rm = lp.Roadmap(ext, weights=w, cost_fn=primitives.get_distance_fn(w, p_norm=2), max_degree=degree)
    rm.add_samples([start, goal] + drawn)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    h = rm.cost_to_go(rm.index(goal), rm.cost_fn)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    sol = lp.lazy_prm(start, goal, sample_fn, ext, col, roadmap=rm, num_samples=num, weights=w, verbose=False)
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    return {"description": "configs[1]: lazy_prm(roadmap=Roadmap(max_degree=%d), num_samples=%d), 6-DoF, 64 boxes, through the "
                           "drop-in signature" % (degree, num),
            "seconds_total": t3 - ts, "sampling_host_s": t0 - ts, "roadmap_build_s": t1 - t0,
            "roadmap_build_note": "batched kNN on the device (a few ms) + the reference-shaped Python edge set and adjacency "
                                  "dictionaries of lazy_prm.Roadmap (about 0.6 M undirected edges inserted one by one in the "
                                  "reference's order, which check_roadmap's iteration order depends on)",
            "cost_to_go_device_s": t2 - t1,
            "lazy_prm_call_s": t3 - t2, "cost_to_go_on_device": h is not None, "solved": sol.path is not None,
            "path_points": 0 if sol.path is None else len(sol.path), "edges": len(rm.edges) // 2,
            "vertices_checked": len(sol.colliding_vertices), "edges_checked": len(sol.colliding_edges) // 2}


def measure_smoothing_c5(n_paths=10000, iters=100, rrt_iters=100):
    """configs[4]'s second half: linear-shortcut smoothing (smoothing.py:51-144) of 10k RRT-Connect paths in the
    3-D world, max_iterations=100 each, every path on its own seeded generators (the result of path i equals a
    seeded smooth_path call on it).  The device does the waypoint and collision batches, the host the per-path
    draws and splices."""
    import random as pyrandom
    import torch
    from motion_planners_b200 import primitives, workloads
    from motion_planners_b200.rrt_connect import rrt_connect_batch
    from motion_planners_b200.smoothing import smooth_paths_batch
    world = workloads.c5_world()
    D = world.dim
    col, ext = primitives.GpuCollisionFn(world), primitives.get_extend_fn()
    dfn = primitives.get_distance_fn(np.ones(D))
    rs = np.random.RandomState(777)
    pool = rs.uniform(size=(8 * n_paths, D))
    free = pool[~col.batch(pool).cpu().numpy()]
    starts, goals = free[:n_paths], free[n_paths:2 * n_paths]
    u = rs.uniform(size=(n_paths, rrt_iters))
    tg = rs.uniform(size=(n_paths, rrt_iters, D))
    t0 = time.perf_counter()
    paths = rrt_connect_batch(starts, goals, dfn, ext, col, u, tg)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    solved = [p for p in paths if p is not None]
    out = smooth_paths_batch(solved, ext, col, dfn, seeds=list(range(len(solved))), max_iterations=iters)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    before = sum(len(p) for p in solved)
    after = sum(len(p) for p in out)
    return {"description": "configs[4]: %d 3-D rrt_connect paths (one launch), then smooth_path(max_iterations=%d) on each, "
                           "in lock step" % (n_paths, iters),
            "paths": len(solved), "rrt_connect_batch_s": t1 - t0, "smoothing_s": t2 - t1,
            "paths_smoothed_per_s": len(solved) / (t2 - t1), "waypoints_before": before, "waypoints_after": after}


def measure_prm_c1(num=2000, degree=10):
    """configs[0], the reference's own CPU-runnable case: prm() on the 2-D demo fixture (3 boxes + 1 disc), 2000
    collision-free samples, target degree 10, seed 0 (the same draws as tests/golden/prm_problem1_s0.npz).  The
    reference needs about 33.5 s for this build on one core."""
    import random as pyrandom
    from motion_planners_b200 import primitives, workloads
    from motion_planners_b200.prm import prm
    world, _, _ = workloads.problem1_world()
    env, obstacles = world.shapes()
    col = primitives.GpuCollisionFn(world)
    ext = primitives.get_extend_fn()
    dfn = primitives.get_distance_fn(np.ones(world.dim))
    pyrandom.seed(0)
    np.random.seed(0)
    cfree = primitives.get_sample_fn(env, obstacles=obstacles)
    start, goal = tuple(cfree()), tuple(cfree())
    t0 = time.perf_counter()
    path = prm(start, goal, dfn, cfree, ext, col, target_degree=degree, num_samples=num)
    sec = time.perf_counter() - t0
    return {"description": "configs[0]: prm() on the 2-D demo fixture, %d c-free samples, target degree %d, "
                           "sampling included; reference: ~33.5 s on one core (SURVEY.md 8a A3)" % (num, degree),
            "seconds": sec, "solved": path is not None, "path_points": 0 if path is None else len(path)}


# ------------------------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("MPB_BENCH_WORKLOAD", "c4"), choices=sorted(WORKLOADS))
    ap.add_argument("--ref-sample", type=int, default=0, help="queries per step of the CPU arm (0: per workload)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra 1M-sample measurement")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        return run_reference_arm(args, rank)

    import torch
    import torch.distributed as dist
    from motion_planners_b200 import ops
    from motion_planners_b200.roadmap import DeviceRoadmap, shard_bounds

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU path); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world_size == args.gpus or world_size == 1, "launch with torchrun for --gpus > 1"

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def measure(workload, steps, warmup, with_e2e):
        world, X, k, desc = make_workload(workload)
        n, D = X.shape
        lo, hi = shard_bounds(n, world_size, rank)
        equal = (n % world_size == 0)
        local = torch.from_numpy(X[lo:hi]).cuda()
        gathered_buf = torch.empty((n, D), dtype=torch.float64, device="cuda") if world_size > 1 else None
        rm = DeviceRoadmap(world, k=k)
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB of L2

        def step():
            if world_size > 1 and equal:
                rm.build(local, gathered=rm.gather_samples_equal(local, gathered_buf))
            else:
                rm.build(local)

        for _ in range(max(warmup, 3)):
            step()
        barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        rm.profile = True
        total_ms, phase = 0.0, {}
        barrier()
        wall0 = time.perf_counter()
        for _ in range(steps):
            flush.zero_()                                           # evict the inputs from L2 (untimed)
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            if world_size > 1:
                dist.barrier()
            e0.record()
            step()
            e1.record()
            e1.synchronize()
            total_ms += e0.elapsed_time(e1)
            for name, ms in rm.phase_ms().items():
                phase[name] = phase.get(name, 0.0) + ms
        barrier()
        wall_s = time.perf_counter() - wall0
        clocks = sampler.stop()
        rm.profile = False
        t = torch.tensor([total_ms], device="cuda", dtype=torch.float64)
        per_rank = None
        if world_size > 1:
            mine = torch.tensor([total_ms / steps] + [phase.get(a, 0.0) / steps for a in ("packed", "knn", "vertices", "edges")],
                                device="cuda", dtype=torch.float64)
            every = torch.empty((world_size, mine.numel()), device="cuda", dtype=torch.float64)
            dist.all_gather_into_tensor(every, mine)
            per_rank = [[round(v, 3) for v in row] for row in every.cpu().tolist()]
            dist.all_reduce(t, op=dist.ReduceOp.MAX)                # max over ranks
        total_ms = float(t.item())
        edges_local = torch.tensor([rm.n_edges_checked()], device="cuda", dtype=torch.int64)
        fb = rm.stats[:1].clone()
        tiles_visited = rm.stats[1:2].clone()
        if world_size > 1:
            dist.all_reduce(edges_local)
            dist.all_reduce(fb)
            dist.all_reduce(tiles_visited)
        edges_total = int(edges_local.item())
        # waypoints per directed edge (samplers.py:119-120), measured on the first rows of this rank (untimed)
        rows = min(hi - lo, 50000)
        own = rm.row_ids[:rows]
        nb = rm.nbr_idx[:rows]
        real = (nb >= 0) & (nb != own.unsqueeze(1))
        q1 = rm.samples[own].unsqueeze(1).expand(-1, nb.shape[1], -1)[real]
        q2 = rm.samples[nb.clamp(min=0)][real]
        wp = ops.edges_count(q1.contiguous(), q2.contiguous(), step=rm.step).double()
        hit_frac = float((rm.edge_hit[:rows][real] >= 0).double().mean().item())
        free_vertices = 1.0 - float(rm.vertex_hit.double().mean().item())
        ms_step = total_ms / steps
        out = {"world": world, "X": X, "k": k, "desc": desc, "n": n, "D": D, "ms_step": ms_step,
               "edges": edges_total, "value": edges_total / (ms_step * 1e-3), "qps": n / (ms_step * 1e-3),
               "phase_ms": {a: b / steps for a, b in phase.items()}, "clocks": clocks, "launches": rm.launches,
               "fallback_queries": int(fb.item()), "n_local": hi - lo, "wall_s": wall_s,
               "tiles_visited": int(tiles_visited.item()), "mean_waypoints": float(wp.mean().item()),
               "edge_hit_frac": hit_frac, "free_vertex_frac": free_vertices, "per_rank": per_rank}

        if with_e2e:
            pinned = torch.from_numpy(np.ascontiguousarray(X[lo:hi])).pin_memory()
            rm2 = DeviceRoadmap(world, k=k)
            if world_size == 1:
                for _ in range(2):
                    rm2.build_from_host(pinned)
                barrier()
                t0 = time.perf_counter()
                for _ in range(steps):
                    res = rm2.build_from_host(pinned)
                torch.cuda.synchronize()
                e2e_s = (time.perf_counter() - t0) / steps
            else:
                def e2e_step():
                    return rm2.build_from_host(
                        pinned, gather=(lambda d: rm2.gather_samples_equal(d, gathered_buf)) if equal else None)
                for _ in range(2):
                    e2e_step()
                barrier()
                t0 = time.perf_counter()
                for _ in range(steps):
                    res = e2e_step()
                barrier()
                e2e_s = (time.perf_counter() - t0) / steps
                ts = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
                dist.all_reduce(ts, op=dist.ReduceOp.MAX)
                e2e_s = float(ts.item())
            out["e2e"] = {"value": edges_total / e2e_s, "unit": UNIT, "ms_per_step": 1e3 * e2e_s,
                          "h2d_bytes_per_step": int(pinned.numel() * 8) * world_size,
                          "d2h_bytes_per_step": int(sum(r.numel() * r.element_size() for r in res)) * world_size,
                          "knn_queries_per_s": n / e2e_s}
        del flush
        return out

    m = measure(args.workload, args.steps, args.warmup, with_e2e=True)

    # ---- rooflines (every number below can be recomputed from the keys beside it)
    peaks_file = {}
    try:
        peaks_file = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    fp32_peak = max(ops.peak_fma(torch.float32), ops.peak_fma(torch.float32, packed=True))
    fp64_peak = ops.peak_fma(torch.float64)              # FLOP/s, FMA = 2
    hbm_peak = peaks_file.get("hbm_gbs")
    bf16_peak = peaks_file.get("bf16_tflops_sustained") or peaks_file.get("bf16_tflops")
    tf32_peak = (bf16_peak / 2.0) if bf16_peak else 1130.0
    from motion_planners_b200 import _lib as _mplib

    def knn_roofline(r):
        """kNN scan, ONE launch = one rank's queries against the whole set.  SURVEY 8d: algorithmic work = 3*D flop per
        (query, sample) pair, all pairs of the launch."""
        n, D, k, n_local = r["n"], r["D"], r["k"], r["n_local"]
        ms = r["phase_ms"].get("knn", float("nan"))
        pairs = float(n_local) * n
        alg = 3.0 * D * pairs
        kind = _mplib.lib().mpb_knn_filter_kind(int(n), int(D), int(k) + 1)
        out = {"kernel": {0: "knn_filter_kernel", 1: "knn_tc_kernel", 2: "knn_exact_kernel"}[kind],
               "kernel_ms": ms, "pairs": pairs, "algorithmic_flops_per_pair": 3 * D,
               "algorithmic_tflops": alg / (ms * 1e-3) / 1e12}
        if kind == 1:
            # tensor-core scan: TF32 pipe.  Executed work = tiles read x (128 queries x 256 candidates) x 2 x 24
            # (every fp32 value enters as two TF32 parts: 24 operand columns).  The scan skips the tiles whose
            # bounding box is provably too far from a query block, so executed << algorithmic and
            # frac (algorithmic / peak) can exceed 1; frac_executed is what the tensor pipe actually did.
            executed = float(r["tiles_visited"]) / world_size * 128 * 256 * 48      # tiles_visited is summed over ranks
            out.update({
                "bound": "tensor", "achieved": alg / (ms * 1e-3) / 1e12, "peak": tf32_peak, "unit": "TFLOP/s",
                "frac": alg / (ms * 1e-3) / 1e12 / tf32_peak,
                "peak_source": ("MEASURED_PEAKS.json bf16_tflops_sustained / 2 (TF32 runs at half the bf16 rate); that figure "
                                "was measured at a %s MHz median SM clock under power capping, this kernel runs at the clock in "
                                "`clocks`" % peaks_file.get("clocks_under_load", {}).get("sm_mhz_median", "?")) if bf16_peak
                else "fallback: nominal dense TF32, 1130 TFLOP/s",
                "executed_flops_per_pair_visited": 48, "tiles_read": r["tiles_visited"] / world_size,
                "tiles_read_per_query_block": r["tiles_visited"] / max(1.0, n_local * world_size / 128.0),
                "tiles_total": (n + 255) // 256, "pairs_visited_frac": executed / 48.0 / pairs,
                "executed_tflops": executed / (ms * 1e-3) / 1e12,
                "frac_executed": executed / (ms * 1e-3) / 1e12 / tf32_peak,
                "ncu_tensor_pipe_active_pct": NCU["knn_tc_kernel"]["tensor_pipe_active_pct"],
                "traffic": NCU["knn_tc_kernel"]["traffic"] if (world_size == 1 and n == 1000000) else None,
                "ncu_source": NCU["knn_tc_kernel"]["source"],
            })
        else:
            out.update({"bound": "fp32", "achieved": alg / (ms * 1e-3) / 1e12, "peak": fp32_peak / 1e12, "unit": "TFLOP/s",
                        "frac": alg / (ms * 1e-3) / fp32_peak, "executed_flops_per_pair": 2 * D,
                        "frac_executed": 2.0 * D * pairs / (ms * 1e-3) / fp32_peak,
                        "peak_source": "measured live: max of the FFMA and FFMA2 micro-benchmarks (mpb_peak_fma_f32*)",
                        "traffic": None})
        return out

    def edges_roofline(r):
        """extend_fn + collision_fn over the directed edges.  SURVEY 8d: per waypoint 3*D flop for the recurrence
        plus the collision test; per edge 16 B of indices + 2*D*8 B of gathered end points + 4 B of verdict."""
        n, D, k = r["n"], r["D"], r["k"]
        world = r["world"]
        ms = r["phase_ms"].get("edges", float("nan"))
        rows_edges = float(r["n_local"]) * (k + 1)       # edges ONE launch walks (this rank's rows, self-edges included)
        wp = r["mean_waypoints"]
        arm = hasattr(world, "model")
        if arm:
            S, B, J = world.n_sph, world.n_box, world.model.n_joints
            fk = 68 * J + 12 * S                                      # chain (sincos, frame turn, link) + sphere centres
            test_all = 17 * S * B                                     # clamp / subtract / fused square-add per axis + compare
            kernel = "arm_edges_refill_kernel"
            model = ("per waypoint: %d (recurrence) + %d (forward kinematics of %d joints, %d sphere centres) + %d "
                     "(%d spheres x %d boxes x 17, un-culled; the kernel tests only the boxes listed for the sphere's grid "
                     "cell and stops at the first hit)" % (3 * D, fk, J, S, test_all, S, B))
        else:
            fk = 0
            test_all = 2 * D * world.n_box + (2 * D + 2) * world.n_sph
            kernel = "edges_check_rows_kernel" if r["n_local"] >= 50000 else "edges_check_kernel"
            model = ("per waypoint: %d (recurrence) + %d (%d boxes x 2*D compares + %d spheres, un-culled; the kernel culls "
                     "obstacles against the row's bounding box first)" % (3 * D, test_all, world.n_box, world.n_sph))
        flop_wp = 3 * D + fk + test_all
        alg = rows_edges * wp * flop_wp
        bytes_edge = 16 + 2 * D * 8 + 4
        gbs = rows_edges * bytes_edge / (ms * 1e-3) / 1e9
        tfl = alg / (ms * 1e-3) / 1e12
        f64 = fp64_peak / 1e12
        frac_f64 = tfl / f64
        frac_hbm = (gbs / hbm_peak) if hbm_peak else None
        ncu = NCU.get(kernel, {})
        return {"kernel": kernel, "kernel_ms": ms, "edges_walked": rows_edges, "mean_waypoints_per_edge": wp,
                "colliding_edge_frac": r["edge_hit_frac"], "free_vertex_frac": r["free_vertex_frac"],
                "algorithmic_flops_per_waypoint": flop_wp, "flop_model": model,
                "bound": "fp64", "achieved": tfl, "peak": f64, "unit": "TFLOP/s", "frac": frac_f64,
                "peak_source": "measured live: DFMA micro-benchmark (mpb_peak_fma_f64), FMA = 2 flop",
                "note": "algorithmic (un-culled, no early exit) flop over time: above 1 means the culling and the early "
                        "exit skipped that much work, not that the pipe ran faster than its peak",
                "bytes_per_edge": bytes_edge, "hbm": {"achieved_gbs": gbs, "peak_gbs": hbm_peak, "frac": frac_hbm},
                "binding": "fp64 pipe / latency (hbm fraction %.4f)" % (frac_hbm or 0.0),
                "ncu_fp64_pipe_active_pct": ncu.get("fp64_pipe_active_pct"),
                "traffic": ncu.get("traffic") if (world_size == 1 and n == 1000000) else None,
                "ncu_source": ncu.get("source")}

    roofline_knn = knn_roofline(m)
    roofline_edges = edges_roofline(m)
    n, D, k = m["n"], m["D"], m["k"]
    knn_ms = m["phase_ms"].get("knn", float("nan"))
    dominant = roofline_knn if knn_ms >= m["phase_ms"].get("edges", 0.0) else roofline_edges
    roofline = dict(dominant)
    roofline["dominant_of"] = {p_: v for p_, v in m["phase_ms"].items()}

    line = {
        "metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": world_size, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": m["ms_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None,
        "dtype": "f64 (fp32 / split-TF32 filter pass inside the kNN, results bit-exact to the fp64 reference)",
        "data": "synthetic", "knn_queries_per_s": m["qps"],
        "config": {"workload": args.workload, "description": m["desc"], "n_samples": n, "dim": D, "k": k,
                   "n_box": m["world"].n_box, "n_sph": m["world"].n_sph, "collision_model": collision_model(m["world"]),
                   "edges_per_step": m["edges"],
                   "sharding": "samples sharded over ranks, one NCCL all-gather per step, edges stay sharded",
                   "l2": "flushed between timed steps (256 MiB write, untimed); inputs are smaller than L2"},
        "phases_ms": m["phase_ms"],
        "per_rank_ms": {"columns": ["step", "packed", "knn", "vertices", "edges"], "rows": m["per_rank"]} if m["per_rank"] else None,
        "kernel_rates": {"knn_queries_per_s": m["n_local"] * world_size / (knn_ms * 1e-3),
                         "edges_per_s": m["edges"] / (m["phase_ms"].get("edges", float("nan")) * 1e-3)},
        "fallback_queries": m["fallback_queries"],
        "roofline": roofline, "roofline_knn": roofline_knn, "roofline_edges": roofline_edges,
        "clocks": m["clocks"], "e2e": m["e2e"], "gpu_launches": m["launches"] * args.steps,
        "wall_s_timed_region": m["wall_s"],
    }

    if rank == 0 and world_size == 1 and not args.no_cpu_baseline:
        from oracle import cpu_baseline
        r = cpu_sample_rate(m["world"], m["X"], k)
        line["cpu_baseline"] = {
            "value": r["n_edges"] / r["seconds"], "unit": UNIT, "cores": cpu_baseline.cores(), "kind": "port",
            "knn_queries_per_s": r["n_queries"] / r["seconds"],
            "sample": "first %d of %d queries and their %d directed edges (scipy cKDTree batched on all cores + "
                      "C oracle under OpenMP); kd-tree build %.2f s untimed" % (r["hi"], n, r["n_edges"],
                                                                                r["tree_build_s"]),
        }
        # the per-call Python path: the unmodified reference itself where its checkout is reachable (the build
        # container), else the port of its per-call structure -- the line says which one ran
        p = cpu_baseline.reference_step(m["world"], m["X"], 0, 40, k)
        which = "unmodified reference (imported from its checkout)"
        if p is None:
            p = cpu_baseline.python_port_step(m["world"], m["X"], 0, 40, k)
            which = "python port of the reference's per-call structure (reference checkout not reachable on this box)"
        line["cpu_baseline"]["python_per_call"] = {
            "value": p["n_edges"] / p["seconds"], "unit": UNIT, "cores": 1, "ran": which,
            "sample": "first 40 queries, %d edges; kd-tree over all samples built inside the timed region "
                      "(nearest.py:54-61 rebuilds it on every add_data)" % p["n_edges"]}

    if not args.no_extra:
        for other in [w_ for w_ in ("c4lite", "c2") if w_ != args.workload]:
            x = measure(other, max(2, min(args.steps, 10 if other == "c2" else 5)), 3, with_e2e=False)
            line["extra_" + other] = {
                "description": x["desc"], "ms_per_step": x["ms_step"], "edges_per_s": x["value"],
                "knn_queries_per_s": x["qps"], "edges_per_step": x["edges"], "phases_ms": x["phase_ms"],
                "roofline_knn": knn_roofline(x), "roofline_edges": edges_roofline(x),
                "fallback_queries": x["fallback_queries"], "clocks": x["clocks"]}

    if not args.no_extra:
        line["extra_c3_rrt_connect_batch"] = measure_rrt_batch(world_size, rank, dist if world_size > 1 else None)
        if rank == 0:
            line["extra_c5_rrt_star"] = measure_rrt_star()
            line["extra_c1_prm"] = measure_prm_c1()
            line["extra_c2_lazy_prm"] = measure_lazy_prm_c2()
            line["extra_c5_smoothing"] = measure_smoothing_c5()

    if rank == 0:
        print(json.dumps(line))
    if world_size > 1:
        dist.barrier()                 # rank 0 may still be inside its single-rank extras
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
