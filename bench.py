#!/usr/bin/env python
"""bench.py -- one JSON line for the kNN + edge-collision hot path (BASELINE.json's metric).

    python bench.py --gpus 1 --steps K --warmup W                       (single GPU)
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...                                (CPU arm: the oracle port)

Default workload: `c4lite` = configs[3], the samples-sharded + all-gathered PRM that BASELINE.json's
"at 1/2/4/8 B200" metric and its target (1M samples, 7-DoF, k=10) are quoted on, with the reference's own
box primitives (the sphere-approximated arm has no counterpart in the reference); it fits one GPU.
`extra_c2` in the same line is configs[1] (100k x 6-DoF, 64 boxes).

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

NCU_DRAM_BYTES_PER_LAUNCH = {"c4lite": 978.515968e6 + 182.364672e6}          # knn_filter_kernel (FFMA2 scan)
NCU_DRAM_BYTES_PER_LAUNCH_TC = {"c4lite": 5.530121e9 + 75.893248e6}        # knn_tc_kernel (tensor-core scan)

METRIC = "PRM edges collision-checked/sec (kNN k=10 + vertex + edge checks per roadmap build step)"
UNIT = "edges/s"

WORKLOADS = {
    # name -> (world factory name, dim, n_samples, k, description)
    "c2": ("c2_world", 6, 100000, 10, "configs[1]: lazy-PRM roadmap, 100k samples, 6-DoF unit cube, 64 random boxes, k=10"),
    "c4lite": ("c4lite_world", 7, 1000000, 10, "configs[3] (the sharded, all-gathered PRM the 1/2/4/8-GPU metric and the "
               "north-star target are quoted on) with reference primitives: 1M samples, 7-DoF, 256 C-space boxes, k=10"),
    "c1": ("c1", 2, 2000, 10, "configs[0]: 2-D demo fixture, 2k samples, k=10"),
}


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
    hi = min(first, n)
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
    sample = min(n, args.ref_sample)
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
                   "n_box": world.n_box, "n_sph": world.n_sph},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu_baseline.cores(), "kind": "port",
                         "sample": "%d of %d queries and their %d directed edges per step; kd-tree built once, "
                                   "untimed" % (sample, n, sample * k)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------- batched RRT extra
def measure_rrt_batch(world_size, rank, dist, total_queries=4096, iters=100):
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


def measure_rrt_star(iters=20000):
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
            "kernel_launches": int(getattr(rrt_star, "last_launches", 0))}


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
    ap.add_argument("--workload", default=os.environ.get("MPB_BENCH_WORKLOAD", "c4lite"), choices=sorted(WORKLOADS))
    ap.add_argument("--ref-sample", type=int, default=100000)
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
        if world_size > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)                # max over ranks
        total_ms = float(t.item())
        edges_local = torch.tensor([rm.n_edges_checked()], device="cuda", dtype=torch.int64)
        fb = rm.stats[:1].clone()
        if world_size > 1:
            dist.all_reduce(edges_local)
            dist.all_reduce(fb)
        edges_total = int(edges_local.item())
        ms_step = total_ms / steps
        out = {"world": world, "X": X, "k": k, "desc": desc, "n": n, "D": D, "ms_step": ms_step,
               "edges": edges_total, "value": edges_total / (ms_step * 1e-3), "qps": n / (ms_step * 1e-3),
               "phase_ms": {a: b / steps for a, b in phase.items()}, "clocks": clocks, "launches": rm.launches,
               "fallback_queries": int(fb.item()), "n_local": hi - lo, "wall_s": wall_s}

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
                    dev = pinned.cuda(non_blocking=True)
                    rm2.build(dev, gathered=rm2.gather_samples_equal(dev, gathered_buf) if equal else None)
                    outs = (rm2.nbr_idx, rm2.vertex_hit, rm2.edge_hit)
                    if not hasattr(rm2, "_host_out"):
                        rm2._host_out = tuple(torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in outs)
                    for hbuf, o in zip(rm2._host_out, outs):
                        hbuf.copy_(o, non_blocking=True)
                    torch.cuda.current_stream().synchronize()
                    return rm2._host_out
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

    # ---- roofline of the dominant kernel (the fp32 filter pass of the kNN)
    peaks_file = {}
    try:
        peaks_file = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    fp32_scalar = ops.peak_fma(torch.float32)
    fp32_packed = ops.peak_fma(torch.float32, packed=True)
    fp32_peak = max(fp32_scalar, fp32_packed)            # the higher of the FFMA and FFMA2 issue peaks
    fp64_peak = ops.peak_fma(torch.float64)
    n, D, k = m["n"], m["D"], m["k"]
    knn_ms = m["phase_ms"].get("knn", float("nan"))
    fma_flops = 2.0 * D * float(m["n_local"]) * n            # one FFMA per dimension per (query, sample) pair
    achieved = fma_flops / (knn_ms * 1e-3)
    hbm_peak = peaks_file.get("hbm_gbs")
    # algorithmic HBM bytes of the step: samples in (fp64), packed tiles out+in once, neighbour matrix
    # (idx + dist) out and idx back in for the edge pass, flags out
    step_bytes = m["n_local"] * (D * 8 + (k + 1) * (8 + 8 + 8 + 4) + 1) + n * (D * 8 + 2 * 32)
    roofline = {
        "kernel": "knn_filter_kernel", "bound": "fp32", "achieved": achieved / 1e12, "peak": fp32_peak / 1e12,
        "unit": "TFLOP/s", "frac": achieved / fp32_peak,
        "peak_source": "measured live on this GPU: max of the FFMA and FFMA2 micro-benchmarks (mpb_peak_fma_f32 / "
                       "mpb_peak_fma_f32x2); MEASURED_PEAKS.json carries only HBM and bf16 tensor peaks; nominal "
                       "fp32 = 148 SM x 128 lanes x 2 x 1.965 GHz = 74.4",
        "peak_ffma_tflops": fp32_scalar / 1e12, "peak_ffma2_tflops": fp32_packed / 1e12,
        "flops_per_pair": 2 * D, "survey_flops_per_pair": 3 * D,
        "frac_on_survey_count": (3.0 * D * m["n_local"] * n / (knn_ms * 1e-3)) / fp32_peak,
        # dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the committed ncu --set full capture
        # (profiles/r1_ncu_full_knn_filter_edges_check_raw.csv); measured for the default workload only
        "traffic": NCU_DRAM_BYTES_PER_LAUNCH.get(args.workload) if world_size == 1 else None,
        "kernel_ms": knn_ms,
        "hbm": {"achieved_gbs": step_bytes / (m["ms_step"] * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                "frac": (step_bytes / (m["ms_step"] * 1e-3) / 1e9 / hbm_peak) if hbm_peak else None,
                "note": "whole step, algorithmic bytes; the path is FMA-pipe bound, not HBM bound"},
        "fp64_fma_peak_tflops": fp64_peak / 1e12,
    }

    from motion_planners_b200 import _lib as _mplib
    if _mplib.lib().mpb_knn_filter_kind(int(n), int(D), int(k) + 1) == 1:
        # The kNN ran on the tensor-core scan: three tcgen05.mma kind::tf32 k-steps (K = 24 operand columns: every
        # fp32 value enters as two TF32 parts) per 128-query x 256-candidate tile.  Executed tensor flops per
        # (query, sample) pair = 2 * 24.  Peak: MEASURED_PEAKS.json has the dense bf16 rate (sustained figure: the
        # kernel is timed inside a long step); TF32 runs at half of it.
        tc_flops = 2.0 * 24 * float(m["n_local"]) * n
        bf16_peak = peaks_file.get("bf16_tflops_sustained") or peaks_file.get("bf16_tflops")
        tf32_peak = (bf16_peak / 2.0) if bf16_peak else 1130.0
        tc_achieved = tc_flops / (knn_ms * 1e-3) / 1e12
        fp32_view = dict(roofline)
        for key in ("kernel", "bound", "traffic", "hbm", "kernel_ms"):
            fp32_view.pop(key, None)
        roofline = {
            "kernel": "knn_tc_kernel", "bound": "tensor", "achieved": tc_achieved, "peak": tf32_peak,
            "unit": "TFLOP/s", "frac": tc_achieved / tf32_peak,
            "peak_source": ("MEASURED_PEAKS.json bf16_tflops_sustained / 2 (TF32 is half the bf16 rate)" if bf16_peak
                            else "fallback: nominal dense TF32, 1130 TFLOP/s"),
            "flops_per_pair": 48, "survey_flops_per_pair": 3 * D,
            "note": "the MMAs occupy about a quarter of the kernel; what bounds it is the CUDA-core epilogue that "
                    "reads every key out of tensor memory (tcgen05.ld), takes a min-tree per 32 keys and inserts the "
                    "takers (profiles/r1_summary.md, r1h)",
            # dram__bytes_read.sum + dram__bytes_write.sum of one launch (profiles/r1j_ncu_full_knn_tc_raw.csv)
            "traffic": NCU_DRAM_BYTES_PER_LAUNCH_TC.get(args.workload) if world_size == 1 else None,
            "kernel_ms": knn_ms, "hbm": roofline["hbm"],
            # the same time expressed on the fp32 formulation's count (2*D flop per pair) against the FFMA2 peak:
            # above 1 means faster than a perfect FFMA2 kernel could be
            "as_fp32_filter": fp32_view,
        }

    line = {
        "metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": world_size, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": m["ms_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None,
        "dtype": "f64 (fp32 / split-TF32 filter pass inside the kNN, results bit-exact to the fp64 reference)",
        "data": "synthetic", "knn_queries_per_s": m["qps"],
        "config": {"workload": args.workload, "description": m["desc"], "n_samples": n, "dim": D, "k": k,
                   "n_box": m["world"].n_box, "n_sph": m["world"].n_sph, "edges_per_step": m["edges"],
                   "sharding": "samples sharded over ranks, one NCCL all-gather per step, edges stay sharded",
                   "l2": "flushed between timed steps (256 MiB write, untimed); inputs are smaller than L2"},
        "phases_ms": m["phase_ms"],
        "kernel_rates": {"knn_queries_per_s": m["n_local"] * world_size / (knn_ms * 1e-3),
                         "edges_per_s": m["edges"] / (m["phase_ms"].get("edges", float("nan")) * 1e-3)},
        "fallback_queries": m["fallback_queries"],
        "roofline": roofline, "clocks": m["clocks"], "e2e": m["e2e"], "gpu_launches": m["launches"] * args.steps,
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
        p = cpu_baseline.python_port_step(m["world"], m["X"], 0, 40, k)
        line["cpu_baseline"]["python_port"] = {
            "value": p["n_edges"] / p["seconds"], "unit": UNIT, "cores": 1,
            "sample": "first 40 queries, %d edges, per-call Python structure of the reference" % p["n_edges"]}

    if not args.no_extra:
        other = "c2" if args.workload != "c2" else "c4lite"
        x = measure(other, max(2, min(args.steps, 10 if other == "c2" else 3)), 3, with_e2e=False)
        x_knn = x["phase_ms"].get("knn", float("nan"))
        line["extra_" + other] = {
            "description": x["desc"], "ms_per_step": x["ms_step"], "edges_per_s": x["value"],
            "knn_queries_per_s": x["qps"], "edges_per_step": x["edges"], "phases_ms": x["phase_ms"],
            "knn_fp32_frac": (2.0 * x["D"] * x["n_local"] * x["n"] / (x_knn * 1e-3)) / fp32_peak,
            "fallback_queries": x["fallback_queries"], "clocks": x["clocks"]}

    if not args.no_extra:
        line["extra_c3_rrt_connect_batch"] = measure_rrt_batch(world_size, rank, dist if world_size > 1 else None)
        if rank == 0:
            line["extra_c5_rrt_star"] = measure_rrt_star()
            line["extra_c1_prm"] = measure_prm_c1()

    if rank == 0:
        print(json.dumps(line))
    if world_size > 1:
        dist.barrier()                 # rank 0 may still be inside its single-rank extras
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
