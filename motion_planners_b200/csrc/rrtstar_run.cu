// RRT* as one persistent cooperative kernel: many iterations per launch, nothing returns to the host in between.
//
// Reference: motion_planners/rrt_star.py:100-137 AS EXECUTED ON PYTHON 3 (the second, "rewire the neighbours"
// loop iterates an exhausted filter object and never runs, so a node's cost is final once its own choose-parent
// pass is over).  Per iteration: goal-biased sample from the caller's pre-drawn streams (the reference draws
// random() only while no solution exists and not at iteration 0, and a configuration only when the goal was not
// chosen: the stream positions advance exactly so), the informed-set rejection (:108-110, a rejected sample does
// not count as an iteration), first arg-min nearest node (:117), the collision-free prefix of the extension
// (:118), one new node (:121), the goal test (:124), and the choose-parent pass over the nodes strictly closer
// than `radius` in index order (:128-137).
//
// The grid is one block per SM; the node scans are split over the blocks, and the short serial parts (the sample,
// the extension, the parent choice, the loop's scalars) are done redundantly by EVERY block on its own registers,
// so that only two grid-wide barriers are left per iteration: after the per-block partials of the nearest scan
// and after the per-block best candidate parents.  (With block 0 doing the sample and the parent choice alone,
// behind a third barrier, an iteration took 23.4 us, 45 % of it waiting at barriers.)  Block 0 writes the new
// node; the other blocks see it in memory after the next barrier and, until then, carry it in registers.  mpb_rrt_star_step (rrtstar.cu) does the same work one iteration per call
// and stays as the cross-check.
#include <cooperative_groups.h>
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#include "mpb_common.cuh"

namespace cg = cooperative_groups;

namespace {

#ifdef MPB_RRTSTAR_TIMING
__device__ __forceinline__ long long gtime()
{
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define RT_MARK(i) if (blockIdx.x == 0 && tid == 0) { const long long now_ = gtime(); S->t_phase[i] += now_ - t_last; t_last = now_; }
#else
#define RT_MARK(i)
#endif

constexpr int kRunThreads = 256;
constexpr int kRunMaxBlocks = 160;
constexpr int kRunWayChunk = 128;        // waypoints of an extension staged in shared memory at a time (<= kRunThreads)
constexpr int kRunLocalCand = 256;       // candidates one block may list per iteration

struct RunPartial {
    double d;
    int32_t i, pad;
};

struct RunCand {
    double d;
    int32_t idx, safe, ok, pad;
};

// everything the kernel keeps between iterations and launches (one per tree, in the caller's workspace)
struct RunState {
    long long n_nodes, iteration, u_pos, s_pos;
    int goal_idx, status, pad0, pad1;
    double goal_cost;
    // per-iteration slots, double-buffered by the parity of the kernel's local trip counter (a block that is still
    // reading trip t's slots cannot meet a block already writing trip t + 2's: a barrier of trip t + 1 lies between)
    RunPartial partial[2][kRunMaxBlocks];  // per block: its nearest node
    RunCand best_cand[2][kRunMaxBlocks];   // per block: its accepted candidate with the smallest cost + d (idx = -1: none)
    long long t_phase[8];   // MPB_RRTSTAR_TIMING builds: nanoseconds block 0 spent per phase, and rejected samples
};

struct RunArgs {
    WorldDev w;
    double *table;          // [cap, D] node configurations
    double *cost;           // [cap]
    int *parent;            // [cap]
    double *path_to;        // [cap, D] target of the extension that created / re-parented the node
    int *path_cnt;          // [cap] waypoints strictly between parent and node
    long long cap;
    const double *u;        // [n_u] random() draws
    long long n_u;
    const double *samples;  // [n_s, D] sample_fn() draws
    long long n_s;
    double start[MPB_MAX_DIM], goal[MPB_MAX_DIM];
    Weights wt;
    int metric;
    double radius, step, eps, goal_probability;
    int informed;
    long long max_iterations;
    RunState *st;
};

template <int D>
__global__ void __launch_bounds__(kRunThreads) rrt_star_run_kernel(RunArgs A)
{
    extern __shared__ __align__(16) double sm[];
    __shared__ double sd[kRunThreads / 32];
    __shared__ int si[kRunThreads / 32], ss[kRunThreads / 32];
    __shared__ int s_first;
    __shared__ double wp[kRunWayChunk * D], cinv[kRunWayChunk];
    __shared__ int lc_n, lc_idx[kRunLocalCand];
    __shared__ double lc_d[kRunLocalCand];
    cg::grid_group grid = cg::this_grid();
    const WorldSmem ws = stage_world(A.w, sm);          // ends with __syncthreads()
    RunState *S = A.st;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int items = 1 + A.w.n_box + A.w.n_sph;        // the environment, the boxes, the spheres

#ifdef MPB_RRTSTAR_TIMING
    long long t_last = gtime();
#endif
    // the loop's scalars, in every thread of every block (block 0 writes them back when the launch ends)
    long long n_nodes = S->n_nodes, iteration = S->iteration, u_pos = S->u_pos, s_pos = S->s_pos;
    int goal_idx = S->goal_idx, status = 0;
    double goal_cost = S->goal_cost;
    bool have_new = false;              // the previous trip appended a node that is not yet visible in memory
    double newest[D];
#pragma unroll
    for (int d = 0; d < D; ++d) newest[d] = 0.;
    for (unsigned trip = 0;; ++trip) {
        const int pb = trip & 1;
        // ---- A (every block, redundantly): the next accepted sample, or the reason to stop
        bool accepted = false, do_goal = false;
        double s[D];
        for (;;) {
            if (iteration >= A.max_iterations) {
                status = MPB_RRTSTAR_DONE;
                break;
            }
            if (n_nodes + 1 >= A.cap) {
                status = MPB_RRTSTAR_FULL;
                break;
            }
            // do_goal = goal_n is None and (iteration == 0 or random() < goal_probability)
            do_goal = false;
            long long u_next = u_pos;
            if (goal_idx < 0) {
                if (iteration == 0) {
                    do_goal = true;
                } else {
                    if (u_next >= A.n_u) {
                        status = MPB_RRTSTAR_NEED_DRAWS;
                        break;
                    }
                    do_goal = A.u[u_next++] < A.goal_probability;
                }
            }
            if (do_goal) {
#pragma unroll
                for (int d = 0; d < D; ++d) s[d] = A.goal[d];
            } else {
                if (s_pos >= A.n_s) {
                    status = MPB_RRTSTAR_NEED_DRAWS;         // the random() draw of this trip is not consumed
                    break;
                }
#pragma unroll
                for (int d = 0; d < D; ++d) s[d] = A.samples[s_pos * D + d];
                s_pos += 1;
            }
            u_pos = u_next;
            if (A.informed && goal_idx >= 0) {
                double a[D], g[D];
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    a[d] = A.start[d];
                    g[d] = A.goal[d];
                }
                if (d_add(metric_dist<D>(a, s, A.wt, A.metric), metric_dist<D>(s, g, A.wt, A.metric)) >= goal_cost) {
#ifdef MPB_RRTSTAR_TIMING
                    if (blockIdx.x == 0 && tid == 0) S->t_phase[7] += 1;
#endif
                    continue;                                   // rrt_star.py:108-110: not an iteration
                }
            }
            iteration += 1;
            accepted = true;
            break;
        }
        RT_MARK(0)
        if (!accepted) break;
        const long long n = n_nodes;

        // ---- B: first arg-min of distance_fn(node, s): per-block partials
        {
            double best = INFINITY;
            int bi = 0x7fffffff;
            const long long n_vis = have_new ? n - 1 : n;       // node n - 1 may not have reached memory yet
            for (long long j = (long long)blockIdx.x * kRunThreads + tid; j < n_vis; j += (long long)gridDim.x * kRunThreads) {
                double a[D];
#pragma unroll
                for (int d = 0; d < D; ++d) a[d] = A.table[j * D + d];
                const double dd = metric_dist<D>(a, s, A.wt, A.metric);
                if (dd < best) {
                    best = dd;
                    bi = (int)j;
                }
            }
            if (have_new && blockIdx.x == 0 && tid == 0) {      // the highest index: it only wins a strictly smaller distance
                const double dd = metric_dist<D>(newest, s, A.wt, A.metric);
                if (dd < best) {
                    best = dd;
                    bi = (int)(n - 1);
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double od = __shfl_down_sync(0xffffffffu, best, o);
                const int oi = __shfl_down_sync(0xffffffffu, bi, o);
                if (od < best || (od == best && oi < bi)) {
                    best = od;
                    bi = oi;
                }
            }
            if (lane == 0) {
                sd[warp] = best;
                si[warp] = bi;
            }
            __syncthreads();
            if (tid == 0) {
                for (int w = 1; w < kRunThreads / 32; ++w)
                    if (sd[w] < best || (sd[w] == best && si[w] < bi)) {
                        best = sd[w];
                        bi = si[w];
                    }
                S->partial[pb][blockIdx.x].d = best;
                S->partial[pb][blockIdx.x].i = bi;
            }
        }
        RT_MARK(2)
        grid.sync();
        RT_MARK(1)

        // ---- C (every block, redundantly): the nearest node and the safe prefix of the extension towards s
        double best = INFINITY;
        int nearest = 0x7fffffff;
        for (int p = tid; p < (int)gridDim.x; p += kRunThreads) {      // one partial per thread, then a block reduction
            const double pd = S->partial[pb][p].d;
            const int pi = S->partial[pb][p].i;
            if (pd < best || (pd == best && pi < nearest)) {
                best = pd;
                nearest = pi;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double od = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, nearest, o);
            if (od < best || (od == best && oi < nearest)) {
                best = od;
                nearest = oi;
            }
        }
        if (lane == 0) {
            sd[warp] = best;
            si[warp] = nearest;
        }
        __syncthreads();
        for (int w = 0; w < kRunThreads / 32; ++w)
            if (sd[w] < best || (sd[w] == best && si[w] < nearest)) {
                best = sd[w];
                nearest = si[w];
            }
        __syncthreads();                                                // sd / si are reused by the next trip
        double a[D];
#pragma unroll
        for (int d = 0; d < D; ++d) a[d] = A.table[(long long)nearest * D + d];
        const int n_way = waypoint_count<D>(a, s, A.step);
        // The waypoints come from a serial recurrence: one thread writes a chunk of them to shared memory, then all
        // threads test (waypoint, obstacle) pairs side by side; the first chunk with a hit ends the walk.
        int safe = n_way;
        double nw[D];
#pragma unroll
        for (int d = 0; d < D; ++d) nw[d] = a[d];
        for (int base = 0; base < n_way; base += kRunWayChunk) {
            const int m = min(kRunWayChunk, n_way - base);
            // c_i = 1 / (n - i) for the chunk, one division per thread; then one thread per coordinate runs
            // q <- fl(fl(c_i * fl(b - q)) + q) (samplers.py:113-127; the coordinates are independent chains)
            if (tid < m) cinv[tid] = __ddiv_rn(1.0, (double)(n_way - (base + tid)));
            if (tid == 0) s_first = 0x7fffffff;
            __syncthreads();
            if (tid < D) {
                double q = (base == 0) ? A.table[(long long)nearest * D + tid] : wp[(kRunWayChunk - 1) * D + tid];
                double b = s[0];
#pragma unroll
                for (int d = 1; d < D; ++d) b = (d == tid) ? s[d] : b;
                for (int i = 0; i < m; ++i) {
                    q = d_add(d_mul(cinv[i], d_sub(b, q)), q);
                    wp[i * D + tid] = q;
                }
            }
            __syncthreads();
            for (int t = tid; t < m * items; t += kRunThreads) {
                const int i = t / items, it = t - i * items;
                if (i >= s_first) continue;                 // a racy read: only skips work that cannot matter
                double q[D];
#pragma unroll
                for (int d = 0; d < D; ++d) q[d] = wp[i * D + d];
                bool hit;
                if (it == 0)
                    hit = outside_env<D>(ws, q);
                else if (it <= A.w.n_box)
                    hit = in_box<D>(ws, it - 1, q);
                else
                    hit = in_sphere<D>(ws, it - 1 - A.w.n_box, q);
                if (hit) atomicMin(&s_first, i);
            }
            __syncthreads();
            const int first = s_first;
            const int last_free = (first == 0x7fffffff ? m : first) - 1;     // within this chunk
            if (last_free >= 0) {
#pragma unroll
                for (int d = 0; d < D; ++d) nw[d] = wp[last_free * D + d];
            }
            __syncthreads();                                // wp and s_first are rewritten by the next chunk
            if (first != 0x7fffffff) {
                safe = base + first;                        // len(safe_path(...)), rrt_star.py:118
                break;
            }
        }
        if (safe == 0) {                                    // every block takes the same branch
            have_new = false;
            continue;
        }
        const double d_near_new = metric_dist<D>(a, nw, A.wt, A.metric);
        const double cost0 = d_add(A.cost[nearest], d_near_new);          // OptimalNode.__init__: parent.cost + d
        RT_MARK(3)

        // ---- D: nodes strictly closer than radius whose cost could beat the new node's = candidate parents.  The scan
        // only lists them (per block, in shared memory); then one WARP per candidate walks the edge candidate -> new:
        // every lane runs the waypoint recurrence, the obstacles are spread over the lanes, a vote finds the first hit.
        if (tid == 0) lc_n = 0;
        __syncthreads();
        for (long long j = (long long)blockIdx.x * kRunThreads + tid; j < n; j += (long long)gridDim.x * kRunThreads) {
            double c[D];
#pragma unroll
            for (int d = 0; d < D; ++d) c[d] = A.table[j * D + d];
            const double dd = metric_dist<D>(c, nw, A.wt, A.metric);
            if (!(dd < A.radius) || !(d_add(A.cost[j], dd) < cost0)) continue;     // rrt_star.py:128, :134
            const int slot = atomicAdd(&lc_n, 1);
            if (slot < kRunLocalCand) {
                lc_idx[slot] = (int)j;
                lc_d[slot] = dd;
            }
        }
        __syncthreads();
        int n_local = lc_n;
        if (n_local > kRunLocalCand) {
            if (tid == 0) S->status = MPB_RRTSTAR_CAND_OVERFLOW;
            n_local = kRunLocalCand;
        }
        double wbest_v = INFINITY;
        int wbest_i = 0x7fffffff, wbest_s = 0;
        for (int c0 = warp; c0 < n_local; c0 += kRunThreads / 32) {
            const int j = lc_idx[c0];
            double c[D], w2[D], prev[D];
#pragma unroll
            for (int d = 0; d < D; ++d) {
                c[d] = A.table[(long long)j * D + d];
                w2[d] = c[d];
                prev[d] = c[d];
            }
            const int nw_way = waypoint_count<D>(c, nw, A.step);
            int sf = nw_way;
            for (int i = 0; i < nw_way; ++i) {
                waypoint_step<D>(w2, nw, nw_way, i);
                bool hit = false;
                for (int it = lane; it < items && !hit; it += 32) {
                    if (it == 0)
                        hit = outside_env<D>(ws, w2);
                    else if (it <= A.w.n_box)
                        hit = in_box<D>(ws, it - 1, w2);
                    else
                        hit = in_sphere<D>(ws, it - 1 - A.w.n_box, w2);
                }
                if (__any_sync(0xffffffffu, hit)) {
                    sf = i;
                    break;
                }
#pragma unroll
                for (int d = 0; d < D; ++d) prev[d] = w2[d];
            }
            if (lane == 0) {
                const bool ok = (sf > 0) && (metric_dist<D>(nw, prev, A.wt, A.metric) < A.eps);   // rrt_star.py:135-136
                if (ok) {
                    const double v = d_add(A.cost[j], lc_d[c0]);
                    if (v < wbest_v || (v == wbest_v && j < wbest_i)) {
                        wbest_v = v;
                        wbest_i = j;
                        wbest_s = sf;
                    }
                }
            }
        }
        // the block's best accepted candidate (smallest cost + d, lowest index among equals)
        if (lane == 0) {
            sd[warp] = wbest_v;
            si[warp] = wbest_i;
            ss[warp] = wbest_s;
        }
        __syncthreads();
        if (tid == 0) {
            double bv = sd[0];
            int bi = si[0], bs = ss[0];
            for (int w = 1; w < kRunThreads / 32; ++w)
                if (sd[w] < bv || (sd[w] == bv && si[w] < bi)) {
                    bv = sd[w];
                    bi = si[w];
                    bs = ss[w];
                }
            RunCand &cd = S->best_cand[pb][blockIdx.x];
            cd.d = bv;
            cd.idx = (bi == 0x7fffffff) ? -1 : bi;
            cd.safe = bs;
            cd.ok = 1;
        }
        __syncthreads();
        RT_MARK(4)
        grid.sync();
        RT_MARK(1)

        // ---- E (every block, redundantly): the parent choice and the goal test; block 0 writes the new node.
        // "for n in neighbors (index order): if (n.cost + d) < new.cost and the edge is free: rewire" ends at the
        // accepted candidate with the smallest n.cost + d, the lowest index among equals (a later equal value is
        // not strictly smaller), provided that beats the cost through the nearest node: an arg-min.
        {
            double bv = INFINITY;
            int bidx = 0x7fffffff, bc = -1;
            for (int c = tid; c < (int)gridDim.x; c += kRunThreads) {
                const RunCand &cd = S->best_cand[pb][c];
                if (cd.idx < 0) continue;
                if (cd.d < bv || (cd.d == bv && cd.idx < bidx)) {
                    bv = cd.d;
                    bidx = cd.idx;
                    bc = c;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
                const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
                if (ov < bv || (ov == bv && oi < bidx)) {
                    bv = ov;
                    bidx = oi;
                    bc = oc;
                }
            }
            if (lane == 0) {
                sd[warp] = bv;
                si[warp] = bidx;
                ss[warp] = bc;
            }
            __syncthreads();
            for (int w = 0; w < kRunThreads / 32; ++w)
                if (sd[w] < bv || (sd[w] == bv && si[w] < bidx)) {
                    bv = sd[w];
                    bidx = si[w];
                    bc = ss[w];
                }
            __syncthreads();                                            // sd / si / ss are reused by the next trip
            double cur = cost0;
            int par = nearest, cnt = safe - 1;
            bool to_sample = true;
            if (bc >= 0 && bv < cost0) {
                cur = bv;
                par = bidx;
                cnt = S->best_cand[pb][bc].safe - 1;
                to_sample = false;
            }
            if (blockIdx.x == 0 && tid == 0) {
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    A.table[n * D + d] = nw[d];
                    A.path_to[n * D + d] = to_sample ? s[d] : nw[d];
                }
                A.cost[n] = cur;
                A.parent[n] = par;
                A.path_cnt[n] = cnt;
            }
            if (do_goal) {
                double g[D];
#pragma unroll
                for (int d = 0; d < D; ++d) g[d] = A.goal[d];
                if (metric_dist<D>(nw, g, A.wt, A.metric) < A.eps) {           // rrt_star.py:124-126
                    goal_idx = (int)n;
                    goal_cost = cur;
                }
            }
            n_nodes = n + 1;
            have_new = true;
#pragma unroll
            for (int d = 0; d < D; ++d) newest[d] = nw[d];
        }
        RT_MARK(5)
    }
    // the launch ends: block 0 leaves the loop's scalars for the next launch and the host (a candidate overflow
    // reported by some block during the run stays)
    if (blockIdx.x == 0 && tid == 0) {
        S->n_nodes = n_nodes;
        S->iteration = iteration;
        S->u_pos = u_pos;
        S->s_pos = s_pos;
        S->goal_idx = goal_idx;
        S->goal_cost = goal_cost;
        if (S->status != MPB_RRTSTAR_CAND_OVERFLOW) S->status = status;
    }
}

template <int D>
int launch_run(const mpb_world *world, RunArgs &A, cudaStream_t st)
{
    A.w = world->dev;
    const size_t smem = world_smem_bytes(world->dev) + 16;
    MPB_CUDA_CHECK(cudaFuncSetAttribute(rrt_star_run_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, sms = 1, per_sm = 0;
    MPB_CUDA_CHECK(cudaGetDevice(&dev));
    MPB_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    MPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rrt_star_run_kernel<D>, kRunThreads, smem));
    MPB_REQUIRE(per_sm >= 1, "mpb_rrt_star_run: the kernel does not fit an SM with this obstacle set");
    int blocks = sms < kRunMaxBlocks ? sms : kRunMaxBlocks;
    if (const char *e = getenv("MPB200_RRTSTAR_BLOCKS")) {     // tuning knob: fewer blocks = cheaper grid barriers
        const int v = atoi(e);
        if (v >= 1 && v < blocks) blocks = v;
    }
    void *args[] = {&A};
    MPB_CUDA_CHECK(cudaLaunchCooperativeKernel((const void *)rrt_star_run_kernel<D>, dim3(blocks), dim3(kRunThreads), args,
                                               smem, st));
    return MPB_OK;
}

}   // namespace

MPB_API size_t mpb_rrt_star_run_state_bytes(void) { return (sizeof(RunState) + 255) & ~(size_t)255; }

MPB_API int mpb_rrt_star_run(const mpb_world *world, const double *start, const double *goal, double radius,
                             const double *weights, int metric, double step, double eps, double goal_probability,
                             int informed, int64_t max_iterations, double *table, double *cost, int32_t *parent,
                             double *path_to, int32_t *path_cnt, int64_t capacity, const double *u_draws, int64_t n_u,
                             const double *sample_draws, int64_t n_s, void *state, int reset, void *stream)
{
    const char *who = "mpb_rrt_star_run";
    MPB_REQUIRE(world != nullptr, "%s: world is NULL", who);
    MPB_REQUIRE_CSPACE(world, who);
    MPB_REQUIRE(start && goal && table && cost && parent && path_to && path_cnt && state, "%s: NULL buffer", who);
    MPB_REQUIRE(capacity >= 2 && capacity < 2147483647LL, "%s: bad capacity", who);
    MPB_REQUIRE(metric >= 0 && metric <= 2, "%s: bad metric %d", who, metric);
    MPB_REQUIRE(step > 0. && radius >= 0. && n_u >= 0 && n_s >= 0 && max_iterations >= 0, "%s: bad argument", who);
    MPB_REQUIRE(n_u == 0 || u_draws, "%s: NULL draws", who);
    MPB_REQUIRE(n_s == 0 || sample_draws, "%s: NULL draws", who);
    MPB_REQUIRE(((uintptr_t)state & 255) == 0, "%s: state must be 256-byte aligned", who);
    MPB_REQUIRE(world_smem_bytes(world->dev) + 16 <= 150 * 1024, "%s: obstacle set too large for shared memory", who);
    const int dim = world->dev.dim;
    cudaStream_t st = (cudaStream_t)stream;
    RunArgs A;
    memset(&A, 0, sizeof(A));
    A.table = table;
    A.cost = cost;
    A.parent = parent;
    A.path_to = path_to;
    A.path_cnt = path_cnt;
    A.cap = capacity;
    A.u = u_draws;
    A.n_u = n_u;
    A.samples = sample_draws;
    A.n_s = n_s;
    for (int d = 0; d < dim; ++d) {
        A.start[d] = start[d];
        A.goal[d] = goal[d];
    }
    A.wt = make_weights(weights, dim);
    A.metric = metric;
    A.radius = radius;
    A.step = step;
    A.eps = eps;
    A.goal_probability = goal_probability;
    A.informed = informed;
    A.max_iterations = max_iterations;
    A.st = reinterpret_cast<RunState *>(state);
    if (reset) {
        // a fresh tree: node 0 = start, cost 0, no parent
        RunState init;
        memset(&init, 0, sizeof(init));
        init.n_nodes = 1;
        init.goal_idx = -1;
        MPB_CUDA_CHECK(cudaMemcpyAsync(state, &init, offsetof(RunState, partial), cudaMemcpyHostToDevice, st));
        MPB_CUDA_CHECK(cudaMemcpyAsync(table, start, sizeof(double) * dim, cudaMemcpyHostToDevice, st));
        MPB_CUDA_CHECK(cudaMemsetAsync(cost, 0, sizeof(double), st));
        const int32_t none = -1;
        MPB_CUDA_CHECK(cudaMemcpyAsync(parent, &none, sizeof(none), cudaMemcpyHostToDevice, st));
        MPB_CUDA_CHECK(cudaMemsetAsync(path_cnt, 0, sizeof(int32_t), st));
        MPB_CUDA_CHECK(cudaStreamSynchronize(st));      // the copies above read host stack memory
    } else {
        // continue: the stream positions restart at the head of the new draw arrays
        MPB_CUDA_CHECK(cudaMemsetAsync(reinterpret_cast<char *>(state) + offsetof(RunState, u_pos), 0, 2 * sizeof(long long), st));
        MPB_CUDA_CHECK(cudaMemsetAsync(reinterpret_cast<char *>(state) + offsetof(RunState, status), 0, sizeof(int), st));
    }
#define CALL(DD) launch_run<DD>(world, A, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    mpb_set_error("%s: unsupported dim %d", who, dim);
    return MPB_EINVAL;
}

// the scalars of the state, for the host: {n_nodes, iteration, u_pos, s_pos, goal_idx, status} as int64 + goal_cost
MPB_API int mpb_rrt_star_run_read(const void *state, int64_t *out6, double *out_goal_cost, void *stream)
{
    MPB_REQUIRE(state && out6 && out_goal_cost, "mpb_rrt_star_run_read: NULL buffer");
    RunState head;
    MPB_CUDA_CHECK(cudaMemcpyAsync(&head, state, offsetof(RunState, partial), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    MPB_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    out6[0] = head.n_nodes;
    out6[1] = head.iteration;
    out6[2] = head.u_pos;
    out6[3] = head.s_pos;
    out6[4] = head.goal_idx;
    out6[5] = head.status;
    *out_goal_cost = head.goal_cost;
    return MPB_OK;
}
