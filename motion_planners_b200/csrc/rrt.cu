// RRT-Connect for a batch of independent queries, one thread block per query, the whole search in
// one launch: the two trees of a query live in global memory, every iteration runs on the block.
//
// Per iteration (motion_planners/rrt_connect.py:56-63), with the query's own pre-drawn random numbers:
//   swap   = u < |T1| / (|T1| + |T2|)                          random_swap, rrt_connect.py:26-29
//   last1  = extend_towards(tree1, target, backward = swap)    primitives.py:19-29
//   last2, success = extend_towards(tree2, last1.config, backward = not swap)
// extend_towards = first arg-min of distance_fn over the tree (block-wide scan), the waypoints of
// extend_fn (tkinter/samplers.py:113-127; forward from the tree node, or -- goal tree -- generated from the
// target towards the tree node and consumed reversed, primitives.py:13-16), collision_fn on every waypoint
// (threads in parallel), one new node per waypoint of the collision-free prefix.
// Same device arithmetic as the stand-alone kernels (mpb_trees_nearest, mpb_edges_check_pairs,
// mpb_trees_append), so the results are identical to composing those.
#include <limits.h>
#include <math.h>

#include "mpb_common.cuh"

#define FULL_MASK 0xffffffffu

namespace {

constexpr int kThreads = 128;
constexpr int kChunk = 256;   // waypoints staged in shared memory at a time

struct Scratch {
    double *way;    // [kChunk, D]
    double *vec;    // from[D], to[D], target[D]
    double *rd;     // [kThreads / 32] reduction
    int *ri;        // [kThreads / 32]
    int *flag;      // [0] first hit, [1] last hit, [2] nearest
};

// first arg-min of the rounded distance over the tree (utils.py:61-65), all threads get the answer
template <int D>
__device__ int block_nearest(const double *base, int n, const double (&tg)[D], const Weights &wt,
                             int metric, const Scratch &sc)
{
    double best = INFINITY;
    int bi = INT_MAX;
    for (int j = threadIdx.x; j < n; j += kThreads) {
        double a[D];
#pragma unroll
        for (int d = 0; d < D; ++d) a[d] = base[(int64_t)j * D + d];
        const double dd = metric_dist<D>(a, tg, wt, metric);
        if (dd < best) {
            best = dd;
            bi = j;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double od = __shfl_xor_sync(FULL_MASK, best, o);
        const int oi = __shfl_xor_sync(FULL_MASK, bi, o);
        if (od < best || (od == best && oi < bi)) {
            best = od;
            bi = oi;
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        sc.rd[warp] = best;
        sc.ri[warp] = bi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kThreads / 32; ++w) {
            if (sc.rd[w] < best || (sc.rd[w] == best && sc.ri[w] < bi)) {
                best = sc.rd[w];
                bi = sc.ri[w];
            }
        }
        sc.flag[2] = bi;
    }
    __syncthreads();
    return sc.flag[2];
}

template <int D>
__device__ __forceinline__ bool config_collides(const WorldDev &w, const WorldSmem &s, const double (&p)[D])
{
    bool hit = outside_env<D>(s, p);
    for (int b = 0; b < w.n_box && !hit; ++b) hit = in_box<D>(s, b, p);
    for (int k = 0; k < w.n_sph && !hit; ++k) hit = in_sphere<D>(s, k, p);
    return hit;
}

// Thread d < D advances coordinate d of the recurrence over waypoints [c0, c0 + len) into shared memory.
template <int D>
__device__ __forceinline__ void generate_chunk(double &qd, double to_d, int n, int c0, int len, double *way)
{
    if (threadIdx.x < D) {
        for (int i = 0; i < len; ++i) {
            const double c = __ddiv_rn(1.0, (double)(n - (c0 + i)));
            qd = d_add(d_mul(c, d_sub(to_d, qd)), qd);
            way[i * D + threadIdx.x] = qd;
        }
    }
}

// extend_towards for one tree.  Returns the index of the tree's new last node; `success` = no waypoint
// collided; `overflow` = the tree has no room (nothing stored).
template <int D>
__device__ int extend_towards(const WorldDev &w, const WorldSmem &ws, double *nodes, int32_t *parents, int &size, int64_t cap, bool backward,
                              const Weights &wt, int metric, double step, const Scratch &sc, bool &success,
                              bool &overflow)
{
    double tg[D];
#pragma unroll
    for (int d = 0; d < D; ++d) tg[d] = sc.vec[2 * D + d];
    const int near = block_nearest<D>(nodes, size, tg, wt, metric, sc);
    if (threadIdx.x < D) {
        const double nv = nodes[(int64_t)near * D + threadIdx.x];
        sc.vec[threadIdx.x] = backward ? tg[threadIdx.x] : nv;          // from
        sc.vec[D + threadIdx.x] = backward ? nv : tg[threadIdx.x];      // to
    }
    if (threadIdx.x == 0) {
        sc.flag[0] = INT_MAX;
        sc.flag[1] = -1;
    }
    __syncthreads();
    double a[D], b[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        a[d] = sc.vec[d];
        b[d] = sc.vec[D + d];
    }
    const int n = waypoint_count<D>(a, b, step);
    success = false;
    if ((int64_t)size + n > cap) {
        overflow = true;
        return near;
    }
    const double from_d = (threadIdx.x < D) ? sc.vec[threadIdx.x] : 0.;
    const double to_d = (threadIdx.x < D) ? sc.vec[D + threadIdx.x] : 0.;
    double qd = from_d;
    int kept = 0;

    // pass A: generate, test; a forward extension stores its safe prefix as it goes
    bool stop = false;
    for (int c0 = 0; c0 < n && !stop; c0 += kChunk) {
        const int len = min(kChunk, n - c0);
        generate_chunk<D>(qd, to_d, n, c0, len, sc.way);
        __syncthreads();
        for (int i = threadIdx.x; i < len; i += kThreads) {
            double p[D];
#pragma unroll
            for (int d = 0; d < D; ++d) p[d] = sc.way[i * D + d];
            if (config_collides<D>(w, ws, p)) {
                atomicMin(&sc.flag[0], c0 + i);
                atomicMax(&sc.flag[1], c0 + i);
            }
        }
        __syncthreads();
        const int first = sc.flag[0];
        if (!backward) {
            const int upto = min(first, c0 + len);
            const int cnt = max(upto - c0, 0);
            for (int e = threadIdx.x; e < cnt * D; e += kThreads) {
                const int i = e / D, d = e - i * D;
                nodes[(int64_t)(size + c0 + i) * D + d] = sc.way[i * D + d];
            }
            for (int i = threadIdx.x; i < cnt; i += kThreads) {
                const int pos = size + c0 + i;
                parents[pos] = (pos == size) ? near : pos - 1;
            }
            kept += cnt;
            stop = (first != INT_MAX);
        }
        __syncthreads();
    }
    if (backward) {
        // the safe prefix of the reversed list = the generated waypoints after the last colliding one
        const int last = sc.flag[1];
        kept = n - 1 - last;
        const bool resident = (n <= kChunk);     // the single chunk is still in shared memory
        qd = from_d;
        for (int c0 = 0; c0 < n; c0 += kChunk) {
            const int len = min(kChunk, n - c0);
            if (!resident) {
                generate_chunk<D>(qd, to_d, n, c0, len, sc.way);
                __syncthreads();
            }
            for (int e = threadIdx.x; e < len * D; e += kThreads) {
                const int i = e / D, d = e - i * D;
                const int g = c0 + i;
                if (g > last) nodes[(int64_t)(size + (n - 1 - g)) * D + d] = sc.way[i * D + d];
            }
            for (int i = threadIdx.x; i < len; i += kThreads) {
                const int g = c0 + i;
                if (g > last) {
                    const int pos = size + (n - 1 - g);
                    parents[pos] = (pos == size) ? near : pos - 1;
                }
            }
            __syncthreads();
        }
    }
    success = (sc.flag[0] == INT_MAX);
    const int last_node = kept > 0 ? size + kept - 1 : near;
    size += kept;
    __threadfence_block();
    __syncthreads();
    return last_node;
}

template <int D>
__global__ void __launch_bounds__(kThreads)
    rrt_connect_kernel(WorldDev w, Weights wt, int metric, double step, const double *__restrict__ starts,
                       const double *__restrict__ goals, const double *__restrict__ swap_u,
                       const double *__restrict__ targets, int iters, double *nodes, int32_t *parents,
                       int32_t *__restrict__ sizes, int64_t cap,
                       uint8_t *__restrict__ out_solved, int32_t *__restrict__ out_last_a,
                       int32_t *__restrict__ out_last_b, uint8_t *__restrict__ out_swapped,
                       int32_t *__restrict__ out_iters, int32_t *__restrict__ out_overflow)
{
    extern __shared__ __align__(16) double sm[];
    const WorldSmem ws = stage_world(w, sm);
    int off = 2 * w.n_box * D + w.n_sph * D + w.n_sph;       // doubles used by stage_world ...
    off += off & 1;
    off += 2 * D;
    off += off & 1;                                            // ... and the scratch starts 16-byte aligned
    Scratch sc;
    sc.way = sm + off;
    sc.vec = sc.way + kChunk * D;
    sc.rd = sc.vec + 3 * D + (D & 1);
    sc.ri = reinterpret_cast<int *>(sc.rd + kThreads / 32);
    sc.flag = sc.ri + kThreads / 32;

    const int64_t q = blockIdx.x;
    double *tree_nodes[2] = {nodes + (2 * q) * cap * D, nodes + (2 * q + 1) * cap * D};
    int32_t *tree_par[2] = {parents + (2 * q) * cap, parents + (2 * q + 1) * cap};
    int size[2] = {1, 1};

    // roots; rrt_connect.py:47: a colliding start or goal fails the query before any iteration
    if (threadIdx.x < D) {
        tree_nodes[0][threadIdx.x] = starts[q * D + threadIdx.x];
        tree_nodes[1][threadIdx.x] = goals[q * D + threadIdx.x];
    }
    if (threadIdx.x == 0) {
        tree_par[0][0] = -1;
        tree_par[1][0] = -1;
        double s[D], g[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            s[d] = starts[q * D + d];
            g[d] = goals[q * D + d];
        }
        sc.flag[3] = (config_collides<D>(w, ws, s) || config_collides<D>(w, ws, g)) ? 1 : 0;
    }
    __threadfence_block();
    __syncthreads();
    const bool blocked = sc.flag[3] != 0;

    bool solved = false, overflow = false, swapped = false;
    int last_a = 0, last_b = 0, it = 0;
    if (!blocked) {
        for (; it < iters; ++it) {
            const double u = swap_u[q * iters + it];
            const bool swap = u < __ddiv_rn((double)size[0], (double)(size[0] + size[1]));
            const int ta = swap ? 1 : 0, tb = 1 - ta;
            if (threadIdx.x < D) sc.vec[2 * D + threadIdx.x] = targets[(q * iters + it) * D + threadIdx.x];
            __syncthreads();
            bool ok1 = false, ok2 = false;
            const int l1 = extend_towards<D>(w, ws, tree_nodes[ta], tree_par[ta], size[ta], cap, swap, wt, metric, step,
                                             sc, ok1, overflow);
            if (overflow) break;
            if (threadIdx.x < D) sc.vec[2 * D + threadIdx.x] = tree_nodes[ta][(int64_t)l1 * D + threadIdx.x];
            __syncthreads();
            const int l2 = extend_towards<D>(w, ws, tree_nodes[tb], tree_par[tb], size[tb], cap, !swap, wt, metric, step,
                                             sc, ok2, overflow);
            if (overflow) break;
            if (ok2) {
                solved = true;
                swapped = swap;
                last_a = l1;
                last_b = l2;
                ++it;
                break;
            }
        }
    }
    if (threadIdx.x == 0) {
        sizes[2 * q] = size[0];
        sizes[2 * q + 1] = size[1];
        out_solved[q] = solved ? 1 : 0;
        out_last_a[q] = last_a;
        out_last_b[q] = last_b;
        out_swapped[q] = swapped ? 1 : 0;
        out_iters[q] = it;
        if (overflow) atomicAdd(out_overflow, 1);
    }
}

template <int D>
int launch_rrt(const mpb_world *world, const double *starts, const double *goals, int64_t B, const Weights &wt,
               int metric, double step, const double *swap_u, const double *targets, int iters, double *nodes,
               int32_t *parents, int32_t *sizes, int64_t cap, uint8_t *out_solved, int32_t *out_last_a,
               int32_t *out_last_b, uint8_t *out_swapped, int32_t *out_iters, int32_t *out_overflow, cudaStream_t st)
{
    const size_t world_b = (world_smem_bytes(world->dev) + 16 + 15) / 16 * 16;   // + the layout's padding
    const size_t smem = world_b + sizeof(double) * ((size_t)kChunk * D + 3 * D + 1 + kThreads / 32) +
                        sizeof(int) * (kThreads / 32 + 8) + 16;
    if (smem > 48 * 1024)
        MPB_CUDA_CHECK(cudaFuncSetAttribute(rrt_connect_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
    rrt_connect_kernel<D><<<(unsigned)B, kThreads, smem, st>>>(world->dev, wt, metric, step, starts, goals, swap_u,
                                                               targets, iters, nodes, parents, sizes, cap, out_solved,
                                                               out_last_a, out_last_b, out_swapped, out_iters,
                                                               out_overflow);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

}   // namespace

MPB_API int mpb_rrt_connect_batch(const mpb_world *world, const double *starts, const double *goals,
                                  int64_t n_queries, const double *weights, int metric, double step,
                                  const double *swap_draws, const double *targets, int iters, double *nodes,
                                  int32_t *parents, int32_t *sizes, int64_t cap, uint8_t *out_solved,
                                  int32_t *out_last_a, int32_t *out_last_b, uint8_t *out_swapped, int32_t *out_iters,
                                  int32_t *out_overflow, void *stream)
{
    MPB_REQUIRE(world != nullptr, "mpb_rrt_connect_batch: world is NULL");
    MPB_REQUIRE_CSPACE(world, "mpb_rrt_connect_batch");
    MPB_REQUIRE(metric >= 0 && metric <= 2, "mpb_rrt_connect_batch: bad metric %d", metric);
    MPB_REQUIRE(n_queries >= 0 && n_queries < 2147483647LL, "mpb_rrt_connect_batch: bad query count");
    MPB_REQUIRE(iters >= 0 && cap >= 1, "mpb_rrt_connect_batch: bad iteration count or capacity");
    MPB_REQUIRE(step > 0., "mpb_rrt_connect_batch: step must be positive");
    if (n_queries == 0) return MPB_OK;
    MPB_REQUIRE(starts && goals && nodes && parents && sizes && out_solved && out_last_a && out_last_b &&
                    out_swapped && out_iters && out_overflow && (iters == 0 || (swap_draws && targets)),
                "mpb_rrt_connect_batch: NULL buffer");
    MPB_REQUIRE(world_smem_bytes(world->dev) <= 150 * 1024,
                "mpb_rrt_connect_batch: obstacle set (%zu bytes) exceeds the shared-memory budget",
                world_smem_bytes(world->dev));
    const int dim = world->dev.dim;
    const Weights wt = make_weights(weights, dim);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_rrt<DD>(world, starts, goals, n_queries, wt, metric, step, swap_draws, targets, iters, nodes, parents, sizes, cap, out_solved, out_last_a, out_last_b, out_swapped, out_iters, out_overflow, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}
