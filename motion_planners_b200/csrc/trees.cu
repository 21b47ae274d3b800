// Tree queries of the RRT family: first arg-min nearest node over many resident trees, and
// the fused nearest + radius scan of RRT*.
//
// Replaces the Python scans
//   motion_planners/primitives.py:21   argmin(lambda n: distance_fn(n.config, target), tree)
//   motion_planners/rrt_star.py:117    argmin over all nodes
//   motion_planners/rrt_star.py:128    filter(distance_fn(n.config, new.config) < radius, nodes)
// with motion_planners/utils.py:61-65 argmin semantics: the first index holding the minimum
// of the ROUNDED distances (after the square root), so two different squared sums that
// round to the same distance tie and the lower index wins.
#include <math.h>

#include "mpb_common.cuh"

#define FULL_MASK 0xffffffffu

namespace {

// (distance, index) lexicographic minimum across a warp
__device__ __forceinline__ void warp_argmin(double &d, int &i)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double od = __shfl_xor_sync(FULL_MASK, d, o);
        const int oi = __shfl_xor_sync(FULL_MASK, i, o);
        if (od < d || (od == d && oi < i)) {
            d = od;
            i = oi;
        }
    }
}

// one warp per tree
template <int D>
__global__ void __launch_bounds__(256)
    trees_nearest_kernel(const double *__restrict__ nodes, const int32_t *__restrict__ sizes, int64_t T, int64_t cap,
                         const double *__restrict__ targets, Weights wt, int metric,
                         const uint8_t *__restrict__ active, int32_t *__restrict__ out_idx,
                         double *__restrict__ out_dist)
{
    const int lane = threadIdx.x & 31;
    const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (t >= T) return;
    if (active && !active[t]) {   // skipped trees report "no node"
        if (lane == 0) {
            out_idx[t] = -1;
            if (out_dist) out_dist[t] = INFINITY;
        }
        return;
    }
    double tg[D];
#pragma unroll
    for (int d = 0; d < D; ++d) tg[d] = targets[t * D + d];
    const int n = sizes[t];
    const double *base = nodes + t * cap * D;
    double best = INFINITY;
    int bi = 0x7fffffff;
    for (int j = lane; j < n; j += 32) {
        double a[D];
#pragma unroll
        for (int d = 0; d < D; ++d) a[d] = base[(int64_t)j * D + d];
        const double dd = metric_dist<D>(a, tg, wt, metric);
        if (dd < best) {   // ascending j per lane: strict < keeps the first minimum
            best = dd;
            bi = j;
        }
    }
    warp_argmin(best, bi);
    if (lane == 0) {
        out_idx[t] = (bi == 0x7fffffff) ? -1 : bi;
        if (out_dist) out_dist[t] = best;
    }
}

template <int D>
__global__ void __launch_bounds__(256)
    scan_dist_kernel(const double *__restrict__ nodes, int64_t n, const double *__restrict__ q, Weights wt, int metric,
                     double *__restrict__ out_dist)
{
    double tg[D];
#pragma unroll
    for (int d = 0; d < D; ++d) tg[d] = q[d];
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        double a[D];
#pragma unroll
        for (int d = 0; d < D; ++d) a[d] = nodes[j * D + d];
        out_dist[j] = metric_dist<D>(a, tg, wt, metric);
    }
}

// single block: first arg-min of dist[] and the ascending list of indices with dist < radius
__global__ void __launch_bounds__(1024)
    scan_select_kernel(const double *__restrict__ dist, int64_t n, double radius, int32_t *__restrict__ out_nearest,
                       int32_t *__restrict__ out_count, int32_t *__restrict__ out_list)
{
    __shared__ double s_d[32];
    __shared__ int s_i[32];
    __shared__ int s_cnt[32];
    __shared__ int s_base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double best = INFINITY;
    int bi = 0x7fffffff;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (int64_t c0 = 0; c0 < n; c0 += blockDim.x) {
        const int64_t j = c0 + threadIdx.x;
        const double dd = (j < n) ? dist[j] : INFINITY;
        if (j < n && dd < best) {
            best = dd;
            bi = (int)j;
        }
        const bool sel = (j < n) && (dd < radius);   // strict, rrt_star.py:128
        const unsigned bal = __ballot_sync(FULL_MASK, sel);
        if (lane == 0) s_cnt[warp] = __popc(bal);
        __syncthreads();
        int off = s_base;
        for (int w = 0; w < warp; ++w) off += s_cnt[w];
        if (sel && out_list) out_list[off + __popc(bal & ((1u << lane) - 1))] = (int)j;
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
            for (int w = 0; w < nw; ++w) tot += s_cnt[w];
            s_base += tot;
        }
        __syncthreads();
    }
    warp_argmin(best, bi);
    if (lane == 0) {
        s_d[warp] = best;
        s_i[warp] = bi;
    }
    __syncthreads();
    if (warp == 0) {
        best = (lane < nw) ? s_d[lane] : INFINITY;
        bi = (lane < nw) ? s_i[lane] : 0x7fffffff;
        warp_argmin(best, bi);
        if (lane == 0) {
            if (out_nearest) *out_nearest = (bi == 0x7fffffff) ? -1 : bi;
            if (out_count) *out_count = s_base;
        }
    }
}

// Append the kept part of an extension to a tree (primitives.py:24-27: one TreeNode per safe waypoint,
// chained parent -> child).  One thread per extension: the waypoints from `from` to `to` are regenerated
// with the reference recurrence (tkinter/samplers.py:121-126) and those with generation index in [lo, hi)
// are stored, in generation order (forward extension) or reversed (backward extension, primitives.py:13-16).
template <int D>
__global__ void __launch_bounds__(128)
    trees_append_kernel(double *__restrict__ nodes, int32_t *__restrict__ parents, int32_t *__restrict__ sizes,
                        int64_t cap, const int32_t *__restrict__ tree_ids, const double *__restrict__ from,
                        const double *__restrict__ to, const int32_t *__restrict__ lo, const int32_t *__restrict__ hi,
                        const uint8_t *__restrict__ reverse, const int32_t *__restrict__ parent0, int64_t B,
                        double step, int32_t *__restrict__ out_last, int32_t *__restrict__ out_overflow)
{
    const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int t = tree_ids[b];
    if (t < 0) {
        out_last[b] = -1;
        return;
    }
    double a[D], q2[D], q[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        a[d] = from[b * D + d];
        q2[d] = to[b * D + d];
        q[d] = a[d];
    }
    double m = 0.;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const double v = fabs(__ddiv_rn(d_sub(q2[d], a[d]), step));
        m = (v > m) ? v : m;
    }
    if (!(m < 2147483000.)) m = 2147483000.;
    const int n = (int)m + 1;
    const int l = lo[b], h = min(hi[b], n);
    const int keep = max(h - l, 0);
    const int size = sizes[t];
    if ((int64_t)size + keep > cap) {
        atomicAdd(out_overflow, 1);
        out_last[b] = parent0[b];
        return;
    }
    double *base = nodes + (int64_t)t * cap * D;
    int32_t *par = parents + (int64_t)t * cap;
    const bool rev = reverse[b] != 0;
    for (int i = 0; i < h; ++i) {
        const double c = __ddiv_rn(1.0, (double)(n - i));
#pragma unroll
        for (int d = 0; d < D; ++d) q[d] = d_add(d_mul(c, d_sub(q2[d], q[d])), q[d]);
        if (i >= l) {
            const int pos = size + (rev ? (h - 1 - i) : (i - l));
#pragma unroll
            for (int d = 0; d < D; ++d) base[(int64_t)pos * D + d] = q[d];
            par[pos] = (pos == size) ? parent0[b] : pos - 1;
        }
    }
    sizes[t] = size + keep;
    out_last[b] = keep > 0 ? size + keep - 1 : parent0[b];
}

template <int D>
int launch_append(double *nodes, int32_t *parents, int32_t *sizes, int64_t cap, const int32_t *tree_ids,
                  const double *from, const double *to, const int32_t *lo, const int32_t *hi, const uint8_t *reverse,
                  const int32_t *parent0, int64_t B, double step, int32_t *out_last, int32_t *out_overflow,
                  cudaStream_t st)
{
    trees_append_kernel<D><<<(unsigned)((B + 127) / 128), 128, 0, st>>>(nodes, parents, sizes, cap, tree_ids, from, to,
                                                                        lo, hi, reverse, parent0, B, step, out_last,
                                                                        out_overflow);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int D>
int launch_trees(const double *nodes, const int32_t *sizes, int64_t T, int64_t cap, const double *targets,
                 const Weights &wt, int metric, const uint8_t *active, int32_t *out_idx, double *out_dist,
                 cudaStream_t st)
{
    const int threads = 256;
    const int64_t blocks = (T * 32 + threads - 1) / threads;
    trees_nearest_kernel<D><<<(unsigned)blocks, threads, 0, st>>>(nodes, sizes, T, cap, targets, wt, metric, active,
                                                                  out_idx, out_dist);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int D>
int launch_scan(const double *nodes, int64_t n, const double *q, const Weights &wt, int metric, double radius,
                int32_t *out_nearest, double *out_dist, int32_t *out_count, int32_t *out_list, cudaStream_t st)
{
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    scan_dist_kernel<D><<<(unsigned)blocks, 256, 0, st>>>(nodes, n, q, wt, metric, out_dist);
    MPB_LAUNCH_CHECK();
    scan_select_kernel<<<1, 1024, 0, st>>>(out_dist, n, radius, out_nearest, out_count, out_list);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

}   // namespace

MPB_API int mpb_trees_nearest(const double *nodes, const int32_t *sizes, int64_t n_trees, int64_t cap, int dim,
                              const double *targets, const double *weights, int metric, const uint8_t *active,
                              int32_t *out_idx, double *out_dist, void *stream)
{
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "mpb_trees_nearest: dim %d outside [1, %d]", dim, MPB_MAX_DIM);
    MPB_REQUIRE(metric >= 0 && metric <= 2, "mpb_trees_nearest: bad metric %d", metric);
    MPB_REQUIRE(n_trees >= 0 && cap >= 0, "mpb_trees_nearest: negative size");
    if (n_trees == 0) return MPB_OK;
    MPB_REQUIRE(nodes && sizes && targets && out_idx, "mpb_trees_nearest: NULL buffer");
    MPB_REQUIRE(n_trees < (1LL << 26), "mpb_trees_nearest: too many trees");
    const Weights wt = make_weights(weights, dim);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_trees<DD>(nodes, sizes, n_trees, cap, targets, wt, metric, active, out_idx, out_dist, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}

MPB_API int mpb_scan_nearest_radius(const double *nodes, int64_t n, int dim, const double *q, const double *weights,
                                    int metric, double radius, int32_t *out_nearest, double *out_dist,
                                    int32_t *out_count, int32_t *out_list, void *stream)
{
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "mpb_scan_nearest_radius: dim %d outside [1, %d]", dim, MPB_MAX_DIM);
    MPB_REQUIRE(metric >= 0 && metric <= 2, "mpb_scan_nearest_radius: bad metric %d", metric);
    MPB_REQUIRE(n >= 0 && n < 2147483647LL, "mpb_scan_nearest_radius: bad node count");
    MPB_REQUIRE(q && out_dist && (n == 0 || nodes), "mpb_scan_nearest_radius: NULL buffer");
    const Weights wt = make_weights(weights, dim);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_scan<DD>(nodes, n, q, wt, metric, radius, out_nearest, out_dist, out_count, out_list, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}

MPB_API int mpb_trees_append(double *nodes, int32_t *parents, int32_t *sizes, int64_t n_trees, int64_t cap, int dim,
                             const int32_t *tree_ids, const double *from, const double *to, const int32_t *lo,
                             const int32_t *hi, const uint8_t *reverse, const int32_t *parent0, int64_t n_ext,
                             double step, int32_t *out_last, int32_t *out_overflow, void *stream)
{
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "mpb_trees_append: dim %d outside [1, %d]", dim, MPB_MAX_DIM);
    MPB_REQUIRE(n_trees >= 0 && cap >= 0 && n_ext >= 0, "mpb_trees_append: negative size");
    MPB_REQUIRE(step > 0., "mpb_trees_append: step must be positive");
    if (n_ext == 0) return MPB_OK;
    MPB_REQUIRE(nodes && parents && sizes && tree_ids && from && to && lo && hi && reverse && parent0 && out_last &&
                    out_overflow,
                "mpb_trees_append: NULL buffer");
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_append<DD>(nodes, parents, sizes, cap, tree_ids, from, to, lo, hi, reverse, parent0, n_ext, step, out_last, out_overflow, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}
