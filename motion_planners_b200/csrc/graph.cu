// Shortest paths over a device-resident kNN roadmap (the graph search either side of check_path:
// motion_planners/lazy_prm.py:41-58 dijkstra, :70-98 wastar_search).
//
// The roadmap is the neighbour matrix the kNN entry points produce (row r = sample row_offset + r, entries =
// neighbour indices with their distances) plus the verdicts of the vertex and edge kernels.  An edge is
// usable when its verdict is "free" and both end points are free; like the reference's REVERSIBLE roadmap
// (lazy_prm.py:141-147) it can be travelled in both directions.
// Distances are relaxed Bellman-Ford style, whole-graph sweeps with atomicMin on the bit pattern of the
// non-negative doubles (order preserving); every sweep relaxes both directions of every listed edge, so no
// reverse adjacency is needed.  The result is the exact fixed point d[v] = min_u fl(d[u] + w(u, v)).
#include <math.h>

#include "mpb_common.cuh"

namespace {

__device__ __forceinline__ unsigned long long dbits(double v) { return (unsigned long long)__double_as_longlong(v); }

__global__ void __launch_bounds__(256)
    roadmap_relax_kernel(const int64_t *__restrict__ nbr, const double *__restrict__ w, const int32_t *__restrict__ hit,
                         const uint8_t *__restrict__ vhit, int64_t n_rows, int kk, int64_t row_offset, double *dist,
                         int32_t *__restrict__ changed)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_rows * kk) return;
    const int64_t r = e / kk;
    const int64_t u = row_offset + r, v = nbr[e];
    if (v < 0 || v == u || hit[e] >= 0) return;
    if (vhit && (vhit[u] || vhit[v])) return;
    const double wt = w[e];
    unsigned long long *du = reinterpret_cast<unsigned long long *>(dist + u);
    unsigned long long *dv = reinterpret_cast<unsigned long long *>(dist + v);
    const double a = dist[u], b = dist[v];
    bool moved = false;
    if (a < INFINITY) {
        const double cand = __dadd_rn(a, wt);
        if (cand < b) moved |= atomicMin(dv, dbits(cand)) > dbits(cand);
    }
    if (b < INFINITY) {
        const double cand = __dadd_rn(b, wt);
        if (cand < a) moved |= atomicMin(du, dbits(cand)) > dbits(cand);
    }
    if (moved) *changed = 1;
}

// parent[v] = a neighbour u with fl(d[u] + w) == d[v] (the smallest such u over the listed edges)
__global__ void __launch_bounds__(256)
    roadmap_parents_kernel(const int64_t *__restrict__ nbr, const double *__restrict__ w,
                           const int32_t *__restrict__ hit, const uint8_t *__restrict__ vhit, int64_t n_rows, int kk,
                           int64_t row_offset, const double *__restrict__ dist, int64_t source, long long *parent)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_rows * kk) return;
    const int64_t r = e / kk;
    const int64_t u = row_offset + r, v = nbr[e];
    if (v < 0 || v == u || hit[e] >= 0) return;
    if (vhit && (vhit[u] || vhit[v])) return;
    const double wt = w[e], a = dist[u], b = dist[v];
    if (a < INFINITY && __dadd_rn(a, wt) == b && v != source) atomicMin(reinterpret_cast<unsigned long long *>(parent + v), (unsigned long long)u);
    if (b < INFINITY && __dadd_rn(b, wt) == a && u != source) atomicMin(reinterpret_cast<unsigned long long *>(parent + u), (unsigned long long)v);
}

// cost of every listed edge under one of the reference's distance procedures (mpb_common.cuh metric_dist)
template <int D>
__global__ void __launch_bounds__(256)
    edge_costs_kernel(const double *__restrict__ x, const int64_t *__restrict__ nbr, int64_t n_rows, int kk,
                      int64_t row_offset, Weights wt, int metric, double *__restrict__ out)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_rows * kk) return;
    const int64_t u = row_offset + e / kk, v = nbr[e];
    if (v < 0) {
        out[e] = INFINITY;
        return;
    }
    double a[D], b[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        a[d] = x[u * D + d];
        b[d] = x[v * D + d];
    }
    out[e] = metric_dist<D>(a, b, wt, metric);
}

template <int D>
int launch_edge_costs(const double *x, const int64_t *nbr, int64_t n_rows, int kk, int64_t row_offset, const Weights &wt,
                      int metric, double *out, cudaStream_t st)
{
    const int64_t total = n_rows * kk;
    edge_costs_kernel<D><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(x, nbr, n_rows, kk, row_offset, wt, metric, out);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

}   // namespace

MPB_API int mpb_edge_costs(const double *samples, int64_t n_samples, int dim, const double *weights, int metric,
                           const int64_t *nbr, int64_t n_rows, int row_len, int64_t row_offset, double *out_cost,
                           void *stream)
{
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "mpb_edge_costs: dim %d outside [1, %d]", dim, MPB_MAX_DIM);
    MPB_REQUIRE(metric >= 0 && metric <= 2, "mpb_edge_costs: bad metric %d", metric);
    MPB_REQUIRE(n_rows >= 0 && row_len >= 1 && row_offset >= 0 && row_offset + n_rows <= n_samples, "mpb_edge_costs: bad rows");
    if (n_rows == 0) return MPB_OK;
    MPB_REQUIRE(samples && nbr && out_cost, "mpb_edge_costs: NULL buffer");
    const Weights wt = make_weights(weights, dim);
#define CALL(DD) launch_edge_costs<DD>(samples, nbr, n_rows, row_len, row_offset, wt, metric, out_cost, (cudaStream_t)stream)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}

MPB_API int mpb_roadmap_relax(const int64_t *nbr, const double *nbr_dist, const int32_t *edge_hit,
                              const uint8_t *vertex_hit, int64_t n_rows, int row_len, int64_t row_offset,
                              double *dist, int32_t *changed, int sweeps, void *stream)
{
    MPB_REQUIRE(n_rows >= 0 && row_len >= 1 && row_offset >= 0 && sweeps >= 1, "mpb_roadmap_relax: bad size");
    if (n_rows == 0) return MPB_OK;
    MPB_REQUIRE(nbr && nbr_dist && edge_hit && dist && changed, "mpb_roadmap_relax: NULL buffer");
    const int64_t total = n_rows * row_len;
    MPB_REQUIRE((total + 255) / 256 < 2147483647LL, "mpb_roadmap_relax: too many edges");
    for (int s = 0; s < sweeps; ++s) {
        roadmap_relax_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
            nbr, nbr_dist, edge_hit, vertex_hit, n_rows, row_len, row_offset, dist, changed);
        MPB_LAUNCH_CHECK();
    }
    return MPB_OK;
}

MPB_API int mpb_roadmap_parents(const int64_t *nbr, const double *nbr_dist, const int32_t *edge_hit,
                                const uint8_t *vertex_hit, int64_t n_rows, int row_len, int64_t row_offset,
                                const double *dist, int64_t source, int64_t *parent, void *stream)
{
    MPB_REQUIRE(n_rows >= 0 && row_len >= 1 && row_offset >= 0, "mpb_roadmap_parents: bad size");
    if (n_rows == 0) return MPB_OK;
    MPB_REQUIRE(nbr && nbr_dist && edge_hit && dist && parent, "mpb_roadmap_parents: NULL buffer");
    const int64_t total = n_rows * row_len;
    roadmap_parents_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        nbr, nbr_dist, edge_hit, vertex_hit, n_rows, row_len, row_offset, dist, source, (long long *)parent);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}
