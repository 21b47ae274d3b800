// Circular (wrap-around) joints: the edge discretiser with wrapped differences and positions, the collision
// sweep over those waypoints, and the exact kNN under the "three images per circular coordinate" metric.
//
// Reference arithmetic reproduced step by step (all IEEE fp64, nothing contracted):
//   utils.py:364-369   wrap_interval(v, (lo, hi))     = (v - lo) % (hi - lo) + lo          (float %, sign of divisor)
//   utils.py:381-384   circular_difference(t2, t1, I) = wrap_interval(t2 - t1, (-E/2, +E/2)),  E = hi - lo
//   primitives.py:74-78  get_difference_fn(circular)    = the above per coordinate, plain t2 - t1 where not circular
//   tkinter/samplers.py:113-127  extend_fn: n = int(max_d |diff(q2, q1)_d / step|) + 1;
//                                q <- c * diff(q2, q) + q, c = 1/(n - i); then every circular coordinate wrapped
//   nearest.py:7-20,74-85  query_neighbors: the query is shifted by {-E, 0, +E} in every circular coordinate
//                          (3^c images), each image queries the kd-tree, an index keeps its smallest distance.
//
// These are separate kernels on purpose: the non-circular edge kernels cull obstacles with the edge's bounding
// box, which a wrapped edge does not have, and the fp32 kNN filter's exactness certificate is stated for one
// image.  Circular workloads are correct first; they take the exact paths.
#include <math.h>

#include "mpb_common.cuh"

namespace {

const size_t kMaxWorldSmem = 200 * 1024;   // same budget as the non-circular edge kernels

struct Circ {
    double lo[MPB_MAX_DIM], hi[MPB_MAX_DIM];
    unsigned mask;   // bit d set: coordinate d is circular
};

// Python / numpy float `%`: fmod, then moved to the sign of the divisor (b > 0 here).
__device__ __forceinline__ double py_mod(double a, double b)
{
    double m = fmod(a, b);
    if (m != 0.) {
        if ((b < 0.) != (m < 0.)) m = d_add(m, b);
    } else {
        m = copysign(0., b);
    }
    return m;
}

__device__ __forceinline__ double wrap_interval(double v, double lo, double hi)
{
    return d_add(py_mod(d_sub(v, lo), d_sub(hi, lo)), lo);
}

// circular_difference(t2, t1): the interval is (-E/2, +E/2) with E = hi - lo computed once.
__device__ __forceinline__ double circ_diff(double t2, double t1, double lo, double hi)
{
    const double E = d_sub(hi, lo);
    const double h = __ddiv_rn(E, 2.0);
    return wrap_interval(d_sub(t2, t1), -h, h);
}

template <int D>
__device__ __forceinline__ double diff_d(const Circ &c, int d, double t2, double t1)
{
    return ((c.mask >> d) & 1u) ? circ_diff(t2, t1, c.lo[d], c.hi[d]) : d_sub(t2, t1);
}

template <int D>
__device__ __forceinline__ int waypoint_count_circ(const Circ &c, const double (&a)[D], const double (&b)[D],
                                                   double step)
{
    double m = 0.;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        double v = fabs(__ddiv_rn(diff_d<D>(c, d, b[d], a[d]), step));
        m = (v > m) ? v : m;
    }
    if (!(m < 2147483000.)) m = 2147483000.;
    return (int)m + 1;
}

template <int D>
__device__ __forceinline__ void waypoint_step_circ(const Circ &c, double (&q)[D], const double (&b)[D], int n, int i)
{
    const double f = __ddiv_rn(1.0, (double)(n - i));
#pragma unroll
    for (int d = 0; d < D; ++d) {
        double v = d_add(d_mul(f, diff_d<D>(c, d, b[d], q[d])), q[d]);
        if ((c.mask >> d) & 1u) v = wrap_interval(v, c.lo[d], c.hi[d]);
        q[d] = v;
    }
}

// ---- collision sweep: one thread per edge, every waypoint against every obstacle ------------
template <int D>
__global__ void __launch_bounds__(128)
    edges_check_circ_kernel(WorldDev w, Circ c, const double *__restrict__ A, const double *__restrict__ B,
                            const int64_t *__restrict__ src, const int64_t *__restrict__ dst, int64_t E, double step,
                            int mode, int skip_last, int32_t *__restrict__ out_hit, int32_t *__restrict__ out_n)
{
    extern __shared__ __align__(16) double sm[];
    const WorldSmem s = stage_world(w, sm);
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
        if (src && (src[e] < 0 || dst[e] < 0)) {
            out_hit[e] = -1;
            if (out_n) out_n[e] = 0;
            continue;
        }
        double a[D], b[D], q[D];
        const double *pa = src ? A + src[e] * D : A + e * D;
        const double *pb = src ? A + dst[e] * D : B + e * D;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            a[d] = pa[d];
            b[d] = pb[d];
            q[d] = a[d];
        }
        const int n = waypoint_count_circ<D>(c, a, b, step);
        const int lim = n - (skip_last ? 1 : 0);
        int best = -1;
        for (int i = 0; i < lim; ++i) {
            waypoint_step_circ<D>(c, q, b, n, i);
            bool hit = outside_env<D>(s, q);
            for (int j = 0; j < w.n_box && !hit; ++j) hit = in_box<D>(s, j, q);
            for (int j = 0; j < w.n_sph && !hit; ++j) hit = in_sphere<D>(s, j, q);
            if (hit) {
                best = i;
                if (mode != MPB_EDGE_LAST) break;
            }
        }
        out_hit[e] = best;
        if (out_n) out_n[e] = n;
    }
}

template <int D>
__global__ void __launch_bounds__(256)
    edges_count_circ_kernel(Circ c, const double *__restrict__ A, const double *__restrict__ B, int64_t E, double step,
                            int32_t *__restrict__ out_n)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    double a[D], b[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        a[d] = A[e * D + d];
        b[d] = B[e * D + d];
    }
    out_n[e] = waypoint_count_circ<D>(c, a, b, step);
}

template <int D>
__global__ void __launch_bounds__(128)
    edges_waypoints_circ_kernel(Circ c, const double *__restrict__ A, const double *__restrict__ B, int64_t E,
                                double step, const int64_t *__restrict__ offsets, double *__restrict__ out)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    double a[D], b[D], q[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        a[d] = A[e * D + d];
        b[d] = B[e * D + d];
        q[d] = a[d];
    }
    const int n = waypoint_count_circ<D>(c, a, b, step);
    const int64_t o = offsets[e];
    const int64_t room = offsets[e + 1] - o;
    for (int i = 0; i < n && i < room; ++i) {
        waypoint_step_circ<D>(c, q, b, n, i);
#pragma unroll
        for (int d = 0; d < D; ++d) out[(o + i) * D + d] = q[d];
    }
}

// ---- exact kNN, three images per circular coordinate ---------------------------------------
// The kd-tree's squared distance is a fixed tree of rounded additions over per-coordinate terms
// fl(fl(e_q - e_x)^2); rounded addition is monotone in each argument, so the smallest image distance of a
// sample is that same sum over the per-coordinate SMALLEST terms: no 3^c loop is needed.
template <int D>
__global__ void __launch_bounds__(256)
    knn_circ_kernel(const double *__restrict__ x, int64_t N, const double *__restrict__ qs, int64_t M, Weights wt,
                    Circ c, int k, double lim, int64_t *__restrict__ out_idx, double *__restrict__ out_dist)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t q = warp0; q < M; q += nwarps) {
        double eq[3][D];   // embedded images: x - E, x, x + E per coordinate (nearest.py:7-20, primitives.py:55-57)
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double v = qs[q * D + d];
            const bool circ = (c.mask >> d) & 1u;
            const double E = circ ? d_sub(c.hi[d], c.lo[d]) : 0.;
            eq[0][d] = d_mul(wt.w[d], d_add(v, -E));
            eq[1][d] = d_mul(wt.w[d], d_add(v, 0.));
            eq[2][d] = d_mul(wt.w[d], d_add(v, E));
        }
        double lk = INFINITY, thr = INFINITY;
        int li = -1;
        for (int64_t j0 = 0; j0 < N; j0 += 32) {
            const int64_t j = j0 + lane;
            double S = INFINITY;
            if (j < N) {
                double t[D];
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    const double ex = d_mul(wt.w[d], x[j * D + d]);
                    double d1 = d_sub(eq[1][d], ex);
                    double best = d_mul(d1, d1);
                    if ((c.mask >> d) & 1u) {
                        double d0 = d_sub(eq[0][d], ex), d2 = d_sub(eq[2][d], ex);
                        best = fmin(best, fmin(d_mul(d0, d0), d_mul(d2, d2)));
                    }
                    t[d] = best;
                }
                // cKDTree's summation order: four running sums over blocks of four, then the tail
                double a0 = 0., a1 = 0., a2 = 0., a3 = 0.;
                constexpr int NB = D / 4;
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    a0 = d_add(a0, t[4 * b + 0]);
                    a1 = d_add(a1, t[4 * b + 1]);
                    a2 = d_add(a2, t[4 * b + 2]);
                    a3 = d_add(a3, t[4 * b + 3]);
                }
                S = d_add(d_add(d_add(a0, a1), a2), a3);
#pragma unroll
                for (int d = 4 * NB; d < D; ++d) S = d_add(S, t[d]);
            }
            unsigned m = __ballot_sync(0xffffffffu, S < thr);
            while (m) {
                const int srcl = __ffs(m) - 1;
                m &= m - 1;
                const double key = __shfl_sync(0xffffffffu, S, srcl);
                if (key < thr) {
                    const int pos = __popc(__ballot_sync(0xffffffffu, lk <= key));
                    const double upk = __shfl_up_sync(0xffffffffu, lk, 1);
                    const int upi = __shfl_up_sync(0xffffffffu, li, 1);
                    if (lane > pos) {
                        lk = upk;
                        li = upi;
                    } else if (lane == pos) {
                        lk = key;
                        li = (int)(j0 + srcl);
                    }
                    thr = __shfl_sync(0xffffffffu, lk, 31);
                }
            }
        }
        if (lane < k) {
            const bool keep = (li >= 0) && (lk < lim);
            out_idx[q * k + lane] = keep ? (int64_t)li : -1;
            if (out_dist) out_dist[q * k + lane] = keep ? __dsqrt_rn(lk) : INFINITY;
        }
    }
}

int make_circ(const char *who, int dim, const double *lo, const double *hi, Circ *c)
{
    MPB_REQUIRE(lo && hi, "%s: NULL circular interval arrays", who);
    c->mask = 0;
    for (int d = 0; d < MPB_MAX_DIM; ++d) {
        c->lo[d] = 0.;
        c->hi[d] = 0.;
    }
    for (int d = 0; d < dim; ++d) {
        const bool unbounded = isinf(lo[d]) && isinf(hi[d]) && lo[d] < 0 && hi[d] > 0;
        if (unbounded) continue;
        MPB_REQUIRE(isfinite(lo[d]) && isfinite(hi[d]) && lo[d] < hi[d],
                    "%s: circular interval %d must be finite with lower < upper (or -inf, +inf for none)", who, d);
        c->lo[d] = lo[d];
        c->hi[d] = hi[d];
        c->mask |= 1u << d;
    }
    return MPB_OK;
}

int sm_count()
{
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
}

template <int D>
int launch_edges_circ(const mpb_world *world, const Circ &c, const double *A, const double *B, const int64_t *src,
                      const int64_t *dst, int64_t E, double step, int mode, int skip_last, int32_t *out_hit,
                      int32_t *out_n, cudaStream_t st)
{
    const size_t smem = world_smem_bytes(world->dev) + 16;
    const int threads = 128;
    int64_t blocks = (E + threads - 1) / threads;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (smem > 48 * 1024)
        MPB_CUDA_CHECK(cudaFuncSetAttribute(edges_check_circ_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
    edges_check_circ_kernel<D><<<(unsigned)blocks, threads, smem, st>>>(world->dev, c, A, B, src, dst, E, step, mode,
                                                                        skip_last, out_hit, out_n);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int D>
int launch_count_circ(const Circ &c, const double *q1, const double *q2, int64_t E, double step, int32_t *out_n,
                      cudaStream_t st)
{
    edges_count_circ_kernel<D><<<(unsigned)((E + 255) / 256), 256, 0, st>>>(c, q1, q2, E, step, out_n);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int D>
int launch_waypoints_circ(const Circ &c, const double *q1, const double *q2, int64_t E, double step,
                          const int64_t *offsets, double *out, cudaStream_t st)
{
    edges_waypoints_circ_kernel<D><<<(unsigned)((E + 127) / 128), 128, 0, st>>>(c, q1, q2, E, step, offsets, out);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int D>
int launch_knn_circ(const double *x, int64_t N, const double *qs, int64_t M, const Weights &wt, const Circ &c, int k,
                    double lim, int64_t *out_idx, double *out_dist, cudaStream_t st)
{
    const int threads = 256;
    int64_t blocks = (M * 32 + threads - 1) / threads;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    knn_circ_kernel<D><<<(unsigned)blocks, threads, 0, st>>>(x, N, qs, M, wt, c, k, lim, out_idx, out_dist);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

}   // namespace

MPB_API int mpb_edges_check_circular(const mpb_world *world, const double *a, const double *b, const int64_t *src,
                                     const int64_t *dst, int64_t n_edges, const double *circ_lo,
                                     const double *circ_hi, double step, int mode, int skip_last, int32_t *out_hit,
                                     int32_t *out_n, void *stream)
{
    const char *who = "mpb_edges_check_circular";
    MPB_REQUIRE(world != nullptr, "%s: world is NULL", who);
    MPB_REQUIRE_CSPACE(world, who);
    MPB_REQUIRE(n_edges >= 0, "%s: negative edge count", who);
    MPB_REQUIRE(mode == MPB_EDGE_ANY || mode == MPB_EDGE_FIRST || mode == MPB_EDGE_LAST, "%s: bad mode %d", who, mode);
    MPB_REQUIRE(step > 0., "%s: step must be positive", who);
    MPB_REQUIRE((src == nullptr) == (dst == nullptr), "%s: src and dst go together", who);
    Circ c;
    int rc = make_circ(who, world->dev.dim, circ_lo, circ_hi, &c);
    if (rc != MPB_OK) return rc;
    if (n_edges == 0) return MPB_OK;
    MPB_REQUIRE(a && out_hit && (src || b), "%s: NULL buffer", who);
    MPB_REQUIRE(world_smem_bytes(world->dev) + 16 <= kMaxWorldSmem,
                "%s: obstacle set (%zu bytes) exceeds the %zu-byte shared-memory budget", who,
                world_smem_bytes(world->dev), kMaxWorldSmem);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_edges_circ<DD>(world, c, a, b, src, dst, n_edges, step, mode, skip_last, out_hit, out_n, st)
    MPB_DISPATCH_DIM(world->dev.dim, CALL)
#undef CALL
    mpb_set_error("%s: unsupported dim %d", who, world->dev.dim);
    return MPB_EINVAL;
}

MPB_API int mpb_edges_count_circular(const double *q1, const double *q2, int64_t n_edges, int dim,
                                     const double *circ_lo, const double *circ_hi, double step, int32_t *out_n,
                                     void *stream)
{
    const char *who = "mpb_edges_count_circular";
    MPB_REQUIRE(n_edges >= 0, "%s: negative edge count", who);
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "%s: dim %d outside [1, %d]", who, dim, MPB_MAX_DIM);
    MPB_REQUIRE(step > 0., "%s: step must be positive", who);
    Circ c;
    int rc = make_circ(who, dim, circ_lo, circ_hi, &c);
    if (rc != MPB_OK) return rc;
    if (n_edges == 0) return MPB_OK;
    MPB_REQUIRE(q1 && q2 && out_n, "%s: NULL buffer", who);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_count_circ<DD>(c, q1, q2, n_edges, step, out_n, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}

MPB_API int mpb_edges_waypoints_circular(const double *q1, const double *q2, int64_t n_edges, int dim,
                                         const double *circ_lo, const double *circ_hi, double step,
                                         const int64_t *offsets, double *out, void *stream)
{
    const char *who = "mpb_edges_waypoints_circular";
    MPB_REQUIRE(n_edges >= 0, "%s: negative edge count", who);
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "%s: dim %d outside [1, %d]", who, dim, MPB_MAX_DIM);
    MPB_REQUIRE(step > 0., "%s: step must be positive", who);
    Circ c;
    int rc = make_circ(who, dim, circ_lo, circ_hi, &c);
    if (rc != MPB_OK) return rc;
    if (n_edges == 0) return MPB_OK;
    MPB_REQUIRE(q1 && q2 && offsets && out, "%s: NULL buffer", who);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_waypoints_circ<DD>(c, q1, q2, n_edges, step, offsets, out, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}

MPB_API int mpb_knn_f64_circular(const double *samples, int64_t n_samples, const double *queries, int64_t n_queries,
                                 int dim, const double *weights, const double *circ_lo, const double *circ_hi, int k,
                                 double max_dist, int64_t *out_idx, double *out_dist, void *stream)
{
    const char *who = "mpb_knn_f64_circular";
    MPB_REQUIRE(n_samples >= 0 && n_queries >= 0, "%s: negative count", who);
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "%s: dim %d outside [1, %d]", who, dim, MPB_MAX_DIM);
    MPB_REQUIRE(k >= 1 && k <= MPB_MAX_K, "%s: k %d outside [1, %d]", who, k, MPB_MAX_K);
    MPB_REQUIRE(n_samples <= 0x7fffffffLL, "%s: too many samples", who);
    MPB_REQUIRE(!(max_dist < 0.), "%s: negative max_dist", who);
    Circ c;
    int rc = make_circ(who, dim, circ_lo, circ_hi, &c);
    if (rc != MPB_OK) return rc;
    if (n_queries == 0) return MPB_OK;
    MPB_REQUIRE(queries && out_idx && (samples || n_samples == 0), "%s: NULL buffer", who);
    const Weights wt = make_weights(weights, dim);
    const double lim = max_dist * max_dist;   // keep iff S < fl(r * r)
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_knn_circ<DD>(samples, n_samples, queries, n_queries, wt, c, k, lim, out_idx, out_dist, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}
