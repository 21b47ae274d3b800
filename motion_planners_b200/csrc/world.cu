// Obstacle set, batched collision_fn and the fused extend_fn + collision_fn edge kernel.
//
// Reference arithmetic reproduced here (paths relative to the reference checkout):
//   motion_planners/tkinter/samplers.py:88-98   get_collision_fn
//   motion_planners/tkinter/viewer.py:69-92     contains_box / contains_circle / point_collides
//   motion_planners/tkinter/samplers.py:113-127 get_extend_fn (circular = {}, norm = inf)
// Every fp64 operation that the reference rounds separately is issued through the
// __d*_rn intrinsics so that ptxas can never contract it into an FMA; the sphere test
// uses an explicit FMA chain because the reference's BLAS dot fuses.
#include <stdarg.h>
#include <string.h>

#include <vector>

#include "mpb_common.cuh"

// ------------------------------------------------------------------------------------------
// error plumbing
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

void mpb_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

MPB_API const char *mpb_last_error(void) { return g_err; }
MPB_API int mpb_version(void) { return MPB_VERSION; }

// ------------------------------------------------------------------------------------------
// world handle
// ------------------------------------------------------------------------------------------
MPB_API int mpb_world_create(const double *box_lo, const double *box_hi, int n_box, const double *sph_c,
                             const double *sph_r, int n_sph, int dim, const double *env_lo,
                             const double *env_hi, double buffer, mpb_world **out)
{
    MPB_REQUIRE(out != nullptr, "mpb_world_create: out is NULL");
    *out = nullptr;
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "mpb_world_create: dim %d outside [1, %d]", dim, MPB_MAX_DIM);
    MPB_REQUIRE(n_box >= 0 && n_sph >= 0, "mpb_world_create: negative obstacle count");
    MPB_REQUIRE(n_box == 0 || (box_lo && box_hi), "mpb_world_create: box arrays are NULL");
    MPB_REQUIRE(n_sph == 0 || (sph_c && sph_r), "mpb_world_create: sphere arrays are NULL");
    MPB_REQUIRE(env_lo && env_hi, "mpb_world_create: environment bounds are NULL");

    const size_t nb = (size_t)n_box * dim, ns = (size_t)n_sph * dim;
    std::vector<double> h(2 * nb + ns + (size_t)n_sph + 2 * (size_t)dim);
    double *p = h.data();
    // viewer.py:74-75: lower - buffer*ones, upper + buffer*ones (buffer*1.0 == buffer exactly)
    for (size_t i = 0; i < nb; ++i) p[i] = box_lo[i] - buffer;
    for (size_t i = 0; i < nb; ++i) p[nb + i] = box_hi[i] + buffer;
    for (size_t i = 0; i < ns; ++i) p[2 * nb + i] = sph_c[i];
    for (int i = 0; i < n_sph; ++i) p[2 * nb + ns + i] = sph_r[i] + buffer;   // viewer.py:82
    for (int i = 0; i < dim; ++i) p[2 * nb + ns + n_sph + i] = env_lo[i];
    for (int i = 0; i < dim; ++i) p[2 * nb + ns + n_sph + dim + i] = env_hi[i];

    mpb_world *w = new mpb_world();
    MPB_CUDA_CHECK(cudaGetDevice(&w->device));
    cudaError_t e = cudaMalloc(&w->blob, h.size() * sizeof(double));
    if (e != cudaSuccess) {
        delete w;
        mpb_set_error("mpb_world_create: cudaMalloc failed: %s", cudaGetErrorString(e));
        return MPB_ENOMEM;
    }
    e = cudaMemcpy(w->blob, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(w->blob);
        delete w;
        mpb_set_error("mpb_world_create: cudaMemcpy failed: %s", cudaGetErrorString(e));
        return MPB_ECUDA;
    }
    const double *d = (const double *)w->blob;
    w->dev.dim = dim;
    w->dev.n_box = n_box;
    w->dev.n_sph = n_sph;
    w->dev.box_lo = d;
    w->dev.box_hi = d + nb;
    w->dev.sph_c = d + 2 * nb;
    w->dev.sph_rb = d + 2 * nb + ns;
    w->dev.env_lo = d + 2 * nb + ns + n_sph;
    w->dev.env_hi = d + 2 * nb + ns + n_sph + dim;
    *out = w;
    return MPB_OK;
}

MPB_API int mpb_world_destroy(mpb_world *w)
{
    if (!w) return MPB_OK;
    mpb_arm_destroy_impl(w->arm);
    cudaFree(w->blob);
    delete w;
    return MPB_OK;
}

// ------------------------------------------------------------------------------------------
// collision_fn over points
// ------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256) points_check_kernel(WorldDev w, const double *__restrict__ q, int64_t n,
                                                           uint8_t *__restrict__ out)
{
    extern __shared__ __align__(16) double sm[];
    const WorldSmem s = stage_world(w, sm);
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double p[D];
#pragma unroll
    for (int d = 0; d < D; ++d) p[d] = q[i * D + d];
    bool hit = outside_env<D>(s, p);
    for (int b = 0; b < w.n_box && !hit; ++b) hit = in_box<D>(s, b, p);
    for (int k = 0; k < w.n_sph && !hit; ++k) hit = in_sphere<D>(s, k, p);
    out[i] = hit ? 1 : 0;
}

template <int D>
static int launch_points(const mpb_world *world, const double *q, int64_t n, uint8_t *out, cudaStream_t st)
{
    const size_t smem = world_smem_bytes(world->dev) + 16;
    if (smem > 48 * 1024)
        MPB_CUDA_CHECK(cudaFuncSetAttribute(points_check_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
    const int threads = 256;
    const int64_t blocks = (n + threads - 1) / threads;
    points_check_kernel<D><<<(unsigned)blocks, threads, smem, st>>>(world->dev, q, n, out);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

static const size_t kMaxWorldSmem = 200 * 1024;

MPB_API int mpb_points_check(const mpb_world *world, const double *q, int64_t n, uint8_t *out, void *stream)
{
    MPB_REQUIRE(world != nullptr, "mpb_points_check: world is NULL");
    MPB_REQUIRE(n >= 0, "mpb_points_check: negative count");
    if (n == 0) return MPB_OK;
    MPB_REQUIRE(q && out, "mpb_points_check: NULL buffer");
    if (world->arm) return mpb_arm_points_check_impl(world->arm, q, n, out, (cudaStream_t)stream);
    MPB_REQUIRE(world_smem_bytes(world->dev) + 16 <= kMaxWorldSmem,
                "mpb_points_check: obstacle set (%zu bytes) exceeds the %zu-byte shared-memory budget",
                world_smem_bytes(world->dev), kMaxWorldSmem);
    MPB_REQUIRE((n + 255) / 256 <= 0x7fffffffLL, "mpb_points_check: too many points");
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_points<DD>(world, q, n, out, st)
    MPB_DISPATCH_DIM(world->dev.dim, CALL)
#undef CALL
    mpb_set_error("mpb_points_check: unsupported dim %d", world->dev.dim);
    return MPB_EINVAL;
}

// ------------------------------------------------------------------------------------------
// fused edge kernel
// ------------------------------------------------------------------------------------------
// One thread per directed edge.  The waypoints of an edge exist only in registers: the
// serial recurrence is replayed once per group of <= 64 obstacles whose (slightly inflated)
// bounding boxes overlap the edge's own bounding box, so a short roadmap edge tests a handful
// of obstacles per waypoint instead of all of them.  Culling cannot change a result: every
// waypoint is a convex combination of the previous one and q2 up to < n*2^-50 relative
// rounding, far inside the 1e-9 relative margin (culling is disabled for n > 10^6).
enum { TEST_ENV = 0, TEST_BOX = 1, TEST_SPH = 2 };

// Replays the waypoints of one edge against one obstacle group; updates `best`.
template <int D, int TEST>
__device__ __forceinline__ void sweep(const WorldSmem &s, const double (&a)[D], const double (&b)[D], int n, int lim,
                                      int mode, int base, unsigned long long mask, int &best)
{
    // FIRST/ANY: only indices below the current best can improve it.
    const int stop = (mode != MPB_EDGE_LAST && best >= 0) ? best : lim;
    double q[D];
#pragma unroll
    for (int d = 0; d < D; ++d) q[d] = a[d];
    for (int i = 0; i < stop; ++i) {
        waypoint_step<D>(q, b, n, i);
        if (mode == MPB_EDGE_LAST && i <= best) continue;
        bool hit = false;
        if (TEST == TEST_ENV) {
            hit = outside_env<D>(s, q);
        } else {
            unsigned long long m = mask;
            while (m && !hit) {
                const int j = __ffsll((long long)m) - 1;
                m &= m - 1;
                hit = (TEST == TEST_BOX) ? in_box<D>(s, base + j, q) : in_sphere<D>(s, base + j, q);
            }
        }
        if (hit) {
            best = i;
            if (mode != MPB_EDGE_LAST) return;
        }
    }
}

template <int D, bool INDEXED>
__global__ void __launch_bounds__(128)
    edges_check_kernel(WorldDev w, const double *__restrict__ A, const double *__restrict__ B,
                       const int64_t *__restrict__ src, const int64_t *__restrict__ dst, int64_t E, double step,
                       int mode, int skip_last, int32_t *__restrict__ out_hit, int32_t *__restrict__ out_n)
{
    extern __shared__ __align__(16) double sm[];
    const WorldSmem s = stage_world(w, sm);
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
    if (INDEXED && (src[e] < 0 || dst[e] < 0)) {   // "no neighbour" slots of a kNN matrix
        out_hit[e] = -1;
        if (out_n) out_n[e] = 0;
        continue;
    }
    double a[D], b[D];
    {
        const double *pa = INDEXED ? A + src[e] * D : A + e * D;
        const double *pb = INDEXED ? A + dst[e] * D : B + e * D;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            a[d] = pa[d];
            b[d] = pb[d];
        }
    }
    const int n = waypoint_count<D>(a, b, step);
    const int lim = n - (skip_last ? 1 : 0);
    const bool cull = n <= 1000000;
    int best = -1;

    // inflated bounding box of the edge, recomputed on the fly to save registers
    auto elo = [&](int d) {
        double lo = fmin(a[d], b[d]), mx = fmax(fabs(a[d]), fabs(b[d]));
        return lo - 1e-9 * (1.0 + mx);
    };
    auto ehi = [&](int d) {
        double hi = fmax(a[d], b[d]), mx = fmax(fabs(a[d]), fabs(b[d]));
        return hi + 1e-9 * (1.0 + mx);
    };

    // environment: skipped when the whole edge box lies inside it
    {
        bool inside = cull;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            double2 ev = s.env[d];
            inside = inside && (ev.x <= elo(d)) && (ehi(d) <= ev.y);
        }
        if (!inside) sweep<D, TEST_ENV>(s, a, b, n, lim, mode, 0, 0ull, best);
    }
    // boxes, 64 at a time
    for (int base = 0; base < w.n_box; base += 64) {
        if (mode == MPB_EDGE_ANY && best >= 0) break;
        const int cnt = min(64, w.n_box - base);
        unsigned long long mask = 0ull;
        for (int j = 0; j < cnt; ++j) {
            const double2 *bb = s.box + (base + j) * D;
            bool ov = true;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                double2 bv = bb[d];
                ov = ov && (bv.x <= ehi(d)) && (elo(d) <= bv.y);
            }
            if (ov || !cull) mask |= 1ull << j;
        }
        if (mask) sweep<D, TEST_BOX>(s, a, b, n, lim, mode, base, mask, best);
    }
    // spheres, 64 at a time (culled by their bounding boxes, inflated by the same margin)
    for (int base = 0; base < w.n_sph; base += 64) {
        if (mode == MPB_EDGE_ANY && best >= 0) break;
        const int cnt = min(64, w.n_sph - base);
        unsigned long long mask = 0ull;
        for (int j = 0; j < cnt; ++j) {
            const double *c = s.sph_c + (base + j) * D;
            const double r = s.sph_rb[base + j] * (1.0 + 1e-9) + 1e-12;
            bool ov = true;
#pragma unroll
            for (int d = 0; d < D; ++d) ov = ov && (c[d] - r <= ehi(d)) && (elo(d) <= c[d] + r);
            if (ov || !cull) mask |= 1ull << j;
        }
        if (mask) sweep<D, TEST_SPH>(s, a, b, n, lim, mode, base, mask, best);
    }
    out_hit[e] = best;
    if (out_n) out_n[e] = n;
    }   // edge loop
}

template <int D>
static int launch_edges(const mpb_world *world, const double *A, const double *B, const int64_t *src,
                        const int64_t *dst, int64_t E, double step, int mode, int skip_last, int32_t *out_hit,
                        int32_t *out_n, cudaStream_t st)
{
    const size_t smem = world_smem_bytes(world->dev) + 16;
    const int threads = 128;
    int64_t blocks = (E + threads - 1) / threads;
    // persistent grid: the obstacle set is staged once per block, edges are strided over it
    int sms = 148;
    {
        int dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    if (blocks > (int64_t)sms * 16) blocks = (int64_t)sms * 16;
    if (src) {
        if (smem > 48 * 1024)
            MPB_CUDA_CHECK(cudaFuncSetAttribute(edges_check_kernel<D, true>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        edges_check_kernel<D, true><<<(unsigned)blocks, threads, smem, st>>>(world->dev, A, B, src, dst, E, step, mode,
                                                                             skip_last, out_hit, out_n);
    } else {
        if (smem > 48 * 1024)
            MPB_CUDA_CHECK(cudaFuncSetAttribute(edges_check_kernel<D, false>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        edges_check_kernel<D, false><<<(unsigned)blocks, threads, smem, st>>>(world->dev, A, B, src, dst, E, step,
                                                                              mode, skip_last, out_hit, out_n);
    }
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

static int edges_common(const char *who, const mpb_world *world, const double *A, const double *B,
                        const int64_t *src, const int64_t *dst, int64_t E, double step, int mode, int skip_last,
                        int32_t *out_hit, int32_t *out_n, void *stream)
{
    MPB_REQUIRE(world != nullptr, "%s: world is NULL", who);
    MPB_REQUIRE(E >= 0, "%s: negative edge count", who);
    MPB_REQUIRE(mode == MPB_EDGE_ANY || mode == MPB_EDGE_FIRST || mode == MPB_EDGE_LAST, "%s: bad mode %d", who, mode);
    MPB_REQUIRE(step > 0., "%s: step must be positive", who);
    if (E == 0) return MPB_OK;
    MPB_REQUIRE(A && out_hit, "%s: NULL buffer", who);
    MPB_REQUIRE((E + 127) / 128 <= 0x7fffffffLL, "%s: too many edges", who);
    if (world->arm)
        return mpb_arm_edges_check_impl(world->arm, A, B, src, dst, 0, 0, E, step, mode, skip_last, out_hit, out_n,
                                        (cudaStream_t)stream);
    MPB_REQUIRE(world_smem_bytes(world->dev) + 16 <= kMaxWorldSmem,
                "%s: obstacle set (%zu bytes) exceeds the %zu-byte shared-memory budget", who,
                world_smem_bytes(world->dev), kMaxWorldSmem);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_edges<DD>(world, A, B, src, dst, E, step, mode, skip_last, out_hit, out_n, st)
    MPB_DISPATCH_DIM(world->dev.dim, CALL)
#undef CALL
    mpb_set_error("%s: unsupported dim %d", who, world->dev.dim);
    return MPB_EINVAL;
}

MPB_API int mpb_edges_check(const mpb_world *world, const double *samples, int64_t n_samples, const int64_t *src,
                            const int64_t *dst, int64_t n_edges, double step, int mode, int skip_last,
                            int32_t *out_hit, int32_t *out_n, void *stream)
{
    MPB_REQUIRE(n_edges == 0 || (src && dst), "mpb_edges_check: NULL index arrays");
    MPB_REQUIRE(n_samples >= 0, "mpb_edges_check: negative sample count");
    return edges_common("mpb_edges_check", world, samples, nullptr, src, dst, n_edges, step, mode, skip_last, out_hit,
                        out_n, stream);
}

MPB_API int mpb_edges_check_pairs(const mpb_world *world, const double *q1, const double *q2, int64_t n_edges,
                                  double step, int mode, int skip_last, int32_t *out_hit, int32_t *out_n,
                                  void *stream)
{
    MPB_REQUIRE(n_edges == 0 || (q1 && q2), "mpb_edges_check_pairs: NULL endpoint arrays");
    return edges_common("mpb_edges_check_pairs", world, q1, q2, nullptr, nullptr, n_edges, step, mode, skip_last,
                        out_hit, out_n, stream);
}

// ------------------------------------------------------------------------------------------
// the same verdicts for a neighbour matrix: one thread per ROW (a sample and its k+1 neighbours)
// ------------------------------------------------------------------------------------------
// All edges of a row leave the same sample and are short, so the obstacles are culled ONCE against the
// bounding box of the sample and all its neighbours; each edge then replays its waypoints against the
// few survivors.  Entry (r, j) equals mpb_edges_check on the pair (row_offset + r, nbr[r, j]).
template <int D>
__global__ void __launch_bounds__(128)
    edges_check_rows_kernel(WorldDev w, const double *__restrict__ X, const int64_t *__restrict__ nbr, int64_t M,
                            int kk, int64_t row_offset, const int64_t *__restrict__ row_ids, double step, int mode,
                            int skip_last,
                            int32_t *__restrict__ out_hit, int32_t *__restrict__ out_n)
{
    extern __shared__ __align__(16) double sm[];
    const WorldSmem s = stage_world(w, sm);
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < M; r += (int64_t)gridDim.x * blockDim.x) {
        double a[D];
        const double *pa = X + (row_ids ? row_ids[r] : row_offset + r) * D;
#pragma unroll
        for (int d = 0; d < D; ++d) a[d] = pa[d];
        // bounding box of the row, inflated like the per-edge boxes of edges_check_kernel
        double lo[D], hi[D];
#pragma unroll
        for (int d = 0; d < D; ++d) lo[d] = hi[d] = a[d];
        bool cull = true;
        const int64_t *row = nbr + r * kk;
        int32_t *hit_row = out_hit + r * kk;
        for (int j = 0; j < kk; ++j) {
            const int64_t v = row[j];
            hit_row[j] = -1;
            if (v < 0) {
                if (out_n) out_n[r * kk + j] = 0;
                continue;
            }
            double b[D];
            const double *pb = X + v * D;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                b[d] = pb[d];
                lo[d] = fmin(lo[d], b[d]);
                hi[d] = fmax(hi[d], b[d]);
            }
            const int n = waypoint_count<D>(a, b, step);
            if (out_n) out_n[r * kk + j] = n;
            cull = cull && (n <= 1000000);
        }
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double m = 1e-9 * (1.0 + fmax(fabs(lo[d]), fabs(hi[d])));
            lo[d] -= m;
            hi[d] += m;
        }
        // one pass per obstacle group: 0 = environment, then boxes, then spheres, 64 at a time
        bool inside = cull;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double2 ev = s.env[d];
            inside = inside && (ev.x <= lo[d]) && (hi[d] <= ev.y);
        }
        const int n_groups = 1 + (w.n_box + 63) / 64 + (w.n_sph + 63) / 64;
        for (int g = 0; g < n_groups; ++g) {
            int kind, base = 0;
            unsigned long long mask = 0ull;
            if (g == 0) {
                kind = TEST_ENV;
                if (inside) continue;
            } else if (g <= (w.n_box + 63) / 64) {
                kind = TEST_BOX;
                base = (g - 1) * 64;
                const int cnt = min(64, w.n_box - base);
                for (int j = 0; j < cnt; ++j) {
                    const double2 *bb = s.box + (base + j) * D;
                    bool ov = true;
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        const double2 bv = bb[d];
                        ov = ov && (bv.x <= hi[d]) && (lo[d] <= bv.y);
                    }
                    if (ov || !cull) mask |= 1ull << j;
                }
                if (!mask) continue;
            } else {
                kind = TEST_SPH;
                base = (g - 1 - (w.n_box + 63) / 64) * 64;
                const int cnt = min(64, w.n_sph - base);
                for (int j = 0; j < cnt; ++j) {
                    const double *c = s.sph_c + (base + j) * D;
                    const double rr = s.sph_rb[base + j] * (1.0 + 1e-9) + 1e-12;
                    bool ov = true;
#pragma unroll
                    for (int d = 0; d < D; ++d) ov = ov && (c[d] - rr <= hi[d]) && (lo[d] <= c[d] + rr);
                    if (ov || !cull) mask |= 1ull << j;
                }
                if (!mask) continue;
            }
            for (int j = 0; j < kk; ++j) {
                const int64_t v = row[j];
                if (v < 0) continue;
                int best = hit_row[j];
                if (mode == MPB_EDGE_ANY && best >= 0) continue;
                double b[D];
                const double *pb = X + v * D;
#pragma unroll
                for (int d = 0; d < D; ++d) b[d] = pb[d];
                const int n = waypoint_count<D>(a, b, step);
                const int lim = n - (skip_last ? 1 : 0);
                if (kind == TEST_ENV)
                    sweep<D, TEST_ENV>(s, a, b, n, lim, mode, 0, 0ull, best);
                else if (kind == TEST_BOX)
                    sweep<D, TEST_BOX>(s, a, b, n, lim, mode, base, mask, best);
                else
                    sweep<D, TEST_SPH>(s, a, b, n, lim, mode, base, mask, best);
                hit_row[j] = best;
            }
        }
    }
}

template <int D>
static int launch_edges_rows(const mpb_world *world, const double *X, const int64_t *nbr, int64_t M, int kk,
                             int64_t row_offset, const int64_t *row_ids, double step, int mode, int skip_last,
                             int32_t *out_hit,
                             int32_t *out_n, cudaStream_t st)
{
    const size_t smem = world_smem_bytes(world->dev) + 16;
    const int threads = 128;
    int64_t blocks = (M + threads - 1) / threads;
    int sms = 148;
    {
        int dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    if (blocks > (int64_t)sms * 16) blocks = (int64_t)sms * 16;
    if (smem > 48 * 1024)
        MPB_CUDA_CHECK(cudaFuncSetAttribute(edges_check_rows_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
    edges_check_rows_kernel<D><<<(unsigned)blocks, threads, smem, st>>>(world->dev, X, nbr, M, kk, row_offset, row_ids, step,
                                                                        mode, skip_last, out_hit, out_n);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

static int edges_knn_common(const mpb_world *world, const double *samples, int64_t n_samples, const int64_t *nbr,
                            int64_t n_rows, int row_len, int64_t row_offset, const int64_t *row_ids, double step,
                            int mode, int skip_last, int32_t *out_hit, int32_t *out_n, void *stream)
{
    MPB_REQUIRE(world != nullptr, "mpb_edges_check_knn: world is NULL");
    MPB_REQUIRE(n_rows >= 0 && row_len >= 1 && row_offset >= 0 && (row_ids || row_offset + n_rows <= n_samples),
                "mpb_edges_check_knn: rows [%lld, %lld) outside the %lld samples", (long long)row_offset,
                (long long)(row_offset + n_rows), (long long)n_samples);
    MPB_REQUIRE(mode == MPB_EDGE_ANY || mode == MPB_EDGE_FIRST || mode == MPB_EDGE_LAST,
                "mpb_edges_check_knn: bad mode %d", mode);
    MPB_REQUIRE(step > 0., "mpb_edges_check_knn: step must be positive");
    if (n_rows == 0) return MPB_OK;
    MPB_REQUIRE(samples && nbr && out_hit, "mpb_edges_check_knn: NULL buffer");
    if (world->arm)
        return mpb_arm_edges_check_impl(world->arm, samples, nullptr, row_ids, nbr, row_len, row_offset,
                                        n_rows * (int64_t)row_len, step, mode, skip_last, out_hit, out_n,
                                        (cudaStream_t)stream);
    MPB_REQUIRE(world_smem_bytes(world->dev) + 16 <= kMaxWorldSmem,
                "mpb_edges_check_knn: obstacle set (%zu bytes) exceeds the %zu-byte shared-memory budget",
                world_smem_bytes(world->dev), kMaxWorldSmem);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_edges_rows<DD>(world, samples, nbr, n_rows, row_len, row_offset, row_ids, step, mode, skip_last, out_hit, out_n, st)
    MPB_DISPATCH_DIM(world->dev.dim, CALL)
#undef CALL
    return MPB_EINVAL;
}

MPB_API int mpb_edges_check_knn(const mpb_world *world, const double *samples, int64_t n_samples, const int64_t *nbr,
                                int64_t n_rows, int row_len, int64_t row_offset, double step, int mode, int skip_last,
                                int32_t *out_hit, int32_t *out_n, void *stream)
{
    return edges_knn_common(world, samples, n_samples, nbr, n_rows, row_len, row_offset, nullptr, step, mode, skip_last,
                            out_hit, out_n, stream);
}

MPB_API int mpb_edges_check_knn_rows(const mpb_world *world, const double *samples, int64_t n_samples,
                                     const int64_t *row_ids, const int64_t *nbr, int64_t n_rows, int row_len, double step,
                                     int mode, int skip_last, int32_t *out_hit, int32_t *out_n, void *stream)
{
    MPB_REQUIRE(n_rows == 0 || row_ids, "mpb_edges_check_knn_rows: row_ids is NULL");
    return edges_knn_common(world, samples, n_samples, nbr, n_rows, row_len, 0, row_ids, step, mode, skip_last, out_hit,
                            out_n, stream);
}

// ------------------------------------------------------------------------------------------
// extend_fn output (the waypoints themselves)
// ------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(128)
    edges_waypoints_kernel(const double *__restrict__ A, const double *__restrict__ B, int64_t E, double step,
                           const int64_t *__restrict__ offsets, double *__restrict__ out)
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
    const int n = waypoint_count<D>(a, b, step);
    const int64_t o = offsets[e];
    const int64_t room = offsets[e + 1] - o;
    for (int i = 0; i < n && i < room; ++i) {
        waypoint_step<D>(q, b, n, i);
#pragma unroll
        for (int d = 0; d < D; ++d) out[(o + i) * D + d] = q[d];
    }
}

template <int D>
__global__ void __launch_bounds__(256)
    edges_count_kernel(const double *__restrict__ A, const double *__restrict__ B, int64_t E, double step,
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
    out_n[e] = waypoint_count<D>(a, b, step);
}

template <int D>
static int launch_count(const double *q1, const double *q2, int64_t E, double step, int32_t *out_n, cudaStream_t st)
{
    edges_count_kernel<D><<<(unsigned)((E + 255) / 256), 256, 0, st>>>(q1, q2, E, step, out_n);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

MPB_API int mpb_edges_count(const double *q1, const double *q2, int64_t n_edges, int dim, double step, int32_t *out_n,
                            void *stream)
{
    MPB_REQUIRE(n_edges >= 0, "mpb_edges_count: negative edge count");
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "mpb_edges_count: dim %d outside [1, %d]", dim, MPB_MAX_DIM);
    MPB_REQUIRE(step > 0., "mpb_edges_count: step must be positive");
    if (n_edges == 0) return MPB_OK;
    MPB_REQUIRE(q1 && q2 && out_n, "mpb_edges_count: NULL buffer");
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_count<DD>(q1, q2, n_edges, step, out_n, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}

template <int D>
static int launch_waypoints(const double *q1, const double *q2, int64_t E, double step, const int64_t *offsets,
                            double *out, cudaStream_t st)
{
    const int threads = 128;
    const int64_t blocks = (E + threads - 1) / threads;
    edges_waypoints_kernel<D><<<(unsigned)blocks, threads, 0, st>>>(q1, q2, E, step, offsets, out);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

MPB_API int mpb_edges_waypoints(const double *q1, const double *q2, int64_t n_edges, int dim, double step,
                                const int64_t *offsets, double *out, void *stream)
{
    MPB_REQUIRE(n_edges >= 0, "mpb_edges_waypoints: negative edge count");
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "mpb_edges_waypoints: dim %d outside [1, %d]", dim, MPB_MAX_DIM);
    MPB_REQUIRE(step > 0., "mpb_edges_waypoints: step must be positive");
    if (n_edges == 0) return MPB_OK;
    MPB_REQUIRE(q1 && q2 && offsets && out, "mpb_edges_waypoints: NULL buffer");
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) launch_waypoints<DD>(q1, q2, n_edges, step, offsets, out, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}
