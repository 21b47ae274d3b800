// Shared helpers for the mpb200 kernels: error reporting, exact-rounding fp64 primitives,
// mbarrier / bulk-copy PTX wrappers (sm_100a).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mpb200.h"

#define MPB_API extern "C" __attribute__((visibility("default")))

void mpb_set_error(const char *fmt, ...);

#define MPB_CUDA_CHECK(expr)                                                                   \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            mpb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,    \
                          __LINE__);                                                           \
            return MPB_ECUDA;                                                                  \
        }                                                                                      \
    } while (0)

#define MPB_REQUIRE(cond, ...)                                                                 \
    do {                                                                                       \
        if (!(cond)) {                                                                         \
            mpb_set_error(__VA_ARGS__);                                                        \
            return MPB_EINVAL;                                                                 \
        }                                                                                      \
    } while (0)

#define MPB_LAUNCH_CHECK()                                                                     \
    do {                                                                                       \
        cudaError_t _e = cudaGetLastError();                                                   \
        if (_e != cudaSuccess) {                                                               \
            mpb_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),         \
                          __FILE__, __LINE__);                                                 \
            return MPB_ECUDA;                                                                  \
        }                                                                                      \
    } while (0)

// Device-side view of an obstacle set.  Bounds are already inflated.
struct WorldDev {
    int dim, n_box, n_sph;
    const double *box_lo;   // [n_box, dim]  lower - buffer
    const double *box_hi;   // [n_box, dim]  upper + buffer
    const double *sph_c;    // [n_sph, dim]
    const double *sph_rb;   // [n_sph]       radius + buffer
    const double *env_lo;   // [dim]
    const double *env_hi;   // [dim]
};

struct mpb_arm_host;   // arm.cu: a serial arm approximated by spheres among workspace boxes

struct mpb_world {
    WorldDev dev;
    void *blob;    // single device allocation backing every array above
    int device;
    mpb_arm_host *arm;   // non-NULL: the collision model is the arm's (dev then only carries dim = joints)
};

// arm.cu
int mpb_arm_points_check_impl(const mpb_arm_host *arm, const double *q, int64_t n, uint8_t *out, cudaStream_t st);
int mpb_arm_edges_check_impl(const mpb_arm_host *arm, const double *Ap, const double *Bp, const int64_t *src,
                             const int64_t *dst, int kk, int64_t row_offset, int64_t E, double step, int mode,
                             int skip_last, int32_t *out_hit, int32_t *out_n, cudaStream_t st);
void mpb_arm_destroy_impl(mpb_arm_host *arm);

// entry points that only know configuration-space boxes and spheres
#define MPB_REQUIRE_CSPACE(world, who) \
    MPB_REQUIRE((world) == nullptr || (world)->arm == nullptr, "%s: not available for an arm world", who)

// ---- exactly rounded fp64 steps (never contracted into FMA) -------------------------------
__device__ __forceinline__ double d_sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double d_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double d_mul(double a, double b) { return __dmul_rn(a, b); }

// cKDTree's squared distance between embedded points held in registers.
template <int D>
__device__ __forceinline__ double sqdist_kdtree(const double (&u)[D], const double (&v)[D])
{
    double a0 = 0., a1 = 0., a2 = 0., a3 = 0.;
    constexpr int NB = D / 4;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        double d0 = d_sub(u[4 * b + 0], v[4 * b + 0]);
        double d1 = d_sub(u[4 * b + 1], v[4 * b + 1]);
        double d2 = d_sub(u[4 * b + 2], v[4 * b + 2]);
        double d3 = d_sub(u[4 * b + 3], v[4 * b + 3]);
        a0 = d_add(a0, d_mul(d0, d0));
        a1 = d_add(a1, d_mul(d1, d1));
        a2 = d_add(a2, d_mul(d2, d2));
        a3 = d_add(a3, d_mul(d3, d3));
    }
    double s = d_add(d_add(d_add(a0, a1), a2), a3);
#pragma unroll
    for (int i = 4 * NB; i < D; ++i) {
        double d = d_sub(u[i], v[i]);
        s = d_add(s, d_mul(d, d));
    }
    return s;
}

// ---- weights and the reference's three distance roundings --------------------------------
struct Weights {
    double w[MPB_MAX_DIM];
};

inline Weights make_weights(const double *weights, int D)
{
    Weights w;
    for (int d = 0; d < MPB_MAX_DIM; ++d) w.w[d] = (weights && d < D) ? weights[d] : 1.0;
    return w;
}

// The reference's three roundings of the weighted Euclidean distance.
template <int D>
__device__ __forceinline__ double metric_dist(const double (&a)[D], const double (&b)[D], const Weights &wt, int metric)
{
    if (metric == MPB_METRIC_KDTREE) {
        double ea[D], eb[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            ea[d] = d_mul(wt.w[d], a[d]);
            eb[d] = d_mul(wt.w[d], b[d]);
        }
        return __dsqrt_rn(sqdist_kdtree<D>(ea, eb));
    }
    double s = 0.;
    if (metric == MPB_METRIC_PRIMITIVES) {
        // primitives.py:60-62: norm(w*q2 - w*q1), norm = sqrt(dot) with a fused sequential dot
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double e = d_sub(d_mul(wt.w[d], b[d]), d_mul(wt.w[d], a[d]));
            s = __fma_rn(e, e, s);
        }
    } else {
        // tkinter/samplers.py:13-15: diff = q2 - q1; sqrt(dot(weights, diff*diff))
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double df = d_sub(b[d], a[d]);
            s = __fma_rn(wt.w[d], d_mul(df, df), s);
        }
    }
    return __dsqrt_rn(s);
}


// ---- packed fp32 pairs (Blackwell FFMA2: one issue slot, two fused multiply-adds) ----------------
typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t pack_f32x2(float lo, float hi)
{
    f32x2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack_f32x2(f32x2_t v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2_t fma_f32x2(f32x2_t a, f32x2_t b, f32x2_t c)
{
    f32x2_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// ---- mbarrier + 1-D bulk async copy (TMA engine, no tensor map) ---------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a barrier that never completes traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins) {
        if (spins > (1u << 26)) __trap();
    }
}
// The same for a whole (converged) warp: the loop condition is a vote, so the wait introduces no divergence
// and the compiler may keep treating the code after it as warp-uniform.
__device__ __forceinline__ void mbar_wait_warp(uint64_t *bar, uint32_t parity)
{
    for (uint32_t spins = 0; !__all_sync(0xffffffffu, mbar_try_wait(bar, parity)); ++spins) {
        if (spins > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void bulk_copy_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ------------------------------------------------------------------------------------------
// world staged in shared memory
// ------------------------------------------------------------------------------------------
// Layout (doubles): box bounds interleaved {lo, hi} per (box, dim); sphere centres; sphere
// radii; environment {lo, hi} per dim.
inline size_t world_smem_bytes(const WorldDev &w)
{
    return sizeof(double) * (2 * (size_t)w.n_box * w.dim + (size_t)w.n_sph * (w.dim + 1) + 2 * (size_t)w.dim);
}

struct WorldSmem {
    const double2 *box;   // [n_box * dim] {lo, hi}
    const double *sph_c;  // [n_sph * dim]
    const double *sph_rb; // [n_sph]
    const double2 *env;   // [dim] {lo, hi}
};

__device__ __forceinline__ WorldSmem stage_world(const WorldDev &w, double *sm)
{
    const int nbd = w.n_box * w.dim, nsd = w.n_sph * w.dim;
    double2 *box = reinterpret_cast<double2 *>(sm);
    double *sc = sm + 2 * nbd;
    double *sr = sc + nsd;
    int env_off = 2 * nbd + nsd + w.n_sph;
    env_off += env_off & 1;   // keep the {lo, hi} pairs 16-byte aligned
    double2 *env = reinterpret_cast<double2 *>(sm + env_off);
    for (int i = threadIdx.x; i < nbd; i += blockDim.x) box[i] = make_double2(w.box_lo[i], w.box_hi[i]);
    for (int i = threadIdx.x; i < nsd; i += blockDim.x) sc[i] = w.sph_c[i];
    for (int i = threadIdx.x; i < w.n_sph; i += blockDim.x) sr[i] = w.sph_rb[i];
    for (int i = threadIdx.x; i < w.dim; i += blockDim.x) env[i] = make_double2(w.env_lo[i], w.env_hi[i]);
    __syncthreads();
    WorldSmem s = {box, sc, sr, env};
    return s;
}

template <int D>
__device__ __forceinline__ bool outside_env(const WorldSmem &s, const double (&q)[D])
{
    bool in = true;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        double2 e = s.env[d];
        in = in && (e.x <= q[d]) && (q[d] <= e.y);
    }
    return !in;
}

template <int D>
__device__ __forceinline__ bool in_box(const WorldSmem &s, int b, const double (&q)[D])
{
    const double2 *bb = s.box + b * D;
    bool in = true;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        double2 e = bb[d];
        in = in && (e.x <= q[d]) && (q[d] <= e.y);
    }
    return in;
}

// ||q - c||_2 <= r + buffer with numpy's norm: sqrt of a sequentially fused dot (viewer.py:80-82)
template <int D>
__device__ __forceinline__ bool in_sphere(const WorldSmem &s, int k, const double (&q)[D])
{
    const double *c = s.sph_c + k * D;
    double acc = 0.;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        double x = d_sub(q[d], c[d]);
        acc = __fma_rn(x, x, acc);
    }
    return __dsqrt_rn(acc) <= s.sph_rb[k];
}


// ---- the edge discretiser (tkinter/samplers.py:113-127, circular = {}, norm = inf) ---------------
template <int D>
__device__ __forceinline__ int waypoint_count(const double (&a)[D], const double (&b)[D], double step)
{
    double m = 0.;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        double v = fabs(__ddiv_rn(d_sub(b[d], a[d]), step));   // |(q2-q1)/step|, samplers.py:119
        m = (v > m) ? v : m;                                    // norm(ord=inf) = max |.|
    }
    if (!(m < 2147483000.)) m = 2147483000.;
    return (int)m + 1;                                          // int() truncates; steps + 1
}

template <int D>
__device__ __forceinline__ void waypoint_step(double (&q)[D], const double (&b)[D], int n, int i)
{
    const double c = __ddiv_rn(1.0, (double)(n - i));   // 1. / (num_steps - i)
#pragma unroll
    for (int d = 0; d < D; ++d) q[d] = d_add(d_mul(c, d_sub(b[d], q[d])), q[d]);
}


// Runtime dimension -> template instantiation.  CALL(D) must be an int-returning expression.
#define MPB_DISPATCH_DIM(dim, CALL)                                                            \
    switch (dim) {                                                                             \
        case 1: return CALL(1);                                                                \
        case 2: return CALL(2);                                                                \
        case 3: return CALL(3);                                                                \
        case 4: return CALL(4);                                                                \
        case 5: return CALL(5);                                                                \
        case 6: return CALL(6);                                                                \
        case 7: return CALL(7);                                                                \
        case 8: return CALL(8);                                                                \
        case 9: return CALL(9);                                                                \
        case 10: return CALL(10);                                                              \
        case 11: return CALL(11);                                                              \
        case 12: return CALL(12);                                                              \
        case 13: return CALL(13);                                                              \
        case 14: return CALL(14);                                                              \
        case 15: return CALL(15);                                                              \
        case 16: return CALL(16);                                                              \
        default: break;                                                                        \
    }
