// collision_fn and fused extend_fn + collision_fn for a sphere-approximated serial arm among workspace boxes
// (BASELINE.json configs[3]).
//
// The reference has no robot model; its collision_fn is a caller-supplied callable (README.md:52-59).  What is
// reproduced from it: the closed environment test on the configuration (tkinter/samplers.py:92), the obstacle
// test `norm(...) <= radius + buffer` (tkinter/viewer.py:80-82, norm = sqrt of a sequentially fused dot) and the
// waypoint recurrence of tkinter/samplers.py:113-127 in joint space.  The model itself (chain, spheres, the
// fixed-polynomial sincos) is defined in oracle/arm_oracle.c, which this file follows operation for operation.
//
// The workspace boxes are staged in shared memory once per block; a 32^3 grid over the workspace lists for every
// cell the boxes a sphere centred there could touch (global memory, read through L1 / L2).  The chain is walked
// joint by joint and the spheres of a joint are tested as soon as its frame is known, so a colliding waypoint
// usually costs a fraction of the forward kinematics.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "arm.cuh"

struct mpb_arm_host {
    ArmDev dev;
    void *blob;
    cudaMemPool_t pool;   // stream-ordered scratch (the edge walk's work cursor); keeps its pages between calls
};

#ifndef MPB_ARM_REFILL_CHUNK
#define MPB_ARM_REFILL_CHUNK 128
#endif
// block shape of the refill kernel, measured on the 1M-sample build (edge phase): 256 threads, 2 blocks per SM (95
// registers) 8.5 ms; 128 x 5 (93 registers, 20 warps per SM) 7.7; 256 x 3 (80 registers, 16 B spilled) 7.7; 128 x 8
// (64 registers, 88 B spilled) 7.7; 64 x 12 8.1 -- the kernel waits on fp64 latency, more warps hide more of it
#ifndef MPB_ARM_REFILL_THREADS
#define MPB_ARM_REFILL_THREADS 128
#endif
#ifndef MPB_ARM_REFILL_MINB
#define MPB_ARM_REFILL_MINB 5
#endif

namespace {

template <int NJ>
__global__ void __launch_bounds__(256) arm_points_kernel(ArmDev A, const double *__restrict__ q, int64_t n,
                                                         uint8_t *__restrict__ out)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const ArmSmem S = arm_stage(A, sm);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double p[NJ];
#pragma unroll
        for (int d = 0; d < NJ; ++d) p[d] = q[i * NJ + d];
        out[i] = arm_collides<NJ>(A, S, p, nullptr) ? 1 : 0;
    }
}

template <int NJ>
__global__ void __launch_bounds__(256) arm_fk_kernel(ArmDev A, const double *__restrict__ q, int64_t n,
                                                     double *__restrict__ out)
{
    const ArmSmem S = {nullptr, nullptr, nullptr};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double p[NJ];
#pragma unroll
        for (int d = 0; d < NJ; ++d) p[d] = q[i * NJ + d];
        arm_collides<NJ>(A, S, p, out + i * 3 * A.n_sph);
    }
}

// Edges: explicit endpoints (Bp != NULL), index pairs (src/dst), or the rows of a neighbour matrix (kk > 0: edge e
// runs from sample row_offset + e / kk to sample dst[e]).  Modes and skip_last as mpb_edges_check.
//
// (Used for mode LAST and small batches; modes ANY / FIRST on large batches go through arm_edges_refill_kernel
// below, see launch_arm_edges.)
// Work is scheduled per LANE and per JOINT.  Edges differ wildly in cost -- most colliding ones end at their first
// waypoints, often at the first joints, free ones walk ten waypoints of seven joints and sixteen spheres -- and
// with one thread per edge 3.5 of a warp's 32 lanes were busy on average (ncu, profiles/r2_summary.md).  Here a
// lane owns a small state machine (edge, waypoint, joint, frame); every trip of the warp's loop advances each lane
// by ONE joint (turn, the joint's spheres, link), and a lane whose edge is finished takes the next edge of its
// warp's share at once.  (Splitting the trip further, into separately voted turn and sphere phases, was measured
// slower: 24.3 against 21.3 ms at 1M samples.)
template <int NJ>
__global__ void __launch_bounds__(256)
    arm_edges_kernel(ArmDev A, const double *__restrict__ Ap, const double *__restrict__ Bp,
                     const int64_t *__restrict__ src, const int64_t *__restrict__ dst, int kk, int64_t row_offset,
                     int64_t E, double step, int mode, int skip_last, int32_t *__restrict__ out_hit,
                     int32_t *__restrict__ out_n)
{
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ ArmTables T;
    arm_fill_tables(A, &T);
    const ArmSmem S = arm_stage(A, sm);                 // ends with __syncthreads()
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    // this warp's share of the edges: every `warps`-th chunk of kChunk consecutive edges.  (Contiguous shares were
    // badly balanced when the rows come in spatial order: whole regions collide, whole regions are free.)
    constexpr int64_t kChunk = 64;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int64_t chunk = wid;                                // this warp's current chunk
    int64_t cursor = chunk * kChunk;                    // next edge of that chunk

    bool active = false;
    int64_t e = 0;
    double q[NJ], b[NJ];
    int n = 0, lim = 0, i = 0, ji = 0, best = -1;
    ArmFrame F;
    for (;;) {
        // ---- lanes without an edge take the next ones of the share
        const unsigned need = __ballot_sync(0xffffffffu, !active);
        if (need) {
            int64_t stop = (chunk + 1) * kChunk < E ? (chunk + 1) * kChunk : E;
            if (cursor >= stop && chunk * kChunk < E) {   // on to this warp's next chunk
                chunk += warps;
                cursor = chunk * kChunk;
                stop = (chunk + 1) * kChunk < E ? (chunk + 1) * kChunk : E;
            }
            const int64_t mine = cursor + __popc(need & lt);
            cursor = cursor + __popc(need) < stop ? cursor + __popc(need) : (cursor > stop ? cursor : stop);
            if (!active && mine < stop) {
                e = mine;
                int64_t ia = e, ib = e;
                const double *pb = Bp;
                if (!Bp) {
                    ia = kk > 0 ? (src ? src[e / kk] : row_offset + e / kk) : src[e];
                    ib = dst[e];
                    pb = Ap;
                }
                if (ia < 0 || ib < 0) {
                    out_hit[e] = -1;
                    if (out_n) out_n[e] = 0;
                } else {
#pragma unroll
                    for (int d = 0; d < NJ; ++d) {
                        q[d] = Ap[ia * NJ + d];
                        b[d] = pb[ib * NJ + d];
                    }
                    n = waypoint_count<NJ>(q, b, step);
                    lim = n - (skip_last ? 1 : 0);
                    i = 0;
                    ji = 0;
                    best = -1;
                    if (lim > i) {
                        active = true;
                    } else {
                        out_hit[e] = -1;
                        if (out_n) out_n[e] = n;
                    }
                }
            }
            if (__ballot_sync(0xffffffffu, active) == 0) {
                if (chunk * kChunk >= E) break;
                continue;
            }
        }
        if (!active) continue;
        // ---- one joint of the lane's current waypoint: turn, the joint's spheres, the link
        bool hit = false;
        if (ji == 0) {
            waypoint_step<NJ>(q, b, n, i);              // samplers.py:113-127: q is the running waypoint
            bool in = true;
#pragma unroll
            for (int d = 0; d < NJ; ++d) in = in && (A.env_lo[d] <= q[d]) && (q[d] <= A.env_hi[d]);
            hit = !in;
            F.X0 = 1.; F.X1 = 0.; F.X2 = 0.;
            F.Y0 = 0.; F.Y1 = 1.; F.Y2 = 0.;
            F.Z0 = 0.; F.Z1 = 0.; F.Z2 = 1.;
            F.p0 = A.base[0]; F.p1 = A.base[1]; F.p2 = A.base[2];
        }
        if (!hit) {
            double qi = q[0];
#pragma unroll
            for (int d = 1; d < NJ; ++d) qi = (d == ji) ? q[d] : qi;
            arm_turn(T, ji, qi, F);
            const int je = T.sph_first[ji + 1];
            for (int j = T.sph_first[ji]; j < je && !hit; ++j) hit = arm_sphere(A, T, S, j, F);
            arm_link(T, ji, F);
            ++ji;
        }
        if (hit || ji == NJ) {
            bool done = false;
            if (hit) {
                best = i;
                done = mode != MPB_EDGE_LAST;
            }
            ++i;
            ji = 0;
            if (done || i >= lim) {
                out_hit[e] = best;
                if (out_n) out_n[e] = n;
                active = false;
            }
        }
    }
}

// Modes ANY and FIRST on large batches: one thread per edge at a time, with lanes that refill.  A lane whose edge
// has ended (collision, or its last waypoint) takes the next edge of its warp's share at once instead of idling
// until the longest edge of the warp is done; a trip of the warp's loop is one waypoint for every lane (whole
// chain, joints unrolled: the configuration stays in registers, unlike arm_edges_kernel's joint-granular trips).
// Warps draw chunks of kChunk consecutive edges from a shared cursor (work_cursor, zero at launch); without one
// (NULL) a warp's share is every `warps`-th chunk, which leaves a tail: 9.0 against 8.6 ms on the 1M-sample build,
// whatever the chunk size from 32 to 256.
template <int NJ>
__global__ void __launch_bounds__(MPB_ARM_REFILL_THREADS, MPB_ARM_REFILL_MINB)
    arm_edges_refill_kernel(ArmDev A, const double *__restrict__ Ap, const double *__restrict__ Bp,
                            const int64_t *__restrict__ src, const int64_t *__restrict__ dst, int kk, int64_t row_offset,
                            int64_t E, double step, int skip_last, int32_t *__restrict__ out_hit,
                            int32_t *__restrict__ out_n, unsigned long long *__restrict__ work_cursor)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const ArmSmem S = arm_stage(A, sm);
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    constexpr int64_t kChunk = MPB_ARM_REFILL_CHUNK;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    int64_t chunk = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int64_t cursor = work_cursor ? 0 : chunk * kChunk;
    int64_t stop = work_cursor ? 0 : (cursor + kChunk < E ? cursor + kChunk : E);
    if (work_cursor && E == 0) return;
    bool active = false;
    int64_t e = 0;
    double q[NJ], b[NJ];
    int n = 0, lim = 0, i = 0;
    for (;;) {
        const unsigned need = __ballot_sync(0xffffffffu, !active);
        if (need) {
            if (cursor >= stop && cursor < E) {          // on to this warp's next chunk
                if (work_cursor) {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(work_cursor, (unsigned long long)kChunk);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    cursor = (int64_t)base < E ? (int64_t)base : E;
                } else {
                    chunk += warps;
                    cursor = chunk * kChunk < E ? chunk * kChunk : E;
                }
                stop = cursor + kChunk < E ? cursor + kChunk : E;
            }
            const int64_t mine = cursor + __popc(need & lt);
            if (!active && mine < stop) {
                e = mine;
                int64_t ia = e, ib = e;
                const double *pb = Bp;
                if (!Bp) {
                    ia = kk > 0 ? (src ? src[e / kk] : row_offset + e / kk) : src[e];
                    ib = dst[e];
                    pb = Ap;
                }
                if (ia < 0 || ib < 0) {
                    out_hit[e] = -1;
                    if (out_n) out_n[e] = 0;
                } else {
#pragma unroll
                    for (int d = 0; d < NJ; ++d) {
                        q[d] = Ap[ia * NJ + d];
                        b[d] = pb[ib * NJ + d];
                    }
                    n = waypoint_count<NJ>(q, b, step);
                    lim = n - (skip_last ? 1 : 0);
                    if (out_n) out_n[e] = n;
                    i = 0;
                    active = lim > 0;
                    if (!active) out_hit[e] = -1;
                }
            }
            cursor = cursor + __popc(need) < stop ? cursor + __popc(need) : stop;
            if (__ballot_sync(0xffffffffu, active) == 0) {
                if (cursor >= E) break;
                continue;
            }
        }
        if (active) {
            waypoint_step<NJ>(q, b, n, i);              // samplers.py:113-127: q is the running waypoint
            const bool hit = arm_collides<NJ>(A, S, q, nullptr);
            if (hit) {
                out_hit[e] = i;
                active = false;
            } else if (++i >= lim) {
                out_hit[e] = -1;
                active = false;
            }
        }
    }
}

int sm_count_arm()
{
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
}

#define ARM_DISPATCH(nj, CALL)   \
    switch (nj) {                \
        case 1: return CALL(1);  \
        case 2: return CALL(2);  \
        case 3: return CALL(3);  \
        case 4: return CALL(4);  \
        case 5: return CALL(5);  \
        case 6: return CALL(6);  \
        case 7: return CALL(7);  \
        case 8: return CALL(8);  \
        default: break;          \
    }

template <int NJ>
int launch_arm_points(const ArmDev &A, const double *q, int64_t n, uint8_t *out, cudaStream_t st)
{
    const size_t smem = arm_smem_bytes(A);
    MPB_CUDA_CHECK(cudaFuncSetAttribute(arm_points_kernel<NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)sm_count_arm() * 8;
    if (blocks > cap) blocks = cap;
    arm_points_kernel<NJ><<<(unsigned)blocks, 256, smem, st>>>(A, q, n, out);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int NJ>
int launch_arm_fk(const ArmDev &A, const double *q, int64_t n, double *out, cudaStream_t st)
{
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)sm_count_arm() * 8;
    if (blocks > cap) blocks = cap;
    arm_fk_kernel<NJ><<<(unsigned)blocks, 256, 0, st>>>(A, q, n, out);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int NJ>
int launch_arm_edges(const ArmDev &A, const double *Ap, const double *Bp, const int64_t *src, const int64_t *dst, int kk,
                     int64_t row_offset, int64_t E, double step, int mode, int skip_last, int32_t *out_hit,
                     int32_t *out_n, cudaMemPool_t pool, cudaStream_t st)
{
    const size_t smem = arm_smem_bytes(A);
    MPB_CUDA_CHECK(cudaFuncSetAttribute(arm_edges_kernel<NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, arm_edges_kernel<NJ>, 256, smem));
    if (per_sm < 1) per_sm = 1;
    int64_t blocks = (E + 255) / 256;
    const int64_t cap = (int64_t)sm_count_arm() * per_sm;
    if (blocks > cap) blocks = cap;
    // Edge phase of the 1M-sample build (10M edges, modes ANY / FIRST): the joint-granular kernel 21.2 ms; first
    // waypoints in one pass and the survivors, one thread per edge, in another 10.7; the same with refilling lanes
    // in the second pass 8.8; refilling lanes from the first waypoint on, in one pass, 8.6.  The work cursor is 256
    // stream-ordered bytes from the world's own pool (its release threshold keeps them mapped: from the device's
    // default pool, which gives its pages back at every synchronisation, the earlier 40 MB survivor list cost each
    // build 4 to 6 ms); without it the kernel falls back to fixed shares.
    // Mode LAST has to look at every waypoint of every edge, and small batches do not fill the refill kernel's grid.
    if (mode != MPB_EDGE_LAST && E >= 65536) {
        MPB_CUDA_CHECK(cudaFuncSetAttribute(arm_edges_refill_kernel<NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
        int per_sm2 = 1;
        MPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm2, arm_edges_refill_kernel<NJ>,
                                                                     MPB_ARM_REFILL_THREADS, smem));
        void *scratch = nullptr;
        if (!pool || cudaMallocFromPoolAsync(&scratch, 256, pool, st) != cudaSuccess) {
            (void)cudaGetLastError();
            scratch = nullptr;
        }
        if (scratch) MPB_CUDA_CHECK(cudaMemsetAsync(scratch, 0, sizeof(unsigned long long), st));
        arm_edges_refill_kernel<NJ><<<(unsigned)(sm_count_arm() * (per_sm2 < 1 ? 1 : per_sm2)), MPB_ARM_REFILL_THREADS, smem, st>>>(
            A, Ap, Bp, src, dst, kk, row_offset, E, step, skip_last, out_hit, out_n,
            reinterpret_cast<unsigned long long *>(scratch));
        MPB_LAUNCH_CHECK();
        if (scratch) MPB_CUDA_CHECK(cudaFreeAsync(scratch, st));
        return MPB_OK;
    }
    arm_edges_kernel<NJ><<<(unsigned)blocks, 256, smem, st>>>(A, Ap, Bp, src, dst, kk, row_offset, E, step, mode,
                                                             skip_last, out_hit, out_n);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

}   // namespace

// ---- entry points used by world.cu's dispatch -------------------------------------------------------------
int mpb_arm_points_check_impl(const mpb_arm_host *arm, const double *q, int64_t n, uint8_t *out, cudaStream_t st)
{
#define CALL(NJ) launch_arm_points<NJ>(arm->dev, q, n, out, st)
    ARM_DISPATCH(arm->dev.n_joints, CALL)
#undef CALL
    return MPB_EINVAL;
}

int mpb_arm_edges_check_impl(const mpb_arm_host *arm, const double *Ap, const double *Bp, const int64_t *src,
                             const int64_t *dst, int kk, int64_t row_offset, int64_t E, double step, int mode,
                             int skip_last, int32_t *out_hit, int32_t *out_n, cudaStream_t st)
{
#define CALL(NJ) \
    launch_arm_edges<NJ>(arm->dev, Ap, Bp, src, dst, kk, row_offset, E, step, mode, skip_last, out_hit, out_n, \
                         arm->pool, st)
    ARM_DISPATCH(arm->dev.n_joints, CALL)
#undef CALL
    return MPB_EINVAL;
}

void mpb_arm_destroy_impl(mpb_arm_host *arm)
{
    if (!arm) return;
    if (arm->pool) cudaMemPoolDestroy(arm->pool);
    cudaFree(arm->blob);
    delete arm;
}

MPB_API int mpb_arm_fk(const mpb_world *world, const double *q, int64_t n, double *out_centres, void *stream)
{
    MPB_REQUIRE(world && world->arm, "mpb_arm_fk: not an arm world");
    MPB_REQUIRE(n >= 0, "mpb_arm_fk: negative count");
    if (n == 0) return MPB_OK;
    MPB_REQUIRE(q && out_centres, "mpb_arm_fk: NULL buffer");
    const mpb_arm_host *arm = world->arm;
#define CALL(NJ) launch_arm_fk<NJ>(arm->dev, q, n, out_centres, (cudaStream_t)stream)
    ARM_DISPATCH(arm->dev.n_joints, CALL)
#undef CALL
    return MPB_EINVAL;
}

MPB_API int mpb_arm_world_create(const mpb_arm_model *m, const double *box_lo, const double *box_hi, int n_box,
                                 const double *env_lo, const double *env_hi, double buffer, mpb_world **out)
{
    MPB_REQUIRE(out != nullptr, "mpb_arm_world_create: out is NULL");
    *out = nullptr;
    MPB_REQUIRE(m != nullptr && env_lo && env_hi, "mpb_arm_world_create: NULL model or bounds");
    MPB_REQUIRE(m->n_joints >= 1 && m->n_joints <= kArmMaxJoints, "mpb_arm_world_create: %d joints outside [1, %d]",
                m->n_joints, kArmMaxJoints);
    MPB_REQUIRE(m->n_sph >= 1 && m->n_sph <= kArmMaxSpheres, "mpb_arm_world_create: %d spheres outside [1, %d]",
                m->n_sph, kArmMaxSpheres);
    MPB_REQUIRE(n_box >= 0 && n_box <= 65535, "mpb_arm_world_create: box count outside [0, 65535]");
    MPB_REQUIRE(n_box == 0 || (box_lo && box_hi), "mpb_arm_world_create: box arrays are NULL");
    for (int i = 0; i < m->n_joints; ++i)
        MPB_REQUIRE(m->axis[i] == 1 || m->axis[i] == 2, "mpb_arm_world_create: joint %d: axis must be 1 (y) or 2 (z)", i);
    for (int j = 0; j < m->n_sph; ++j)
        MPB_REQUIRE(m->sph_link[j] >= 0 && m->sph_link[j] < m->n_joints && m->sph_r[j] >= 0.,
                    "mpb_arm_world_create: sphere %d: bad joint or radius", j);

    mpb_arm_host *arm = new mpb_arm_host();
    ArmDev &A = arm->dev;
    memset(&A, 0, sizeof(A));
    A.n_joints = m->n_joints;
    A.n_sph = m->n_sph;
    A.n_box = n_box;
    for (int i = 0; i < m->n_joints; ++i) {
        A.axis[i] = m->axis[i];
        A.theta_lo[i] = m->theta_lo[i];
        A.theta_span[i] = m->theta_span[i];
        A.a[i] = m->a[i];
        A.d[i] = m->d[i];
        A.env_lo[i] = env_lo[i];
        A.env_hi[i] = env_hi[i];
    }
    for (int k = 0; k < 3; ++k) A.base[k] = m->base[k];
    // spheres grouped by the joint that carries them (the verdict is an `any`, so the order is free)
    int pos = 0;
    double rmax = 0.;
    for (int i = 0; i < m->n_joints; ++i) {
        A.sph_first[i] = pos;
        for (int j = 0; j < m->n_sph; ++j) {
            if (m->sph_link[j] != i) continue;
            A.sph_orig[pos] = j;
            A.sph_u[pos] = m->sph_u[j];
            A.sph_w[pos] = m->sph_w[j];
            const double rb = m->sph_r[j] + buffer;          // viewer.py:82 radius + buffer
            A.sph_rb[pos] = rb;
            A.sph_lo2[pos] = rb * rb * (1.0 - 1e-12);
            A.sph_hi2[pos] = rb * rb * (1.0 + 1e-12);
            rmax = std::max(rmax, rb);
            ++pos;
        }
    }
    for (int i = m->n_joints; i <= kArmMaxJoints; ++i) A.sph_first[i] = pos;

    // the grid: it spans every box inflated by the largest radius; cell lists hold every box whose inflated
    // bounds (plus a margin far above any rounding of the cell index) overlap the cell
    const int G = kArmGrid;
    double glo[3] = {0., 0., 0.}, ghi[3] = {1., 1., 1.};
    const double infl = rmax * (1.0 + 1e-6) + 1e-9;
    for (int k = 0; k < 3; ++k) {
        double lo = INFINITY, hi = -INFINITY;
        for (int b = 0; b < n_box; ++b) {
            lo = std::min(lo, box_lo[3 * b + k] - infl);
            hi = std::max(hi, box_hi[3 * b + k] + infl);
        }
        if (!(lo < hi)) {
            lo = 0.;
            hi = 1.;
        }
        const double pad = 1e-6 * (hi - lo) + 1e-9;
        glo[k] = lo - pad;
        ghi[k] = hi + pad;
        A.grid_lo[k] = glo[k];
        A.grid_inv[k] = (double)G / (ghi[k] - glo[k]);
    }
    std::vector<std::vector<unsigned short>> cells((size_t)G * G * G);
    for (int b = 0; b < n_box; ++b) {
        int c0[3], c1[3];
        for (int k = 0; k < 3; ++k) {
            const double h = (ghi[k] - glo[k]) / G;
            const double margin = infl + 1e-6 * h;
            c0[k] = std::max(0, std::min(G - 1, (int)floor((box_lo[3 * b + k] - margin - glo[k]) / h)));
            c1[k] = std::max(0, std::min(G - 1, (int)floor((box_hi[3 * b + k] + margin - glo[k]) / h)));
        }
        for (int x = c0[0]; x <= c1[0]; ++x)
            for (int y = c0[1]; y <= c1[1]; ++y)
                for (int z = c0[2]; z <= c1[2]; ++z) cells[((size_t)x * G + y) * G + z].push_back((unsigned short)b);
    }
    std::vector<int> cell_start((size_t)G * G * G + 1);
    std::vector<unsigned short> items;
    for (size_t c = 0; c < cells.size(); ++c) {
        cell_start[c] = (int)items.size();
        items.insert(items.end(), cells[c].begin(), cells[c].end());
    }
    cell_start[cells.size()] = (int)items.size();
    A.n_items = (int)items.size();
    if (arm_smem_bytes(A) > 200 * 1024) {
        delete arm;
        mpb_set_error("mpb_arm_world_create: %d boxes with %d grid entries exceed the shared-memory budget", n_box,
                      A.n_items);
        return MPB_EINVAL;
    }
    std::vector<double> boxes((size_t)n_box * 6);
    for (int b = 0; b < n_box; ++b)
        for (int k = 0; k < 3; ++k) {
            boxes[(size_t)b * 6 + 2 * k] = box_lo[3 * b + k];
            boxes[(size_t)b * 6 + 2 * k + 1] = box_hi[3 * b + k];
        }
    const size_t box_bytes = (boxes.size() * sizeof(double) + 255) & ~(size_t)255;
    const size_t cs_bytes = (cell_start.size() * sizeof(int) + 255) & ~(size_t)255;
    const size_t it_bytes = (items.size() * sizeof(unsigned short) + 255) & ~(size_t)255;
    cudaError_t e = cudaMalloc(&arm->blob, box_bytes + cs_bytes + it_bytes + 256);
    if (e != cudaSuccess) {
        delete arm;
        mpb_set_error("mpb_arm_world_create: cudaMalloc failed: %s", cudaGetErrorString(e));
        return MPB_ENOMEM;
    }
    char *base = reinterpret_cast<char *>(arm->blob);
    e = cudaMemcpy(base, boxes.data(), boxes.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(base + box_bytes, cell_start.data(), cell_start.size() * sizeof(int), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !items.empty())
        e = cudaMemcpy(base + box_bytes + cs_bytes, items.data(), items.size() * sizeof(unsigned short), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(arm->blob);
        delete arm;
        mpb_set_error("mpb_arm_world_create: cudaMemcpy failed: %s", cudaGetErrorString(e));
        return MPB_ECUDA;
    }
    A.boxes = reinterpret_cast<const double *>(base);
    A.cell_start = reinterpret_cast<const int *>(base + box_bytes);
    A.cell_items = reinterpret_cast<const unsigned short *>(base + box_bytes + cs_bytes);

    mpb_world *w = new mpb_world();
    memset(&w->dev, 0, sizeof(w->dev));
    w->dev.dim = m->n_joints;
    w->blob = nullptr;
    w->arm = arm;
    cudaGetDevice(&w->device);
    // a pool of this world's own (without one the edge walk uses fixed shares)
    cudaMemPoolProps props;
    memset(&props, 0, sizeof(props));
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = w->device;
    arm->pool = nullptr;
    if (cudaMemPoolCreate(&arm->pool, &props) == cudaSuccess) {
        unsigned long long keep = ~0ULL;
        cudaMemPoolSetAttribute(arm->pool, cudaMemPoolAttrReleaseThreshold, &keep);
    } else {
        arm->pool = nullptr;
        (void)cudaGetLastError();
    }
    *out = w;
    return MPB_OK;
}
