// k nearest neighbours: exact fp64 brute force, and the tiled fp32 filter + fp64 re-rank.
//
// Replaces motion_planners/nearest.py:36-85 (KDNeighbors over scipy's cKDTree, p = 2,
// eps = 0) as driven by motion_planners/lazy_prm.py:148-160, and the distance sort of
// motion_planners/prm.py:244-246.  Ranking key = cKDTree's squared distance between the
// embedded points e = fl(w*x) (motion_planners/primitives.py:55-57): four fp64 accumulators
// over blocks of four dimensions, no FMA, ties to the smaller index.
//
// fp32 filter: samples are packed once into fp32 tiles (centred on the bounding-box
// midpoint, plus the squared norm), streamed through shared memory with 1-D bulk async
// copies (TMA engine) in a multi-stage mbarrier ring, and each warp scores 32 candidates per
// step for Q register-resident queries with the expansion key = |x|^2 - 2 q.x: one packed
// FFMA2 per coordinate serves two queries.  Each query keeps its 16 (k <= 12) or 32 best
// candidates in a warp-distributed sorted list (shuffle insert).  At the end the exact fp64
// keys of the survivors are ranked and a rounding-error bound certifies that nothing outside
// the list can enter the top k; uncertified queries are queued for the exact kernel.
#include <math.h>

#include <mutex>

#include "mpb_common.cuh"

#define FULL_MASK 0xffffffffu

namespace {

#ifndef MPB_KNN_WARPS
#define MPB_KNN_WARPS 8
#endif
constexpr int kWarps = MPB_KNN_WARPS;   // warps per block
constexpr int kBlocksPerSM = 16 / MPB_KNN_WARPS;   // 16 warps of 128 registers fill an SM
#ifndef MPB_KNN_STAGES
#define MPB_KNN_STAGES 3
#endif
#ifndef MPB_KNN_TILE2
#define MPB_KNN_TILE2 1024
#endif
// Two builds of the filter kernel, chosen by the number of candidates (measured, profiles/r1_summary.md):
//  * LONG scans spend their time in the multiply-add stream: 8 groups per loop body and every plane of the next
//    group prefetched keep the FMA pipe fed;
//  * SHORT scans (up to kShortScanMax candidates) still insert into the lists at most steps, the instruction
//    stream runs through the insertion code all the time, and a body that fits the instruction cache wins:
//    2 groups per body, no prefetch registers.
#ifndef MPB_KNN_UNROLL
#define MPB_KNN_UNROLL 8
#endif
#ifndef MPB_KNN_PREFETCH
#define MPB_KNN_PREFETCH 8   // planes of the next group fetched one step ahead (clamped to the planes there are)
#endif
#ifndef MPB_KNN_UNROLL_SHORT
#define MPB_KNN_UNROLL_SHORT 2
#endif
#ifndef MPB_KNN_PREFETCH_SHORT
#define MPB_KNN_PREFETCH_SHORT 0
#endif
#ifndef MPB_KNN_SHORT_MAX
#define MPB_KNN_SHORT_MAX 262144
#endif
constexpr int64_t kShortScanMax = MPB_KNN_SHORT_MAX;
constexpr int kStages = MPB_KNN_STAGES;   // shared-memory ring depth
constexpr int kSmallListMaxK = 12; // k up to which 16 survivors per query are kept (else 32)
constexpr int kFilterMaxK = 28;    // beyond this the 32-entry lists cannot certify anything: exact scan directly
constexpr int kHeaderBytes = 1024;

struct KnnHeader {
    unsigned long long mm[2 * MPB_MAX_DIM];   // order-encoded min [0,16) and max [16,32)
    double center[MPB_MAX_DIM];
    double R;                                 // >= ||e - c|| and >= ||fl32(e - c)|| for every sample
    int fb_count;                             // queries queued for the exact kernel
    int pad;
    int dim_order[MPB_MAX_DIM];               // coordinates by extent, widest first (kdorder.cuh)
    unsigned long long tiles_visited;         // candidate tiles read by the tensor-core scan, summed over its CTAs
    unsigned long long dbg[16];               // MPB_TC_STATS builds only
};
static_assert(sizeof(KnnHeader) <= kHeaderBytes, "header overflows its slot");

__host__ __device__ constexpr int planes_for(int D) { return (D + 1 + 3) / 4; }
__host__ __device__ constexpr int tile_for(int D) { return planes_for(D) == 1 ? 1024 : (planes_for(D) == 2 ? MPB_KNN_TILE2 : 256); }

__device__ __forceinline__ unsigned long long enc_f64(double v)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(v);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double dec_f64(unsigned long long u)
{
    u = (u >> 63) ? (u & 0x7fffffffffffffffull) : ~u;
    return __longlong_as_double((long long)u);
}

// ------------------------------------------------------------------------------------------
// prepare: bounding box -> centre and radius -> packed fp32 tiles
// ------------------------------------------------------------------------------------------
__global__ void knn_init_kernel(KnnHeader *h)
{
    const int i = threadIdx.x;
    if (i < MPB_MAX_DIM) {
        h->mm[i] = enc_f64(INFINITY);
        h->mm[MPB_MAX_DIM + i] = enc_f64(-INFINITY);
        h->center[i] = 0.;
    }
    if (i == 0) {
        h->R = 0.;
        h->fb_count = 0;
    }
}

template <int D>
__global__ void __launch_bounds__(256) knn_minmax_kernel(const double *__restrict__ x, int64_t N, Weights wt, KnnHeader *h)
{
    double lo[D], hi[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        lo[d] = INFINITY;
        hi[d] = -INFINITY;
    }
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < N; j += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double e = d_mul(wt.w[d], x[j * D + d]);
            lo[d] = fmin(lo[d], e);
            hi[d] = fmax(hi[d], e);
        }
    }
    // warp, then block, then one atomic per block and bound (one per warp had 130k atomics queue on 14 addresses:
    // 100 us of a 1.2 ms prepare)
    __shared__ double s_lo[8][D], s_hi[8][D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[d] = fmin(lo[d], __shfl_xor_sync(FULL_MASK, lo[d], o));
            hi[d] = fmax(hi[d], __shfl_xor_sync(FULL_MASK, hi[d], o));
        }
        if ((threadIdx.x & 31) == 0) {
            s_lo[threadIdx.x >> 5][d] = lo[d];
            s_hi[threadIdx.x >> 5][d] = hi[d];
        }
    }
    __syncthreads();
    if (threadIdx.x < D) {
        const int d = threadIdx.x;
        double l = s_lo[0][d], u = s_hi[0][d];
        for (int w = 1; w < 8; ++w) {
            l = fmin(l, s_lo[w][d]);
            u = fmax(u, s_hi[w][d]);
        }
        atomicMin(&h->mm[d], enc_f64(l));
        atomicMax(&h->mm[MPB_MAX_DIM + d], enc_f64(u));
    }
}

__global__ void knn_set_stat_kernel(int64_t *stats, int64_t v) { stats[0] = v; }

// Where the answer of query q goes: row q, or -- when the queries are a slice of the samples' spatial order
// (mpb_knn_search_part) -- the query's position in that order, counted from the slice's first position.
// kdorder.cuh (included further down)
__host__ __device__ inline int64_t kd_leaf_start(int64_t j, int64_t n, int64_t L);
__host__ __device__ inline int64_t kd_leaf_of(int64_t pos, int64_t n, int64_t L);

struct RowMap {
    const int32_t *inv;        // sample index -> position in the spatial order (NULL: row = q)
    int64_t pos0;              // rowstart == NULL: the rows are the positions from pos0 on
    const int32_t *rowstart;   // a dealt share (kd_share_leaf): first row of each of its blocks
    int parts;
    int64_t n, leaves;
    __device__ __forceinline__ int64_t row(int64_t q) const
    {
        if (!inv) return q;
        const int64_t pos = inv[q];
        if (!rowstart) return pos - pos0;
        const int64_t l = kd_leaf_of(pos, n, leaves);
        return (int64_t)rowstart[l / parts] + pos - kd_leaf_start(l, n, leaves);
    }
};

__global__ void knn_finalize_kernel(KnnHeader *h, int D)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double r2 = 0.;
    for (int d = 0; d < D; ++d) {
        const double lo = dec_f64(h->mm[d]), hi = dec_f64(h->mm[MPB_MAX_DIM + d]);
        double c = 0.5 * lo + 0.5 * hi;
        if (!isfinite(c)) c = 0.;
        h->center[d] = c;
        const double ext = fmax(fabs(hi - c), fabs(c - lo));
        r2 += ext * ext;
    }
    // 1e-6 covers the rounding of c, ext, the sum, and of fl32(e - c) (2^-24)
    h->R = sqrt(r2) * (1.0 + 1e-6) + 1e-300;
}

// Tile t holds TILE candidates, 4P floats each: the centred fp32 coordinates, at index D the squared norm,
// zeros after it.  Slots past N carry an infinite norm so that they never qualify.
template <int D>
__global__ void __launch_bounds__(256)
    knn_pack_kernel(const double *__restrict__ x, int64_t N, int64_t Npad, Weights wt, const KnnHeader *__restrict__ h,
                    float4 *__restrict__ packed)
{
    constexpr int P = planes_for(D), TILE = tile_for(D);
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= Npad) return;
    float v[4 * P];
#pragma unroll
    for (int i = 0; i < 4 * P; ++i) v[i] = 0.f;
    if (j < N) {
        double n2 = 0.;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double e = d_mul(wt.w[d], x[j * D + d]);
            const float f = (float)d_sub(e, h->center[d]);
            v[d] = f;
            n2 += (double)f * (double)f;
        }
        v[D] = (float)n2;
    } else {
        v[D] = INFINITY;
    }
    // Tile layout for the pair-wise scoring of the filter kernel: within a group of 64 candidates lane l owns
    // candidates l ("a") and 32 + l ("b"); value i of both sits side by side, two values per float4:
    //   plane m (0 .. 2P-1), entry g*32 + l  =  (a[2m], b[2m], a[2m+1], b[2m+1])
    // so that one LDS.128 delivers two ready-made (a, b) register pairs for FFMA2.
    const int64_t t = j / TILE;
    const int c = (int)(j - t * TILE);
    const int g = c >> 6, r = c & 63, l = r & 31, half = r >> 5;
    float *base = reinterpret_cast<float *>(packed) + (size_t)t * TILE * P * 4;
#pragma unroll
    for (int i = 0; i < 4 * P; ++i) {
        const int m = i >> 1;
        base[((size_t)m * (TILE / 2) + (size_t)(g * 32 + l)) * 4 + (i & 1) * 2 + half] = v[i];
    }
}

// ------------------------------------------------------------------------------------------
// exact kernel: one warp per query, sorted (key, index) list distributed over the lanes
// ------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256)
    knn_exact_kernel(const double *__restrict__ x, int64_t N, const double *__restrict__ qs, int64_t M, Weights wt,
                     int metric, int k, double lim, const int *__restrict__ list, const int *__restrict__ list_count,
                     const unsigned long long *__restrict__ visited, RowMap rows, int64_t *__restrict__ out_idx, double *__restrict__ out_dist, int64_t *__restrict__ stats)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t total = list ? (int64_t)*list_count : M;
    if (stats && warp0 == 0 && lane == 0) {
        stats[0] = list ? total : 0;
        stats[1] = visited ? (int64_t)*visited : 0;      // candidate tiles the tensor-core scan read
    }

    for (int64_t it = warp0; it < total; it += nwarps) {
        const int64_t q = list ? (int64_t)list[it] : it;
        // ranking key: the kd-tree's squared distance, or (other metrics) the rounded distance itself,
        // because the reference sorts / arg-mins on distance_fn's return value
        double eq[D], qr[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            qr[d] = qs[q * D + d];
            eq[d] = d_mul(wt.w[d], qr[d]);
        }
        double lk = INFINITY, thr = INFINITY;
        int li = -1;
        for (int64_t j0 = 0; j0 < N; j0 += 32) {
            const int64_t j = j0 + lane;
            double S = INFINITY;
            if (j < N) {
                double ex[D];
                if (metric == MPB_METRIC_KDTREE) {
#pragma unroll
                    for (int d = 0; d < D; ++d) ex[d] = d_mul(wt.w[d], x[j * D + d]);
                    S = sqdist_kdtree<D>(eq, ex);
                } else {
#pragma unroll
                    for (int d = 0; d < D; ++d) ex[d] = x[j * D + d];
                    S = metric_dist<D>(qr, ex, wt, metric);
                }
            }
            unsigned m = __ballot_sync(FULL_MASK, S < thr);
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1;
                const double key = __shfl_sync(FULL_MASK, S, src);
                if (key < thr) {
                    // equal keys: the newcomer (larger index) goes after the incumbents
                    const int pos = __popc(__ballot_sync(FULL_MASK, lk <= key));
                    const double upk = __shfl_up_sync(FULL_MASK, lk, 1);
                    const int upi = __shfl_up_sync(FULL_MASK, li, 1);
                    if (lane > pos) {
                        lk = upk;
                        li = upi;
                    } else if (lane == pos) {
                        lk = key;
                        li = (int)(j0 + src);
                    }
                    thr = __shfl_sync(FULL_MASK, lk, 31);
                }
            }
        }
        if (lane < k) {
            const bool kd = (metric == MPB_METRIC_KDTREE);
            const bool keep = (li >= 0) && (kd ? (lk < lim) : (lk <= lim));
            const int64_t r = rows.row(q);
            out_idx[r * k + lane] = keep ? (int64_t)li : -1;
            if (out_dist) out_dist[r * k + lane] = keep ? (kd ? __dsqrt_rn(lk) : lk) : INFINITY;
        }
    }
}

// ------------------------------------------------------------------------------------------
// fp32 filter kernel
// ------------------------------------------------------------------------------------------
template <int D, int Q>
struct FilterSmem {
    static constexpr int P = planes_for(D);
    static constexpr int TILE = tile_for(D);
    static constexpr int STAGE_BYTES = TILE * P * 16;
    static constexpr int RING_BYTES = kStages * STAGE_BYTES;
    static constexpr int BYTES = RING_BYTES + 128;   // + full barriers and release counters
};
static_assert(kStages * 8 + kStages * 4 <= 128, "sync area too small");

// 128-bit shared-memory load by 32-bit shared-space address (no generic->shared conversion in the loop).
// volatile: keeps its order relative to the mbarrier waits (also volatile asm).
__device__ __forceinline__ float4 lds128(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

// SEG = candidates kept per query (16 or 32).  A 32-lane register holds 32/SEG query lists side by side:
// query q of the warp lives in register q / (32/SEG), lanes [SEG*(q % (32/SEG)), +SEG), sorted ascending.
// (Measured alternatives that lost, see profiles/r1_summary.md: a shared-memory append buffer with periodic
// compaction; a fixed producer warp; staggered tile order between blocks.)
template <int D, int Q, int SEG, int UNROLL, int PREFETCH>
__global__ void __launch_bounds__(kWarps * 32, kBlocksPerSM)
    knn_filter_kernel(const double *__restrict__ x, int64_t N, const double *__restrict__ qs, int64_t M, Weights wt,
                      int k, double r2, const float4 *__restrict__ packed, int n_tiles, KnnHeader *__restrict__ h,
                      int *__restrict__ fb_list, int64_t *__restrict__ out_idx, double *__restrict__ out_dist)
{
    using SM = FilterSmem<D, Q>;
    constexpr int P = SM::P, TILE = SM::TILE;
    constexpr int PER = 32 / SEG;          // query lists per register
    constexpr int NREG = Q / PER;          // list registers per warp
    constexpr unsigned SEGMASK = (SEG == 32) ? 0xffffffffu : ((1u << (SEG & 31)) - 1u);
    static_assert(Q % PER == 0, "Q must be a multiple of the lists per register");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *stage_base = reinterpret_cast<float4 *>(smem_raw);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + SM::RING_BYTES);
    int *released = reinterpret_cast<int *>(full + kStages);      // warps done with each ring slot

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int seg = lane / SEG;            // which list of a register this lane belongs to
    const int sl = lane % SEG;             // position inside that list

    auto request_tile = [&](int i) {   // one lane: arm the slot's barrier and start the bulk copy of tile i
        const int slot = i % kStages;
        mbar_arrive_expect_tx(&full[slot], SM::STAGE_BYTES);
        bulk_copy_g2s(stage_base + (size_t)slot * TILE * P, packed + (size_t)i * TILE * P, SM::STAGE_BYTES,
                      &full[slot]);
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full[s], 1);
            released[s] = 0;
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages && i < n_tiles; ++i) request_tile(i);
    }

    // ---- this warp's Q queries, as -2*fl32(e_q - c), uniform across the lanes
    const int64_t q0 = ((int64_t)blockIdx.x * kWarps + warp) * Q;
    // Every lane computes the same value; passing it through a warp reduction lands it in a UNIFORM register
    // (REDUX writes one), and FFMA2 takes a uniform scalar as its broadcast operand.  With the coefficients off
    // the vector register file, and the candidate pair reused by the Q chains of a coordinate, each FFMA2 reads
    // a single fresh vector operand (its accumulator) and the FMA pipe runs at its full rate; with a second
    // fresh vector operand it sustains only ~63 % of that (operand-pattern micro-benchmarks, profiles/).
    float qm[Q][D];
#pragma unroll
    for (int qi = 0; qi < Q; ++qi) {
        int64_t q = q0 + qi;
        if (q >= M) q = M - 1;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double e = d_mul(wt.w[d], qs[q * D + d]);
            const float v = -2.f * (float)d_sub(e, h->center[d]);
            qm[qi][d] = __uint_as_float(__reduce_or_sync(FULL_MASK, __float_as_uint(v)));
        }
    }
    float thr[Q], lk[NREG];
    int li[NREG];
#pragma unroll
    for (int qi = 0; qi < Q; ++qi) thr[qi] = INFINITY;
#pragma unroll
    for (int r = 0; r < NREG; ++r) {
        lk[r] = INFINITY;
        li[r] = -1;
    }

    const uint32_t smem0 = smem_u32(stage_base) + (uint32_t)lane * 16u;

    // Insertion rounds for one group of 32 candidates (lane <-> candidate cbase + lane) whose keys are `key32`.
    // Every round gives each query at most one of its passing candidates, so the shuffle chains of the list
    // registers that have a taker overlap.  A candidate that no longer beats its list (or "no candidate":
    // key = +inf) lands at position SEG, i.e. nowhere.
    auto insert_group = [&](const float (&key32)[Q], int cbase) {
        unsigned m[Q];
#pragma unroll
        for (int qi = 0; qi < Q; ++qi) m[qi] = __ballot_sync(FULL_MASK, key32[qi] < thr[qi]);
        bool more = false;
#pragma unroll
        for (int qi = 0; qi < Q; ++qi) more = more || (m[qi] != 0);
        while (more) {
            more = false;
#pragma unroll
            for (int r = 0; r < NREG; ++r) {
                // late in the scan a group has a taker for one query at most: skip the list registers with none
                // (the masks are ballots, so the branch is uniform)
                unsigned any_r = 0;
#pragma unroll
                for (int hs = 0; hs < PER; ++hs) any_r |= m[r * PER + hs];
                if (any_r == 0) continue;
                float key = INFINITY;
                int cidx = -1;
#pragma unroll
                for (int hs = 0; hs < PER; ++hs) {
                    const int qi = r * PER + hs;
                    const int src = (__ffs(m[qi]) - 1) & 31;
                    const float kq = __shfl_sync(FULL_MASK, key32[qi], src);
                    if (seg == hs) {
                        key = m[qi] ? kq : INFINITY;
                        cidx = cbase + src;
                    }
                    m[qi] &= m[qi] - 1;
                    more = more || (m[qi] != 0);
                }
                const unsigned le = __ballot_sync(FULL_MASK, lk[r] <= key);
                const int pos = __popc((le >> ((SEG * seg) & 31)) & SEGMASK);
                const float upk = __shfl_up_sync(FULL_MASK, lk[r], 1, SEG);
                const int upi = __shfl_up_sync(FULL_MASK, li[r], 1, SEG);
                if (sl > pos) {
                    lk[r] = upk;
                    li[r] = upi;
                } else if (sl == pos) {
                    lk[r] = key;
                    li[r] = cidx;
                }
#pragma unroll
                for (int hs = 0; hs < PER; ++hs)
                    thr[r * PER + hs] = __shfl_sync(FULL_MASK, lk[r], SEG * hs + SEG - 1);
            }
        }
    };
    // Both halves of a group of 64 go through ONE inlined copy of the insertion code: the scan loop is unrolled,
    // and every copy costs instruction-cache space (fetch stalls are visible in the profile of this loop).
    auto insert_halves = [&](const float (&ka)[Q], const float (&kb)[Q], int cbase, bool any_a, bool any_b) {
#pragma unroll 1
        for (int hh = 0; hh < 2; ++hh) {
            if (!(hh ? any_b : any_a)) continue;
            float key32[Q];
#pragma unroll
            for (int qi = 0; qi < Q; ++qi) key32[qi] = hh ? kb[qi] : ka[qi];
            insert_group(key32, cbase + 32 * hh);
        }
    };
    float prev_a[Q], prev_b[Q];   // keys of the 64 candidates scored one step ago, not yet compared
    int prev_base = 0;
#pragma unroll
    for (int qi = 0; qi < Q; ++qi) prev_a[qi] = prev_b[qi] = INFINITY;

    // ---- stream every tile through the ring
    for (int t = 0; t < n_tiles; ++t) {
        const int s = t % kStages;
        const uint32_t ph = (uint32_t)(t / kStages) & 1u;
        mbar_wait_warp(&full[s], ph);

        const uint32_t tbase = smem0 + (uint32_t)s * (uint32_t)SM::STAGE_BYTES;
        const int cbase0 = t * TILE;
        // Each step scores 64 candidates: lane l owns candidates l and 32 + l of the group as a register PAIR per
        // coordinate, and one FFMA2 advances both keys of one query.  Software pipeline: while the FFMA2 chains
        // of this group run, the keys of the PREVIOUS group (complete long ago) are compared with the thresholds
        // and, if any of them passes, inserted -- so the compare/vote/branch tail never waits for the arithmetic
        // it follows.  The keys in flight cross tile boundaries in registers (`prev_a`, `prev_b`, `prev_base`);
        // the last group is drained after the tile loop.
        constexpr int NV = 2 * P;        // float4 per lane per group
        static_assert(TILE % 64 == 0, "tiles hold whole groups of 64 candidates");
        // Planes 0 .. NPF-1 of a group are fetched one step ahead (their values are needed the moment the step
        // starts); the remaining planes are loaded at the top of their own step, where the FFMA2 chains of the
        // earlier planes cover the shared-memory latency.  Prefetching everything costs 4 registers per plane and
        // pushes the loop over the 128-register budget.
        constexpr int NPF = (PREFETCH < NV) ? PREFETCH : NV;
        float4 nxt[NPF > 0 ? NPF : 1];
#pragma unroll
        for (int m = 0; m < NPF; ++m) nxt[m] = lds128(tbase + (uint32_t)(m * (TILE / 2)) * 16u);
#pragma unroll UNROLL
        for (int c0 = 0; c0 < TILE; c0 += 64) {
            f32x2_t xv[4 * P];           // xv[i] = (value i of candidate a, value i of candidate b)
#pragma unroll
            for (int m = 0; m < NV; ++m) {
                const float4 f = (m < NPF) ? nxt[m] : lds128(tbase + (uint32_t)(m * (TILE / 2) + c0 / 2) * 16u);
                xv[2 * m + 0] = pack_f32x2(f.x, f.y);
                xv[2 * m + 1] = pack_f32x2(f.z, f.w);
            }
            if (c0 + 64 < TILE) {
#pragma unroll
                for (int m = 0; m < NPF; ++m)
                    nxt[m] = lds128(tbase + (uint32_t)(m * (TILE / 2) + (c0 + 64) / 2) * 16u);
            }
            bool pass_a = false, pass_b = false;
#pragma unroll
            for (int qi = 0; qi < Q; ++qi) {
                pass_a = pass_a || (prev_a[qi] < thr[qi]);
                pass_b = pass_b || (prev_b[qi] < thr[qi]);
            }
            float acc_a[Q], acc_b[Q];
#pragma unroll
            for (int qi = 0; qi < Q; ++qi) {
                f32x2_t a = xv[D];
#pragma unroll
                for (int d = 0; d < D; ++d) a = fma_f32x2(xv[d], pack_f32x2(qm[qi][d], qm[qi][d]), a);
                unpack_f32x2(a, acc_a[qi], acc_b[qi]);
            }
            const bool any_a = __any_sync(FULL_MASK, pass_a), any_b = __any_sync(FULL_MASK, pass_b);
            if (any_a || any_b) insert_halves(prev_a, prev_b, prev_base, any_a, any_b);
#pragma unroll
            for (int qi = 0; qi < Q; ++qi) {
                prev_a[qi] = acc_a[qi];
                prev_b[qi] = acc_b[qi];
            }
            prev_base = cbase0 + c0;
        }
        // The last warp to release a slot requests its next tenant: nobody waits for anybody.
        __syncwarp();
        if (lane == 0) {
            __threadfence_block();
            const int done = atomicAdd(&released[s], 1);
            if (done == kWarps - 1) {
                released[s] = 0;
                __threadfence_block();
                if (t + kStages < n_tiles) request_tile(t + kStages);
            }
        }
        __syncwarp();
    }

    {   // drain the software pipeline: the last group scored has not been compared yet
        insert_halves(prev_a, prev_b, prev_base, true, true);
    }

    // ---- exact re-rank of the survivors and the certificate; the PER lists of a register are handled
    //      side by side, each lane working for the query its segment belongs to
    const double R = h->R;
    constexpr double U32 = 5.9604644775390625e-08;   // 2^-24
#pragma unroll
    for (int r = 0; r < NREG; ++r) {
        const int64_t q = q0 + r * PER + seg;
        const bool qvalid = q < M;
        const int64_t qc = qvalid ? q : (M - 1);
        double eq[D];
        double nq2 = 0.;
        float T32 = thr[r * PER];
#pragma unroll
        for (int hs = 1; hs < PER; ++hs) T32 = (seg == hs) ? thr[r * PER + hs] : T32;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            eq[d] = d_mul(wt.w[d], qs[qc * D + d]);
            // the query's fp32 coordinate, recomputed exactly as it was for the scan (eq[d] is this lane's query)
            const float qmd = -2.f * (float)d_sub(eq[d], h->center[d]);
            const double qh = -0.5 * (double)qmd;
            nq2 += qh * qh;
        }
        const int ci = li[r];
        double S = INFINITY;
        if (ci >= 0) {
            double ex[D];
#pragma unroll
            for (int d = 0; d < D; ++d) ex[d] = d_mul(wt.w[d], x[(int64_t)ci * D + d]);
            S = sqdist_kdtree<D>(eq, ex);
        }
        int rank = 0;
#pragma unroll 4
        for (int o = 0; o < SEG; ++o) {
            const double So = __shfl_sync(FULL_MASK, S, o, SEG);
            const int io = __shfl_sync(FULL_MASK, ci, o, SEG);
            rank += ((So < S) || (So == S && io < ci)) ? 1 : 0;
        }
        // k-th smallest exact key among this segment's survivors (inf when fewer than k exist)
        const unsigned mk_all = __ballot_sync(FULL_MASK, ci >= 0 && rank == k - 1);
        const unsigned mk = (mk_all >> ((SEG * seg) & 31)) & SEGMASK;
        const double Sk_src = __shfl_sync(FULL_MASK, S, mk ? (__ffs(mk) - 1) : 0, SEG);
        const double Sk = mk ? Sk_src : INFINITY;
        const unsigned valid_all = __ballot_sync(FULL_MASK, ci >= 0);
        const int nvalid = __popc((valid_all >> ((SEG * seg) & 31)) & SEGMASK);
        const double T = (double)T32;
        // list never filled: every sample should be a survivor; fewer means a key was not comparable (overflow,
        // NaN) and the query goes to the exact kernel
        bool ok = isinf(T) && T > 0 && (int64_t)nvalid >= N;
        if (!ok) {
            // every sample outside the list has computed key >= T, hence exact squared
            // distance >= (sqrt(T - eta + |q|^2) - delta)^2  ("kNN certificate")
            const double nq = sqrt(nq2);
            const double delta = U32 * (R + nq) * 1.0001 + 1e-30;
            const double eta = (double)(D + 2) * U32 * 1.001 * (R * R + 2.0 * nq * R) + 1e-37;
            double slow = T - eta + nq2;
            slow -= 1e-12 * (fabs(T) + eta + nq2);
            double bound = 0.;
            if (slow > 0.) {
                const double rho = sqrt(slow) * (1.0 - 1e-15) - delta;
                if (rho > 0.) bound = rho * rho * (1.0 - 1e-12);
            }
            ok = (Sk < bound) || (r2 <= bound);
        }
        if (qvalid) {
            if (ok) {
                if (ci >= 0 && rank < k) {
                    const bool keep = S < r2;
                    out_idx[q * k + rank] = keep ? (int64_t)ci : -1;
                    if (out_dist) out_dist[q * k + rank] = keep ? __dsqrt_rn(S) : INFINITY;
                }
                if (sl < k && sl >= nvalid) {
                    out_idx[q * k + sl] = -1;
                    if (out_dist) out_dist[q * k + sl] = INFINITY;
                }
            } else if (sl == 0) {
                const int pos = atomicAdd(&h->fb_count, 1);
                fb_list[pos] = (int)q;
            }
        }
    }
}

// kernels launched by the kNN entry points on this host thread (bench.py reports the count it claims)
static thread_local int g_last_launches = 0;

#include "kdorder.cuh"
#include "knn_tc.cuh"

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
// Which filter serves a search.  The tensor-core scan (knn_tc.cuh) reads only the tiles near a query block, the
// FFMA2 kernel scans every pair; the scan's prepare (k-d order, operand tiles) costs 0.2 to 0.7 ms against 0.04.
// Measured on B200, uniform samples, k = 11, set against itself, prepare + search, tensor-core scan vs FFMA2
// kernel (tools/tc_policy_sweep.py, profiles/r2_tc_policy_sweep.txt): 7-DoF 20k 0.61 vs 0.52 ms, 50k 1.02 vs 1.46,
// 100k 1.77 vs 4.09, 1M 14.6 vs 244; 6-DoF 20k 0.54 vs 0.51, 50k 0.89 vs 1.38, 100k 1.53 vs 3.85, 1M 10.8 vs 226;
// 5-DoF 20k 0.47 vs 0.49, 50k 0.69 vs 1.30, 1M 8.9 vs 207; 4-DoF 20k 0.35 vs 0.47, 100k 0.79 vs 3.32, 1M 24.7 vs
// 180.  In 2 and 3 dimensions large sets are so dense (neighbours 1e-3 apart in a unit box) that the TF32
// certificate cannot separate the k-th from the next key and queries fall through to the exact kernel (3-D: 33 of
// 200k, 23k of 1M; 2-D: all of 1M): those stay with the FFMA2 kernel.  (Round 1's boundaries, 160k to 500k, were
// measured before the scan pruned tiles.)  Default: from kTcAutoMin[dim] samples on.  MPB200_KNN_TC=0 switches
// the tensor-core scan off, MPB200_KNN_TC=1 forces it for
// every set it can serve (dimension <= 7, at least MPB200_KNN_TC_MIN samples, default 8192); both are read once,
// when the library is first used.
constexpr int64_t kTcNever = INT64_MAX;
constexpr int64_t kTcAutoMin[kTcMaxDim + 1] = {kTcNever, kTcNever, kTcNever, kTcNever, 16384, 24576, 32768, 32768};

struct TcPolicy {
    int mode;            // 0 off, 1 forced, 2 automatic
    int64_t forced_min;
};

const TcPolicy &tc_policy()
{
    static const TcPolicy p = [] {
        TcPolicy q = {2, 8192};
        const char *e = getenv("MPB200_KNN_TC");
        if (e && e[0] == '0') q.mode = 0;
        if (e && e[0] == '1') q.mode = 1;
        const char *m = getenv("MPB200_KNN_TC_MIN");
        if (m) {
            const long long v = atoll(m);
            if (v >= kTcTile) q.forced_min = v;
        }
        return q;
    }();
    return p;
}

bool use_tc(int64_t N, int D)
{
    const TcPolicy &p = tc_policy();
    if (p.mode == 0 || D > kTcMaxDim) return false;
    if (p.mode == 1) return N >= p.forced_min;
    return N >= kTcAutoMin[D];
}

// k-d partitions of the tensor-core scan: samples in leaves of at most 128 (two per candidate tile), a foreign query
// set in leaves of at most 128 (one per query block)
inline KdPlan tc_sample_plan(int64_t N, int D) { return kd_plan(N, 2 * ((N + kTcTile - 1) / kTcTile), D, true); }
inline KdPlan tc_query_plan(int64_t M, int D) { return kd_plan(M, (M + kTcQueries - 1) / kTcQueries, D, false); }

struct WsLayout {
    size_t packed_off, packed_bytes, tc_off, tc_bytes, slot_off, slot_bytes, box_off, box_bytes, perm_off, perm_bytes, inv_off, inv_bytes, fb_off,
        fb_bytes, surv_off,
        surv_bytes, thr_off, thr_bytes, qord_off, qord_bytes, sort_off, sort_bytes, total;
    int64_t n_tiles, n_pad, tc_tiles, tc_slots;
    int box_stride;
};

// Workspace: [header][packed fp32 tiles][tensor-core tiles][tile slot -> sample index][tile boxes] -- written by
// prepare, their offsets depend on the sample set only -- then [fallback list][survivors][thresholds][query
// order][sort scratch], which belong to one search (prepare uses the sort scratch too, at its M = 0 offset).
WsLayout ws_layout(int64_t N, int64_t M, int D, bool allow_tc = true)
{
    WsLayout L;
    const int P = planes_for(D), TILE = tile_for(D);
    const bool tc = allow_tc && use_tc(N, D);
    const size_t m = (size_t)(M > 0 ? M : 1);
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    L.n_tiles = (N + TILE - 1) / TILE;
    L.n_pad = L.n_tiles * TILE;
    L.packed_off = kHeaderBytes;
    L.packed_bytes = (size_t)L.n_pad * P * 16;
    L.tc_tiles = tc ? tc_sample_plan(N, D).leaves / 2 : 0;
    L.tc_off = (L.packed_off + L.packed_bytes + 1023) & ~(size_t)1023;
    L.tc_bytes = (size_t)L.tc_tiles * kTcTileBytes;
    L.slot_off = up(L.tc_off + L.tc_bytes);
    L.slot_bytes = up((size_t)L.tc_tiles * kTcTile * sizeof(int));
    L.box_stride = (int)((L.tc_tiles + 31) & ~(int64_t)31);
    L.box_off = L.slot_off + L.slot_bytes;
    L.box_bytes = up((size_t)16 * L.box_stride * sizeof(float));
    L.perm_off = L.box_off + L.box_bytes;
    L.perm_bytes = tc ? up((size_t)(N > 0 ? N : 1) * sizeof(uint32_t)) : 0;     // the samples' k-d leaf order, kept
    L.inv_off = L.perm_off + L.perm_bytes;
    L.inv_bytes = L.perm_bytes;                                                  // ... and its inverse
    L.fb_off = L.inv_off + L.inv_bytes;
    L.fb_bytes = up(m * sizeof(int));
    // query slots: blocks of 128; the sample set as its own query set uses the tiles' own slots
    const int64_t own = (M == N) ? L.tc_tiles * kTcTile : 0;
    const int64_t other = tc ? tc_query_plan(M > 0 ? M : 1, D).leaves * kTcQueries : 0;
    L.tc_slots = tc ? (own > other ? own : other) : 0;
    L.surv_off = L.fb_off + L.fb_bytes;
    L.surv_bytes = tc ? up((size_t)L.tc_slots * kTcSurvivors * sizeof(int)) : 0;
    L.thr_off = L.surv_off + L.surv_bytes;
    L.thr_bytes = tc ? up((size_t)L.tc_slots * kTcLists * sizeof(float)) : 0;
    L.qord_off = L.thr_off + L.thr_bytes;
    L.qord_bytes = tc ? up((size_t)(other > 0 ? other : kTcQueries) * sizeof(int)) : 0;
    L.sort_off = L.qord_off + L.qord_bytes;
    L.sort_bytes = tc ? kd_scratch_bytes(N > M ? N : M) : 0;
    L.total = L.sort_off + L.sort_bytes;
    return L;
}

int check_common(const char *who, const double *samples, int64_t N, const double *queries, int64_t M, int D, int k,
                 const int64_t *out_idx)
{
    MPB_REQUIRE(D >= 1 && D <= MPB_MAX_DIM, "%s: dim %d outside [1, %d]", who, D, MPB_MAX_DIM);
    MPB_REQUIRE(k >= 1 && k <= MPB_MAX_K, "%s: k %d outside [1, %d]", who, k, MPB_MAX_K);
    MPB_REQUIRE(N >= 0 && M >= 0, "%s: negative size", who);
    MPB_REQUIRE(N < 2147483647LL && M < 2147483647LL, "%s: more than 2^31-1 samples or queries", who);
    MPB_REQUIRE(M == 0 || (queries && out_idx), "%s: NULL query or output buffer", who);
    MPB_REQUIRE(N == 0 || samples, "%s: NULL sample buffer", who);
    return MPB_OK;
}

int sm_count()
{
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
}

template <int D>
int launch_exact(const double *x, int64_t N, const double *qs, int64_t M, const Weights &wt, int metric, int k,
                 double lim, const int *list, const int *list_count, int64_t *out_idx, double *out_dist,
                 int64_t *stats, cudaStream_t st, const unsigned long long *visited = nullptr,
                 RowMap rows = RowMap{nullptr, 0})
{
    const int threads = 256;
    int64_t blocks = list ? (int64_t)sm_count() * 4 : (M * 32 + threads - 1) / threads;
    if (blocks > (int64_t)sm_count() * 32) blocks = (int64_t)sm_count() * 32;
    if (blocks < 1) blocks = 1;
    knn_exact_kernel<D><<<(unsigned)blocks, threads, 0, st>>>(x, N, qs, M, wt, metric, k, lim, list, list_count,
                                                              visited, rows, out_idx, out_dist, stats);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int D>
int launch_prepare(const double *x, int64_t N, const Weights &wt, void *ws, bool allow_tc, cudaStream_t st)
{
    const WsLayout L = ws_layout(N, 0, D, allow_tc);
    KnnHeader *h = reinterpret_cast<KnnHeader *>(ws);
    float4 *packed = reinterpret_cast<float4 *>(reinterpret_cast<char *>(ws) + L.packed_off);
    knn_init_kernel<<<1, 32, 0, st>>>(h);
    MPB_LAUNCH_CHECK();
    if (N > 0) {
        int64_t blocks = (N + 255) / 256;
        if (blocks > (int64_t)sm_count() * 4) blocks = (int64_t)sm_count() * 4;
        knn_minmax_kernel<D><<<(unsigned)blocks, 256, 0, st>>>(x, N, wt, h);
        MPB_LAUNCH_CHECK();
    }
    knn_finalize_kernel<<<1, 32, 0, st>>>(h, D);
    MPB_LAUNCH_CHECK();
    if (L.n_pad > 0) {
        knn_pack_kernel<D><<<(unsigned)((L.n_pad + 255) / 256), 256, 0, st>>>(x, N, L.n_pad, wt, h, packed);
        MPB_LAUNCH_CHECK();
    }
    g_last_launches += 2 + (N > 0 ? 1 : 0) + (L.n_pad > 0 ? 1 : 0);
    if constexpr (D <= kTcMaxDim) {
        if (L.tc_tiles > 0) {
            // k-d leaf order of the samples (two leaves of at most 128 per tile), then the operand tiles, the
            // slot -> sample table and the tile boxes in that order
            char *base = reinterpret_cast<char *>(ws);
            unsigned char *tiles = reinterpret_cast<unsigned char *>(base + L.tc_off);
            int *slot_idx = reinterpret_cast<int *>(base + L.slot_off);
            float *boxes = reinterpret_cast<float *>(base + L.box_off);
            kd_dim_order_kernel<<<1, 32, 0, st>>>(h, D);
            MPB_LAUNCH_CHECK();
            const KdPlan plan = tc_sample_plan(N, D);
            const int64_t leaves = plan.leaves;
            const KdScratch scratch = kd_scratch_at(base + L.sort_off, N);
            const uint32_t *perm = nullptr;
            int rc = kd_order<D>(x, N, wt, h, plan, scratch, &perm, st, &g_last_launches);
            if (rc != MPB_OK) return rc;
            knn_tc_pack_kernel<D><<<(unsigned)L.tc_tiles, kTcTile, 0, st>>>(x, N, leaves, perm, wt, h, tiles, slot_idx,
                                                                         boxes, L.box_stride);
            MPB_LAUNCH_CHECK();
            // the order outlives the sort scratch (mpb_knn_sample_order hands it to callers that shard by it)
            MPB_CUDA_CHECK(cudaMemcpyAsync(base + L.perm_off, perm, (size_t)N * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
            kd_invert_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(perm, N, reinterpret_cast<int32_t *>(base + L.inv_off));
            MPB_LAUNCH_CHECK();
            g_last_launches += 1;
            g_last_launches += 2;
        }
    }
    return MPB_OK;
}

template <int D, int Q, int SEG, int UNROLL, int PREFETCH>
int launch_filter_as(const double *x, int64_t N, const double *qs, int64_t M, const Weights &wt, int k, double r2,
                     const float4 *packed, int n_tiles, KnnHeader *h, int *fb, int64_t *out_idx, double *out_dist,
                     cudaStream_t st)
{
    using SM = FilterSmem<D, Q>;
    auto kernel = knn_filter_kernel<D, Q, SEG, UNROLL, PREFETCH>;
    MPB_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM::BYTES));
    const int64_t blocks = (M + (int64_t)kWarps * Q - 1) / ((int64_t)kWarps * Q);
    kernel<<<(unsigned)blocks, kWarps * 32, SM::BYTES, st>>>(x, N, qs, M, wt, k, r2, packed, n_tiles, h, fb, out_idx,
                                                             out_dist);
    MPB_LAUNCH_CHECK();
    return MPB_OK;
}

template <int D, int Q, int SEG>
int launch_filter(const double *x, int64_t N, const double *qs, int64_t M, const Weights &wt, int k, double r2,
                  const float4 *packed, int n_tiles, KnnHeader *h, int *fb, int64_t *out_idx, double *out_dist,
                  cudaStream_t st)
{
    if (N <= kShortScanMax)
        return launch_filter_as<D, Q, SEG, MPB_KNN_UNROLL_SHORT, MPB_KNN_PREFETCH_SHORT>(
            x, N, qs, M, wt, k, r2, packed, n_tiles, h, fb, out_idx, out_dist, st);
    return launch_filter_as<D, Q, SEG, MPB_KNN_UNROLL, MPB_KNN_PREFETCH>(x, N, qs, M, wt, k, r2, packed, n_tiles, h,
                                                                         fb, out_idx, out_dist, st);
}

template <int D>
int launch_search(const double *x, int64_t N, const double *qs, int64_t M, const Weights &wt, int k, double r2,
                  void *ws, bool allow_tc, int64_t *out_idx, double *out_dist, int64_t *stats, cudaStream_t st,
                  int part = -1, int parts = 1, int64_t *out_rows = nullptr)
{
#ifndef MPB_KNN_Q
#define MPB_KNN_Q 8
#endif
    constexpr int Q = (D <= 8) ? MPB_KNN_Q : 4;
    const WsLayout L = ws_layout(N, M, D, allow_tc);
    KnnHeader *h = reinterpret_cast<KnnHeader *>(ws);
    const float4 *packed = reinterpret_cast<const float4 *>(reinterpret_cast<char *>(ws) + L.packed_off);
    int *fb = reinterpret_cast<int *>(reinterpret_cast<char *>(ws) + L.fb_off);
    if (k > kFilterMaxK) {
        // A 32-entry survivor list has no slack left to certify more than kFilterMaxK neighbours (measured: every
        // query of a k = 32 search fell through to the exact scan after paying for the filter pass): scan exactly.
        int rc0 = launch_exact<D>(x, N, qs, M, wt, MPB_METRIC_KDTREE, k, r2, nullptr, nullptr, out_idx, out_dist,
                                  nullptr, st);
        if (rc0 != MPB_OK) return rc0;
        if (stats) {
            knn_set_stat_kernel<<<1, 1, 0, st>>>(stats, M);      // every query went to the exact scan
            MPB_LAUNCH_CHECK();
        }
        g_last_launches += 2;
        return MPB_OK;
    }
    MPB_CUDA_CHECK(cudaMemsetAsync(&h->fb_count, 0, sizeof(int), st));
    MPB_CUDA_CHECK(cudaMemsetAsync(&h->tiles_visited, 0, 17 * sizeof(unsigned long long), st));
    if constexpr (D <= kTcMaxDim) {
        if (L.tc_tiles > 0 && k <= kSmallListMaxK) {
            // tensor-core scan over the tiles near each query block -> 16 survivors per query -> exact re-rank +
            // certificate -> exact re-do of the rest
            char *base = reinterpret_cast<char *>(ws);
            const unsigned char *tiles = reinterpret_cast<const unsigned char *>(base + L.tc_off);
            const int *slot_idx = reinterpret_cast<const int *>(base + L.slot_off);
            const float *boxes = reinterpret_cast<const float *>(base + L.box_off);
            int *surv = reinterpret_cast<int *>(base + L.surv_off);
            float *sthr = reinterpret_cast<float *>(base + L.thr_off);
            const int *qorder = slot_idx;                  // the sample set queried against itself: block = half tile
            int64_t blocks = 2 * L.tc_tiles;
            int64_t own_block0 = 0;
            int own_part = 0, own_parts = 1;
            RowMap rows = RowMap{nullptr, 0, nullptr, 1, 0, 0};
            const bool own = (qs == x && M == N);
            if (part >= 0) {
                // the queries are the samples of share `part` of the leaves (kd_share_leaf): the same blocks a
                // whole-set search has, a share of them; block i's answers go to rows rowstart[i] onwards.  The two
                // small tables live in the sort scratch, which a search of the set against itself does not use.
                const KdPlan plan = tc_sample_plan(N, D);
                blocks = kd_share_blocks(plan.leaves, part, parts);
                if (blocks == 0) return MPB_OK;
                int *share_q = reinterpret_cast<int *>(base + L.sort_off);
                int32_t *rowstart = reinterpret_cast<int32_t *>(base + L.sort_off) + blocks * kTcQueries;
                kd_share_rowstart_kernel<<<1, 1024, 0, st>>>(N, plan.leaves, part, parts, blocks, rowstart);
                MPB_LAUNCH_CHECK();
                kd_share_slots_kernel<<<(unsigned)blocks, 128, 0, st>>>(
                    slot_idx, reinterpret_cast<const uint32_t *>(base + L.perm_off), N, plan.leaves, part, parts, rowstart,
                    share_q, out_rows);
                MPB_LAUNCH_CHECK();
                g_last_launches += 2;
                qorder = share_q;
                own_block0 = 0;
                own_part = part;
                own_parts = parts;
                rows = RowMap{reinterpret_cast<const int32_t *>(base + L.inv_off), 0, rowstart, parts, N, plan.leaves};
            } else if (!own) {
                own_block0 = -1;
                // any other query set gets its own k-d order, one leaf of at most 128 queries per block
                const KdPlan plan = tc_query_plan(M, D);
                blocks = plan.leaves;
                const KdScratch scratch = kd_scratch_at(base + L.sort_off, N > M ? N : M);
                const uint32_t *perm = nullptr;
                int rc = kd_order<D>(qs, M, wt, h, plan, scratch, &perm, st, &g_last_launches);
                if (rc != MPB_OK) return rc;
                int *qo = reinterpret_cast<int *>(base + L.qord_off);
                knn_tc_qorder_kernel<<<(unsigned)blocks, kTcQueries, 0, st>>>(M, blocks, perm, qo);
                MPB_LAUNCH_CHECK();
                g_last_launches += 1;
                qorder = qo;
            }
            MPB_CUDA_CHECK(cudaFuncSetAttribute(knn_tc_kernel<D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)TcSmem::BYTES));
            knn_tc_kernel<D, false><<<(unsigned)blocks, kTcThreads, TcSmem::BYTES, st>>>(
                qs, qorder, wt, tiles, (int)L.tc_tiles, boxes, L.box_stride, h, surv, sthr, nullptr, nullptr, nullptr, own_block0, own_part, own_parts);
            MPB_LAUNCH_CHECK();
            const int64_t n_slots = blocks * kTcQueries;
            const int64_t rthreads = n_slots * (kTcSurvivors <= 16 ? 16 : 32);      // a half warp per query slot
            knn_tc_rerank_kernel<D><<<(unsigned)((rthreads + 255) / 256), 256, 0, st>>>(
                x, N, qs, M, wt, k, r2, h, qorder, n_slots, slot_idx, surv, sthr, fb, rows, out_idx, out_dist);
            MPB_LAUNCH_CHECK();
            g_last_launches += 3;
            return launch_exact<D>(x, N, qs, M, wt, MPB_METRIC_KDTREE, k, r2, fb, &h->fb_count, out_idx, out_dist, stats,
                                   st, &h->tiles_visited, rows);
        }
    }
    // 16 survivors per query leave a comfortable certificate margin up to k = 12 (a PRM's k + 1 = 11) and
    // halve the insertion work; larger k keeps 32
    g_last_launches += 2;
    int rc = (k <= kSmallListMaxK)
                 ? launch_filter<D, Q, 16>(x, N, qs, M, wt, k, r2, packed, (int)L.n_tiles, h, fb, out_idx, out_dist, st)
                 : launch_filter<D, Q, 32>(x, N, qs, M, wt, k, r2, packed, (int)L.n_tiles, h, fb, out_idx, out_dist, st);
    if (rc != MPB_OK) return rc;
    return launch_exact<D>(x, N, qs, M, wt, MPB_METRIC_KDTREE, k, r2, fb, &h->fb_count, out_idx, out_dist, stats, st);
}

// measuring build of the tensor-core scan (mpb_knn_tc_key_error): prepared workspace in, two counters out
template <int D>
int launch_tc_diag(const double *x, int64_t N, const double *qs, int64_t M, const Weights &wt, void *ws,
                   unsigned long long *out, cudaStream_t st)
{
    if constexpr (D <= kTcMaxDim) {
        const WsLayout L = ws_layout(N, M, D, true);
        if (L.tc_tiles == 0) return MPB_EINVAL;
        KnnHeader *h = reinterpret_cast<KnnHeader *>(ws);
        char *base = reinterpret_cast<char *>(ws);
        const KdPlan plan = tc_query_plan(M, D);
        const int64_t blocks = plan.leaves;
        const KdScratch scratch = kd_scratch_at(base + L.sort_off, N > M ? N : M);
        const uint32_t *perm = nullptr;
        int rc = kd_order<D>(qs, M, wt, h, plan, scratch, &perm, st, &g_last_launches);
        if (rc != MPB_OK) return rc;
        int *qo = reinterpret_cast<int *>(base + L.qord_off);
        knn_tc_qorder_kernel<<<(unsigned)blocks, kTcQueries, 0, st>>>(M, blocks, perm, qo);
        MPB_LAUNCH_CHECK();
        MPB_CUDA_CHECK(cudaMemsetAsync(out, 0, 2 * sizeof(unsigned long long), st));
        MPB_CUDA_CHECK(cudaFuncSetAttribute(knn_tc_kernel<D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)TcSmem::BYTES));
        knn_tc_kernel<D, true><<<(unsigned)blocks, kTcThreads, TcSmem::BYTES, st>>>(
            qs, qo, wt, reinterpret_cast<const unsigned char *>(base + L.tc_off), (int)L.tc_tiles,
            reinterpret_cast<const float *>(base + L.box_off), L.box_stride, h, nullptr, nullptr, x,
            reinterpret_cast<const int *>(base + L.slot_off), out, -1, 0, 1);
        MPB_LAUNCH_CHECK();
        return MPB_OK;
    } else {
        return MPB_EINVAL;
    }
}

int dispatch_exact(int D, const double *x, int64_t N, const double *qs, int64_t M, const Weights &wt, int metric, int k,
                   double lim, int64_t *out_idx, double *out_dist, cudaStream_t st)
{
#define CALL(DD) launch_exact<DD>(x, N, qs, M, wt, metric, k, lim, nullptr, nullptr, out_idx, out_dist, nullptr, st)
    MPB_DISPATCH_DIM(D, CALL)
#undef CALL
    return MPB_EINVAL;
}
int dispatch_prepare(int D, const double *x, int64_t N, const Weights &wt, void *ws, bool allow_tc, cudaStream_t st)
{
#define CALL(DD) launch_prepare<DD>(x, N, wt, ws, allow_tc, st)
    MPB_DISPATCH_DIM(D, CALL)
#undef CALL
    return MPB_EINVAL;
}
int dispatch_search(int D, const double *x, int64_t N, const double *qs, int64_t M, const Weights &wt, int k, double r2,
                    void *ws, bool allow_tc, int64_t *out_idx, double *out_dist, int64_t *stats, cudaStream_t st,
                    int part = -1, int parts = 1, int64_t *out_rows = nullptr)
{
#define CALL(DD) launch_search<DD>(x, N, qs, M, wt, k, r2, ws, allow_tc, out_idx, out_dist, stats, st, part, parts, out_rows)
    MPB_DISPATCH_DIM(D, CALL)
#undef CALL
    return MPB_EINVAL;
}

}   // namespace

MPB_API int mpb_knn_f64(const double *samples, int64_t n_samples, const double *queries, int64_t n_queries, int dim,
                        const double *weights, int k, double max_dist, int64_t *out_idx, double *out_dist,
                        void *stream)
{
    int rc = check_common("mpb_knn_f64", samples, n_samples, queries, n_queries, dim, k, out_idx);
    if (rc != MPB_OK) return rc;
    MPB_REQUIRE(max_dist >= 0., "mpb_knn_f64: max_dist must be >= 0");
    if (n_queries == 0) return MPB_OK;
    return dispatch_exact(dim, samples, n_samples, queries, n_queries, make_weights(weights, dim), MPB_METRIC_KDTREE,
                          k, max_dist * max_dist, out_idx, out_dist, (cudaStream_t)stream);
}

MPB_API int mpb_knn_f64_metric(const double *samples, int64_t n_samples, const double *queries, int64_t n_queries,
                               int dim, const double *weights, int metric, int k, double max_dist, int64_t *out_idx,
                               double *out_dist, void *stream)
{
    int rc = check_common("mpb_knn_f64_metric", samples, n_samples, queries, n_queries, dim, k, out_idx);
    if (rc != MPB_OK) return rc;
    MPB_REQUIRE(metric >= 0 && metric <= 2, "mpb_knn_f64_metric: bad metric %d", metric);
    MPB_REQUIRE(max_dist >= 0., "mpb_knn_f64_metric: max_dist must be >= 0");
    if (n_queries == 0) return MPB_OK;
    const double lim = (metric == MPB_METRIC_KDTREE) ? max_dist * max_dist : max_dist;
    return dispatch_exact(dim, samples, n_samples, queries, n_queries, make_weights(weights, dim), metric, k, lim,
                          out_idx, out_dist, (cudaStream_t)stream);
}

MPB_API int mpb_knn_filter_kind(int64_t n_samples, int dim, int k)
{
    if (dim < 1 || dim > MPB_MAX_DIM || n_samples < 0 || k < 1 || k > MPB_MAX_K) return MPB_EINVAL;
    if (k > kFilterMaxK) return MPB_KNN_KIND_EXACT;
    if (k <= kSmallListMaxK && use_tc(n_samples, dim)) return MPB_KNN_KIND_TENSOR;
    return MPB_KNN_KIND_FP32;
}

MPB_API size_t mpb_knn_workspace_bytes(int64_t n_samples, int64_t n_queries, int dim)
{
    if (dim < 1 || dim > MPB_MAX_DIM || n_samples < 0 || n_queries < 0) return 0;
    return ws_layout(n_samples, n_queries, dim).total;
}

static int check_ws(const char *who, void *ws, size_t ws_bytes, int64_t N, int64_t M, int D, bool allow_tc = true)
{
    MPB_REQUIRE(ws != nullptr, "%s: workspace is NULL", who);
    MPB_REQUIRE(((uintptr_t)ws & 255) == 0, "%s: workspace must be 256-byte aligned", who);
    const size_t need = ws_layout(N, M, D, allow_tc).total;
    if (ws_bytes < need) {
        mpb_set_error("%s: workspace of %zu bytes is smaller than the %zu required", who, ws_bytes, need);
        return MPB_ENOMEM;
    }
    return MPB_OK;
}

MPB_API int mpb_knn_prepare(const double *samples, int64_t n_samples, int dim, const double *weights, void *workspace,
                            size_t workspace_bytes, void *stream)
{
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM, "mpb_knn_prepare: dim %d outside [1, %d]", dim, MPB_MAX_DIM);
    MPB_REQUIRE(n_samples >= 0 && n_samples < 2147483647LL, "mpb_knn_prepare: bad sample count");
    MPB_REQUIRE(n_samples == 0 || samples, "mpb_knn_prepare: NULL sample buffer");
    int rc = check_ws("mpb_knn_prepare", workspace, workspace_bytes, n_samples, 0, dim);
    if (rc != MPB_OK) return rc;
    return dispatch_prepare(dim, samples, n_samples, make_weights(weights, dim), workspace, true, (cudaStream_t)stream);
}

MPB_API int mpb_knn_search(const double *samples, int64_t n_samples, const double *queries, int64_t n_queries, int dim,
                           const double *weights, int k, double max_dist, void *workspace, size_t workspace_bytes,
                           int64_t *out_idx, double *out_dist, int64_t *stats, void *stream)
{
    int rc = check_common("mpb_knn_search", samples, n_samples, queries, n_queries, dim, k, out_idx);
    if (rc != MPB_OK) return rc;
    MPB_REQUIRE(max_dist >= 0., "mpb_knn_search: max_dist must be >= 0");
    rc = check_ws("mpb_knn_search", workspace, workspace_bytes, n_samples, n_queries, dim);
    if (rc != MPB_OK) return rc;
    if (n_queries == 0) return MPB_OK;
    return dispatch_search(dim, samples, n_samples, queries, n_queries, make_weights(weights, dim), k,
                           max_dist * max_dist, workspace, true, out_idx, out_dist, stats, (cudaStream_t)stream);
}

MPB_API int mpb_knn_tc_key_error(const double *samples, int64_t n_samples, const double *queries, int64_t n_queries,
                                 int dim, const double *weights, void *workspace, size_t workspace_bytes,
                                 unsigned long long *out, void *stream)
{
    MPB_REQUIRE(dim >= 1 && dim <= kTcMaxDim, "mpb_knn_tc_key_error: the tensor-core scan serves dim 1..%d", kTcMaxDim);
    MPB_REQUIRE(samples && queries && out && n_samples > 0 && n_queries > 0, "mpb_knn_tc_key_error: empty input");
    MPB_REQUIRE(n_samples < 2147483647LL && n_queries < 2147483647LL, "mpb_knn_tc_key_error: too many points");
    MPB_REQUIRE(use_tc(n_samples, dim), "mpb_knn_tc_key_error: the tensor-core scan does not serve %lld samples at "
                "dim %d under the current policy (MPB200_KNN_TC)", (long long)n_samples, dim);
    int rc = check_ws("mpb_knn_tc_key_error", workspace, workspace_bytes, n_samples, n_queries, dim);
    if (rc != MPB_OK) return rc;
    const Weights wt = make_weights(weights, dim);
    rc = dispatch_prepare(dim, samples, n_samples, wt, workspace, true, (cudaStream_t)stream);
    if (rc != MPB_OK) return rc;
#define CALL(DD) launch_tc_diag<DD>(samples, n_samples, queries, n_queries, wt, workspace, out, (cudaStream_t)stream)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    return MPB_EINVAL;
}

MPB_API int64_t mpb_knn_part_rows(int64_t n_samples, int dim, int part, int parts)
{
    if (dim < 1 || dim > kTcMaxDim || n_samples <= 0 || parts < 1 || part < 0 || part >= parts) return -1;
    if (!use_tc(n_samples, dim)) return -1;
    const KdPlan plan = tc_sample_plan(n_samples, dim);
    int64_t rows = 0;
    const int64_t blocks = kd_share_blocks(plan.leaves, part, parts);
    for (int64_t i = 0; i < blocks; ++i) {
        const int64_t l = kd_share_leaf(i, part, parts);
        rows += kd_leaf_start(l + 1, n_samples, plan.leaves) - kd_leaf_start(l, n_samples, plan.leaves);
    }
    return rows;
}

MPB_API int mpb_knn_search_part(const double *samples, int64_t n_samples, int dim, const double *weights, int k,
                                double max_dist, void *workspace, size_t workspace_bytes, int part, int parts,
                                int64_t *out_rows, int64_t *out_idx, double *out_dist, int64_t *stats, void *stream)
{
    int rc = check_common("mpb_knn_search_part", samples, n_samples, samples, n_samples, dim, k, out_idx);
    if (rc != MPB_OK) return rc;
    MPB_REQUIRE(max_dist >= 0., "mpb_knn_search_part: max_dist must be >= 0");
    MPB_REQUIRE(parts >= 1 && part >= 0 && part < parts, "mpb_knn_search_part: part %d of %d", part, parts);
    MPB_REQUIRE(k <= kSmallListMaxK && mpb_knn_part_rows(n_samples, dim, part, parts) >= 0,
                "mpb_knn_search_part: needs a set the tensor-core scan serves (mpb_knn_filter_kind) and k <= %d",
                kSmallListMaxK);
    rc = check_ws("mpb_knn_search_part", workspace, workspace_bytes, n_samples, n_samples, dim);
    if (rc != MPB_OK) return rc;
    return dispatch_search(dim, samples, n_samples, samples, n_samples, make_weights(weights, dim), k,
                           max_dist * max_dist, workspace, true, out_idx, out_dist, stats, (cudaStream_t)stream, part,
                           parts, out_rows);
}

MPB_API int mpb_knn_sample_order(const void *workspace, size_t workspace_bytes, int64_t n_samples, int dim,
                                 int32_t *out_order, void *stream)
{
    MPB_REQUIRE(dim >= 1 && dim <= MPB_MAX_DIM && n_samples > 0 && workspace && out_order, "mpb_knn_sample_order: bad argument");
    const WsLayout L = ws_layout(n_samples, 0, dim, true);
    if (L.tc_tiles == 0) {
        mpb_set_error("mpb_knn_sample_order: a set of %lld samples at dim %d is not spatially ordered (the tensor-core "
                      "scan does not serve it)", (long long)n_samples, dim);
        return MPB_EINVAL;
    }
    MPB_REQUIRE(workspace_bytes >= L.total, "mpb_knn_sample_order: workspace too small");
    MPB_CUDA_CHECK(cudaMemcpyAsync(out_order, reinterpret_cast<const char *>(workspace) + L.perm_off,
                                   (size_t)n_samples * sizeof(int32_t), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return MPB_OK;
}

MPB_API int mpb_knn_launch_count(void) { return g_last_launches; }

// prepare + search on one workspace; `allow_tc` = the workspace was sized with room for the tensor-core tiles
static int filter_on_workspace(const char *who, const double *samples, int64_t n_samples, const double *queries,
                               int64_t n_queries, int dim, const double *weights, int k, double max_dist,
                               void *workspace, size_t workspace_bytes, bool allow_tc, int64_t *out_idx,
                               double *out_dist, int64_t *stats, void *stream)
{
    int rc = check_common(who, samples, n_samples, queries, n_queries, dim, k, out_idx);
    if (rc != MPB_OK) return rc;
    MPB_REQUIRE(max_dist >= 0., "%s: max_dist must be >= 0", who);
    MPB_REQUIRE(n_samples < 2147483647LL, "%s: bad sample count", who);
    rc = check_ws(who, workspace, workspace_bytes, n_samples, n_queries, dim, allow_tc);
    if (rc != MPB_OK) return rc;
    const Weights wt = make_weights(weights, dim);
    rc = dispatch_prepare(dim, samples, n_samples, wt, workspace, allow_tc, (cudaStream_t)stream);
    if (rc != MPB_OK) return rc;
    if (n_queries == 0) return MPB_OK;
    return dispatch_search(dim, samples, n_samples, queries, n_queries, wt, k, max_dist * max_dist, workspace, allow_tc,
                           out_idx, out_dist, stats, (cudaStream_t)stream);
}

MPB_API int mpb_knn_f32filter(const double *samples, int64_t n_samples, const double *queries, int64_t n_queries,
                              int dim, const double *weights, int k, double max_dist, void *workspace,
                              size_t workspace_bytes, int64_t *out_idx, double *out_dist, int64_t *stats, void *stream)
{
    return filter_on_workspace("mpb_knn_f32filter", samples, n_samples, queries, n_queries, dim, weights, k, max_dist,
                               workspace, workspace_bytes, true, out_idx, out_dist, stats, stream);
}

// ------------------------------------------------------------------------------------------
// kNN on scratch memory that lives inside the library
// ------------------------------------------------------------------------------------------
// Some users of this ABI have no way to hand us a workspace: a planner thread inside a robot's
// control loop must not touch the driver's allocator (cudaMalloc takes a process-wide lock, may
// synchronise the device and has no latency bound); a process whose device memory was handed out
// in full at start-up to another pool (a simulator, an inference runtime) has nothing left to ask
// the driver for; and a thin foreign-language binding that wraps only the functions of this header
// owns no device allocator at all.  For them the filter's scratch -- header, packed fp32 tiles,
// fallback list -- sits in the fixed array below.  Every GPU that loads the library carries these
// MPB_KNN_STATIC_BYTES until the process exits, which is why the array is kept small.
static __device__ __align__(256) unsigned char g_knn_static_scratch[MPB_KNN_STATIC_BYTES];

// In host code `g_knn_static_scratch` is only a handle for the runtime, not an address a kernel could
// use; the device address is looked up once per device and remembered here.
static void *g_knn_static_ptr[64];
static std::mutex g_knn_static_mutex;

MPB_API int mpb_knn_f32filter_static(const double *samples, int64_t n_samples, const double *queries,
                                     int64_t n_queries, int dim, const double *weights, int k, double max_dist,
                                     int64_t *out_idx, double *out_dist, int64_t *stats, void *stream)
{
    int rc = check_common("mpb_knn_f32filter_static", samples, n_samples, queries, n_queries, dim, k, out_idx);
    if (rc != MPB_OK) return rc;
    MPB_REQUIRE(max_dist >= 0., "mpb_knn_f32filter_static: max_dist must be >= 0");
    // sized without the tensor-core tiles (96 bytes per sample would not leave room for much in this array)
    const size_t need = ws_layout(n_samples, n_queries, dim, false).total;
    if (need > (size_t)MPB_KNN_STATIC_BYTES) {
        mpb_set_error("mpb_knn_f32filter_static: %lld samples x %lld queries at dim %d need %zu bytes of scratch, "
                      "the built-in array holds %zu; use mpb_knn_f32filter with a caller-owned workspace",
                      (long long)n_samples, (long long)n_queries, dim, need, (size_t)MPB_KNN_STATIC_BYTES);
        return MPB_ENOMEM;
    }
    int dev = 0;
    MPB_CUDA_CHECK(cudaGetDevice(&dev));
    MPB_REQUIRE(dev >= 0 && dev < 64, "mpb_knn_f32filter_static: device ordinal %d not supported", dev);
    void *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_knn_static_mutex);
        if (g_knn_static_ptr[dev] == nullptr) {
            // first call on this device: resolve the array's device address (the device is current)
            void *p = nullptr;
            MPB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_knn_static_scratch));
            g_knn_static_ptr[dev] = p;
        }
        scratch = g_knn_static_ptr[dev];
    }
    // prepare re-initialises the header (bounding box, fallback counter) and rewrites every tile slot it
    // will read, so nothing left behind by an earlier call is ever consumed
    return filter_on_workspace("mpb_knn_f32filter_static", samples, n_samples, queries, n_queries, dim, weights, k,
                               max_dist, scratch, (size_t)MPB_KNN_STATIC_BYTES, false, out_idx, out_dist, stats, stream);
}
