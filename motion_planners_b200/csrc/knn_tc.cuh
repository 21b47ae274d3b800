// Tensor-core form of the kNN filter pass (included by knn.cu, inside its anonymous namespace).
//
// The filter key  ||x^||^2 - 2 q^.x^  of a (query, candidate) pair is one row-by-row product when every fp32
// value is split into two TF32 parts, v = v_hi + v_lo (v_hi = v rounded to 10 mantissa bits, v_lo = the rest
// rounded the same way; what is dropped is below 2^-22 |v|):
//
//   column    0..6     7      8..14     15     16..22    23
//   candidate x_hi[d]  n_hi   x_lo[d]   n_lo   x_hi[d]   0          n = fl32(||x^||^2)
//   query     q_hi[d]  1      q_hi[d]   1      q_lo[d]   0          q = -2 x^_q
//
// Products of two TF32 numbers are exact in fp32, so three tcgen05.mma kind::tf32 steps (K = 8 each) of a
// 128-query x 256-candidate tile give the key to within a few 2^-22 of its terms -- fp32-grade, which is what the
// exactness certificate needs (its eta is widened accordingly; the fp64 re-rank stays the arbiter).
// One CTA per 128 queries: thread t of the four epilogue warps IS query t (TMEM lane t), so its threshold is a
// plain register and the 256 keys of a tile come to it through tcgen05.ld; a producer thread streams candidate
// tiles (prepared in the operand's canonical shared-memory image, so a tile is one 1-D bulk copy) through a
// ring, one thread issues the MMAs into a double-buffered TMEM accumulator (2 x 256 columns).
#pragma once

constexpr int kTcK = 24;                       // operand columns (D <= 7)
constexpr int kTcTile = 256;                   // candidates per tile = MMA N
constexpr int kTcQueries = 128;                // queries per CTA = MMA M = TMEM lanes
#ifndef MPB_TC_SLEEP_ROLE
#define MPB_TC_SLEEP_ROLE 40      // ns between polls of the producer / MMA threads
#endif
#ifndef MPB_TC_SLEEP_EPI
#define MPB_TC_SLEEP_EPI 32       // ns between polls of an epilogue warp
#endif
#ifndef MPB_TC_PARTS
#define MPB_TC_PARTS 4
#endif
#ifndef MPB_TC_STAGES
#define MPB_TC_STAGES 4
#endif
constexpr int kTcStages = MPB_TC_STAGES;       // shared-memory ring depth
constexpr int kTcKeep = 16;                    // survivors per list
constexpr int kTcParts = MPB_TC_PARTS;          // epilogue warps per TMEM lane quarter
constexpr int kTcEpiWarps = 4 * kTcParts;      // warp w: lane quarter w & 3, part w >> 2 (8 / kTcParts chunks of 32 keys per tile)
constexpr int kTcThreads = 32 * (kTcEpiWarps + 2);
#ifndef MPB_TC_SHARED_LIST
#define MPB_TC_SHARED_LIST 1      // 1: the warps of a lane quarter share one list per query (per-query lock)
#endif
constexpr int kTcLists = MPB_TC_SHARED_LIST ? 1 : kTcParts;   // lists per query
constexpr int kTcSurvivors = kTcKeep * kTcLists;   // survivors per query handed to the exact re-rank
constexpr uint32_t kTcLbo = 128;               // bytes between the two 16-byte k-chunks of a k-step
constexpr uint32_t kTcSbo = (kTcK / 4) * 128;  // bytes between 8-row groups
constexpr int kTcTileBytes = (kTcTile / 8) * (int)kTcSbo;      // 24576
constexpr int kTcQueryBytes = (kTcQueries / 8) * (int)kTcSbo;  // 12288
constexpr int kTcMaxDim = 7;

// byte offset of (row r, column kk) in an operand image: K-major, no swizzle (8 x 16-byte core matrices)
__host__ __device__ inline uint32_t tc_offset(int r, int kk)
{
    return (uint32_t)(r >> 3) * kTcSbo + (uint32_t)(kk >> 2) * kTcLbo + (uint32_t)(r & 7) * 16u + (uint32_t)(kk & 3) * 4u;
}

// round to TF32 (10 explicit mantissa bits), nearest-even; the tensor core ignores the 13 low bits
__device__ __forceinline__ float tf32_rn(float x)
{
    uint32_t u = __float_as_uint(x);
    u += 0x00000fffu + ((u >> 13) & 1u);
    u &= 0xffffe000u;
    return __uint_as_float(u);
}

// shared-memory matrix descriptor: start address, leading / stride byte offsets (all >> 4), sm_100 version bit
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((kTcLbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((kTcSbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;   // layout type 0 = no swizzle
    return d;
}

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// bounded wait of a single-thread role: sleeps between polls (the polling thread shares a scheduler with an
// epilogue warp), traps instead of hanging
__device__ __forceinline__ void mbar_wait_backoff(uint64_t *bar, uint32_t parity)
{
    for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins) {
        __nanosleep(MPB_TC_SLEEP_ROLE);
        if (spins > (1u << 26)) __trap();
    }
}

// the same for a whole converged warp (the vote keeps the code after it warp-uniform); sixteen warps polling
// without a pause took a third of the SM's issue slots
__device__ __forceinline__ void mbar_wait_warp_backoff(uint64_t *bar, uint32_t parity)
{
    for (uint32_t spins = 0; !__all_sync(0xffffffffu, mbar_try_wait(bar, parity)); ++spins) {
        __nanosleep(MPB_TC_SLEEP_EPI);
        if (spins > (1u << 26)) __trap();
    }
}

// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t (&v)[32], uint32_t taddr)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}

// ---- candidate tiles in the operand image ------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256)
    knn_tc_pack_kernel(const double *__restrict__ x, int64_t N, int64_t Npad, Weights wt, const KnnHeader *__restrict__ h,
                       unsigned char *__restrict__ tiles)
{
    static_assert(D <= kTcMaxDim, "three k-steps hold up to seven coordinates");
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= Npad) return;
    float row[kTcK];
#pragma unroll
    for (int i = 0; i < kTcK; ++i) row[i] = 0.f;
    if (j < N) {
        double n2 = 0.;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double e = d_mul(wt.w[d], x[j * D + d]);
            const float f = (float)d_sub(e, h->center[d]);
            n2 += (double)f * (double)f;
            const float hi = tf32_rn(f);
            row[d] = hi;
            row[8 + d] = tf32_rn(f - hi);      // f - hi is exact in fp32
            row[16 + d] = hi;
        }
        const float n = (float)n2;
        const float nh = tf32_rn(n);
        row[7] = nh;
        row[15] = tf32_rn(n - nh);
    } else {
        row[7] = 3.0e38f;                      // padding never qualifies (finite: no inf - inf inside the MMA)
    }
    const int64_t t = j / kTcTile;
    const int r = (int)(j - t * kTcTile);
    unsigned char *base = tiles + (size_t)t * kTcTileBytes;
#pragma unroll
    for (int c = 0; c < kTcK / 4; ++c)
        *reinterpret_cast<float4 *>(base + tc_offset(r, 4 * c)) =
            make_float4(row[4 * c], row[4 * c + 1], row[4 * c + 2], row[4 * c + 3]);
}

struct TcSmem {
    static constexpr int A_OFF = 0;
    static constexpr int B_OFF = A_OFF + kTcQueryBytes;
    static constexpr int LK_OFF = B_OFF + kTcStages * kTcTileBytes;
    static constexpr int LI_OFF = LK_OFF + kTcSurvivors * kTcQueries * 4;
    static constexpr int SCR_OFF = LI_OFF + kTcSurvivors * kTcQueries * 4;   // [epilogue warp][32 lanes][33]
    static constexpr int THR_OFF = SCR_OFF + kTcEpiWarps * 32 * 33 * 4;        // [part][128] published thresholds
    static constexpr int LOCK_OFF = THR_OFF + kTcLists * kTcQueries * 4;       // [128] list locks (shared-list build)
    static constexpr int BAR_OFF = LOCK_OFF + kTcQueries * 4;
    static constexpr int BYTES = BAR_OFF + 256;
};

// ---- the scan: 128 queries per CTA against every candidate tile --------------------------------------
// Warps 0..15: epilogue (lane = query of TMEM lane quarter w & 3; part w >> 2 takes two of a tile's eight 32-key
// chunks; the four warps of a quarter insert into one 16-entry list per query, one warp at a time per query).
// Warp 16, one lane: MMA issue.  Warp 17, one lane: tile producer.
template <int D>
__global__ void __launch_bounds__(kTcThreads, 1)
    knn_tc_kernel(const double *__restrict__ qs, int64_t M, Weights wt, const unsigned char *__restrict__ tiles,
                  int n_tiles, const KnnHeader *__restrict__ h, int *__restrict__ surv_idx, float *__restrict__ surv_thr)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char *sA = smem + TcSmem::A_OFF;
    unsigned char *sB = smem + TcSmem::B_OFF;
    float *lk = reinterpret_cast<float *>(smem + TcSmem::LK_OFF);      // [part][128][kTcKeep] sorted keys per query
    int *li = reinterpret_cast<int *>(smem + TcSmem::LI_OFF);          // [part][128][kTcKeep] their candidate indices
    float *scr = reinterpret_cast<float *>(smem + TcSmem::SCR_OFF);
    float *sthr = reinterpret_cast<float *>(smem + TcSmem::THR_OFF);
    int *locks = reinterpret_cast<int *>(smem + TcSmem::LOCK_OFF);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + TcSmem::BAR_OFF);
    uint64_t *empty = full + kTcStages;
    uint64_t *tfull = empty + kTcStages;
    uint64_t *tempty = tfull + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t q0 = (int64_t)blockIdx.x * kTcQueries;

    if (tid < kTcQueries) {            // this thread's query row of the A operand, and its empty list
        int64_t q = q0 + tid;
        if (q >= M) q = M - 1;
        float row[kTcK];
#pragma unroll
        for (int i = 0; i < kTcK; ++i) row[i] = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float f = -2.f * (float)d_sub(d_mul(wt.w[d], qs[q * D + d]), h->center[d]);
            const float hi = tf32_rn(f);
            row[d] = hi;
            row[8 + d] = hi;
            row[16 + d] = tf32_rn(f - hi);
        }
        row[7] = 1.f;
        row[15] = 1.f;
#pragma unroll
        for (int c = 0; c < kTcK / 4; ++c)
            *reinterpret_cast<float4 *>(sA + tc_offset(tid, 4 * c)) =
                make_float4(row[4 * c], row[4 * c + 1], row[4 * c + 2], row[4 * c + 3]);
    }
    for (int i = tid; i < kTcSurvivors * kTcQueries; i += kTcThreads) {
        lk[i] = INFINITY;
        li[i] = -1;
    }
    for (int i = tid; i < kTcLists * kTcQueries; i += kTcThreads) sthr[i] = INFINITY;
    for (int i = tid; i < kTcQueries; i += kTcThreads) locks[i] = 0;
    if (tid == 0) {
        for (int s = 0; s < kTcStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull[b], 1);
            mbar_init(&tempty[b], kTcEpiWarps);
        }
        mbar_fence_init();
    }
    if (warp == kTcEpiWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the query rows were written by plain stores
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;

    if (warp == kTcEpiWarps + 1) {
        // ---- producer: one bulk copy per tile
        if (lane == 0) {
            for (int i = 0; i < n_tiles; ++i) {
                const int s = i % kTcStages;
                if (i >= kTcStages) mbar_wait_backoff(&empty[s], (uint32_t)((i / kTcStages - 1) & 1));
                mbar_arrive_expect_tx(&full[s], kTcTileBytes);
                bulk_copy_g2s(sB + (size_t)s * kTcTileBytes, tiles + (size_t)i * kTcTileBytes, kTcTileBytes, &full[s]);
            }
        }
    } else if (warp == kTcEpiWarps) {
        // ---- MMA issue: three k-steps per tile into accumulator i & 1
        if (lane == 0) {
            // D = f32, A = B = tf32, both K-major, N = 256, M = 128
            constexpr uint32_t idesc =
                (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTcTile >> 3) << 17) | ((uint32_t)(kTcQueries >> 4) << 24);
            const uint32_t a_addr = smem_u32(sA);
            for (int i = 0; i < n_tiles; ++i) {
                const int s = i % kTcStages, b = i & 1;
                mbar_wait_backoff(&full[s], (uint32_t)((i / kTcStages) & 1));
                if (i >= 2) mbar_wait_backoff(&tempty[b], (uint32_t)((i / 2 - 1) & 1));
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t b_addr = smem_u32(sB + (size_t)s * kTcTileBytes);
                const uint32_t d_addr = tmem + (uint32_t)(b * kTcTile);
#pragma unroll
                for (int ks = 0; ks < kTcK / 8; ++ks) {
                    const uint64_t da = tc_desc(a_addr + ks * 2 * kTcLbo);
                    const uint64_t db = tc_desc(b_addr + ks * 2 * kTcLbo);
                    const uint32_t acc = ks > 0 ? 1u : 0u;
                    asm volatile(
                        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                        " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}"
                        ::"r"(d_addr), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
                }
                // both commits track every MMA issued so far: the ring slot is free, the accumulator is ready
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                             ::"r"(smem_u32(&empty[s])) : "memory");
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                             ::"r"(smem_u32(&tfull[b])) : "memory");
            }
        }
    } else {
        // ---- epilogue: lane = query; this warp's part of every tile is two chunks of 32 keys
        const int quarter = warp & 3, part = warp >> 2;
        const int q = quarter * 32 + lane;
        float thr = INFINITY;
        const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
        const int my_list = MPB_TC_SHARED_LIST ? 0 : part;
        float *lkp = lk + my_list * kTcKeep * kTcQueries;
        int *lip = li + my_list * kTcKeep * kTcQueries;
        // The 32 keys a lane just received are examined with one min-tree.  When some lane's minimum beats its
        // threshold, the flagged lanes park their keys in the warp's scratch rows, and each flagged lane is then
        // served by the WHOLE warp: lane j reads key j of that row, one ballot finds the takers, and each taker is
        // inserted into the query's shared-memory list by lanes 0..15 side by side.  (A lane-private walk over
        // the 32 keys makes every lane of the warp wait for it, and at 1M candidates one chunk in five has a
        // flagged lane: profiles/r1_summary.md.)
        float *rows = scr + warp * 32 * 33;
        auto examine = [&](const uint32_t (&v)[32], int cbase) {
            float g[4];        // minima of the four groups of eight keys
#pragma unroll
            for (int gi = 0; gi < 4; ++gi) {
                g[gi] = __uint_as_float(v[8 * gi]);
#pragma unroll
                for (int j = 1; j < 8; ++j) g[gi] = fminf(g[gi], __uint_as_float(v[8 * gi + j]));
            }
            const float m = fminf(fminf(g[0], g[1]), fminf(g[2], g[3]));
            unsigned flagged = __ballot_sync(FULL_MASK, m < thr);
            if (flagged == 0) return;
            // A list that already holds 16 keys below t proves that nothing at or above t is among the query's
            // 16 best, whichever list it would go to: before any work is spent on a flagged chunk, every lane
            // lowers its threshold to the smallest one published for its query (read without synchronisation:
            // a stale value is only a larger one) and the chunk is tested again.
#pragma unroll
            for (int pp = 0; pp < kTcLists; ++pp) thr = fminf(thr, sthr[pp * kTcQueries + q]);
            flagged = __ballot_sync(FULL_MASK, m < thr);
            if (flagged == 0) return;
            // only the groups of eight that hold a key below the threshold are parked
            unsigned gm = 0;
#pragma unroll
            for (int gi = 0; gi < 4; ++gi) {
                if (g[gi] < thr) {
                    gm |= 1u << gi;
#pragma unroll
                    for (int j = 0; j < 8; ++j) rows[lane * 33 + 8 * gi + j] = __uint_as_float(v[8 * gi + j]);
                }
            }
            __syncwarp();
            while (flagged) {
                const int L = __ffs(flagged) - 1;
                flagged &= flagged - 1;
                float thrL = __shfl_sync(FULL_MASK, thr, L);
                const int qL = quarter * 32 + L;
#if MPB_TC_SHARED_LIST
                // the four warps of this lane quarter insert into the same list: one warp at a time per query
                if (lane == 0) {
                    for (uint32_t spins = 0; atomicCAS(&locks[qL], 0, 1) != 0; ++spins)
                        if (spins > (1u << 24)) __trap();
                }
                __syncwarp();
                __threadfence_block();
                thrL = fminf(thrL, lkp[qL * kTcKeep + kTcKeep - 1]);
#endif
                const unsigned gmL = __shfl_sync(FULL_MASK, gm, L);
                const float mine = rows[L * 33 + lane];      // stale where the group was not parked: masked out
                unsigned takers = __ballot_sync(FULL_MASK, ((gmL >> (lane >> 3)) & 1u) && mine < thrL);
                while (takers) {
                    const int j = __ffs(takers) - 1;
                    takers &= takers - 1;
                    const float key = __shfl_sync(FULL_MASK, mine, j);
                    if (key < thrL) {
                        const bool in_list = lane < kTcKeep;
                        const float e = in_list ? lkp[qL * kTcKeep + lane] : INFINITY;
                        const int ei = in_list ? lip[qL * kTcKeep + lane] : -1;
                        // equal keys: the newcomer goes after the incumbents
                        const int pos = __popc(__ballot_sync(FULL_MASK, in_list && e <= key));
                        if (lane >= pos && lane < kTcKeep - 1) {
                            lkp[qL * kTcKeep + lane + 1] = e;
                            lip[qL * kTcKeep + lane + 1] = ei;
                        }
                        if (lane == pos) {
                            lkp[qL * kTcKeep + pos] = key;
                            lip[qL * kTcKeep + pos] = cbase + j;
                        }
                        __syncwarp();
                        thrL = lkp[qL * kTcKeep + kTcKeep - 1];
                    }
                }
                if (lane == L) {
                    thr = thrL;
                    sthr[my_list * kTcQueries + q] = thrL;
                }
#if MPB_TC_SHARED_LIST
                __syncwarp();
                __threadfence_block();
                if (lane == 0) atomicExch(&locks[qL], 0);
#endif
            }
            __syncwarp();
        };
        for (int i = 0; i < n_tiles; ++i) {
            const int b = i & 1;
            mbar_wait_warp_backoff(&tfull[b], (uint32_t)((i / 2) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            constexpr int CH = (kTcTile / 32) / kTcParts;      // chunks per warp per tile
            static_assert(CH % 2 == 0, "chunks are processed in pairs");
            uint32_t va[32], vb[32];
            const int c0 = CH * part;
            const uint32_t tile_base = lane_base + (uint32_t)(b * kTcTile + c0 * 32);
            if constexpr (CH == 2) {
                // both chunks go to registers first and the accumulator is handed back at once: the next MMAs
                // never wait for a warp that is busy inserting
                __syncwarp();
                tmem_ld32(va, tile_base);
                tmem_ld32(vb, tile_base + 32u);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[b]);
                examine(va, i * kTcTile + c0 * 32);
                examine(vb, i * kTcTile + (c0 + 1) * 32);
            } else {
                // two register sets: the next chunk's load is in flight while the current chunk is examined; the
                // accumulator is handed back as soon as this warp's last chunk is in registers
                __syncwarp();
                tmem_ld32(va, tile_base);
#pragma unroll 1
                for (int c = 0; c < CH; c += 2) {
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    __syncwarp();
                    tmem_ld32(vb, tile_base + (uint32_t)((c + 1) * 32));
                    examine(va, i * kTcTile + (c0 + c) * 32);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    __syncwarp();
                    if (c + 2 < CH) {
                        tmem_ld32(va, tile_base + (uint32_t)((c + 2) * 32));
                    } else {
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        if (lane == 0) mbar_arrive(&tempty[b]);
                    }
                    examine(vb, i * kTcTile + (c0 + c + 1) * 32);
                }
            }
        }
#if MPB_TC_SHARED_LIST
        asm volatile("bar.sync 1, %0;" ::"r"(kTcEpiWarps * 32) : "memory");
#endif
        // this part's survivors of the query (ascending fp32 key) and its final threshold
        const int64_t qq = q0 + q;
        if (qq < M && (!MPB_TC_SHARED_LIST || part == 0)) {
#pragma unroll
            for (int p = 0; p < kTcKeep; ++p) surv_idx[qq * kTcSurvivors + my_list * kTcKeep + p] = lip[q * kTcKeep + p];
            surv_thr[qq * kTcLists + my_list] = MPB_TC_SHARED_LIST ? lkp[q * kTcKeep + kTcKeep - 1] : thr;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == kTcEpiWarps) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

// ---- exact re-rank of the survivors of every query, and the certificate ---------------------------------
// One warp per query, kTcSurvivors / 32 survivors per lane (kTcParts lists of kTcKeep).  Same argument as in
// knn_filter_kernel: every sample outside list p has computed key >= T_p, so every sample outside all lists has
// computed key >= T = min_p T_p, hence exact squared distance >= (sqrt(T - eta + |q^|^2) - delta)^2; here eta also
// covers the TF32 split (three dropped or rounded 2^-22 terms) and the tensor core's accumulation, budgeted at
// 2^-22 per term.
template <int D>
__global__ void __launch_bounds__(256)
    knn_tc_rerank_kernel(const double *__restrict__ x, int64_t N, const double *__restrict__ qs, int64_t M, Weights wt,
                         int k, double r2, KnnHeader *__restrict__ h, const int *__restrict__ surv_idx,
                         const float *__restrict__ surv_thr, int *__restrict__ fb_list, int64_t *__restrict__ out_idx,
                         double *__restrict__ out_dist)
{
    constexpr int NU = (kTcSurvivors + 31) / 32;   // survivors per lane
    const int lane = threadIdx.x & 31;
    const int64_t q = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= M) return;                         // whole warps leave together
    double eq[D];
    double nq2 = 0.;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        eq[d] = d_mul(wt.w[d], qs[q * D + d]);
        const double qh = (double)(float)d_sub(eq[d], h->center[d]);
        nq2 += qh * qh;
    }
    int ci[NU];
    double S[NU];
#pragma unroll
    for (int u = 0; u < NU; ++u) {
        ci[u] = (u * 32 + lane < kTcSurvivors) ? surv_idx[q * kTcSurvivors + u * 32 + lane] : -1;
        if (ci[u] >= N) ci[u] = -1;             // a padding slot (only when a part saw fewer than 16 real candidates)
        S[u] = INFINITY;
        if (ci[u] >= 0) {
            double ex[D];
#pragma unroll
            for (int d = 0; d < D; ++d) ex[d] = d_mul(wt.w[d], x[(int64_t)ci[u] * D + d]);
            S[u] = sqdist_kdtree<D>(eq, ex);
        }
    }
    int rank[NU];
#pragma unroll
    for (int u = 0; u < NU; ++u) rank[u] = 0;
    for (int o = 0; o < 32; ++o) {
#pragma unroll
        for (int w = 0; w < NU; ++w) {
            const double So = __shfl_sync(FULL_MASK, S[w], o);
            const int io = __shfl_sync(FULL_MASK, ci[w], o);
#pragma unroll
            for (int u = 0; u < NU; ++u) rank[u] += ((So < S[u]) || (So == S[u] && io < ci[u])) ? 1 : 0;
        }
    }
    // k-th smallest exact key among the survivors (inf when fewer than k exist)
    double Sk = INFINITY;
#pragma unroll
    for (int u = 0; u < NU; ++u) {
        const unsigned mk = __ballot_sync(FULL_MASK, ci[u] >= 0 && rank[u] == k - 1);
        const double cand = __shfl_sync(FULL_MASK, S[u], mk ? (__ffs(mk) - 1) : 0);
        if (mk) Sk = cand;
    }
    int nvalid = 0;
#pragma unroll
    for (int u = 0; u < NU; ++u) nvalid += __popc(__ballot_sync(FULL_MASK, ci[u] >= 0));
    float T32 = INFINITY;
#pragma unroll
    for (int p = 0; p < kTcLists; ++p) {
        const float tp = surv_thr[q * kTcLists + p];
        if (tp < 1.0e38f) T32 = fminf(T32, tp);  // a list that never filled (or filled with padding) excludes nothing
    }
    const double T = (double)T32;
    bool ok = isinf(T) && T > 0;                 // no list ever filled: every sample is a survivor
    if (!ok) {
        const double R = h->R;
        constexpr double U24 = 5.9604644775390625e-08;   // 2^-24: rounding of the centred fp32 coordinates
        constexpr double U22 = 2.384185791015625e-07;    // 2^-22: TF32 split and accumulation budget per term
        const double nq = sqrt(nq2);
        const double delta = U24 * (R + nq) * 1.0001 + 1e-30;
        const double eta = (double)(kTcK + 8) * U22 * 1.01 * (R * R + 2.0 * nq * R) + 1e-37;
        double slow = T - eta + nq2;
        slow -= 1e-12 * (fabs(T) + eta + nq2);
        double bound = 0.;
        if (slow > 0.) {
            const double rho = sqrt(slow) * (1.0 - 1e-15) - delta;
            if (rho > 0.) bound = rho * rho * (1.0 - 1e-12);
        }
        ok = (Sk < bound) || (r2 <= bound);
    }
    if (ok) {
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            if (ci[u] >= 0 && rank[u] < k) {
                const bool keep = S[u] < r2;
                out_idx[q * k + rank[u]] = keep ? (int64_t)ci[u] : -1;
                if (out_dist) out_dist[q * k + rank[u]] = keep ? __dsqrt_rn(S[u]) : INFINITY;
            }
        }
        if (lane < k && lane >= nvalid) {
            out_idx[q * k + lane] = -1;
            if (out_dist) out_dist[q * k + lane] = INFINITY;
        }
    } else if (lane == 0) {
        const int pos = atomicAdd(&h->fb_count, 1);
        fb_list[pos] = (int)q;
    }
}
