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
//
// Which tiles a CTA scans: samples and queries are first brought into the leaf order of a balanced k-d partition
// (kdorder.cuh), so a tile is a compact cell with a known fp32 bounding box and the 128 queries of a CTA are
// neighbours in space.  A CTA visits its nearest tile first, then, in rounds, the tiles whose box lies closer to
// the block of queries than the largest current k-th distance in the block -- nearest box first, re-checked
// against the shrinking distances just before each copy is issued.  A skipped tile cannot hold a neighbour of any
// query of the block, and the exactness certificate (knn_tc_rerank_kernel) is unchanged: the results stay bit
// for bit those of the full scan.  At 1M uniform 7-DoF samples a CTA reads about 6 % of the tiles.
#pragma once

#ifdef MPB_TC_STATS
#define TC_STAT(i, v) atomicAdd(&const_cast<KnnHeader *>(h)->dbg[i], (unsigned long long)(v))
#define TC_CNT(v) ++v
#define TC_CLK(var) const long long var = clock64()
#define TC_ACC(acc, since) acc += clock64() - (since)
#else
#define TC_CNT(v) ((void)0)
#define TC_CLK(var) ((void)0)
#define TC_ACC(acc, since) ((void)0)
#define TC_STAT(i, v) ((void)0)
#endif

constexpr int kTcK = 24;                       // operand columns (D <= 7)
constexpr int kTcTile = 256;                   // candidates per tile = MMA N
constexpr int kTcQueries = 128;                // queries per CTA = MMA M = TMEM lanes
#ifndef MPB_TC_SLEEP_ROLE
#define MPB_TC_SLEEP_ROLE 40      // ns between polls of the producer / MMA threads
#endif
#ifndef MPB_TC_SLEEP_EPI
#define MPB_TC_SLEEP_EPI 32       // ns between polls of an epilogue warp
#endif
#ifndef MPB_TC_WAIT_HINT
#define MPB_TC_WAIT_HINT 4000    // ns an epilogue warp may stay suspended inside one mbarrier.try_wait
#endif
#ifndef MPB_TC_MERGER_SLEEP
#define MPB_TC_MERGER_SLEEP 0     // ns the threshold merger pauses between lane quarters
#endif
#ifndef MPB_TC_REFRESH_EVERY
#define MPB_TC_REFRESH_EVERY 0    // 1: every epilogue thread re-reads its query's published threshold at every tile
#endif
#ifndef MPB_TC_BULK_MIN
#define MPB_TC_BULK_MIN 5         // flagged lanes of a chunk from which the whole warp takes the sorting-network path
#endif
#ifndef MPB_TC_FIRST_BOUND
#define MPB_TC_FIRST_BOUND 0.5f
#endif
#ifndef MPB_TC_PARTS
#define MPB_TC_PARTS 4
#endif
#ifndef MPB_TC_STAGES
#define MPB_TC_STAGES 3
#endif
constexpr int kTcStages = MPB_TC_STAGES;       // shared-memory ring depth
constexpr int kTcKeep = 16;                    // survivors per list
constexpr int kTcParts = MPB_TC_PARTS;          // epilogue warps per TMEM lane quarter
constexpr int kTcEpiWarps = 4 * kTcParts;      // warp w: lane quarter w & 3, part w >> 2 (8 / kTcParts chunks of 32 keys per tile)
constexpr int kTcThreads = 32 * (kTcEpiWarps + 3);   // + MMA issue, tile producer, threshold merger
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
constexpr int kTcListCap = 512;                 // candidate tiles planned per round (shared-memory list)
constexpr int kTcHalf = kTcTile / 2;            // points per k-d leaf: a tile is two leaves, a query block one
static_assert(kTcHalf == kTcQueries, "a query block of the sample set itself is one half tile");

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
__device__ __forceinline__ bool mbar_try_wait_hint(uint64_t *bar, uint32_t parity, uint32_t ns)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(ns)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_warp_backoff(uint64_t *bar, uint32_t parity)
{
    // the try_wait itself suspends the thread (up to the hinted time) and is woken by the phase change
    for (uint32_t spins = 0; !__all_sync(0xffffffffu, mbar_try_wait_hint(bar, parity, MPB_TC_WAIT_HINT)); ++spins) {
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



// ---- sorting networks over a thread's registers (every index is a compile-time constant after unrolling) ------
// compare-exchange on keys alone
#define TC_CE(x, y)                     \
    {                                   \
        const float lo_ = fminf(x, y);  \
        const float hi_ = fmaxf(x, y);  \
        x = lo_;                        \
        y = hi_;                        \
    }
// ascending bitonic sort of a[OFF .. OFF + N)
template <int OFF, int N, int LEN>
__device__ __forceinline__ void tc_sort_asc(float (&a)[LEN])
{
#pragma unroll
    for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                const int l = i ^ j;
                if (l > i) {
                    if ((i & k) == 0) {
                        TC_CE(a[OFF + i], a[OFF + l]);
                    } else {
                        TC_CE(a[OFF + l], a[OFF + i]);
                    }
                }
            }
        }
    }
}
// a[0 .. N) is bitonic: make it ascending (keys alone)
template <int N, int LEN>
__device__ __forceinline__ void tc_merge_asc(float (&a)[LEN])
{
#pragma unroll
    for (int j = N >> 1; j > 0; j >>= 1) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const int l = i ^ j;
            if (l > i) TC_CE(a[i], a[l]);
        }
    }
}
// the same with a payload riding along (equal keys keep their order of arrival: the lower position wins)
template <int N>
__device__ __forceinline__ void tc_merge_asc_kv(float (&a)[N], int (&b)[N])
{
#pragma unroll
    for (int j = N >> 1; j > 0; j >>= 1) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const int l = i ^ j;
            if (l > i) {
                const bool sw = a[l] < a[i];
                const float ka = sw ? a[l] : a[i], kb = sw ? a[i] : a[l];
                const int va = sw ? b[l] : b[i], vb = sw ? b[i] : b[l];
                a[i] = ka;
                a[l] = kb;
                b[i] = va;
                b[l] = vb;
            }
        }
    }
}

// ---- candidate tiles in the operand image ------------------------------------------------------------
// One block per tile.  Tile t holds k-d leaves 2t and 2t+1 of the sample order (at most 128 points each, slot
// 128 * half + i); unused slots carry a huge norm and slot_idx = -1.  The block also writes the tile's bounding
// box over the centred fp32 coordinates it packs (exactly the values the scan multiplies), plane-major:
// tile_box[d * stride + t] = min, tile_box[(8 + d) * stride + t] = max.
template <int D>
__global__ void __launch_bounds__(kTcTile)
    knn_tc_pack_kernel(const double *__restrict__ x, int64_t N, int64_t leaves, const uint32_t *__restrict__ perm,
                       Weights wt, const KnnHeader *__restrict__ h, unsigned char *__restrict__ tiles,
                       int *__restrict__ slot_idx, float *__restrict__ tile_box, int box_stride)
{
    static_assert(D <= kTcMaxDim, "three k-steps hold up to seven coordinates");
    __shared__ float red[kTcTile / 32][2][8];
    const int64_t t = blockIdx.x;
    const int r = threadIdx.x, lane = r & 31, warp = r >> 5;
    const int64_t leaf = 2 * t + (r >> 7);
    const int i = r & (kTcHalf - 1);
    int64_t j = -1;
    if (leaf < leaves) {
        const int64_t start = kd_leaf_start(leaf, N, leaves);
        const int64_t cnt = kd_leaf_start(leaf + 1, N, leaves) - start;
        // slot i of the half holds leaf position 37 i mod 128: the 32-column chunks an epilogue warp reads are
        // then fair samples of the leaf instead of slices of its (sorted) last coordinate
        const int src = (i * 37) & (kTcHalf - 1);
        if (src < cnt) j = (int64_t)perm[start + src];
    }
    float row[kTcK], f[D];
#pragma unroll
    for (int c = 0; c < kTcK; ++c) row[c] = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) f[d] = 0.f;
    if (j >= 0) {
        double n2 = 0.;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const double e = d_mul(wt.w[d], x[j * D + d]);
            f[d] = (float)d_sub(e, h->center[d]);
            n2 += (double)f[d] * (double)f[d];
            const float hi = tf32_rn(f[d]);
            row[d] = hi;
            row[8 + d] = tf32_rn(f[d] - hi);      // f - hi is exact in fp32
            row[16 + d] = hi;
        }
        const float n = (float)n2;
        const float nh = tf32_rn(n);
        row[7] = nh;
        row[15] = tf32_rn(n - nh);
    } else {
        row[7] = 3.0e38f;                          // padding never qualifies (finite: no inf - inf inside the MMA)
    }
    unsigned char *base = tiles + (size_t)t * kTcTileBytes;
#pragma unroll
    for (int c = 0; c < kTcK / 4; ++c)
        *reinterpret_cast<float4 *>(base + tc_offset(r, 4 * c)) =
            make_float4(row[4 * c], row[4 * c + 1], row[4 * c + 2], row[4 * c + 3]);
    slot_idx[t * kTcTile + r] = (int)j;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        float lo = (j >= 0) ? f[d] : INFINITY, hi = (j >= 0) ? f[d] : -INFINITY;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(FULL_MASK, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(FULL_MASK, hi, o));
        }
        if (lane == 0) {
            red[warp][0][d] = lo;
            red[warp][1][d] = hi;
        }
    }
    __syncthreads();
    if (r < 16 && (r & 7) < D) {
        const int side = r >> 3, d = r & 7;
        float v = red[0][side][d];
#pragma unroll
        for (int w = 1; w < kTcTile / 32; ++w) v = side ? fmaxf(v, red[w][1][d]) : fminf(v, red[w][0][d]);
        tile_box[(size_t)r * box_stride + t] = v;
    }
}

// query blocks of a query set that is not the sample set: block g = leaf g of the queries' own k-d order
__global__ void __launch_bounds__(kTcQueries)
    knn_tc_qorder_kernel(int64_t M, int64_t leaves, const uint32_t *__restrict__ perm, int *__restrict__ qorder)
{
    const int64_t g = blockIdx.x;
    const int64_t start = kd_leaf_start(g, M, leaves);
    const int64_t cnt = kd_leaf_start(g + 1, M, leaves) - start;
    const int i = threadIdx.x;
    qorder[g * kTcQueries + i] = (i < cnt) ? (int)perm[start + i] : -1;
}

struct TcCtl {                 // per-CTA control block in shared memory
    float qlo[8], qhi[8];      // bounding box of the block's centred fp32 queries
    float red_f[32];
    unsigned long long red_u[32];
    unsigned long long lo_key; // every tile with (lower bound, index) <= lo_key has been handled
    int slot_tile[kTcStages];  // tile in each ring slot (-1: end of round)
    int acc_tile[2];           // tile in each accumulator (-1: end of round)
    int list_n;                // tiles planned for the coming round
    int list_all;              // candidates counted by the last collection pass (may exceed the list)
    int home;                  // the tile nearest to the block
    int cursor;                // tiles below this index were all visited (rounds before any threshold existed)
    int has_lo;
    int final_round;           // the planned round covers everything that is left
    int round_over;            // the epilogue has reached the end of the round (tells the threshold merger to stop)
    float rmax2;
};

struct TcSmem {
    static constexpr int A_OFF = 0;
    static constexpr int B_OFF = A_OFF + kTcQueryBytes;
    static constexpr int PK_OFF = B_OFF + kTcStages * kTcTileBytes;            // [epilogue warp][32 lanes][kTcKeep + 1] keys
    static constexpr int PI_OFF = PK_OFF + (kTcKeep + 1) * kTcEpiWarps * 32 * 4;   // the same shape: their candidate slots
    static constexpr int SCR_OFF = PI_OFF + (kTcKeep + 1) * kTcEpiWarps * 32 * 4;  // [epilogue warp][32 lanes][33]
    static constexpr int THR_OFF = SCR_OFF + kTcEpiWarps * 32 * 33 * 4;        // [128] the queries' published thresholds
    static constexpr int NQ_OFF = THR_OFF + kTcQueries * 4;                    // [128] |q^|^2, rounded up
    static constexpr int LIST_OFF = NQ_OFF + kTcQueries * 4;                   // [kTcListCap] (lower bound, tile) keys
    static constexpr int CTL_OFF = LIST_OFF + kTcListCap * 8;
    static constexpr int BAR_OFF = CTL_OFF + 640;
    static constexpr int BYTES = BAR_OFF + 256;
};
static_assert(sizeof(TcCtl) <= 640, "control block overflows its slot");
static_assert(TcSmem::LIST_OFF % 8 == 0 && TcSmem::CTL_OFF % 8 == 0 && TcSmem::BAR_OFF % 8 == 0, "alignment");
static_assert(TcSmem::BYTES <= 227 * 1024, "shared memory per CTA");
static_assert(kTcParts == 4 && kTcEpiWarps == 16, "four epilogue warps per TMEM lane quarter");

// squared k-th distance implied by a list threshold (a key is d^2 - |q^|^2), rounded up; +inf stays +inf
__device__ __forceinline__ float tc_r2_of(float thr, float nq2_up)
{
    return fmaxf(thr + nq2_up, 0.f) * 1.000001f + 1e-37f;
}

// centred fp32 image of query slot `slot` (an unused slot of the block repeats the block's first query)
template <int D>
__device__ __forceinline__ bool tc_query_hat(const double *__restrict__ qs, const int *__restrict__ qorder, int64_t slot0,
                                             int q, const Weights &wt, const KnnHeader *__restrict__ h, float (&f)[D])
{
    int qi = qorder[slot0 + q];
    const bool valid = qi >= 0;
    if (!valid) qi = qorder[slot0];
#pragma unroll
    for (int d = 0; d < D; ++d) f[d] = (float)d_sub(d_mul(wt.w[d], qs[(int64_t)qi * D + d]), h->center[d]);
    return valid;
}

// ---- the scan: one block of (at most) 128 queries per CTA against the tiles that can matter to it ------------
// Warps 0..15: epilogue (lane = query of TMEM lane quarter w & 3; part w >> 2 takes two of a tile's eight 32-key
// chunks; the four warps of a quarter insert into one 16-entry list per query, one warp at a time per query).
// Warp 16, one lane: MMA issue.  Warp 17: tile producer (all lanes follow the list, lane 0 issues the copies).
// Between rounds every warp of the CTA plans the next round together.
// DIAG = true is the measuring build behind mpb_knn_tc_key_error: every tile is visited, nothing is inserted,
// and each key read from tensor memory is compared with the same expression evaluated in fp64.
template <int D, bool DIAG>
__global__ void __launch_bounds__(kTcThreads, 1)
    knn_tc_kernel(const double *__restrict__ qs, const int *__restrict__ qorder, Weights wt,
                  const unsigned char *__restrict__ tiles, int n_tiles, const float *__restrict__ tile_box,
                  int box_stride, const KnnHeader *__restrict__ h, int *__restrict__ surv_idx,
                  float *__restrict__ surv_thr, const double *__restrict__ x, const int *__restrict__ slot_idx,
                  unsigned long long *__restrict__ diag, int64_t own_block0, int own_part, int own_parts)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char *sA = smem + TcSmem::A_OFF;
    unsigned char *sB = smem + TcSmem::B_OFF;
    float *pk = reinterpret_cast<float *>(smem + TcSmem::PK_OFF);      // private sorted lists: keys
    int *pi = reinterpret_cast<int *>(smem + TcSmem::PI_OFF);          // ... candidate slots
    float *scr = reinterpret_cast<float *>(smem + TcSmem::SCR_OFF);
    float *sthr = reinterpret_cast<float *>(smem + TcSmem::THR_OFF);
    float *nq2s = reinterpret_cast<float *>(smem + TcSmem::NQ_OFF);
    unsigned long long *list = reinterpret_cast<unsigned long long *>(smem + TcSmem::LIST_OFF);
    TcCtl *ctl = reinterpret_cast<TcCtl *>(smem + TcSmem::CTL_OFF);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + TcSmem::BAR_OFF);
    uint64_t *empty = full + kTcStages;
    uint64_t *tfull = empty + kTcStages;
    uint64_t *tempty = tfull + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t slot0 = (int64_t)blockIdx.x * kTcQueries;

    if (tid < kTcQueries) {            // this thread's query row of the A operand, its norm, the block's box
        float f[D];
        tc_query_hat<D>(qs, qorder, slot0, tid, wt, h, f);
        float row[kTcK];
#pragma unroll
        for (int i = 0; i < kTcK; ++i) row[i] = 0.f;
        double n2 = 0.;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            n2 += (double)f[d] * (double)f[d];
            const float m2 = -2.f * f[d];
            const float hi = tf32_rn(m2);
            row[d] = hi;
            row[8 + d] = hi;
            row[16 + d] = tf32_rn(m2 - hi);
        }
        row[7] = 1.f;
        row[15] = 1.f;
#pragma unroll
        for (int c = 0; c < kTcK / 4; ++c)
            *reinterpret_cast<float4 *>(sA + tc_offset(tid, 4 * c)) =
                make_float4(row[4 * c], row[4 * c + 1], row[4 * c + 2], row[4 * c + 3]);
        nq2s[tid] = (float)n2 * 1.000001f;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            float lo = f[d], hi = f[d];            // an unused slot repeats a real query: the box is unchanged
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lo = fminf(lo, __shfl_xor_sync(FULL_MASK, lo, o));
                hi = fmaxf(hi, __shfl_xor_sync(FULL_MASK, hi, o));
            }
            if (lane == 0) {
                scr[warp * 16 + d] = lo;
                scr[warp * 16 + 8 + d] = hi;
            }
        }
    }
    for (int i = tid; i < (kTcKeep + 1) * kTcEpiWarps * 32; i += kTcThreads) pk[i] = INFINITY;
    for (int i = tid; i < (kTcKeep + 1) * kTcEpiWarps * 32; i += kTcThreads) pi[i] = -1;
    for (int i = tid; i < kTcQueries; i += kTcThreads) sthr[i] = INFINITY;
    if (tid == 0) {
        for (int s = 0; s < kTcStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull[b], 2);               // the MMAs' commit + the issuing thread's own arrival
            mbar_init(&tempty[b], kTcEpiWarps);
        }
        mbar_fence_init();
        ctl->cursor = 0;
        ctl->has_lo = 0;
        ctl->lo_key = 0ull;
        ctl->final_round = 0;
        ctl->round_over = 0;
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

    // ---- the block's box, and the lower bound of the squared distance from it to a tile's box
    float qlo[D], qhi[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        qlo[d] = fminf(fminf(scr[d], scr[16 + d]), fminf(scr[32 + d], scr[48 + d]));
        qhi[d] = fmaxf(fmaxf(scr[8 + d], scr[24 + d]), fmaxf(scr[40 + d], scr[56 + d]));
    }
    // computed in fp32 and scaled down past its own rounding (D + 3 operations of 2^-24 each), so it never
    // exceeds the true box-to-box distance; the boxes are exact (min / max of the very values the MMA reads)
    auto tile_lb = [&](int t) -> float {
        float s = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float blo = tile_box[(size_t)d * box_stride + t], bhi = tile_box[(size_t)(8 + d) * box_stride + t];
            const float gap = fmaxf(fmaxf(blo - qhi[d], qlo[d] - bhi), 0.f);
            s = __fmaf_rn(gap, gap, s);
        }
        return s * 0.999998f;
    };
    auto key_of = [](float lb, int t) -> unsigned long long {
        return ((unsigned long long)__float_as_uint(lb) << 32) | (unsigned long long)(uint32_t)t;
    };

    // ---- round 0: the nearest tile alone (ties: lowest index); it gives every query of the block a first
    // threshold.  own_block0 >= 0: the queries are samples in their own leaf order, block b is leaf b of share
    // own_part of own_parts (kd_share_leaf; 0 of 1 = every leaf), one half of tile leaf / 2.
    if (own_block0 >= 0) {
        if (tid == 0) {
            const int64_t leaf = own_block0 + kd_share_leaf((int64_t)blockIdx.x, own_part, own_parts);
            ctl->home = (int)(leaf >> 1);
            list[0] = key_of(0.f, (int)(leaf >> 1));
            ctl->list_n = 1;
        }
        __syncthreads();
    } else {
        unsigned long long best = ~0ull;
#pragma unroll 4
        for (int t = tid; t < n_tiles; t += kTcThreads) best = min(best, key_of(tile_lb(t), t));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(FULL_MASK, best, o));
        if (lane == 0) ctl->red_u[warp] = best;
        __syncthreads();
        if (tid == 0) {
            unsigned long long b = ctl->red_u[0];
            for (int w = 1; w < kTcThreads / 32; ++w) b = min(b, ctl->red_u[w]);
            ctl->home = (int)(uint32_t)b;
            list[0] = b;
            ctl->list_n = 1;
        }
        __syncthreads();
    }
    const int home = ctl->home;

    // ---- planning of the next round, by the whole CTA (uniform control flow; ends with list / list_n ready)
    auto plan_round = [&]() {
        // largest squared k-th distance in the block right now
        if (tid < kTcQueries) {
            float r = tc_r2_of(sthr[tid], nq2s[tid]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(FULL_MASK, r, o));
            if (lane == 0) ctl->red_f[warp] = r;
        }
        __syncthreads();
        const float rmax2 = fmaxf(fmaxf(ctl->red_f[0], ctl->red_f[1]), fmaxf(ctl->red_f[2], ctl->red_f[3]));
        const int cursor = ctl->cursor;
        const bool has_lo = ctl->has_lo != 0;
        const unsigned long long lo_key = ctl->lo_key;
        int n = 0;
        if (!(rmax2 < 3.0e38f)) {
            // some list is not full yet (or holds padding): no bound to prune with; take the next tiles by index
            const int end = min(n_tiles, cursor + kTcListCap);
            if (tid == 0) ctl->list_all = 0;
            __syncthreads();
            for (int t = cursor + tid; t < end; t += kTcThreads)
                if (t != home) list[atomicAdd(&ctl->list_all, 1)] = key_of(0.f, t);
            __syncthreads();
            n = ctl->list_all;
            if (tid == 0) {
                ctl->cursor = end;
                ctl->final_round = (end >= n_tiles) ? 1 : 0;
                ctl->red_f[4] = 1.f;
            }
        } else {
            // tiles whose box is closer than that, not handled yet.  More than the list holds (a block whose
            // first tile gave loose distances, or a box that touches thousands of tiles): the nearer ones first,
            // by shrinking the bound; if that does not separate them, the next kTcListCap tiles by index.
            // (the first such round stops at half the bound: the distances a block has after its own tile alone are
            // loose, most of the tiles they admit are dropped again once the nearest ones have been read, and a
            // short list is cheaper to sort; what is left is collected against the tightened bound afterwards)
            float bound = has_lo ? rmax2 : rmax2 * MPB_TC_FIRST_BOUND;
            int tries = 0, end = n_tiles;
            bool window = false;
            for (;;) {
                if (tid == 0) ctl->list_all = 0;
                __syncthreads();
#pragma unroll 4
                for (int t = cursor + tid; t < end; t += kTcThreads) {
                    if (t == home) continue;
                    const float lb = tile_lb(t);
                    const unsigned long long key = key_of(lb, t);
                    if (lb < bound && (!has_lo || key > lo_key)) {
                        const int pos = atomicAdd(&ctl->list_all, 1);
                        if (pos < kTcListCap) list[pos] = key;
                    }
                }
                __syncthreads();
                n = ctl->list_all;
                if (n <= kTcListCap && (n > 0 || window || bound == rmax2)) break;
                if (n > kTcListCap && !window && ++tries <= 3) {
                    bound *= 0.25f;
                } else if (n == 0 && !window && bound < rmax2) {
                    bound = rmax2;                 // nothing that near: everything below the real bound
                } else {
                    window = true;
                    bound = rmax2;
                    end = min(n_tiles, cursor + kTcListCap);
                }
                __syncthreads();
            }
            if (tid == 0) {
                ctl->final_round = (bound == rmax2 && end == n_tiles) ? 1 : 0;
                if (window) ctl->cursor = end;
                ctl->red_f[4] = window ? 1.f : 0.f;
            }
        }
        // nearest box first (bitonic sort of the keys; the pad sorts to the end)
        int n2 = 2;
        while (n2 < n) n2 <<= 1;
        if (n > 1) {
            for (int i = n + tid; i < n2; i += kTcThreads) list[i] = ~0ull;
            __syncthreads();
            for (int k2 = 2; k2 <= n2; k2 <<= 1) {
                for (int j = k2 >> 1; j > 0; j >>= 1) {
                    for (int i = tid; i < n2; i += kTcThreads) {
                        const int p = i ^ j;
                        if (p > i) {
                            const unsigned long long a = list[i], b = list[p];
                            if ((a > b) == ((i & k2) == 0)) {
                                list[i] = b;
                                list[p] = a;
                            }
                        }
                    }
                    __syncthreads();
                }
            }
        }
        if (tid == 0) {
            ctl->round_over = 0;
            ctl->list_n = n;
            if (n > 0 && rmax2 < 3.0e38f && ctl->red_f[4] == 0.f) {     // a round ordered by distance, not by index
                ctl->lo_key = list[n - 1];
                ctl->has_lo = 1;
            }
        }
        __syncthreads();
    };

    // ---- epilogue state (warps 0..15)
    const int quarter = warp & 3, part = warp >> 2;
    const int q = quarter * 32 + lane;             // this thread's query (epilogue warps only)
    float thr = INFINITY;
#ifdef MPB_TC_STATS
    unsigned n_ev = 0, n_tk = 0, n_bulk = 0;
#endif
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    float *rows = scr + (warp < kTcEpiWarps ? warp : 0) * 32 * 33;
    // ---- what a query keeps.  Every epilogue thread (part p of query q) owns a PRIVATE sorted list of the 16
    // smallest keys seen in its part's columns; the warp that owns it is the only writer, so inserting needs no
    // lock and no fence.  The query's threshold is the 16th smallest key of the union of its four lists.  Lists only
    // ever decrease, entry by entry, so that value may be computed from plain racy reads at any time -- whatever a
    // reader sees is at least as large as the truth, i.e. a valid, slightly stale threshold: one warp of each lane
    // quarter (taking turns) merges the four sorted lists of its 32 queries every fourth tile and publishes the
    // result.  Candidates are dealt to the four parts like cards (the pack kernel strides the slots), so each part
    // sees a fair quarter of every tile.
    constexpr int KS = kTcKeep + 1;                // key lists are padded: lane-strided reads stay conflict-free
    auto kbase = [&](int w, int l) -> float * { return pk + (w * 32 + l) * KS; };
    auto ibase = [&](int w, int l) -> int * { return pi + (w * 32 + l) * KS; };
    // 16th smallest key of the union of this lane's query's four sorted lists (inf while fewer than 16 keys)
    auto union16 = [&](int quarter) -> float {
        int hd[kTcParts];
        float a[kTcParts];
#pragma unroll
        for (int p = 0; p < kTcParts; ++p) {
            hd[p] = 0;
            a[p] = *(volatile float *)(kbase(p * 4 + quarter, lane));
        }
        // The largest of 16 values taken from 16 distinct list positions.  At rest that is the union's 16th
        // smallest key.  While other warps insert, a position may be read before or after a shift (so the same key
        // can turn up twice, or the pops may not come in order), but whatever is read from a position is at least
        // that position's final value, and the largest of any 16 positions' final values is at least the union's
        // 16th smallest: the result is a valid threshold in every interleaving.
        float top = -INFINITY;
#pragma unroll 1
        for (int s = 0; s < kTcKeep; ++s) {
            int best = 0;
            float bv = a[0];
#pragma unroll
            for (int p = 1; p < kTcParts; ++p)
                if (a[p] < bv) {
                    bv = a[p];
                    best = p;
                }
            top = fmaxf(top, bv);
#pragma unroll
            for (int p = 0; p < kTcParts; ++p)
                if (p == best) {
                    ++hd[p];
                    a[p] = hd[p] < kTcKeep ? *(volatile float *)(kbase(p * 4 + quarter, lane) + hd[p]) : INFINITY;
                }
        }
        return top;
    };

    // Many lanes of a chunk flagged (the first tiles a block reads): instead of serving them one by one, every lane
    // takes the 16 smallest of its 32 keys through a sorting network over its own registers and merges them with
    // its list, all 32 lanes side by side -- about 700 instructions for the warp whatever the number of takers,
    // where one sorted insert costs a hundred.  The column of a key rides through the network in its five lowest
    // mantissa bits (a perturbation of at most 31 ulp, which the certificate's eta covers).
    auto bulk = [&](const uint32_t (&v)[32], int cbase) {
        float a[32];
#pragma unroll
        for (int j = 0; j < 32; ++j)
            a[j] = __uint_as_float(v[j]) < thr ? __uint_as_float((v[j] & 0xffffffe0u) | (uint32_t)j) : INFINITY;
        tc_sort_asc<0, 16, 32>(a);
        tc_sort_asc<16, 16, 32>(a);
        float c[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) c[i] = fminf(a[i], a[31 - i]);      // the 16 smallest, as a bitonic sequence
        tc_merge_asc<16, 16>(c);
        float *kb = kbase(warp, lane);
        int *ib = ibase(warp, lane);
        float d[16];
        int di[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float mine = kb[i], cand = c[15 - i];
            const bool take = cand < mine;                                // the incumbent wins a tie
            d[i] = take ? cand : mine;
            di[i] = take ? cbase + (int)(__float_as_uint(cand) & 31u) : ib[i];
        }
        tc_merge_asc_kv<16>(d, di);
        // every position only ever gets a smaller key (the new list is the 16 smallest of old list + chunk), which
        // is what the unsynchronised readers of union16 rely on
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            kb[i] = d[i];
            ib[i] = di[i];
        }
        thr = fminf(thr, d[15]);
        TC_CNT(n_bulk);
    };

    // The 32 keys a lane just received are examined with one min-tree.  When some lane's minimum beats its
    // threshold, the flagged lanes park the groups of eight that hold such a key in the warp's scratch rows, and
    // each flagged lane is then served by the WHOLE warp: lane j reads key j of that row, one ballot finds the
    // takers, and each taker is inserted into the lane's private list by lanes 0..15 side by side.
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
        // the union threshold published last for the query; then the chunk is tested again
        thr = fminf(thr, *(volatile float *)&sthr[q]);
        flagged = __ballot_sync(FULL_MASK, m < thr);
        if (flagged == 0) return;
        if (__popc(flagged) >= MPB_TC_BULK_MIN) {
            bulk(v, cbase);
            return;
        }
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
            TC_CNT(n_ev);
            float thrL = __shfl_sync(FULL_MASK, thr, L);
            const unsigned gmL = __shfl_sync(FULL_MASK, gm, L);
            const float mine = rows[L * 33 + lane];      // stale where the group was not parked: masked out
            unsigned takers = __ballot_sync(FULL_MASK, ((gmL >> (lane >> 3)) & 1u) && mine < thrL);
            float *kb = kbase(warp, L);
            int *ib = ibase(warp, L);
            while (takers) {
                const int j = __ffs(takers) - 1;
                takers &= takers - 1;
                const float key = __shfl_sync(FULL_MASK, mine, j);
                if (key < thrL) {
                    TC_CNT(n_tk);
                    const bool in_list = lane < kTcKeep;
                    const float e = in_list ? kb[lane] : INFINITY;
                    const int ei = in_list ? ib[lane] : -1;
                    // equal keys: the newcomer goes after the incumbents
                    const int pos = __popc(__ballot_sync(FULL_MASK, in_list && e <= key));
                    if (lane >= pos && lane < kTcKeep - 1) {
                        kb[lane + 1] = e;
                        ib[lane + 1] = ei;
                    }
                    if (lane == pos) {
                        kb[pos] = key;
                        ib[pos] = cbase + j;
                    }
                    // the list's new 16th: what was its 15th, or the key itself
                    const float e14 = __shfl_sync(FULL_MASK, e, kTcKeep - 2);
                    thrL = fminf(thrL, pos == kTcKeep - 1 ? key : e14);
                    __syncwarp();
                }
            }
            if (lane == L) thr = thrL;
        }
        __syncwarp();
    };

    // DIAG: this thread's query in fp32 and fp64 terms, and the running maxima of the key error
    float dq[D];
    double d_scale = 1., d_max = 0.;
    unsigned long long d_pairs = 0;
    bool d_valid = false;
    if constexpr (DIAG) {
        if (warp < kTcEpiWarps) {
            d_valid = tc_query_hat<D>(qs, qorder, slot0, q, wt, h, dq);
            double n2 = 0.;
#pragma unroll
            for (int d = 0; d < D; ++d) n2 += (double)dq[d] * (double)dq[d];
            d_scale = h->R * h->R + 2.0 * sqrt(n2) * h->R;
        }
    }
    // every key of the chunk against the same expression in fp64: n + sum_d (-2 q^_d) x^_d over the fp32 values
    auto measure = [&](const uint32_t (&v)[32], int cbase) {
        if constexpr (DIAG) {
            for (int i = 0; i < 32; ++i) {
                const int j = slot_idx[cbase + i];
                if (j < 0 || !d_valid) continue;
                double n2 = 0., dot = 0.;
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    const float f = (float)d_sub(d_mul(wt.w[d], x[(int64_t)j * D + d]), h->center[d]);
                    n2 += (double)f * (double)f;
                    dot += (double)(-2.f * dq[d]) * (double)f;
                }
                const double exact = (double)(float)n2 + dot;
                const double err = fabs((double)__uint_as_float(v[i]) - exact) / d_scale;
                d_max = fmax(d_max, err);
                ++d_pairs;
            }
        }
    };

    // ---- rounds: scan the planned tiles, then plan again, until a round has covered everything that is left
    uint32_t it = 0;                               // tiles and end-of-round marks this thread's role has handled
    uint32_t visited = 0;
#ifdef MPB_TC_STATS
    long long c_start = clock64(), c_plan = 0, c_r0 = 0;
    long long w_a = 0, w_b = 0;
    int round_no = 0;
#endif
    for (;;) {
#ifdef MPB_TC_STATS
        const long long c_scan0 = clock64();
#endif
        if (warp == kTcEpiWarps + 1) {
            // ---- producer: follows the list; a tile is dropped (and with it the rest of the round: the list is
            // sorted) when its box is no longer closer than the block's largest k-th distance
            const int n = ctl->list_n;
            for (int i = 0; i <= n; ++i) {
                int t = -1;
                if (i < n) {
                    const unsigned long long e = list[i];
                    const float lb = __uint_as_float((uint32_t)(e >> 32));
                    float r = 0.f;
#pragma unroll
                    for (int u = 0; u < kTcQueries / 32; ++u) {
                        const int qq = lane + 32 * u;
                        r = fmaxf(r, tc_r2_of(*(volatile float *)&sthr[qq], nq2s[qq]));
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(FULL_MASK, r, o));
                    if (lb >= r) i = n;            // -> the end-of-round mark
                    else t = (int)(uint32_t)e;
                    visited += (t >= 0);
                }
                const int s = it % kTcStages;
                if (lane == 0) {
                    TC_CLK(c0);
                    if (it >= kTcStages) mbar_wait_backoff(&empty[s], (uint32_t)((it / kTcStages - 1) & 1));
                    TC_ACC(w_a, c0);
                    *(volatile int *)&ctl->slot_tile[s] = t;
                    if (t >= 0) {
                        mbar_arrive_expect_tx(&full[s], kTcTileBytes);
                        bulk_copy_g2s(sB + (size_t)s * kTcTileBytes, tiles + (size_t)t * kTcTileBytes, kTcTileBytes, &full[s]);
                    } else {
                        mbar_arrive(&full[s]);
                    }
                }
                ++it;
                __syncwarp();
            }
        } else if (warp == kTcEpiWarps + 2) {
            // ---- threshold merger: keeps publishing the union threshold of every query (one lane quarter per
            // trip) from plain reads of the private lists, until the epilogue has seen the end of the round
            if constexpr (!DIAG) {
                for (int qt = 0; *(volatile int *)&ctl->round_over == 0; qt = (qt + 1) & 3) {
                    const float u = union16(qt);
                    float *dst = &sthr[qt * 32 + lane];
                    if (u < *(volatile float *)dst) *(volatile float *)dst = u;
                    if (MPB_TC_MERGER_SLEEP) __nanosleep(MPB_TC_MERGER_SLEEP);
                }
            }
        } else if (warp == kTcEpiWarps) {
            // ---- MMA issue: three k-steps per tile into accumulator it & 1
            if (lane == 0) {
                // D = f32, A = B = tf32, both K-major, N = 256, M = 128
                constexpr uint32_t idesc =
                    (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTcTile >> 3) << 17) | ((uint32_t)(kTcQueries >> 4) << 24);
                const uint32_t a_addr = smem_u32(sA);
                for (;;) {
                    const int s = it % kTcStages, b = it & 1;
                    TC_CLK(c0);
                    mbar_wait_backoff(&full[s], (uint32_t)((it / kTcStages) & 1));
                    TC_ACC(w_a, c0);
                    const int t = *(volatile int *)&ctl->slot_tile[s];
                    TC_CLK(c1);
                    if (it >= 2) mbar_wait_backoff(&tempty[b], (uint32_t)((it / 2 - 1) & 1));
                    TC_ACC(w_b, c1);
                    ++it;
                    *(volatile int *)&ctl->acc_tile[b] = t;
                    if (t < 0) {
                        mbar_arrive(&empty[s]);
                        mbar_arrive(&tfull[b]);
                        mbar_arrive(&tfull[b]);
                        break;
                    }
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
                    // both commits track every MMA issued so far: the ring slot is free, the accumulator is ready;
                    // the plain arrival publishes acc_tile[b] to the epilogue warps
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                 ::"r"(smem_u32(&empty[s])) : "memory");
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                 ::"r"(smem_u32(&tfull[b])) : "memory");
                    mbar_arrive(&tfull[b]);
                }
            }
            __syncwarp();
        } else {
            // ---- epilogue: lane = query; this warp's part of every tile is two chunks of 32 keys
            for (;;) {
                const int b = it & 1;
                TC_CLK(cw0);
                mbar_wait_warp_backoff(&tfull[b], (uint32_t)((it / 2) & 1));
                TC_ACC(w_a, cw0);
                ++it;
                const int t = *(volatile int *)&ctl->acc_tile[b];
                if (t < 0) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty[b]);
                    if (tid == 0) *(volatile int *)&ctl->round_over = 1;
                    break;
                }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                constexpr int CH = (kTcTile / 32) / kTcParts;      // chunks per warp per tile
                static_assert(CH == 2, "both chunks of a warp go to registers at once");
                uint32_t va[32], vb[32];
                const int c0 = CH * part;
                const uint32_t tile_base = lane_base + (uint32_t)(b * kTcTile + c0 * 32);
                // both chunks go to registers first and the accumulator is handed back at once: the next MMAs
                // never wait for a warp that is busy inserting
                __syncwarp();
                tmem_ld32(va, tile_base);
                tmem_ld32(vb, tile_base + 32u);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[b]);
                if (MPB_TC_REFRESH_EVERY) thr = fminf(thr, *(volatile float *)&sthr[q]);
                if constexpr (DIAG) {
                    measure(va, t * kTcTile + c0 * 32);
                    measure(vb, t * kTcTile + (c0 + 1) * 32);
                } else {
                    TC_CLK(c2);
                    examine(va, t * kTcTile + c0 * 32);
                    examine(vb, t * kTcTile + (c0 + 1) * 32);
                    TC_ACC(w_b, c2);
                }
            }
        }
        __syncwarp();
        __syncthreads();
#ifdef MPB_TC_STATS
        if (round_no++ == 0) c_r0 = clock64() - c_scan0;
        if (tid == 0) TC_STAT(3, 1);
#endif
        if (ctl->final_round) break;
#ifdef MPB_TC_STATS
        const long long c_p0 = clock64();
#endif
        // exact thresholds for the planning (every list is at rest here)
        if (!DIAG && warp < kTcParts) sthr[q] = union16(quarter);
        __syncthreads();
        plan_round();
#ifdef MPB_TC_STATS
        c_plan += clock64() - c_p0;
        if (tid == 0) TC_STAT(7, ctl->list_n);
#endif
    }
    if (warp == kTcEpiWarps + 1 && lane == 0) atomicAdd(&const_cast<KnnHeader *>(h)->tiles_visited, (unsigned long long)visited);
#ifdef MPB_TC_STATS
    if (tid == 0) {
        TC_STAT(0, 1);
        TC_STAT(4, clock64() - c_start);
        TC_STAT(5, c_plan);
        TC_STAT(6, c_r0);
    }
    if (lane == 0) {
        if (warp == kTcEpiWarps + 1) TC_STAT(8, w_a);
        else if (warp == kTcEpiWarps) { TC_STAT(9, w_a); TC_STAT(10, w_b); }
        else { TC_STAT(11, w_a); TC_STAT(12, w_b); TC_STAT(1, n_ev); TC_STAT(2, n_tk); TC_STAT(13, n_bulk); }
    }
#endif

    // the 16 survivors of every query (ascending fp32 key: the 4-way merge of its lists) and its final threshold
    if (!DIAG && warp < kTcParts) {
        int hd[kTcParts];
        float a[kTcParts];
#pragma unroll
        for (int p = 0; p < kTcParts; ++p) {
            hd[p] = 0;
            a[p] = kbase(p * 4 + quarter, lane)[0];
        }
        float last = INFINITY;
#pragma unroll 1
        for (int s = 0; s < kTcKeep; ++s) {
            int best = 0;
            float bv = a[0];
#pragma unroll
            for (int p = 1; p < kTcParts; ++p)
                if (a[p] < bv) {
                    bv = a[p];
                    best = p;
                }
            last = bv;
            int slot = -1;
#pragma unroll
            for (int p = 0; p < kTcParts; ++p)
                if (p == best) {
                    slot = ibase(p * 4 + quarter, lane)[hd[p]];
                    ++hd[p];
                    a[p] = hd[p] < kTcKeep ? kbase(p * 4 + quarter, lane)[hd[p]] : INFINITY;
                }
            surv_idx[(slot0 + q) * kTcKeep + s] = (bv < INFINITY) ? slot : -1;
        }
        surv_thr[slot0 + q] = last;
    }
    if constexpr (DIAG) {
        if (warp < kTcEpiWarps && d_pairs > 0) {
            atomicMax(&diag[0], (unsigned long long)__double_as_longlong(d_max));   // non-negative doubles order as integers
            atomicAdd(&diag[1], d_pairs);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == kTcEpiWarps) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

// ---- exact re-rank of the survivors of every query, and the certificate ---------------------------------
// One warp per query slot.  Same argument as in knn_filter_kernel: every sample of a VISITED tile that is not in
// the list has computed key >= T (the list's final threshold), hence exact squared distance
// >= (sqrt(T - eta + |q^|^2) - delta)^2; here eta also covers the TF32 split (three dropped or rounded 2^-22
// terms) and the tensor core's accumulation, budgeted at 2^-22 per term (measured: mpb_knn_tc_key_error).
// Every sample of a SKIPPED tile lies in a box whose squared distance to the block's box, rounded down, was at
// least the block's largest T_q + |q^|^2, rounded up, when the tile was dropped -- and thresholds only fall --
// so its fp32-space squared distance is >= T + |q^|^2 and the same bound holds for it a fortiori.
template <int D>
__global__ void __launch_bounds__(256)
    knn_tc_rerank_kernel(const double *__restrict__ x, int64_t N, const double *__restrict__ qs, int64_t M, Weights wt,
                         int k, double r2, KnnHeader *__restrict__ h, const int *__restrict__ qorder, int64_t n_slots,
                         const int *__restrict__ slot_idx, const int *__restrict__ surv_idx,
                         const float *__restrict__ surv_thr, int *__restrict__ fb_list, RowMap rows,
                         int64_t *__restrict__ out_idx, double *__restrict__ out_dist)
{
    constexpr int NU = (kTcSurvivors + 31) / 32;   // survivors per lane
    // a query is served by a group of G lanes: the whole warp, or -- with 16 survivors -- half of it, two queries
    // per warp (0.92 -> about 0.6 ms at 1M queries: most of the kernel is per-warp bookkeeping, not the 16 distances)
    constexpr int G = (kTcSurvivors <= 16) ? 16 : 32;
    const int lane = threadIdx.x & 31;
    const int gl = lane & (G - 1), gbase = lane & ~(G - 1);
    const unsigned gmask = (G == 32) ? FULL_MASK : (0xffffu << gbase);
    const int64_t slot = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 / G) + lane / G;
    if (slot - lane / G >= n_slots) return;     // whole warps leave together (n_slots is a multiple of 128)
    const int64_t q_raw = slot < n_slots ? (int64_t)qorder[slot] : -1;
    const bool live = q_raw >= 0 && q_raw < M;  // not an unused slot of a query block
    if (!__any_sync(FULL_MASK, live)) return;
    const int64_t q = live ? q_raw : 0;
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
        ci[u] = (live && u * 32 + gl < kTcSurvivors) ? surv_idx[slot * kTcSurvivors + u * 32 + gl] : -1;
        if (ci[u] >= 0) ci[u] = slot_idx[ci[u]];   // tile slot -> sample index (-1: a padding slot)
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
    // Lanes past the last survivor hold (inf, -1): they come before an entry only if that entry's own key is inf
    // (then they all do), so the loop stops at the survivors and such an entry gets them added at once.
    constexpr int kLive = (kTcSurvivors < 32) ? kTcSurvivors : 32;
    for (int o = 0; o < kLive; ++o) {
#pragma unroll
        for (int w = 0; w < NU; ++w) {
            const double So = __shfl_sync(FULL_MASK, S[w], gbase + o);
            const int io = __shfl_sync(FULL_MASK, ci[w], gbase + o);
#pragma unroll
            for (int u = 0; u < NU; ++u) rank[u] += ((So < S[u]) || (So == S[u] && io < ci[u])) ? 1 : 0;
        }
    }
    if (NU == 1 && kLive < 32) {
        if (ci[0] >= 0 && S[0] == INFINITY) rank[0] += 32 - kLive;
    }
    // k-th smallest exact key among the survivors (inf when fewer than k exist)
    double Sk = INFINITY;
#pragma unroll
    for (int u = 0; u < NU; ++u) {
        const unsigned mk = __ballot_sync(FULL_MASK, ci[u] >= 0 && rank[u] == k - 1) & gmask;
        const double cand = __shfl_sync(FULL_MASK, S[u], mk ? (__ffs(mk) - 1) : 0);
        if (mk) Sk = cand;
    }
    int nvalid = 0;
#pragma unroll
    for (int u = 0; u < NU; ++u) nvalid += __popc(__ballot_sync(FULL_MASK, ci[u] >= 0) & gmask);
    float T32 = INFINITY;
#pragma unroll
    for (int p = 0; p < kTcLists; ++p) {
        const float tp = surv_thr[slot * kTcLists + p];
        if (tp < 1.0e38f) T32 = fminf(T32, tp);  // a list that never filled (or filled with padding) excludes nothing
    }
    const double T = (double)T32;
    // A list that never filled saw every tile (nothing is skipped without a threshold), so all N samples should
    // be among the survivors; fewer means that some key was not comparable (overflow, NaN): not certified.
    bool ok = isinf(T) && T > 0 && (int64_t)nvalid >= N;
    if (!ok) {
        const double R = h->R;
        constexpr double U24 = 5.9604644775390625e-08;   // 2^-24: rounding of the centred fp32 coordinates
        constexpr double U22 = 2.384185791015625e-07;    // 2^-22: TF32 split and accumulation budget per term
        const double nq = sqrt(nq2);
        const double delta = U24 * (R + nq) * 1.0001 + 1e-30;
        // 2^-22 per operand column and accumulation step (measured: under 2 x 2^-22 in all, mpb_knn_tc_key_error), plus
        // 16 x 2^-22 = 2^-18 for the column bits the sorting-network path writes into a key
        const double eta = (double)(kTcK + 8 + 16) * U22 * 1.01 * (R * R + 2.0 * nq * R) + 1e-37;
        double slow = T - eta + nq2;
        slow -= 1e-12 * (fabs(T) + eta + nq2);
        double bound = 0.;
        if (slow > 0.) {
            const double rho = sqrt(slow) * (1.0 - 1e-15) - delta;
            if (rho > 0.) bound = rho * rho * (1.0 - 1e-12);
        }
        ok = (Sk < bound) || (r2 <= bound);
    }
    if (!live) return;                          // (no warp-wide operation below this line)
    if (ok) {
        const int64_t row = rows.row(q);
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            if (ci[u] >= 0 && rank[u] < k) {
                const bool keep = S[u] < r2;
                out_idx[row * k + rank[u]] = keep ? (int64_t)ci[u] : -1;
                if (out_dist) out_dist[row * k + rank[u]] = keep ? __dsqrt_rn(S[u]) : INFINITY;
            }
        }
        if (gl < k && gl >= nvalid) {
            out_idx[row * k + gl] = -1;
            if (out_dist) out_dist[row * k + gl] = INFINITY;
        }
    } else if (gl == 0) {
        const int pos = atomicAdd(&h->fb_count, 1);
        fb_list[pos] = (int)q;
    }
}
