// Spatial ordering of a point set for the tensor-core kNN scan (included by knn.cu, inside its anonymous
// namespace, after KnnHeader).
//
// The scan of knn_tc.cuh skips every candidate tile whose bounding box is provably farther from a block of
// queries than the block's current k-th distances.  That only pays when a tile (256 consecutive candidates)
// and a query block (128 consecutive queries) are compact in space, so both sets are first brought into the
// leaf order of a balanced k-d partition: `leaves` cells of floor/ceil(n / leaves) points each, obtained level
// by level -- every level sorts the points of each current cell along one coordinate and cuts the cell into
// 2^bits parts of equal leaf counts.  One level = one stable LSD radix sort of (cell id, coordinate) keys over
// the whole array, so cells stay where they are and only their contents are reordered.  Everything is
// stream-ordered device work on caller-provided scratch: no host round trip, no allocation, deterministic.
//
// The reference has no counterpart (scipy's cKDTree builds its own tree inside nearest.py:59-61); the order
// only decides which tiles are visited, never what the search returns.
#pragma once

constexpr int kRsThreads = 256;                    // threads per block of the radix passes
constexpr int kRsItems = 16;                       // keys per thread
constexpr int kRsChunk = kRsThreads * kRsItems;    // keys per block
constexpr int kKdMaxLevels = 8;
constexpr int kKdCoordBits = 24;                   // leading bits of the order-preserving fp32 image kept in a key

struct KdPlan {
    int64_t n;            // points
    int64_t leaves;       // cells at the end (each holds floor or ceil of n / leaves points)
    int levels;           // sorts
    int ways[kKdMaxLevels];   // level l cuts every cell ways[l] ways along coordinate dim_order[l]
    int leaf_bits;        // bits needed for a cell id
    int coord_bits;       // leading bits of the order-preserving fp32 image kept in a key
};

inline int ceil_log2_i64(int64_t v)
{
    int b = 0;
    while (((int64_t)1 << b) < v) ++b;
    return b;
}

// The smallest leaf count >= min_leaves that is a product of one cut per coordinate (at most kKdMaxLevels of
// them) with all cuts within a factor two of each other -- cells as close to cubes as the count allows, and every
// cell cut the same way (a count that is not such a product leaves some cells uncut along the last coordinates,
// and elongated cells prune badly: measured 324 against 233 tiles per query block at 1M x 7-DoF).  `even`: the
// count must be even (a candidate tile is two consecutive leaves).  The widest coordinate gets the largest cut.
inline KdPlan kd_plan(int64_t n, int64_t min_leaves, int D, bool even)
{
    KdPlan p;
    p.n = n;
    if (min_leaves < 1) min_leaves = 1;
    if (even && min_leaves < 2) min_leaves = 2;
    const int dims = D < kKdMaxLevels ? D : kKdMaxLevels;
    int64_t best = -1;
    int best_lo = 1, best_hi_count = 0, best_hi = 1;
    // cuts take two adjacent-ish values: `hc` coordinates are cut `hi` ways, the others `lo` ways, lo <= hi <= 2 lo
    for (int lo = 1; lo <= 64; ++lo) {
        for (int hi = lo; hi <= 2 * lo; ++hi) {
            for (int hc = 0; hc <= dims; ++hc) {
                if (hi == lo && hc > 0) continue;
                int64_t prod = 1;
                for (int d = 0; d < dims; ++d) prod *= (d < hc ? hi : lo);
                if (prod < min_leaves || (even && (prod & 1))) continue;
                if (best < 0 || prod < best) {
                    best = prod;
                    best_lo = lo;
                    best_hi = hi;
                    best_hi_count = hc;
                }
            }
        }
        int64_t all_lo = 1;
        for (int d = 0; d < dims; ++d) all_lo *= lo;
        if (best > 0 && all_lo >= best) break;      // larger cuts can only give larger counts
    }
    if (best < 0) {                                  // more than 64^dims leaves: plain powers of two
        best = (int64_t)1 << ceil_log2_i64(min_leaves);
        best_lo = best_hi = 2;
        best_hi_count = 0;
    }
    p.leaves = best;
    p.levels = 0;
    for (int l = 0; l < kKdMaxLevels; ++l) p.ways[l] = 1;
    if (best_lo == 2 && best_hi == 2 && best > 1 && ((int64_t)1 << ceil_log2_i64(best)) == best && best_hi_count == 0 &&
        ceil_log2_i64(best) > dims) {
        // fallback shape: distribute the bits over the coordinates
        const int T = ceil_log2_i64(best);
        for (int l = 0; l < dims; ++l) p.ways[l] = 1 << (T / dims + (l < T % dims ? 1 : 0));
        p.levels = dims;
    } else {
        for (int d = 0; d < dims; ++d) {
            const int w = d < best_hi_count ? best_hi : best_lo;
            if (w > 1) p.ways[p.levels++] = w;
        }
    }
    p.leaf_bits = ceil_log2_i64(p.leaves);
    // a key = (cell << coord_bits) | coordinate bits: with at most 32 significant bits a level takes four radix
    // passes instead of five; 16 coordinate bits (sign, exponent, 7 mantissa bits) still order a cell's points well
    // enough for a split into a handful of parts, and ties keep their previous order (the sort is stable)
    p.coord_bits = 32 - p.leaf_bits;
    if (p.coord_bits > kKdCoordBits) p.coord_bits = kKdCoordBits;
    if (p.coord_bits < 16) p.coord_bits = 16;
    return p;
}

// first position of leaf j, and the leaf that holds position pos (leaf j = positions [floor(j n / L), floor((j+1) n / L)))
__host__ __device__ inline int64_t kd_leaf_start(int64_t j, int64_t n, int64_t L) { return (j * n) / L; }
__host__ __device__ inline int64_t kd_leaf_of(int64_t pos, int64_t n, int64_t L) { return ((pos + 1) * L - 1) / n; }

// first leaf of the cell that holds leaf j after `depth` levels of cuts
__device__ inline int64_t kd_cell_start(const KdPlan &p, int64_t j, int depth)
{
    int64_t s = 0, c = p.leaves;
    for (int l = 0; l < depth; ++l) {
        const int64_t w = p.ways[l];
        const int64_t r = j - s;
        const int64_t i = ((r + 1) * w - 1) / c;          // child i covers leaves [floor(i c / w), floor((i+1) c / w))
        const int64_t a = (i * c) / w, b = ((i + 1) * c) / w;
        s += a;
        c = b - a;
    }
    return s;
}

__device__ __forceinline__ uint32_t ord_f32(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u >> 31) ? ~u : (u | 0x80000000u);
}

// coordinates ranked by extent of the sample set's bounding box, widest first (ties: lower coordinate)
__global__ void kd_dim_order_kernel(KnnHeader *h, int D)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double ext[MPB_MAX_DIM];
    for (int d = 0; d < D; ++d) {
        const double e = dec_f64(h->mm[MPB_MAX_DIM + d]) - dec_f64(h->mm[d]);
        ext[d] = (e == e) ? e : 0.;
    }
    bool used[MPB_MAX_DIM];
    for (int d = 0; d < D; ++d) used[d] = false;
    for (int r = 0; r < D; ++r) {
        int best = -1;
        for (int d = 0; d < D; ++d)
            if (!used[d] && (best < 0 || ext[d] > ext[best])) best = d;
        used[best] = true;
        h->dim_order[r] = best;
    }
}

__global__ void __launch_bounds__(256) kd_invert_kernel(const uint32_t *__restrict__ perm, int64_t n, int32_t *__restrict__ inv)
{
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) inv[perm[p]] = (int32_t)p;
}

// ---- shares of the leaves for G ranks --------------------------------------------------------------------
// Leaves are dealt in rotation: leaf l = G a + b goes to share (a + b) mod G, so block i of share r is leaf
// G i + ((r - i) mod G).  Every share gets leaves from all over the partition.  (Contiguous slices of the leaf
// order gave the shares that lie inside the sampled region 28 % more tiles to read than the ones at its rim:
// 4.09 against 3.21 ms at four GPUs on the 1M-sample set.)
__host__ __device__ inline int64_t kd_share_leaf(int64_t i, int part, int parts)
{
    const int64_t b = ((((int64_t)part - i) % parts) + parts) % parts;
    return (int64_t)parts * i + b;
}
__host__ __device__ inline int64_t kd_share_blocks(int64_t leaves, int part, int parts)
{
    const int64_t a = leaves / parts, rem = leaves % parts;
    const int64_t b = ((((int64_t)part - a) % parts) + parts) % parts;
    return a + (b < rem ? 1 : 0);
}

// first row of every block of a share (rows = the share's samples, block after block): an exclusive scan of the
// leaf sizes, one block of 1024 threads; rowstart[blocks] = the share's row count
__global__ void __launch_bounds__(1024) kd_share_rowstart_kernel(int64_t n, int64_t leaves, int part, int parts, int64_t blocks,
                                                                 int32_t *__restrict__ rowstart)
{
    __shared__ int32_t part_sum[1024];
    const int t = threadIdx.x;
    const int64_t per = (blocks + 1023) / 1024;
    const int64_t b0 = t * per, b1 = (b0 + per < blocks) ? b0 + per : blocks;
    int32_t sum = 0;
    for (int64_t i = b0; i < b1; ++i) {
        const int64_t l = kd_share_leaf(i, part, parts);
        sum += (int32_t)(kd_leaf_start(l + 1, n, leaves) - kd_leaf_start(l, n, leaves));
    }
    part_sum[t] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int32_t v = (t >= o) ? part_sum[t - o] : 0;
        __syncthreads();
        part_sum[t] += v;
        __syncthreads();
    }
    int32_t run = part_sum[t] - sum;                 // exclusive
    for (int64_t i = b0; i < b1; ++i) {
        rowstart[i] = run;
        const int64_t l = kd_share_leaf(i, part, parts);
        run += (int32_t)(kd_leaf_start(l + 1, n, leaves) - kd_leaf_start(l, n, leaves));
    }
    if (t == 1023) rowstart[blocks] = part_sum[1023];
}

// the share's query slots (128 per block: the tile slots of the block's leaf) and, optionally, the sample of every row
__global__ void __launch_bounds__(128) kd_share_slots_kernel(const int *__restrict__ slot_idx, const uint32_t *__restrict__ perm,
                                                             int64_t n, int64_t leaves, int part, int parts,
                                                             const int32_t *__restrict__ rowstart, int *__restrict__ qorder,
                                                             int64_t *__restrict__ out_rows)
{
    const int64_t i = blockIdx.x;
    const int t = threadIdx.x;
    const int64_t l = kd_share_leaf(i, part, parts);
    qorder[i * 128 + t] = slot_idx[l * 128 + t];
    if (out_rows) {
        const int64_t start = kd_leaf_start(l, n, leaves);
        const int64_t cnt = kd_leaf_start(l + 1, n, leaves) - start;
        if (t < cnt) out_rows[rowstart[i] + t] = (int64_t)perm[start + t];
    }
}

// keys of one level: (number of the point's current cell) << coord_bits | leading bits of its centred coordinate
template <int D>
__global__ void __launch_bounds__(256)
    kd_key_kernel(const double *__restrict__ x, Weights wt, const KnnHeader *__restrict__ h, KdPlan p, int level,
                  int64_t leaves_per_cell, const uint32_t *__restrict__ perm_in, uint64_t *__restrict__ keys,
                  uint32_t *__restrict__ vals)
{
    const int64_t pos = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= p.n) return;
    const uint32_t j = perm_in ? perm_in[pos] : (uint32_t)pos;
    const int d = h->dim_order[level];
    // dim_order is data; the coordinate is fetched by a runtime index from global memory, the weight from a
    // small switch (wt lives in the kernel's parameter space)
    double w = wt.w[0];
#pragma unroll
    for (int dd = 1; dd < D; ++dd)
        if (dd == d) w = wt.w[dd];
    const float f = (float)d_sub(d_mul(w, x[(int64_t)j * D + d]), h->center[d]);
    const uint64_t cell = (uint64_t)(kd_cell_start(p, kd_leaf_of(pos, p.n, p.leaves), level) / leaves_per_cell);
    keys[pos] = (cell << p.coord_bits) | (uint64_t)(ord_f32(f) >> (32 - p.coord_bits));
    vals[pos] = j;
}

// One level of the ordering for cells that fit a block: cell c (a run of consecutive positions) is sorted by
// the level's coordinate in shared memory, one block per cell -- a bitonic network over 32-bit keys
// (the leading 32 - log2 P bits of the coordinate's order-preserving image, then the point's place in the cell,
// which makes the keys distinct and the result that of a stable sort).  Replaces the key kernel and the four radix
// passes of the level (13 launches, 170 us at 1M points) where the cells hold at most kKdBlockSortMax points: the
// last three of seven levels at 1M x 7-DoF (prepare 1.08 -> 0.75 ms; 16384-point cells, a fourth level: 0.74).
constexpr int kKdBlockSortMax = 4096;

template <int D>
__global__ void __launch_bounds__(1024)
    kd_cell_sort_kernel(const double *__restrict__ x, Weights wt, const KnnHeader *__restrict__ h, KdPlan p, int level,
                        int64_t leaves_per_cell, int P, const uint32_t *__restrict__ perm_in, uint32_t *__restrict__ perm_out)
{
    extern __shared__ uint32_t sk[];
    const int tid = threadIdx.x;
    const int64_t c = blockIdx.x;
    const int64_t start = kd_leaf_start(c * leaves_per_cell, p.n, p.leaves);
    const int cnt = (int)(kd_leaf_start((c + 1) * leaves_per_cell, p.n, p.leaves) - start);
    const int d = h->dim_order[level];
    double w = wt.w[0];
#pragma unroll
    for (int dd = 1; dd < D; ++dd)
        if (dd == d) w = wt.w[dd];
    const double ctr = h->center[d];
    const int idx_bits = 31 - __clz(P);                          // P is a power of two
    const uint32_t top_max = (0xffffffffu >> idx_bits) - 1u;
    for (int i = tid; i < P; i += blockDim.x) {
        uint32_t key = 0xffffffffu;                              // padding: after every real key
        if (i < cnt) {
            const int64_t j = perm_in ? (int64_t)perm_in[start + i] : start + i;
            const float f = (float)d_sub(d_mul(w, x[j * D + d]), ctr);
            uint32_t top = ord_f32(f) >> idx_bits;
            top = top < top_max ? top : top_max;
            key = (top << idx_bits) | (uint32_t)i;
        }
        sk[i] = key;
    }
    __syncthreads();
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (P >> 1); t += blockDim.x) {
                const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int hi = lo | j;
                const uint32_t a = sk[lo], b = sk[hi];
                const bool up = (lo & k) == 0;
                if ((a > b) == up) {
                    sk[lo] = b;
                    sk[hi] = a;
                }
            }
            __syncthreads();
        }
    }
    for (int i = tid; i < cnt; i += blockDim.x) {
        const int64_t li = (int64_t)(sk[i] & (uint32_t)(P - 1));
        perm_out[start + i] = perm_in ? perm_in[start + li] : (uint32_t)(start + li);
    }
}

// ---- stable LSD radix sort of (64-bit key, 32-bit value) pairs, one 8-bit digit per pass -----------------
// Pass = digit histogram per block -> exclusive scan over (digit, block) -> stable scatter.  A block owns
// kRsChunk consecutive keys; inside it warp w owns a contiguous run of them and walks it 32 keys at a time, so
// "earlier in the array" is "earlier warp, earlier step, lower lane" and the rank of a key among the equal
// digits before it comes from match.any + popcount.
__global__ void __launch_bounds__(kRsThreads)
    rs_hist_kernel(const uint64_t *__restrict__ keys, int64_t n, int shift, uint32_t *__restrict__ hist, int nblocks)
{
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kRsChunk;
#pragma unroll 4
    for (int i = 0; i < kRsItems; ++i) {
        const int64_t idx = base + (int64_t)i * kRsThreads + threadIdx.x;
        if (idx < n) atomicAdd(&h[(uint32_t)(keys[idx] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// per digit (one block each): exclusive scan of the digit's per-block counts in place, and the digit's total
__global__ void __launch_bounds__(256) rs_scan_kernel(uint32_t *__restrict__ hist, int nblocks, uint32_t *__restrict__ totals)
{
    __shared__ uint32_t warp_sums[8];
    __shared__ uint32_t carry_s;
    uint32_t *row = hist + (size_t)blockIdx.x * nblocks;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += 256) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < nblocks ? row[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        uint32_t before = carry_s;
        for (int w = 0; w < warp; ++w) before += warp_sums[w];
        if (i < nblocks) row[i] = before + inc - v;
        __syncthreads();
        if (threadIdx.x == 255) carry_s = before + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry_s;
}

__global__ void __launch_bounds__(kRsThreads)
    rs_scatter_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ vals, int64_t n, int shift,
                      const uint32_t *__restrict__ hist, int nblocks, const uint32_t *__restrict__ totals,
                      uint64_t *__restrict__ out_keys, uint32_t *__restrict__ out_vals)
{
    constexpr int W = kRsThreads / 32;
    constexpr int STEPS = kRsChunk / kRsThreads;           // 32-key steps per warp
    __shared__ uint32_t wh[W][256];
    __shared__ uint32_t digit_base[256], wsum[W];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < W * 256; i += kRsThreads) (&wh[0][0])[i] = 0;
    {
        // first output position of every digit: exclusive scan of the 256 digit totals
        const uint32_t v = totals[threadIdx.x];
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        uint32_t before = 0;
        for (int w = 0; w < warp; ++w) before += wsum[w];
        digit_base[threadIdx.x] = before + inc - v;
    }
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kRsChunk + (int64_t)warp * (STEPS * 32);
    uint64_t k[STEPS];
    uint32_t v[STEPS];
#pragma unroll
    for (int s = 0; s < STEPS; ++s) {
        const int64_t idx = base + s * 32 + lane;
        const bool in = idx < n;
        k[s] = in ? keys[idx] : 0ull;
        v[s] = in ? vals[idx] : 0u;
        const uint32_t dg = in ? ((uint32_t)(k[s] >> shift) & 255u) : 256u;
        const unsigned m = __match_any_sync(0xffffffffu, dg);
        if (in && lane == __ffs(m) - 1) wh[warp][dg] += __popc(m);
        __syncwarp();
    }
    __syncthreads();
    {
        const int d = threadIdx.x;                        // kRsThreads == 256 digits
        uint32_t run = digit_base[d] + hist[(size_t)d * nblocks + blockIdx.x];
#pragma unroll
        for (int w = 0; w < W; ++w) {
            const uint32_t t = wh[w][d];
            wh[w][d] = run;
            run += t;
        }
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < STEPS; ++s) {
        const int64_t idx = base + s * 32 + lane;
        const bool in = idx < n;
        const uint32_t dg = in ? ((uint32_t)(k[s] >> shift) & 255u) : 256u;
        const unsigned m = __match_any_sync(0xffffffffu, dg);
        if (in) {
            const uint32_t pos = wh[warp][dg] + __popc(m & ((1u << lane) - 1u));
            out_keys[pos] = k[s];
            out_vals[pos] = v[s];
        }
        __syncwarp();
        if (in && lane == __ffs(m) - 1) wh[warp][dg] += __popc(m);
        __syncwarp();
    }
}
static_assert(kRsThreads == 256, "one thread per digit in the scatter kernel");

struct KdScratch {
    uint64_t *keys[2];
    uint32_t *vals[2];
    uint32_t *hist;
};

inline size_t kd_scratch_bytes(int64_t n)
{
    const size_t nn = (size_t)(n > 0 ? n : 1);
    const size_t nblocks = (nn + kRsChunk - 1) / kRsChunk;
    return 2 * ((nn * 8 + 255) & ~(size_t)255) + 2 * ((nn * 4 + 255) & ~(size_t)255) + (((nblocks + 1) * 256 * 4 + 255) & ~(size_t)255);
}

inline KdScratch kd_scratch_at(void *base, int64_t n)
{
    const size_t nn = (size_t)(n > 0 ? n : 1);
    const size_t kb = (nn * 8 + 255) & ~(size_t)255, vb = (nn * 4 + 255) & ~(size_t)255;
    char *p = reinterpret_cast<char *>(base);
    KdScratch s;
    s.keys[0] = reinterpret_cast<uint64_t *>(p);
    s.keys[1] = reinterpret_cast<uint64_t *>(p + kb);
    s.vals[0] = reinterpret_cast<uint32_t *>(p + 2 * kb);
    s.vals[1] = reinterpret_cast<uint32_t *>(p + 2 * kb + vb);
    s.hist = reinterpret_cast<uint32_t *>(p + 2 * kb + 2 * vb);
    return s;
}

// Sorts in place of the scratch: the pairs start in buffer `cur` and end in the returned buffer index.
inline int rs_sort(const KdScratch &s, int cur, int64_t n, int key_bits, cudaStream_t st, int *launches)
{
    const int nblocks = (int)((n + kRsChunk - 1) / kRsChunk);
    for (int shift = 0; shift < key_bits; shift += 8) {
        rs_hist_kernel<<<nblocks, kRsThreads, 0, st>>>(s.keys[cur], n, shift, s.hist, nblocks);
        rs_scan_kernel<<<256, 256, 0, st>>>(s.hist, nblocks, s.hist + (size_t)256 * nblocks);
        rs_scatter_kernel<<<nblocks, kRsThreads, 0, st>>>(s.keys[cur], s.vals[cur], n, shift, s.hist, nblocks,
                                                          s.hist + (size_t)256 * nblocks, s.keys[cur ^ 1], s.vals[cur ^ 1]);
        cur ^= 1;
        if (launches) *launches += 3;
    }
    return cur;
}

// The whole ordering: after the call, *perm_out (one of the scratch value buffers) lists the point indices in
// leaf order.  n >= 1.
template <int D>
int kd_order(const double *x, int64_t n, const Weights &wt, const KnnHeader *h, const KdPlan &plan, const KdScratch &s,
             const uint32_t **perm_out, cudaStream_t st, int *launches)
{
    int cur = 0;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    if (plan.levels == 0) {
        // a single leaf: the identity order (written through the same kernel for uniformity)
        KdPlan p1 = plan;
        kd_key_kernel<D><<<blocks, 256, 0, st>>>(x, wt, h, p1, 0, p1.leaves, nullptr, s.keys[0], s.vals[0]);
        MPB_LAUNCH_CHECK();
        if (launches) *launches += 1;
        *perm_out = s.vals[0];
        return MPB_OK;
    }
    int64_t cells = 1;                                    // cells at the current level = product of the cuts so far
    for (int l = 0; l < plan.levels; ++l) {
        const uint32_t *perm_in = (l == 0) ? nullptr : s.vals[cur];
        const int64_t leaves_per_cell = plan.leaves / cells;
        const int64_t max_cnt = (n * leaves_per_cell + plan.leaves - 1) / plan.leaves + 1;
        if (max_cnt <= kKdBlockSortMax) {
            int P = 2;
            while (P < max_cnt) P <<= 1;
            const int threads = P / 2 < 32 ? 32 : (P / 2 > 1024 ? 1024 : P / 2);
            if ((size_t)P * sizeof(uint32_t) > 48 * 1024)
                MPB_CUDA_CHECK(cudaFuncSetAttribute(kd_cell_sort_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                    kKdBlockSortMax * (int)sizeof(uint32_t)));
            kd_cell_sort_kernel<D><<<(unsigned)cells, threads, (size_t)P * sizeof(uint32_t), st>>>(
                x, wt, h, plan, l, leaves_per_cell, P, perm_in, s.vals[cur ^ 1]);
            MPB_LAUNCH_CHECK();
            if (launches) *launches += 1;
            cur ^= 1;
        } else {
            // the keys are written into the buffer the sort will read; the values ride along
            kd_key_kernel<D><<<blocks, 256, 0, st>>>(x, wt, h, plan, l, leaves_per_cell, perm_in, s.keys[cur ^ 1],
                                                     s.vals[cur ^ 1]);
            MPB_LAUNCH_CHECK();
            if (launches) *launches += 1;
            cur ^= 1;
            // the cell number takes log2(cells) bits, not the leaf number's: one radix pass fewer on the upper levels
            const int bits = plan.coord_bits + ceil_log2_i64(cells);
            cur = rs_sort(s, cur, n, bits, st, launches);
            MPB_LAUNCH_CHECK();
        }
        cells *= plan.ways[l];
    }
    *perm_out = s.vals[cur];
    return MPB_OK;
}
