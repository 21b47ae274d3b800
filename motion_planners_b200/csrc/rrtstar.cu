// One RRT* iteration's device work in a single call (reference: rrt_star.py:100-137 as executed on Python 3).
//
// The reference's iteration is: first arg-min nearest node to the sample (rrt_star.py:117), the collision-free
// prefix of the extension towards it (safe_path, :118 / :76-82), one new node at the end of that prefix (:121),
// then for every node strictly closer than `radius` to the new one (:128, new node included) the distance and
// the collision-free prefix of the edge node -> new, accepted when it ends within EPSILON of the new node
// (:133-137).  The cost bookkeeping that decides WHICH of those candidates becomes the parent is sequential
// (each rewire changes new.cost) and stays with the caller; everything the caller needs for it comes back in
// one small host record.  Four launches and two small device->host copies per iteration, no allocation.
#include <math.h>

#include "mpb_common.cuh"

namespace {

const size_t kMaxWorldSmem = 200 * 1024;
const int kScanBlocks = 148;     // partial arg-mins of the node scan

struct Sample {
    double v[MPB_MAX_DIM];
};

// device copy of the iteration record; the first MPB_RRTSTAR_HEADER_BYTES travel to the host
struct StepHeader {
    int32_t nearest, first, cnt, safe, n_cand, pad;
    double d_near_new;
    double new_cfg[MPB_MAX_DIM];
};
static_assert(sizeof(StepHeader) == MPB_RRTSTAR_HEADER_BYTES, "header layout is part of the ABI");

struct Cand {
    int32_t idx, safe, ok, pad;
    double d;
};
static_assert(sizeof(Cand) == MPB_RRTSTAR_CAND_BYTES, "candidate layout is part of the ABI");

struct Partial {
    double d;
    int32_t i, pad;
};

const int kInlineCands = 256;   // candidates copied to the host together with the header

// header + candidate records, rounded up so that the partials behind them stay 16-byte aligned
inline size_t cand_region_bytes(int64_t capacity)
{
    return (sizeof(StepHeader) + sizeof(Cand) * (size_t)(capacity + 1) + 255) & ~(size_t)255;
}

// ---- (1) first arg-min of distance_fn(node, sample) over the nodes: per-block partials -----------
template <int D>
__global__ void __launch_bounds__(256)
    nearest_partials_kernel(const double *__restrict__ table, int64_t n, Sample s, Weights wt, int metric,
                            Partial *__restrict__ partial)
{
    __shared__ double sd[8];
    __shared__ int si[8];
    double tg[D];
#pragma unroll
    for (int d = 0; d < D; ++d) tg[d] = s.v[d];
    double best = INFINITY;
    int bi = 0x7fffffff;
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        double a[D];
#pragma unroll
        for (int d = 0; d < D; ++d) a[d] = table[j * D + d];
        const double dd = metric_dist<D>(a, tg, wt, metric);
        if (dd < best) {    // j ascends within a thread: the first minimum is kept
            best = dd;
            bi = (int)j;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double od = __shfl_down_sync(0xffffffffu, best, o);
        const int oi = __shfl_down_sync(0xffffffffu, bi, o);
        if (od < best || (od == best && oi < bi)) {
            best = od;
            bi = oi;
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        sd[warp] = best;
        si[warp] = bi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
            if (sd[w] < best || (sd[w] == best && si[w] < bi)) {
                best = sd[w];
                bi = si[w];
            }
        partial[blockIdx.x].d = best;
        partial[blockIdx.x].i = bi;
    }
}

// ---- (2) the extension nearest -> sample: one block, obstacles spread over the threads ------------
template <int D>
__global__ void __launch_bounds__(128)
    extend_kernel(WorldDev w, double *__restrict__ table, int64_t n, Sample s, Weights wt, int metric, double step,
                  const Partial *__restrict__ partial, int n_partial, StepHeader *__restrict__ h)
{
    extern __shared__ __align__(16) double sm[];
    __shared__ int s_first;
    const WorldSmem ws = stage_world(w, sm);   // ends with __syncthreads()
    // reduce the partial arg-mins (every thread redundantly: n_partial is small)
    double best = INFINITY;
    int bi = 0x7fffffff;
    for (int p = 0; p < n_partial; ++p) {
        const double pd = partial[p].d;
        const int pi = partial[p].i;
        if (pd < best || (pd == best && pi < bi)) {
            best = pd;
            bi = pi;
        }
    }
    if (threadIdx.x == 0) s_first = 0x7fffffff;
    __syncthreads();
    double a[D], b[D], q[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        a[d] = table[(int64_t)bi * D + d];
        b[d] = s.v[d];
        q[d] = a[d];
    }
    const int n_way = waypoint_count<D>(a, b, step);
    // item 0 = the environment, then the boxes, then the spheres
    const int items = 1 + w.n_box + w.n_sph;
    int mine = 0x7fffffff;
    if ((int)threadIdx.x < items) {
        for (int i = 0; i < n_way; ++i) {
            waypoint_step<D>(q, b, n_way, i);
            bool hit = false;
            for (int it = threadIdx.x; it < items && !hit; it += blockDim.x) {
                if (it == 0)
                    hit = outside_env<D>(ws, q);
                else if (it <= w.n_box)
                    hit = in_box<D>(ws, it - 1, q);
                else
                    hit = in_sphere<D>(ws, it - 1 - w.n_box, q);
            }
            if (hit) {
                mine = i;
                break;
            }
        }
    }
    if (mine != 0x7fffffff) atomicMin(&s_first, mine);
    __syncthreads();
    if (threadIdx.x == 0) {
        const int first = (s_first == 0x7fffffff) ? -1 : s_first;
        const int safe = (first < 0) ? n_way : first;      // len(safe_path(...)), rrt_star.py:118
        h->nearest = bi;
        h->first = first;
        h->cnt = n_way;
        h->safe = safe;
        h->n_cand = 0;
        h->pad = 0;
        double c[D];
#pragma unroll
        for (int d = 0; d < D; ++d) c[d] = a[d];
        for (int i = 0; i < safe; ++i) waypoint_step<D>(c, b, n_way, i);
        h->d_near_new = (safe > 0) ? metric_dist<D>(a, c, wt, metric) : 0.;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            h->new_cfg[d] = c[d];
            if (safe > 0) table[n * D + d] = c[d];          // the new node takes slot n
        }
    }
}

// ---- (3) nodes strictly closer than radius to the new node (the new node included) -----------------
template <int D>
__global__ void __launch_bounds__(256)
    radius_kernel(const double *__restrict__ table, int64_t n1, Weights wt, int metric, double radius,
                  StepHeader *__restrict__ h, Cand *__restrict__ cand)
{
    if (h->safe <= 0) return;
    double nw[D];
#pragma unroll
    for (int d = 0; d < D; ++d) nw[d] = h->new_cfg[d];
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n1; j += (int64_t)gridDim.x * blockDim.x) {
        double a[D];
#pragma unroll
        for (int d = 0; d < D; ++d) a[d] = table[j * D + d];
        const double dd = metric_dist<D>(a, nw, wt, metric);
        if (dd < radius) {                                  // strict, rrt_star.py:128
            const int slot = atomicAdd(&h->n_cand, 1);      // order restored by the caller (sort by idx)
            cand[slot].idx = (int)j;
            cand[slot].d = dd;
        }
    }
}

// ---- (4) the edge candidate -> new node for every candidate: one thread each ----------------------------
template <int D>
__global__ void __launch_bounds__(128)
    candidates_kernel(WorldDev w, const double *__restrict__ table, Weights wt, int metric, double step, double eps,
                      const StepHeader *__restrict__ h, Cand *__restrict__ cand)
{
    extern __shared__ __align__(16) double sm[];
    const WorldSmem ws = stage_world(w, sm);
    if (h->safe <= 0) return;
    const int nc = h->n_cand;
    double nw[D];
#pragma unroll
    for (int d = 0; d < D; ++d) nw[d] = h->new_cfg[d];
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < nc; j += gridDim.x * blockDim.x) {
        const int idx = cand[j].idx;
        double a[D], q[D], prev[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            a[d] = table[(int64_t)idx * D + d];
            q[d] = a[d];
            prev[d] = a[d];
        }
        const int n_way = waypoint_count<D>(a, nw, step);
        int safe = n_way;
        for (int i = 0; i < n_way; ++i) {
            waypoint_step<D>(q, nw, n_way, i);
            bool hit = outside_env<D>(ws, q);
            for (int k = 0; k < w.n_box && !hit; ++k) hit = in_box<D>(ws, k, q);
            for (int k = 0; k < w.n_sph && !hit; ++k) hit = in_sphere<D>(ws, k, q);
            if (hit) {
                safe = i;
                break;
            }
#pragma unroll
            for (int d = 0; d < D; ++d) prev[d] = q[d];
        }
        // accepted iff the safe prefix is non-empty and ends within eps of the new node (rrt_star.py:135-136)
        const bool ok = (safe > 0) && (metric_dist<D>(nw, prev, wt, metric) < eps);
        cand[j].safe = safe;
        cand[j].ok = ok ? 1 : 0;
        cand[j].pad = 0;
    }
}

template <int D>
int launch_step(const mpb_world *world, double *table, int64_t n, int64_t capacity, const Sample &s, double radius,
                const Weights &wt, int metric, double step, double eps, void *workspace, void *out_host,
                cudaStream_t st)
{
    unsigned char *base = static_cast<unsigned char *>(workspace);
    StepHeader *h = reinterpret_cast<StepHeader *>(base);
    Cand *cand = reinterpret_cast<Cand *>(base + sizeof(StepHeader));
    Partial *partial = reinterpret_cast<Partial *>(base + cand_region_bytes(capacity));
    const size_t smem = world_smem_bytes(world->dev) + 16;
    int64_t blocks = (n + 255) / 256;
    if (blocks > kScanBlocks) blocks = kScanBlocks;
    nearest_partials_kernel<D><<<(unsigned)blocks, 256, 0, st>>>(table, n, s, wt, metric, partial);
    MPB_LAUNCH_CHECK();
    if (smem > 48 * 1024) {
        MPB_CUDA_CHECK(cudaFuncSetAttribute(extend_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MPB_CUDA_CHECK(
            cudaFuncSetAttribute(candidates_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    extend_kernel<D><<<1, 128, smem, st>>>(world->dev, table, n, s, wt, metric, step, partial, (int)blocks, h);
    MPB_LAUNCH_CHECK();
    int64_t rblocks = (n + 1 + 255) / 256;
    if (rblocks > kScanBlocks * 2) rblocks = kScanBlocks * 2;
    radius_kernel<D><<<(unsigned)rblocks, 256, 0, st>>>(table, n + 1, wt, metric, radius, h, cand);
    MPB_LAUNCH_CHECK();
    candidates_kernel<D><<<kScanBlocks, 128, smem, st>>>(world->dev, table, wt, metric, step, eps, h, cand);
    MPB_LAUNCH_CHECK();
    // The header and the first kInlineCands candidates sit next to each other on the device too, so the
    // common case (a handful of candidates) is one copy and one synchronisation.
    unsigned char *out = static_cast<unsigned char *>(out_host);
    size_t first_bytes = sizeof(StepHeader) + sizeof(Cand) * (size_t)kInlineCands;
    const size_t all_bytes = sizeof(StepHeader) + sizeof(Cand) * (size_t)(capacity + 1);
    if (first_bytes > all_bytes) first_bytes = all_bytes;
    MPB_CUDA_CHECK(cudaMemcpyAsync(out, h, first_bytes, cudaMemcpyDeviceToHost, st));
    MPB_CUDA_CHECK(cudaStreamSynchronize(st));
    const StepHeader *hh = reinterpret_cast<const StepHeader *>(out);
    if (hh->safe > 0 && hh->n_cand > kInlineCands) {
        MPB_CUDA_CHECK(cudaMemcpyAsync(out + first_bytes, reinterpret_cast<unsigned char *>(h) + first_bytes,
                                       sizeof(Cand) * (size_t)(hh->n_cand - kInlineCands), cudaMemcpyDeviceToHost,
                                       st));
        MPB_CUDA_CHECK(cudaStreamSynchronize(st));
    }
    return MPB_OK;
}

}   // namespace

MPB_API size_t mpb_rrt_star_workspace_bytes(int64_t capacity)
{
    if (capacity < 0) return 0;
    return cand_region_bytes(capacity) + sizeof(Partial) * (size_t)kScanBlocks;
}

MPB_API size_t mpb_rrt_star_record_bytes(int64_t capacity)
{
    if (capacity < 0) return 0;
    return sizeof(StepHeader) + sizeof(Cand) * (size_t)(capacity + 1);
}

MPB_API int mpb_rrt_star_step(const mpb_world *world, double *table, int64_t n_nodes, int64_t capacity,
                              const double *sample, double radius, const double *weights, int metric, double step,
                              double eps, void *workspace, size_t workspace_bytes, void *out_record,
                              size_t record_bytes, void *stream)
{
    const char *who = "mpb_rrt_star_step";
    MPB_REQUIRE(world != nullptr, "%s: world is NULL", who);
    MPB_REQUIRE_CSPACE(world, who);
    MPB_REQUIRE(table && sample && workspace && out_record, "%s: NULL buffer", who);
    MPB_REQUIRE(n_nodes >= 1 && n_nodes < capacity && capacity < 2147483647LL,
                "%s: need 1 <= n_nodes < capacity (the new node takes slot n_nodes)", who);
    MPB_REQUIRE(metric >= 0 && metric <= 2, "%s: bad metric %d", who, metric);
    MPB_REQUIRE(step > 0. && radius >= 0., "%s: step must be positive and radius non-negative", who);
    MPB_REQUIRE(workspace_bytes >= mpb_rrt_star_workspace_bytes(capacity), "%s: workspace too small", who);
    MPB_REQUIRE(record_bytes >= mpb_rrt_star_record_bytes(capacity), "%s: host record too small", who);
    MPB_REQUIRE(((uintptr_t)workspace & 255) == 0, "%s: workspace must be 256-byte aligned", who);
    MPB_REQUIRE(world_smem_bytes(world->dev) + 16 <= kMaxWorldSmem,
                "%s: obstacle set (%zu bytes) exceeds the %zu-byte shared-memory budget", who,
                world_smem_bytes(world->dev), kMaxWorldSmem);
    const int dim = world->dev.dim;
    Sample s;
    for (int d = 0; d < MPB_MAX_DIM; ++d) s.v[d] = d < dim ? sample[d] : 0.;
    const Weights wt = make_weights(weights, dim);
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(DD) \
    launch_step<DD>(world, table, n_nodes, capacity, s, radius, wt, metric, step, eps, workspace, out_record, st)
    MPB_DISPATCH_DIM(dim, CALL)
#undef CALL
    mpb_set_error("%s: unsupported dim %d", who, dim);
    return MPB_EINVAL;
}
