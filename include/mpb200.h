/*
 * mpb200.h -- C ABI of the B200-native kNN + edge-collision path for sampling-based
 * motion planners (drop-in for the hot path of caelan/motion-planners).
 *
 * The reference has no FFI layer: its operator API is four Python callables
 * (README.md:52-59).  This ABI is what a binding for the BATCHED forms of those
 * callables binds; each entry point cites the reference code it replaces
 * (paths relative to the reference checkout, motion_planners/...).
 *
 * Conventions
 *  - Plain C: pointers and sizes only.  `stream` is a cudaStream_t passed as void*
 *    (NULL = the legacy default stream).  Every call is asynchronous and ordered on
 *    `stream`; nothing synchronises the device unless stated.
 *  - "dev" pointers are device pointers on the CURRENT device, "host" pointers are
 *    host pointers.  The caller owns every buffer; the library allocates nothing
 *    except inside mpb_world_create / mpb_arm_world_create (and, for an arm world, 256
 *    stream-ordered bytes per large edge check from a pool that world owns: see there).
 *  - Configurations are row-major [n, D] IEEE binary64, contiguous, D <= MPB_MAX_DIM.
 *  - Return value: 0 on success, a negative MPB_E* code otherwise; mpb_last_error()
 *    returns a thread-local message.  Nothing throws, nothing aborts.
 *  - All results are bit-exact to the reference's fp64 arithmetic.
 */
#ifndef MPB200_H
#define MPB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPB_VERSION 1
#define MPB_MAX_DIM 16
#define MPB_MAX_K 32

#define MPB_OK 0
#define MPB_EINVAL (-22)   /* bad argument (size, dim, k, null pointer, misalignment) */
#define MPB_ENOMEM (-12)   /* workspace too small / allocation failed                */
#define MPB_ECUDA (-5)     /* a CUDA runtime call failed; see mpb_last_error()        */

typedef struct mpb_world mpb_world;

int mpb_version(void);
const char *mpb_last_error(void);

/* ---- obstacle set --------------------------------------------------------------------
 * Replaces the closure built by tkinter/samplers.py:88-98 get_collision_fn(environment,
 * obstacles) over tkinter/viewer.py:16-17 Box / Circle.  All array arguments are HOST
 * pointers and are copied to the current device (this is the one call that allocates).
 * box_lo/box_hi: [n_box, dim] raw bounds; sph_c: [n_sph, dim]; sph_r: [n_sph];
 * env_lo/env_hi: [dim]; buffer: obstacle inflation (viewer.py:70 MIN_PROXIMITY = 1e-3).
 * The inflated bounds lower-buffer, upper+buffer and radius+buffer are rounded once in
 * fp64, exactly as viewer.py:72-82 does on every call. */
int mpb_world_create(const double *box_lo, const double *box_hi, int n_box, const double *sph_c,
                     const double *sph_r, int n_sph, int dim, const double *env_lo,
                     const double *env_hi, double buffer, mpb_world **out);
int mpb_world_destroy(mpb_world *world);

/* ---- collision_fn over a batch of configurations ---------------------------------------
 * out[i] = collision_fn(q[i]) (tkinter/samplers.py:90-96): 1 if q is outside the closed
 * environment box or inside any inflated obstacle, else 0.  Replaces the per-vertex
 * calls at lazy_prm.py:173-178, rrt_connect.py:47, meta.py:24.
 * q: dev [n, dim] f64; out: dev [n] u8. */
int mpb_points_check(const mpb_world *world, const double *q, int64_t n, uint8_t *out, void *stream);

/* ---- a serial arm approximated by spheres, among boxes in the 3-D workspace ----------------
 * BASELINE.json configs[3]'s collision model.  The reference has no robot model -- its collision_fn is a
 * callable the caller supplies (README.md:52-59) -- so this entry point has no counterpart there; it builds a
 * collision_fn in the reference's terms: closed bounds on the configuration (tkinter/samplers.py:92) and, for
 * every sphere k and box b, ||clamp(c_k, lo_b, hi_b) - c_k|| <= r_k + buffer (tkinter/viewer.py:80-82: norm =
 * sqrt of a sequentially fused dot).  The model -- chain, spheres and the fixed-polynomial sincos of the forward
 * kinematics -- is stated in oracle/arm_oracle.c, which the kernels follow operation for operation.
 * Configuration q in R^n_joints; angle_i = fl(fl(q_i * theta_span[i]) + theta_lo[i]); frame i = frame i-1 turned
 * about its local axis[i] (1 = y, 2 = z); link i = a[i] along the turned x + d[i] along the turned z; sphere j
 * sits on frame sph_link[j] at sph_u[j] along x and sph_w[j] along z from the frame's origin.
 * The returned handle is an mpb_world: mpb_points_check, mpb_edges_check, mpb_edges_check_pairs and
 * mpb_edges_check_knn take it (configurations have n_joints coordinates); the entry points that are stated for
 * configuration-space obstacles only (circular joints, the resident-tree planners) refuse it.
 * box_lo/box_hi: HOST [n_box, 3] raw bounds; env_lo/env_hi: HOST [n_joints].
 * Memory: the world owns a CUDA memory pool; an edge check of 65 536 edges or more (modes ANY / FIRST) takes 256
 * stream-ordered bytes from it for the duration of its kernel (cudaMallocFromPoolAsync / cudaFreeAsync on the
 * caller's stream: no synchronisation, nothing from the device's default pool), and runs without them when the
 * pool cannot give any. */
#define MPB_ARM_MAX_JOINTS 8
#define MPB_ARM_MAX_SPHERES 32
typedef struct mpb_arm_model {
    int n_joints;
    int axis[MPB_ARM_MAX_JOINTS];
    double theta_lo[MPB_ARM_MAX_JOINTS];
    double theta_span[MPB_ARM_MAX_JOINTS];
    double a[MPB_ARM_MAX_JOINTS];
    double d[MPB_ARM_MAX_JOINTS];
    double base[3];
    int n_sph;
    int sph_link[MPB_ARM_MAX_SPHERES];
    double sph_u[MPB_ARM_MAX_SPHERES];
    double sph_w[MPB_ARM_MAX_SPHERES];
    double sph_r[MPB_ARM_MAX_SPHERES];
} mpb_arm_model;
int mpb_arm_world_create(const mpb_arm_model *model, const double *box_lo, const double *box_hi, int n_box,
                         const double *env_lo, const double *env_hi, double buffer, mpb_world **out);
/* Forward kinematics alone: out_centres dev [n, n_sph, 3] f64 = the sphere centres of q dev [n, n_joints]
 * (what the collision test of an arm world compares with the boxes; for tests and visualisation). */
int mpb_arm_fk(const mpb_world *world, const double *q, int64_t n, double *out_centres, void *stream);

/* ---- extend_fn + collision_fn fused over a batch of directed edges ----------------------
 * For each edge q1 -> q2: discretise with tkinter/samplers.py:113-127 get_extend_fn
 * (circular = {}, norm = inf): n = int(max_d |(q2-q1)/step|) + 1 waypoints from the serial
 * recurrence q <- (1/(n-i))*(q2-q) + q, each operation rounded separately; test every
 * waypoint i < n - skip_last with collision_fn.
 *   mode MPB_EDGE_ANY   out_hit = index of SOME colliding waypoint, -1 if none
 *                       (lazy_prm.py:189-195 check_edge; prm.py:250-251 with skip_last=1)
 *   mode MPB_EDGE_FIRST out_hit = smallest colliding index (primitives.py:23 takewhile,
 *                       rrt_star.py:76-82 safe_path)
 *   mode MPB_EDGE_LAST  out_hit = largest colliding index (backward extension,
 *                       primitives.py:13-16: the prefix of the reversed list)
 * out_n (may be NULL) receives n.  Both outputs are dev [E] i32. */
#define MPB_EDGE_ANY 0
#define MPB_EDGE_FIRST 1
#define MPB_EDGE_LAST 2
/* edges as index pairs into a sample table: samples dev [N, dim] f64, src/dst dev [E] i64.
 * A pair with a negative index (the "no neighbour" slot of a kNN row) is skipped:
 * out_hit = -1, out_n = 0. */
int mpb_edges_check(const mpb_world *world, const double *samples, int64_t n_samples,
                    const int64_t *src, const int64_t *dst, int64_t n_edges, double step, int mode,
                    int skip_last, int32_t *out_hit, int32_t *out_n, void *stream);
/* edges as the rows of a neighbour matrix (the output of the kNN entry points, consumed in place): entry
 * (r, j) is the edge from sample row_offset + r to sample nbr[r, j]; nbr dev [n_rows, row_len] i64, outputs
 * dev [n_rows, row_len] i32.  Same verdicts as mpb_edges_check on those pairs (negative = no neighbour:
 * out_hit = -1, out_n = 0), but the obstacle culling is done once per row instead of once per edge.
 * This is lazy_prm.py:211-215 check_roadmap over the edges lazy_prm.py:148-160 add_samples creates. */
int mpb_edges_check_knn(const mpb_world *world, const double *samples, int64_t n_samples,
                        const int64_t *nbr, int64_t n_rows, int row_len, int64_t row_offset, double step,
                        int mode, int skip_last, int32_t *out_hit, int32_t *out_n, void *stream);
/* mpb_edges_check_knn for rows that are not a contiguous range of samples: row r leaves sample row_ids[r]
 * (dev [n_rows] i64) -- the rows of one rank of a roadmap sharded by spatial order (mpb_knn_search_part). */
int mpb_edges_check_knn_rows(const mpb_world *world, const double *samples, int64_t n_samples,
                             const int64_t *row_ids, const int64_t *nbr, int64_t n_rows, int row_len, double step,
                             int mode, int skip_last, int32_t *out_hit, int32_t *out_n, void *stream);
/* edges as explicit endpoint arrays: q1, q2 dev [E, dim] f64 */
int mpb_edges_check_pairs(const mpb_world *world, const double *q1, const double *q2,
                          int64_t n_edges, double step, int mode, int skip_last, int32_t *out_hit,
                          int32_t *out_n, void *stream);
/* Waypoint counts only: out_n[e] = int(max_d |(q2-q1)/step|) + 1 (tkinter/samplers.py:119-120). */
int mpb_edges_count(const double *q1, const double *q2, int64_t n_edges, int dim, double step,
                    int32_t *out_n, void *stream);
/* The waypoints themselves (extend_fn output) for one edge batch, for callers that need
 * the path (lazy_prm.py:223-230 get_path, prm.py:51-58 Edge.path): offsets dev [E+1] i64 is
 * the exclusive prefix sum of the counts from out_n; out dev [offsets[E], dim] f64. */
int mpb_edges_waypoints(const double *q1, const double *q2, int64_t n_edges, int dim, double step,
                        const int64_t *offsets, double *out, void *stream);

/* ---- circular (wrap-around) coordinates ---------------------------------------------------
 * The same three operations for configurations with circular joints: tkinter/samplers.py:113-127
 * get_extend_fn(circular=...) with primitives.py:74-78 get_difference_fn and utils.py:364-369,381-384
 * wrap_interval / circular_difference (Python float %: result takes the sign of the divisor).
 * circ_lo, circ_hi: HOST [dim]; a coordinate with (-inf, +inf) is not circular, otherwise the interval
 * must be finite with lower < upper.  Waypoints: q <- c*diff(q2, q) + q, then every circular coordinate
 * wrapped into its interval; n = int(max_d |diff(q2, q1)_d / step|) + 1.
 * mpb_edges_check_circular: src/dst both NULL = explicit endpoint arrays a, b dev [E, dim]; otherwise
 * a = samples dev [N, dim], b unused, src/dst dev [E] i64 (negative = no edge: out_hit -1, out_n 0).
 * Modes, skip_last and outputs as mpb_edges_check. */
int mpb_edges_check_circular(const mpb_world *world, const double *a, const double *b,
                             const int64_t *src, const int64_t *dst, int64_t n_edges,
                             const double *circ_lo, const double *circ_hi, double step, int mode,
                             int skip_last, int32_t *out_hit, int32_t *out_n, void *stream);
int mpb_edges_count_circular(const double *q1, const double *q2, int64_t n_edges, int dim,
                             const double *circ_lo, const double *circ_hi, double step,
                             int32_t *out_n, void *stream);
int mpb_edges_waypoints_circular(const double *q1, const double *q2, int64_t n_edges, int dim,
                                 const double *circ_lo, const double *circ_hi, double step,
                                 const int64_t *offsets, double *out, void *stream);

/* ---- k nearest neighbours ---------------------------------------------------------------
 * Replaces nearest.py:36-85 KDNeighbors.query_neighbors (scipy cKDTree, p = 2, eps = 0,
 * no circular dimensions) as called per sample by lazy_prm.py:148-160, and the full
 * distance sort of prm.py:244-246.  Points are embedded as e = fl(weights * x)
 * (primitives.py:55-57); the ranking key is cKDTree's squared distance (4-way unrolled
 * fp64 sum, no FMA), ties broken by the smaller index; out_dist = sqrt(key).  Only
 * neighbours with key < fl(max_dist*max_dist) are returned (scipy's distance_upper_bound);
 * missing slots hold idx = -1, dist = +inf.  Rows are sorted ascending.
 * samples dev [N, dim], queries dev [M, dim] f64; weights HOST [dim] (NULL = ones);
 * out_idx dev [M, k] i64; out_dist dev [M, k] f64 (may be NULL); 1 <= k <= MPB_MAX_K. */

/* Exact fp64 brute force: one warp per query, no scratch memory. */
int mpb_knn_f64(const double *samples, int64_t n_samples, const double *queries, int64_t n_queries,
                int dim, const double *weights, int k, double max_dist, int64_t *out_idx,
                double *out_dist, void *stream);

/* The same exact scan under any of the reference's distance procedures (MPB_METRIC_* below).  For
 * the non-kd-tree metrics the ranking key is the ROUNDED distance distance_fn returns (the reference
 * sorts on that value: prm.py:244-246 sorted(...), nearest.py:102-109 BruteForceNeighbors), ties to the
 * smaller index (Python's sort is stable), and a neighbour is kept iff distance <= max_dist
 * (prm.py:244 connect_distance filter).  metric = MPB_METRIC_KDTREE is mpb_knn_f64. */
int mpb_knn_f64_metric(const double *samples, int64_t n_samples, const double *queries,
                       int64_t n_queries, int dim, const double *weights, int metric, int k,
                       double max_dist, int64_t *out_idx, double *out_dist, void *stream);

/* mpb_knn_f64 with circular coordinates: nearest.py:7-20 expand_circular + nearest.py:74-85
 * query_neighbors shift the query by {-E, 0, +E} (E = upper - lower) in every circular coordinate,
 * query the kd-tree once per image and keep each index's smallest distance.  Here every sample's
 * key is the kd-tree sum over the per-coordinate smallest of the three image terms (equal to the
 * smallest image key, because rounded addition is monotone).  circ_lo / circ_hi as above. */
int mpb_knn_f64_circular(const double *samples, int64_t n_samples, const double *queries,
                         int64_t n_queries, int dim, const double *weights, const double *circ_lo,
                         const double *circ_hi, int k, double max_dist, int64_t *out_idx,
                         double *out_dist, void *stream);

/* Tiled fp32 filter + fp64 re-rank, same results bit for bit: an fp32 pass over
 * shared-memory tiles keeps 32 candidates per query, the exact fp64 keys of those
 * candidates are ranked, and a rigorous rounding-error bound proves that no excluded
 * sample can belong to the answer; queries for which the proof fails are redone by the
 * exact kernel inside the same call.
 * workspace: dev, 256-byte aligned, at least mpb_knn_workspace_bytes(N, M, dim) bytes.
 * stats (dev [2] i64, may be NULL): [0] = queries re-done exactly, [1] = candidate tiles read by the
 * tensor-core scan, summed over its query blocks (0 on the other paths). */
size_t mpb_knn_workspace_bytes(int64_t n_samples, int64_t n_queries, int dim);
int mpb_knn_prepare(const double *samples, int64_t n_samples, int dim, const double *weights,
                    void *workspace, size_t workspace_bytes, void *stream);
int mpb_knn_search(const double *samples, int64_t n_samples, const double *queries,
                   int64_t n_queries, int dim, const double *weights, int k, double max_dist,
                   void *workspace, size_t workspace_bytes, int64_t *out_idx, double *out_dist,
                   int64_t *stats, void *stream);
/* prepare + search in one call */
int mpb_knn_f32filter(const double *samples, int64_t n_samples, const double *queries,
                      int64_t n_queries, int dim, const double *weights, int k, double max_dist,
                      void *workspace, size_t workspace_bytes, int64_t *out_idx, double *out_dist,
                      int64_t *stats, void *stream);

/* The same search for callers that cannot allocate device memory -- a planner thread in a control
 * loop that must stay away from the driver's allocator (process-wide lock, possible device
 * synchronisation, unbounded latency), a process whose device memory was handed out in full to
 * another pool at start-up, or a thin foreign-language binding that owns no device allocator.  The
 * scratch is a fixed array of MPB_KNN_STATIC_BYTES inside the library, resident on every GPU that
 * loads it until the process exits; inputs and outputs are still the caller's.  No counterpart in
 * the reference (it has no device memory to manage).
 *  - This entry point always runs the packed-fp32 filter (never the tensor-core scan), and its scratch
 *    need is that filter's alone (smaller than mpb_knn_workspace_bytes for sets the tensor-core scan
 *    serves).  Sizes whose need exceeds MPB_KNN_STATIC_BYTES are refused with MPB_ENOMEM and a
 *    message; 4 MiB covers e.g. 100k samples x 100k queries at dim <= 7.
 *  - The first call on each device looks up the array's device address; make that call before
 *    starting any stream capture on that device.
 *  - The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).
 *  - All calls share the one array and it keeps the previous call's bytes; each call rewrites what
 *    it reads (header, counters, tile slots), so results do not depend on history -- but two calls
 *    that overlap on the GPU (different streams, or threads) race on it.  Serialise them. */
#define MPB_KNN_STATIC_BYTES (4u << 20)
int mpb_knn_f32filter_static(const double *samples, int64_t n_samples, const double *queries,
                             int64_t n_queries, int dim, const double *weights, int k,
                             double max_dist, int64_t *out_idx, double *out_dist, int64_t *stats,
                             void *stream);

/* ---- tree queries (RRT family) ----------------------------------------------------------
 * metric selects the reference's rounding procedure for the weighted Euclidean distance:
 *   MPB_METRIC_KDTREE     the kNN key above
 *   MPB_METRIC_PRIMITIVES primitives.py:60-62 get_distance_fn (sequential-FMA dot)
 *   MPB_METRIC_SAMPLERS   tkinter/samplers.py:11-16 get_distance_fn (sequential-FMA dot) */
#define MPB_METRIC_KDTREE 0
#define MPB_METRIC_PRIMITIVES 1
#define MPB_METRIC_SAMPLERS 2
/* For T independent trees stored as nodes dev [T, cap, dim] with sizes dev [T] i32 and one
 * target per tree (targets dev [T, dim]): out_idx[t] = first arg-min of distance_fn(node,
 * target) over the tree's nodes (utils.py:61-65 argmin as used by primitives.py:21 and
 * rrt_star.py:117), out_dist[t] its distance (may be NULL).  `active` (dev [T] u8, may be
 * NULL) skips trees whose flag is 0. */
int mpb_trees_nearest(const double *nodes, const int32_t *sizes, int64_t n_trees, int64_t cap,
                      int dim, const double *targets, const double *weights, int metric,
                      const uint8_t *active, int32_t *out_idx, double *out_dist, void *stream);
/* Append the kept part of one extension per entry to resident trees (primitives.py:24-27: one TreeNode per
 * safe waypoint, each the child of the previous one, the first the child of parent0).  Entry b extends tree
 * tree_ids[b] (negative = skip; at most one entry per tree per call): the waypoints of extend_fn(from[b],
 * to[b]) with generation index in [lo[b], hi[b]) are stored at the end of the tree, in generation order, or
 * reversed when reverse[b] != 0 (the backward extension of primitives.py:13-16).  nodes dev [T, cap, dim] f64,
 * parents dev [T, cap] i32, sizes dev [T] i32 (updated); out_last[b] = index of the last node stored (parent0[b]
 * when nothing was kept); *out_overflow is incremented for every entry that did not fit in cap (left unstored). */
int mpb_trees_append(double *nodes, int32_t *parents, int32_t *sizes, int64_t n_trees, int64_t cap, int dim,
                     const int32_t *tree_ids, const double *from, const double *to, const int32_t *lo,
                     const int32_t *hi, const uint8_t *reverse, const int32_t *parent0, int64_t n_ext,
                     double step, int32_t *out_last, int32_t *out_overflow, void *stream);
/* rrt_connect (rrt_connect.py:31-72) for n_queries independent queries in ONE launch, one thread block per
 * query: per iteration the swap decision (rrt_connect.py:26-29) from swap_draws[q, it], the target
 * targets[q, it, :] (the value sample_fn() returned), and two extend_towards steps (primitives.py:19-29) --
 * block-wide first arg-min over the tree, the extend_fn waypoints (forward for the start tree, generated from
 * the target and consumed reversed for the goal tree), collision_fn on every waypoint, one node per safe
 * waypoint.  A query whose start or goal collides fails at once (rrt_connect.py:47).
 * starts, goals dev [Q, dim]; swap_draws dev [Q, iters]; targets dev [Q, iters, dim]; weights HOST [dim];
 * nodes dev [2Q, cap, dim] f64 / parents dev [2Q, cap] i32 / sizes dev [2Q] i32 receive the trees (tree 2q =
 * start tree, 2q+1 = goal tree; contents on entry are ignored).  Per query: out_solved u8, out_iters i32
 * (iterations consumed), out_swapped u8 (the swap flag of the connecting iteration) and out_last_a / out_last_b
 * i32 = last node of the first- / second-extended tree of that iteration, from which the path is
 * configs(path1[:-1] + path2[::-1]) (rrt_connect.py:65-71).  *out_overflow counts queries stopped because a
 * tree reached cap.  Results equal composing mpb_trees_nearest, mpb_edges_check_pairs and mpb_trees_append. */
int mpb_rrt_connect_batch(const mpb_world *world, const double *starts, const double *goals,
                          int64_t n_queries, const double *weights, int metric, double step,
                          const double *swap_draws, const double *targets, int iters, double *nodes,
                          int32_t *parents, int32_t *sizes, int64_t cap, uint8_t *out_solved,
                          int32_t *out_last_a, int32_t *out_last_b, uint8_t *out_swapped,
                          int32_t *out_iters, int32_t *out_overflow, void *stream);
/* One tree: nearest node to q and every node with distance_fn(node, q) < radius
 * (rrt_star.py:117,128).  out_nearest dev [1] i32; out_dist dev [n] f64 receives all
 * distances; out_count dev [1] i32; out_list dev [n] i32 the qualifying indices ascending. */
int mpb_scan_nearest_radius(const double *nodes, int64_t n, int dim, const double *q,
                            const double *weights, int metric, double radius, int32_t *out_nearest,
                            double *out_dist, int32_t *out_count, int32_t *out_list, void *stream);

/* ---- one RRT* iteration ---------------------------------------------------------------------
 * The device work of one iteration of rrt_star.py:100-137 (as executed on Python 3) in one call:
 *   nearest   = first arg-min over nodes [0, n_nodes) of distance_fn(node, sample)        (rrt_star.py:117)
 *   safe      = len(safe_path(extend_fn(nearest, sample), collision_fn))                  (:118, :76-82)
 *   new       = waypoint safe-1 of that extension, stored in table[n_nodes] when safe > 0 (:121)
 *   candidates = every node j in [0, n_nodes] with distance_fn(node_j, new) < radius      (:128, strict)
 *   per candidate: d = that distance, safe_j = len(safe_path(extend_fn(node_j, new))), and
 *   ok = safe_j > 0 and distance_fn(new, last safe waypoint) < eps                        (:133-136)
 * The sequential cost / rewire bookkeeping (:134-137, OptimalNode) stays with the caller.
 * table dev [capacity, dim] f64 (dim from the world); sample, weights HOST; metric = MPB_METRIC_*;
 * workspace dev, 256-byte aligned, >= mpb_rrt_star_workspace_bytes(capacity);
 * out_record HOST (pinned for speed), >= mpb_rrt_star_record_bytes(capacity), written as
 *   { i32 nearest, first (-1 = none), cnt (waypoints of the extension), safe, n_cand, pad;
 *     f64 d_near_new; f64 new_cfg[MPB_MAX_DIM]; }                 = MPB_RRTSTAR_HEADER_BYTES
 *   followed by n_cand records { i32 idx, safe, ok, pad; f64 d; } = MPB_RRTSTAR_CAND_BYTES each,
 *   in no particular order (sort by idx for the reference's iteration order).
 * The call synchronises the stream (it returns host data). */
#define MPB_RRTSTAR_HEADER_BYTES 160
#define MPB_RRTSTAR_CAND_BYTES 24
size_t mpb_rrt_star_workspace_bytes(int64_t capacity);
size_t mpb_rrt_star_record_bytes(int64_t capacity);
int mpb_rrt_star_step(const mpb_world *world, double *table, int64_t n_nodes, int64_t capacity,
                      const double *sample, double radius, const double *weights, int metric,
                      double step, double eps, void *workspace, size_t workspace_bytes,
                      void *out_record, size_t record_bytes, void *stream);

/* ---- RRT*: many iterations per launch ----------------------------------------------------------
 * rrt_star.py:100-137 as executed on Python 3 (see mpb_rrt_star_step), with the tree, the costs and the parent
 * links resident on the device and the whole loop inside ONE cooperative kernel: per iteration the goal-biased
 * sample from the caller's pre-drawn streams, the informed-set rejection, nearest node, safe extension, new node,
 * goal test and the choose-parent pass.  The caller pre-draws what the reference would draw one by one:
 * u_draws dev [n_u] f64 = consecutive random.random() values; sample_draws dev [n_s, dim] f64 =
 * consecutive sample_fn() configurations.  The kernel consumes them in the reference's order (a random() only
 * while no solution exists and not at iteration 0; a configuration only when the goal was not chosen) and stops
 * when max_iterations iterations have been counted (status MPB_RRTSTAR_DONE), a stream runs dry
 * (MPB_RRTSTAR_NEED_DRAWS: call again with reset = 0 and fresh arrays that continue the streams; positions
 * consumed so far come back through mpb_rrt_star_run_read), or the node arrays are full (MPB_RRTSTAR_FULL: grow
 * them, copy, call again).  Node i: table[i] (configuration), cost[i], parent[i] (-1 for the root), and the
 * recipe of the waypoints strictly between parent and node: the first path_cnt[i] waypoints of
 * extend_fn(table[parent[i]], path_to[i]).  All arrays dev, capacity rows; state dev, 256-byte aligned,
 * mpb_rrt_star_run_state_bytes() bytes, owned by the caller for the life of the tree.  reset = 1 starts a new
 * tree at `start` (HOST [dim]); goal HOST [dim].  Synchronises the stream when reset = 1. */
#define MPB_RRTSTAR_RUNNING 0
#define MPB_RRTSTAR_DONE 1
#define MPB_RRTSTAR_NEED_DRAWS 2
#define MPB_RRTSTAR_FULL 3
#define MPB_RRTSTAR_CAND_OVERFLOW 4
size_t mpb_rrt_star_run_state_bytes(void);
int mpb_rrt_star_run(const mpb_world *world, const double *start, const double *goal, double radius,
                     const double *weights, int metric, double step, double eps, double goal_probability,
                     int informed, int64_t max_iterations, double *table, double *cost, int32_t *parent,
                     double *path_to, int32_t *path_cnt, int64_t capacity, const double *u_draws, int64_t n_u,
                     const double *sample_draws, int64_t n_s, void *state, int reset, void *stream);
/* out6 HOST [6] i64 = {n_nodes, iterations counted, random() draws consumed, configurations consumed, index of the
 * goal node or -1, status}; out_goal_cost HOST.  Synchronises the stream. */
int mpb_rrt_star_run_read(const void *state, int64_t *out6, double *out_goal_cost, void *stream);

/* Edge costs of a neighbour matrix under one of the reference's distance procedures: out_cost[r, j] =
 * distance_fn(samples[row_offset + r], samples[nbr[r, j]]) with the rounding of `metric` (MPB_METRIC_*), +inf for
 * a missing neighbour.  lazy_prm.py:216-222 Roadmap.get_cost evaluates cost_fn per edge on demand; the graph
 * search below wants them all, bit for bit what cost_fn returns.  out_cost dev [n_rows, row_len] f64. */
int mpb_edge_costs(const double *samples, int64_t n_samples, int dim, const double *weights, int metric,
                   const int64_t *nbr, int64_t n_rows, int row_len, int64_t row_offset, double *out_cost,
                   void *stream);

/* ---- shortest paths over a device-resident roadmap ------------------------------------------
 * The graph search either side of check_path (lazy_prm.py:41-58 dijkstra, :70-98 wastar_search) for the
 * roadmap the entry points above produce: nbr / nbr_dist dev [n_rows, row_len] (kNN output; row r = sample
 * row_offset + r), edge_hit dev [n_rows, row_len] i32 (mpb_edges_check_knn output: >= 0 = colliding),
 * vertex_hit dev [N] u8 indexed by sample (may be NULL).  Usable edges are travelled in both directions
 * (lazy_prm.py:141-147 REVERSIBLE).  dist dev [N] f64: +inf everywhere except 0 at the source(s) on entry.
 * mpb_roadmap_relax runs `sweeps` whole-graph relaxation sweeps, d[v] = min(d[v], fl(d[u] + w)), and sets
 * *changed (dev i32) to 1 if any distance moved; call until it stays 0.  mpb_roadmap_parents then stores, for
 * every reached vertex but the source, the smallest neighbour index u with fl(d[u] + w) == d[v]
 * (parent dev [N] i64, pre-filled with INT64_MAX by the caller). */
int mpb_roadmap_relax(const int64_t *nbr, const double *nbr_dist, const int32_t *edge_hit,
                      const uint8_t *vertex_hit, int64_t n_rows, int row_len, int64_t row_offset,
                      double *dist, int32_t *changed, int sweeps, void *stream);
int mpb_roadmap_parents(const int64_t *nbr, const double *nbr_dist, const int32_t *edge_hit,
                        const uint8_t *vertex_hit, int64_t n_rows, int row_len, int64_t row_offset,
                        const double *dist, int64_t source, int64_t *parent, void *stream);

/* Which filter mpb_knn_search / mpb_knn_f32filter run for a sample set of this size and dimension at this k
 * (for reporting; results are identical): the packed-fp32 FFMA2 scan, the tensor-core scan (tcgen05.mma
 * kind::tf32 on operands split into two TF32 parts, accumulator in tensor memory), or the exact fp64 scan alone.
 * The tensor-core scan is chosen for k <= 12 and sets of at least 32 768 / 32 768 / 24 576 / 16 384 samples in
 * 7 / 6 / 5 / 4 dimensions (where it was measured faster);
 * MPB200_KNN_TC=0 / =1 in the environment (read once) switches it off / forces it wherever it applies. */
#define MPB_KNN_KIND_FP32 0
#define MPB_KNN_KIND_TENSOR 1
#define MPB_KNN_KIND_EXACT 2
int mpb_knn_filter_kind(int64_t n_samples, int dim, int k);

/* Diagnostic for the tensor-core filter (no counterpart in the reference; scipy's cKDTree, nearest.py:81, has
 * no filter pass).  The exactness certificate of mpb_knn_search budgets the rounding of a filter key --
 * the TF32 split of the operands plus whatever tcgen05.mma kind::tf32 does inside its accumulator -- at
 * 32 * 2^-22 * (R^2 + 2 |q^| R) (R: radius of the centred sample set, q^: centred query).  This call measures it:
 * it prepares the workspace for `samples`, runs the scan's measuring build for `queries` against EVERY tile and
 * compares each key read from tensor memory with the same expression evaluated in fp64 on the same fp32
 * operands.  out (dev, 2 x u64): [0] = bit pattern of the double max |key_tc - key_fp64| / (R^2 + 2 |q^| R),
 * [1] = pairs measured.  The budget holds iff out[0] <= 32 * 2^-22.  Cost: one fp64 key per pair.
 * Needs a set the tensor-core scan serves (mpb_knn_filter_kind == MPB_KNN_KIND_TENSOR at k <= 12). */
int mpb_knn_tc_key_error(const double *samples, int64_t n_samples, const double *queries, int64_t n_queries,
                         int dim, const double *weights, void *workspace, size_t workspace_bytes,
                         unsigned long long *out, void *stream);

/* The spatial order of a prepared sample set: out_order dev [n_samples] i32 = the sample indices in the leaf order
 * of the balanced k-d partition mpb_knn_prepare built for the tensor-core scan (consecutive entries are close in
 * space).  mpb_knn_search_part forms a rank's query blocks from whole leaves of this order, so that each rank's
 * blocks are as compact as a single GPU's (roadmap.py).  MPB_EINVAL when the set is not served by
 * the tensor-core scan (mpb_knn_filter_kind), which builds no order.  No counterpart in the reference (scipy's
 * cKDTree keeps its tree to itself, nearest.py:59-61). */
int mpb_knn_sample_order(const void *workspace, size_t workspace_bytes, int64_t n_samples, int dim,
                         int32_t *out_order, void *stream);

/* mpb_knn_search for one share of the sample set queried against itself -- the multi-GPU form of
 * lazy_prm.py:148-160 add_samples (every sample asks for its neighbours among all samples): after the all-gather
 * every rank holds all samples, prepares the same workspace, and answers part `part` of `parts` of the queries.
 * A part is a share of the set's k-d leaves (whole leaves, dealt in rotation: leaf parts * a + b belongs to part
 * (a + b) mod parts), so every rank's 128-query blocks are exactly the blocks a single-GPU search forms, a share
 * of them from all over the sampled region, and prune as well.  Rows: the part's leaves in ascending order, the
 * samples of a leaf in the order of mpb_knn_sample_order.
 * out_rows dev [rows] i64 (may be NULL): the sample index of every row; out_idx dev [rows, k], out_dist dev
 * [rows, k] (may be NULL); rows = mpb_knn_part_rows(n_samples, dim, part, parts) (host arithmetic; -1 when the
 * tensor-core scan does not serve the set: use mpb_knn_search on an index shard then).  The union of all parts'
 * rows is the result of mpb_knn_search(samples, samples), bit for bit.  k <= 12.  Workspace as for
 * mpb_knn_workspace_bytes(n_samples, n_samples, dim), prepared by mpb_knn_prepare. */
int64_t mpb_knn_part_rows(int64_t n_samples, int dim, int part, int parts);
int mpb_knn_search_part(const double *samples, int64_t n_samples, int dim, const double *weights, int k,
                        double max_dist, void *workspace, size_t workspace_bytes, int part, int parts,
                        int64_t *out_rows, int64_t *out_idx, double *out_dist, int64_t *stats, void *stream);

/* Kernels launched so far by the mpb_knn_* entry points on the calling host thread (a monotone counter;
 * callers report differences).  No counterpart in the reference. */
int mpb_knn_launch_count(void);

/* ---- measurement helpers (not part of the planner path) ---------------------------------
 * Issue-rate micro-benchmarks used by bench.py for the roofline denominators: each thread
 * runs `iters` rounds of 8 independent dependent-FMA chains.  flops = 2 * 8 * iters *
 * blocks * threads.  sink: dev, at least blocks*threads elements. */
int mpb_peak_fma_f32(int blocks, int threads, int iters, float *sink, void *stream);
int mpb_peak_fma_f64(int blocks, int threads, int iters, double *sink, void *stream);
/* The same with packed fp32 pairs (fma.rn.f32x2, SASS FFMA2): flops = 2 * 2 * 8 * iters * blocks * threads. */
int mpb_peak_fma_f32x2(int blocks, int threads, int iters, float *sink, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MPB200_H */
