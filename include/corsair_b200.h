/* corsair_b200 -- C ABI of the B200-native correspondence-and-registration hot path.
 *
 * The reference (ExistentialRobotics/CORSAIR) has no FFI of its own: its boundary for this
 * path is a set of plain Python functions (utils/find_nn.py, utils/eval_pose.py,
 * utils/symmetry.py, utils/preprocess.py, utils/retrieval.py, utils/pc_dist.py,
 * utils/_transform_estimation.py).  Each entry point below names the reference function(s)
 * (file:line under the reference tree) whose arithmetic it replaces; the Python modules in
 * corsair_b200/ keep the reference signatures and call these through ctypes.
 *
 * Conventions
 *   - every data pointer is a DEVICE pointer owned by the caller, unless the parameter name
 *     starts with h_ (host).  The library never frees or retains caller memory.
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*; NULL = legacy
 *     default stream) and never synchronises, unless documented otherwise.
 *   - a cb_ctx owns a device workspace arena; one ctx serves one stream at a time.
 *   - return value: 0 on success, negative cb_status otherwise; text via cb_last_error().
 *   - matrices are row-major; poses are row-major 4x4.
 */
#ifndef CORSAIR_B200_H
#define CORSAIR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cb_ctx cb_ctx;

enum cb_status {
    CB_OK = 0,
    CB_EINVAL = -1,       /* bad argument (shape, null pointer, unsupported k ...) */
    CB_ECUDA = -2,        /* CUDA runtime error (message holds cudaGetErrorString) */
    CB_ENOMEM = -3,       /* workspace allocation failed */
    CB_EUNSUPPORTED = -4, /* valid request the kernels do not cover (e.g. feature width) */
    CB_ENCCL = -5,
    CB_ENOTRUN = -6       /* cb_sym_pose_many: the pair was never started because an earlier one failed */
};

/* metric selectors */
enum cb_knn_metric {
    CB_KNN_SQL2 = 0,   /* squared L2, "SquareL2" of utils/find_nn.py:28-29 */
    CB_KNN_L2 = 1,     /* sqrt(d2 + 1e-7), "L2" of utils/find_nn.py:25-27 */
    CB_KNN_MAXDOT = 2  /* largest inner product, utils/_transform_estimation.py:149-150 */
};
enum cb_retrieval_metric {
    CB_RET_EUCLID = 0, /* scipy cdist default, utils/retrieval.py:175 */
    CB_RET_COSINE = 1  /* scipy pdist "cosine", utils/retrieval.py:24,84,131 */
};

int cb_version(void);
int cb_create(cb_ctx** out, int device, size_t workspace_bytes);
int cb_destroy(cb_ctx* ctx);
const char* cb_last_error(const cb_ctx* ctx);
/* number of kernels this ctx has launched since creation (bench.py's gpu_launches) */
int64_t cb_launch_count(const cb_ctx* ctx);
/* name of the kNN candidate kernel variant the ctx dispatches to: "tcgen05" or "simt" */
const char* cb_knn_variant(const cb_ctx* ctx);
/* force a kNN candidate variant (0 = auto, 1 = simt fp32, 2 = tcgen05 bf16x3); for tests/bench */
int cb_set_knn_variant(cb_ctx* ctx, int variant);
/* RANSAC scoring variant: "tcgen05" (split-bf16 GEMM, default) or "simt" (packed fp32);
 * cb_set_ransac_variant: 0 = auto, 1 = simt, 2 = tcgen05; for tests/bench */
const char* cb_ransac_variant(const cb_ctx* ctx);
/* columns of the scoring GEMM's widened K axis (bf16 flop issued per evaluation = 2 x this) */
int cb_ransac_gemm_k(void);
int cb_set_ransac_variant(cb_ctx* ctx, int variant);
/* test hook: the next cb_ransac calls of this ctx (tcgen05 variant) also store the raw
 * tensor-core values U[h, c] = d^2 - max_dist^2 (before the exact in-band re-decision) into the
 * device array u_out [h_total, ld] (ld >= m); NULL switches the dump off */
int cb_ransac_debug_dump(cb_ctx* ctx, float* u_out, int ld);
/* diagnostic: rows of the most recent E1 kNN problem of this ctx that failed the re-rank's proof
 * obligation and were recomputed by the exact brute-force kernel (synchronises the stream);
 * -1 if no E1 problem has run since the last arena growth */
int cb_knn_last_fallback_rows(cb_ctx* ctx, void* stream);

/* ---- instrumentation for bench.py ---------------------------------------------------------
 * With timing enabled every launch of a tracked kernel class is bracketed by CUDA events on the
 * launching stream.  cb_timing_read synchronises the device and returns, per class, the summed
 * device time (ms), the algorithmic work launched (flop for the kNN candidate kernel; for RANSAC
 * scoring the (hypothesis, correspondence) evaluations of the CTAs that ran, counted on the device)
 * and the launch count. */
enum cb_timing_class {
    CB_T_KNN_CAND = 0,   /* kNN candidate pass (distance GEMM + running top-k) */
    CB_T_KNN_RERANK = 1, /* exact float64 re-rank + fallback */
    CB_T_RANSAC_SCORE = 2,
    CB_T_RANSAC_FIT = 3,
    CB_T_CHAMFER_NN = 4,
    CB_T_ROWS_SELECT = 5, /* exact rows + radix select (retrieval, 50-NN rows) */
    CB_T_KMEANS = 6,
    CB_T_RANSAC_WAVES = 7, /* work counter only: scoring launches that found hypotheses below the budget */
    CB_TIMING_CLASSES = 8
};
/* bytes of host<->device copies the library itself has issued on this context since cb_create (the
 * staging of cb_sym_pose: seed rows and k-means++ draws up, segment offsets and poses back) */
int cb_copy_bytes(const cb_ctx* ctx, int64_t* h2d, int64_t* d2h);
int cb_timing_enable(cb_ctx* ctx, int on);
int cb_timing_read(cb_ctx* ctx, int cls, double* ms, double* work, int64_t* launches);
int cb_timing_reset(cb_ctx* ctx);

/* ---- K1: feature-space kNN -------------------------------------------------------------
 * Replaces find_nn_cpu / find_knn_cpu (utils/find_nn.py:34-49, scipy KDTree in float64, the
 * kNN the shipped pipeline runs at utils/eval_pose.py:68,71) and find_nn_gpu / find_knn_gpu
 * (utils/find_nn.py:52-119, broadcast pdist + min/topk).
 * q [nq,d], r [nr,d] float32.  idx [nq,k] int64, dist [nq,k] float32 (may be NULL), rows
 * sorted ascending by (exact float64 difference-form distance, index).  For CB_KNN_MAXDOT the
 * order is descending inner product and dist holds the products.
 * mutual (may be NULL) [nq] uint8: 1 iff the query is also the nearest reference's nearest
 * query (k-th column 0 only; new, not in the reference).
 * 1 <= k <= 64, k <= nr, 1 <= d <= 256. */
int cb_knn(cb_ctx* ctx, const float* q, int nq, const float* r, int nr, int d, int k, int metric,
           int64_t* idx, float* dist, uint8_t* mutual, void* stream);

/* Ragged batch of independent kNN problems sharing d,k,metric: problem p uses query rows
 * [h_q_off[p], h_q_off[p+1]) of q and reference rows [h_r_off[p], h_r_off[p+1]) of r; idx
 * holds indices LOCAL to the problem's reference block.  Replaces the per-segment loop of
 * split_corr (utils/symmetry.py:145-179).  h_q_off/h_r_off are HOST arrays of n_problems+1. */
int cb_knn_ragged(cb_ctx* ctx, const float* q, const float* r, int d, int k, int metric,
                  const int32_t* h_q_off, const int32_t* h_r_off, int n_problems, int64_t* idx,
                  float* dist, void* stream);

/* Dense pairwise distance tile, pdist (utils/find_nn.py:24-31): out [n,m] float32,
 * dist_type 0 = SquareL2, 1 = L2 (+1e-7 under the root). */
int cb_pdist(cb_ctx* ctx, const float* a, int n, const float* b, int m, int d, int dist_type,
             float* out, void* stream);

/* Dense inner-product matrix F0 F1^T (utils/_transform_estimation.py:149, return_corr=True). */
int cb_dot_matrix(cb_ctx* ctx, const float* a, int n, const float* b, int m, int d, float* out,
                  void* stream);

/* ---- K2: global-descriptor retrieval ----------------------------------------------------
 * Replaces scipy cdist + np.argsort (utils/retrieval.py:139-141,175; evaluation-scan2cad.py
 * :322-323) and cosine pdist + argsort of get_rank / retrieval_eval (utils/retrieval.py:84,
 * 129-136).  q [nq,d], db [ndb,d] float32.  For each query the k best database rows in
 * ascending (float64 distance, index) order: idx [nq,k] int32, dist [nq,k] float64 (may be
 * NULL; Euclidean distance, not squared, or cosine distance).  exclude_self != 0 drops column
 * j == query row + self_offset (get_rank's positional self-exclusion).  1 <= k <= ndb. */
int cb_retrieval_topk(cb_ctx* ctx, const float* q, int nq, const float* db, int ndb, int d, int k,
                      int metric, int exclude_self, int self_offset, int32_t* idx, double* dist,
                      void* stream);

/* ---- K3: correspondence RANSAC ----------------------------------------------------------
 * Replaces registration_based_on_corr (utils/eval_pose.py:82-100), i.e. Open3D's
 * registration_ransac_based_on_correspondence(ransac_n=10, point-to-point, no scaling,
 * max_iteration 100000, confidence 0.999) on an identity correspondence set.
 * src_xyz [n_src,3], tgt_xyz [n_tgt,3] float32; correspondence i pairs
 * src_xyz[src_idx[i]] with tgt_xyz[tgt_idx[i]] (int64 index arrays of length m; either may be
 * NULL = identity).  Hypothesis h fits minimal set h: minimal [h_total, ransac_n] int32 if
 * given, else drawn on device by cb_draw_minimal_sets' counter-based rule from `seed`.
 * adaptive != 0 applies Open3D's early-termination rule in hypothesis order.
 * Outputs (device): T [16] float64, stats [4] float64 = {best_h, inliers, sum_d2,
 * n_evaluated}; counts [h_total] int32 and sums [h_total] float64 (may be NULL) receive the
 * per-hypothesis inlier count / sum of squared inlier distances for h < n_evaluated waves. */
int cb_ransac(cb_ctx* ctx, const float* src_xyz, int n_src, const float* tgt_xyz, int n_tgt,
              const int64_t* src_idx, const int64_t* tgt_idx, int m, const int32_t* minimal,
              uint64_t seed, int h_total, int ransac_n, double max_dist, int adaptive,
              double confidence, double* T, double* stats, int32_t* counts, double* sums,
              void* stream);

/* The library's seeded minimal-set rule, exported so the oracle consumes identical sets:
 * set[h][j] = (hi32(splitmix64(seed*0x9E3779B97F4A7C15 + h*ransac_n + j)) * m) >> 32.
 * out [h_total, ransac_n] int32 (device). */
int cb_draw_minimal_sets(cb_ctx* ctx, uint64_t seed, int m, int h_total, int ransac_n,
                         int32_t* out, void* stream);

/* Batched Umeyama fits only (Open3D TransformationEstimationPointToPoint, a10 in SURVEY §8):
 * T_all [h_total,16] float64. */
int cb_fit_minimal(cb_ctx* ctx, const float* src_xyz, const float* tgt_xyz, int m,
                   const int32_t* minimal, int h_total, int ransac_n, double* T_all, void* stream);

/* ---- K4: Chamfer ------------------------------------------------------------------------
 * One-direction: mean over a of the Euclidean distance to the nearest b, after a <- T a
 * (T [16] float64 row-major 4x4, may be NULL): chamfer_kdtree_1direction(apply_transform(..))
 * (utils/preprocess.py:39-48,67-70) and chamfer_gpu_1direction (:59-64).
 * two_dir != 0 adds the reverse direction: chamfer_gpu (:51-56) == pc_dist.chamfer
 * (utils/pc_dist.py:45-52).  a [na,3], b [nb,3] float32; out [1] float64 (device).
 * nn_idx (may be NULL) [na] int32 receives the nearest-b index of each a. */
int cb_chamfer(cb_ctx* ctx, const float* a, int na, const float* b, int nb, const double* T,
               int two_dir, double* out, int32_t* nn_idx, void* stream);

/* ---- Chamfer without any device allocation ------------------------------------------------------
 * No counterpart in the reference: this entry point exists for the hosts of this library, not for
 * the algorithm.  It computes exactly what cb_chamfer computes (same kernels, same launch geometry,
 * same results) but takes no cb_ctx, because a cb_ctx owns a cudaMalloc'ed arena.  It is meant for
 * callers that may not allocate device memory themselves: plugins inside an application whose own
 * pool has already claimed the card (caching allocators, inference servers with a fixed per-model
 * budget), real-time loops that cannot afford cudaMalloc's device-wide synchronisation, processes
 * whose memory accounting must see every allocation.  Its scratch is a fixed array in the library's
 * device image, CB_CHAMFER_STATIC_SCRATCH_BYTES (about 3 MB) on every GPU that loads the library,
 * held until the process exits.
 *   - limit: na, nb <= CB_CHAMFER_STATIC_MAX_POINTS each; larger clouds are refused with
 *     CB_EUNSUPPORTED (use cb_chamfer).  No cudaMalloc / cudaFree / allocating library call is made.
 *   - device: the CURRENT device of the calling thread (cudaSetDevice), which must own a, b, T, out.
 *   - the first call on each device looks up the array's address there (cudaGetSymbolAddress) and
 *     the SM count; make that first call outside any stream capture.  Later calls only launch.
 *   - every call on a device uses the same array and finds the previous call's contents in it; the
 *     entry point resets what it needs, but two calls that can run on the GPU at the same time
 *     (different streams without an ordering between them, or two host threads) corrupt each other.
 *     Order them on one stream or with events.
 * a, b, T, two_dir, out, nn_idx, stream as for cb_chamfer.  h_err (HOST, may be NULL) receives the
 * error text of a failing call, truncated to h_err_len bytes (there is no ctx for cb_last_error). */
#define CB_CHAMFER_STATIC_MAX_POINTS 65536
#define CB_CHAMFER_STATIC_SCRATCH_BYTES 3153920
int cb_chamfer_static(const float* a, int na, const float* b, int nb, const double* T, int two_dir,
                      double* out, int32_t* nn_idx, char* h_err, size_t h_err_len, void* stream);

/* All-pairs two-direction Chamfer rows of compute_dist (utils/pc_dist.py:61-99):
 * pts [L,P,3] float32; out [(row1-row0), L] float64 rows row0..row1-1 of the full symmetric
 * table; the diagonal is 200 (eye*100 doubled by `table += table.T`, :66,:98). */
int cb_chamfer_matrix(cb_ctx* ctx, const float* pts, int L, int P, int row0, int row1, double* out,
                      void* stream);

/* The same table for row-sharded runs without redundant work: of rows row0..row1-1 only the entries
 * of the circulant half are computed — row i owns the pairs (i, (i + t) mod L), t = 1 .. floor((L-1)/2),
 * plus t = L/2 for i < L/2 when L is even — every unordered pair exactly once over all rows and the
 * same number in every row; all other entries (diagonal included) are written as 0.  The full table
 * is H + H^T with the diagonal set to 200 (corsair_b200/dist.py: chamfer_matrix_sharded).  P <= 4266. */
int cb_chamfer_matrix_half(cb_ctx* ctx, const float* pts, int L, int P, int row0, int row1, double* out,
                           void* stream);

/* Cross table of two-direction Chamfer distances (utils/precompute_chamfer_splats.py:113-149, the
 * table evaluation-scan2cad.py:311-317 reads back): a [la,pa,3], b [lb,pb,3] float32; out [la,lb]
 * float64 (DEVICE), out[i,j] = mean NN distance A_i -> B_j plus mean NN distance B_j -> A_i. */
int cb_chamfer_cross(cb_ctx* ctx, const float* a, int la, int pa, const float* b, int lb, int pb, double* out,
                     void* stream);

/* Hausdorff-style max of NN distances (evaluation-shapenet.py:122-155 chamfer_max), same
 * arguments as cb_chamfer one-direction; out [1] float64. */
int cb_nn_max(cb_ctx* ctx, const float* a, int na, const float* b, int nb, const double* T,
              double* out, void* stream);

/* chamfer_max of a cloud against moved copies of itself — the test behind get_symmetry_label
 * (evaluation-shapenet.py:122-155): for each of n_poses rigid poses T [n_poses,16] float64 (DEVICE,
 * row-major 4x4), out[r] (DEVICE float64) = max over both directions of the nearest-neighbour distance
 * between xyz [n,3] and T_r applied to xyz. */
int cb_self_hausdorff(cb_ctx* ctx, const float* xyz, int n, const double* T, int n_poses, double* out,
                      void* stream);

/* ---- K3': symmetric_cut4 pieces (utils/symmetry.py:182-259) -------------------------------
 * feature rows + 50-NN (:201-214): for each of n_samples rows sample_idx[s] of feat [n,d], the
 * n_nn nearest rows (itself included) as ASCENDING row indices: nn_idx [n_samples,n_nn] int32. */
int cb_feature_rows_topk(cb_ctx* ctx, const float* feat, int n, int d, const int32_t* sample_idx,
                         int n_samples, int n_nn, int32_t* nn_idx, void* stream);

/* Batched tiny k-means (sklearn KMeans(n_clusters=K, random_state=0, n_init=10) on the n_nn
 * gathered xyz, :216) + the acceptance statistics of :218-236.  rng_first [n_init] int32 and
 * rng_unif [n_init,K-1,n_trials] float64 are the data-independent draws of sklearn's
 * k-means++ (generated by the host shim from numpy RandomState).  Outputs per sample:
 * centers [n_samples,K,3] float32, stats [n_samples,4] float64 = {min centre distance, max
 * mean intra-cluster error, std of label ratios over all n points, inertia}. */
int cb_kmeans_batch(cb_ctx* ctx, const float* xyz, int n, const int32_t* nn_idx, int n_samples,
                    int n_nn, int K, const int32_t* rng_first, const double* rng_unif, int n_init,
                    int n_trials, float* centers, double* stats, void* stream);

/* KMeans.predict over all n points (:224,:240): labels [n] int32 nearest of K centres
 * (centers [K,3] float32), counts [K] int32 histogram (may be NULL). */
int cb_assign_labels(cb_ctx* ctx, const float* xyz, int n, const float* centers, int K,
                     int32_t* labels, int32_t* counts, void* stream);

/* Model selection (utils/symmetry.py:233-235), KMeans.predict of the winner over all n points
 * (:246), the K = 4 centre ranking (:239-244) and the resulting segment permutation, without a host round trip.
 * centers [n_samples,K,3] f32 and stats [n_samples,4] f64 as written by cb_kmeans_batch.
 * perm [n] int64: ascending point indices of segment 0, then segment 1, ... in the order of the masks
 * symmetric_cut4 returns (:248-256); labels [n] int32 (may be NULL); seg_off [K+2] int32 (DEVICE):
 * K+1 segment offsets, then the accepted sample (-1: none — the reference raises there, :237).
 * K must be 2 or 4 (:257-258). */
int cb_symmetry_segments(cb_ctx* ctx, const float* xyz, int n, const float* centers, const double* stats,
                         int n_samples, int K, double min_center, int64_t* perm, int32_t* labels,
                         int32_t* seg_off, void* stream);

/* ---- sym_pose body in one call (utils/symmetry.py:262-358) -----------------------------------
 * The vanilla registration (:267-283: feature kNN over the whole clouds -> correspondence RANSAC
 * with ransac_n = 10 and the seeded counter-based draws -> one-direction Chamfer of the
 * float32-rounded pose) followed by n_pairings symmetry pairings (:303-356, split_corr per
 * pairing: per-segment kNN, concatenated correspondences -> RANSAC -> Chamfer), all enqueued on
 * `stream` without host synchronisation.
 * base_f [n0,d], base_xyz [n0,3], pos_f [n1,d], pos_xyz [n1,3] float32 (device).
 * a_perm [n0] / b_perm [n1] int64 (device): point indices sorted by segment (ascending index
 * inside a segment); h_q_off / h_r_off [n_seg+1] int32 (HOST): segment offsets into them.
 * h_pair_map [n_pairings, n_seg] int32 (HOST): pairing p matches base segment s with pos segment
 * h_pair_map[p][s].  n_pairings = 0 runs the vanilla registration only (perms may be NULL).
 * Outputs (device): T_out [(1+n_pairings),16] float64 holding float32-rounded poses (vanilla
 * first), chamfer_out [(1+n_pairings)] float64, stats_out (may be NULL) [(1+n_pairings),4] as
 * cb_ransac's stats. */
int cb_register_pairings(cb_ctx* ctx, const float* base_f, const float* base_xyz, int n0,
                         const float* pos_f, const float* pos_xyz, int n1, int d, int k_nn,
                         const int64_t* a_perm, const int32_t* h_q_off, const int64_t* b_perm,
                         const int32_t* h_r_off, int n_seg, const int32_t* h_pair_map, int n_pairings,
                         double max_dist, uint64_t seed, int h_total, int adaptive, double confidence,
                         double* T_out, double* chamfer_out, double* stats_out, void* stream);

/* HOST ONLY, no device work: replays `rs = np.random.RandomState(seed)` followed by `count` successive
 * `rs.choice(n[c], size, replace=False)` (legacy MT19937 seeding, Fisher-Yates permutation with the
 * masked-rejection bounded draw) into out [count, size] int32.  CB_EINVAL when some n[c] < size
 * (NumPy raises there). */
int cb_np_seeded_choices(uint32_t seed, const int32_t* n, int count, int size, int32_t* out);

/* Voxel de-duplication (quantize_pc, evaluation-shapenet.py:97-107; utils/Info/CADLib.py:106-121): for
 * xyz [n,3] float32 the indices of the points that are the FIRST (lowest index) in their voxel
 * floor(p / voxel), ascending: idx_out [<= n] int64, their integer voxel coordinates grid_out
 * [<= n, 3] int32 (may be NULL), and *count (DEVICE int32) = how many (-1: a coordinate beyond
 * +-2^20 voxels or not finite).  float32 division and floor, like NumPy on float32 input. */
int cb_voxel_unique(cb_ctx* ctx, const float* xyz, int n, float voxel, int64_t* idx_out, int32_t* grid_out,
                    int32_t* count, void* stream);

/* ---- sym_pose end to end (utils/symmetry.py:262-358) ------------------------------------------
 * One call per query-CAD pair: the vanilla registration, the symmetry segmentation of both clouds
 * (symmetric_cut4 with K = 4 when pos_sym >= 2, else 2: :289-298) and, when both accept a model,
 * every pairing of the sweeps (:303-356).  The 1 + P registrations are independent and run on a
 * few internal streams (CB_SYM_STREAMS, default 3; 0 = all on `stream`); the call returns after its
 * results are on the host.  Device inputs as for cb_register_pairings.
 * HOST inputs: h_samples_base / h_samples_pos [n_samples] int32 — the seed rows the caller drew with
 * choice(n, 100, replace=False) (:193; n_samples = 0 skips the segmentation) — or both NULL with
 * np_seed in [0, 2^32): the call draws them itself as np.random.seed(np_seed) followed by the two
 * choices would (cb_np_seeded_choices); h_km_first [n_init]
 * int32 and h_km_unif [n_init, K-1, n_trials] float64 — the data-independent RandomState(0) draws of
 * sklearn's k-means++ (cb_kmeans_batch).
 * h_out [40] float64 (HOST): [0,16) best pose (float32 values), [16] its Chamfer distance, [17,33)
 * the vanilla pose, [33] its Chamfer distance, [34] status (0 symmetry search ran; 1 / 2 the base /
 * pos segmentation accepted no model — the reference raises and falls back, :299-301; 3 not
 * attempted), [35] index of the winning registration (0 vanilla), [36] pairings evaluated. */
int cb_sym_pose(cb_ctx* ctx, const float* base_f, const float* base_xyz, int n0, const float* pos_f,
                const float* pos_xyz, int n1, int d, int pos_sym, int k_nn, const int32_t* h_samples_base,
                const int32_t* h_samples_pos, int n_samples, int64_t np_seed, int n_nn, const int32_t* h_km_first,
                const double* h_km_unif, int n_init, int n_trials, double min_center, double max_dist,
                uint64_t seed, int h_total, int adaptive, double confidence, double* h_out, void* stream);

/* ---- the registration loop of evaluation.py:297-313 in one call ----------------------------------
 * One record per query-CAD pair: the arguments of cb_sym_pose (device pointers first, host pointers
 * where that call takes host pointers; every pair brings its own k-means++ draws because they depend
 * on n_nn = min(50, n0, n1) and K). */
typedef struct cb_sym_job {
    const float* base_f;    /* device [n0, d] */
    const float* base_xyz;  /* device [n0, 3] */
    const float* pos_f;     /* device [n1, d] */
    const float* pos_xyz;   /* device [n1, 3] */
    int32_t n0, n1, d, pos_sym, k_nn, n_samples, n_nn, n_init, n_trials, h_total, adaptive, reserved;
    int64_t np_seed;
    const int32_t* h_samples_base; /* host, or NULL with np_seed >= 0 */
    const int32_t* h_samples_pos;
    const int32_t* h_km_first;     /* host [n_init] */
    const double* h_km_unif;       /* host [n_init, K-1, n_trials] */
    double min_center, max_dist, confidence;
    uint64_t seed;
    double* h_out;                 /* host [40], as cb_sym_pose */
} cb_sym_job;
/* Registers n_jobs independent pairs with up to `concurrency` of them in flight (each on its own
 * internal stream set), driven by the calling thread alone: it enqueues a pair's next stage when the
 * event of the previous one has completed and never sleeps in the driver while another pair has work
 * to enqueue.  The inputs must be complete on `stream`; on return every h_out is on the host.
 * Results are those of n_jobs cb_sym_pose calls.  h_status [n_jobs] (may be NULL) receives each
 * pair's status; the first failure stops new pairs from starting and is returned: the pairs that
 * never started get h_status CB_ENOTRUN and an all-zero h_out with h_out[34] = -1.  At most 32 pairs
 * are in flight whatever `concurrency` says.  concurrency <= 1 (or timing enabled) runs the pairs one
 * after the other on `stream`. */
int cb_sym_pose_many(cb_ctx* ctx, const cb_sym_job* jobs, int n_jobs, int concurrency, int32_t* h_status,
                     void* stream);

/* ---- K5: IRLS pose (utils/_transform_estimation.py:110-137) ------------------------------
 * est_quad_linear_robust: iters (reference: 20) iterations of the linearised weighted solve.
 * p0,p1 [n,3] float32, w [n] float32 (NULL = ones); T [16] float32 (device). */
int cb_irls_pose(cb_ctx* ctx, const float* p0, const float* p1, const float* w, int n, int iters,
                 float* T, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CORSAIR_B200_H */
