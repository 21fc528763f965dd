// K3': the device pieces of symmetric_cut4 (utils/symmetry.py:182-259 of the reference).
//   cb_kmeans_batch   one block per sampled seed point: sklearn-style KMeans(n_clusters=K,
//                     random_state=0, n_init=10) on the n_nn (=50) gathered xyz -- one warp per
//                     initialisation (k-means++ with the data-independent RandomState(0) draws
//                     supplied by the host shim, then Lloyd with sklearn's stopping rules) --
//                     followed by the acceptance statistics of :218-236 over all n points.
//   cb_assign_labels  KMeans.predict over all points for the winning model (:224,:240).
// The arithmetic follows sklearn 1.x (float32 data, distances for the seeding computed in float64
// and rounded, sequential float32 centre sums).  BLAS summation orders inside sklearn are not
// reproducible, so agreement is exact except where a decision sits on a float32 rounding tie;
// tests report the mask-equality rate against the reference.
#include "common.cuh"
#include "knn_internal.cuh"

#include <math.h>

constexpr int kKmMaxPts = 64;   // n_nn <= 64: two points per lane
constexpr int kKmMaxK = 8;
constexpr int kKmMaxInit = 16;
constexpr int kKmMaxIter = 300;  // sklearn default max_iter

struct KmWarp {
    float closest[kKmMaxPts];
    float cs[kKmMaxPts];
    float cand[4][kKmMaxPts];
    int label[kKmMaxPts];
    int label_old[kKmMaxPts];
    float c[kKmMaxK][3];
    float cn[kKmMaxK][3];
    float inertia;
};

__device__ __forceinline__ float km_d2_upcast(const float* a, const float* b) {
    // sklearn _euclidean_distances_upcast: float64 arithmetic, clipped at 0, rounded to float32
    double dx = (double)a[0] - (double)b[0], dy = (double)a[1] - (double)b[1], dz = (double)a[2] - (double)b[2];
    double d = dx * dx + dy * dy + dz * dz;
    return (float)fmax(d, 0.0);
}

__device__ __forceinline__ float km_warp_sum(float v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// nearest centre in the GEMM form sklearn's Lloyd / predict use: argmin_c |c|^2 - 2 x.c (first min)
__device__ __forceinline__ int km_nearest(const float* x, const float (*c)[3], int K) {
    int best = 0;
    float bv = INFINITY;
    for (int k = 0; k < K; ++k) {
        float cc = fmaf(c[k][2], c[k][2], fmaf(c[k][1], c[k][1], c[k][0] * c[k][0]));
        float dot = fmaf(x[2], c[k][2], fmaf(x[1], c[k][1], x[0] * c[k][0]));
        float v = fmaf(-2.f, dot, cc);
        if (v < bv) { bv = v; best = k; }
    }
    return best;
}

__global__ void __launch_bounds__(32 * kKmMaxInit)
kmeans_batch_kernel(const float* __restrict__ xyz, int n, const int32_t* __restrict__ nn_idx, int n_nn, int K,
                    const int32_t* __restrict__ rng_first, const double* __restrict__ rng_unif, int n_init,
                    int n_trials, float* __restrict__ centers_out, double* __restrict__ stats_out) {
    extern __shared__ __align__(16) unsigned char km_smem[];
    KmWarp* W = reinterpret_cast<KmWarp*>(km_smem);
    __shared__ float X[kKmMaxPts][3];   // centred neighbourhood
    __shared__ float Xraw[kKmMaxPts][3];
    __shared__ float mean[3];
    __shared__ float tol_s;
    __shared__ int best_init;
    __shared__ int hist[kKmMaxK];
    __shared__ float fin_c[kKmMaxK][3];
    __shared__ float err_sum[kKmMaxK];
    __shared__ int err_cnt[kKmMaxK];
    const int s = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid < n_nn) {
        int j = nn_idx[(size_t)s * n_nn + tid];
        Xraw[tid][0] = xyz[3 * (size_t)j];
        Xraw[tid][1] = xyz[3 * (size_t)j + 1];
        Xraw[tid][2] = xyz[3 * (size_t)j + 2];
    }
    __syncthreads();
    if (tid < 3) {
        double m = 0.0;
        for (int i = 0; i < n_nn; ++i) m += (double)Xraw[i][tid];
        mean[tid] = (float)(m / n_nn);
    }
    __syncthreads();
    if (tid < n_nn)
        for (int a = 0; a < 3; ++a) X[tid][a] = Xraw[tid][a] - mean[a];
    __syncthreads();
    if (tid == 0) {
        // tol = mean(var(X, axis=0)) * 1e-4  (sklearn _tolerance, tol=1e-4)
        double v = 0.0;
        for (int a = 0; a < 3; ++a) {
            double mu = 0.0;
            for (int i = 0; i < n_nn; ++i) mu += (double)X[i][a];
            mu /= n_nn;
            double q = 0.0;
            for (int i = 0; i < n_nn; ++i) { double d = (double)X[i][a] - mu; q += d * d; }
            v += q / n_nn;
        }
        tol_s = (float)(v / 3.0 * 1e-4);
    }
    __syncthreads();

    if (warp < n_init) {
        KmWarp& w = W[warp];
        // ---- k-means++ seeding
        const int first = rng_first[warp];
        if (lane < 3) w.c[0][lane] = X[first][lane];
        __syncwarp();
        float pot_part = 0.f;
        for (int i = lane; i < n_nn; i += 32) {
            float d = km_d2_upcast(X[i], w.c[0]);
            w.closest[i] = d;
            pot_part += d;
        }
        float pot = km_warp_sum(pot_part);
        __syncwarp();
        for (int c = 1; c < K; ++c) {
            if (lane == 0) {
                float run = 0.f;
                for (int i = 0; i < n_nn; ++i) { run += w.closest[i]; w.cs[i] = run; }
            }
            __syncwarp();
            float best_pot = INFINITY;
            int best_t = 0, best_id = 0;
            for (int t = 0; t < n_trials; ++t) {
                const double rv = rng_unif[((size_t)warp * (K - 1) + (c - 1)) * n_trials + t] * (double)pot;
                int id = 0;  // searchsorted(cs, rv, side='left')
                for (int i = 0; i < n_nn; ++i) id += ((double)w.cs[i] < rv) ? 1 : 0;
                id = min(id, n_nn - 1);
                float part = 0.f;
                for (int i = lane; i < n_nn; i += 32) {
                    float d = fminf(w.closest[i], km_d2_upcast(X[i], X[id]));
                    w.cand[t][i] = d;
                    part += d;
                }
                float p = km_warp_sum(part);
                if (p < best_pot) { best_pot = p; best_t = t; best_id = id; }
            }
            __syncwarp();
            for (int i = lane; i < n_nn; i += 32) w.closest[i] = w.cand[best_t][i];
            if (lane < 3) w.c[c][lane] = X[best_id][lane];
            pot = best_pot;
            __syncwarp();
        }
        // ---- Lloyd
        for (int i = lane; i < n_nn; i += 32) { w.label[i] = -1; w.label_old[i] = -1; }
        __syncwarp();
        bool strict = false;
        for (int it = 0; it < kKmMaxIter; ++it) {
            for (int i = lane; i < n_nn; i += 32) w.label[i] = km_nearest(X[i], w.c, K);
            __syncwarp();
            if (lane < K) {
                float sx = 0.f, sy = 0.f, sz = 0.f;
                int cnt = 0;
                for (int i = 0; i < n_nn; ++i)
                    if (w.label[i] == lane) { sx += X[i][0]; sy += X[i][1]; sz += X[i][2]; ++cnt; }
                if (cnt > 0) {
                    w.cn[lane][0] = sx / (float)cnt; w.cn[lane][1] = sy / (float)cnt; w.cn[lane][2] = sz / (float)cnt;
                } else {  // sklearn relocates empty clusters; with k-means++ on 50 points this is rare
                    w.cn[lane][0] = w.c[lane][0]; w.cn[lane][1] = w.c[lane][1]; w.cn[lane][2] = w.c[lane][2];
                }
            }
            __syncwarp();
            float sh = 0.f;
            if (lane < K) {
                float dx = w.cn[lane][0] - w.c[lane][0], dy = w.cn[lane][1] - w.c[lane][1],
                      dz = w.cn[lane][2] - w.c[lane][2];
                sh = dx * dx + dy * dy + dz * dz;
                w.c[lane][0] = w.cn[lane][0]; w.c[lane][1] = w.cn[lane][1]; w.c[lane][2] = w.cn[lane][2];
            }
            sh = km_warp_sum(sh);
            bool same = true;
            for (int i = lane; i < n_nn; i += 32) same = same && (w.label[i] == w.label_old[i]);
            same = __all_sync(0xffffffffu, same);
            __syncwarp();
            if (same) { strict = true; break; }
            if (sh <= tol_s) break;
            for (int i = lane; i < n_nn; i += 32) w.label_old[i] = w.label[i];
            __syncwarp();
        }
        if (!strict) {
            for (int i = lane; i < n_nn; i += 32) w.label[i] = km_nearest(X[i], w.c, K);
            __syncwarp();
        }
        float part = 0.f;
        for (int i = lane; i < n_nn; i += 32) {
            const float* cc = w.c[w.label[i]];
            float dx = X[i][0] - cc[0], dy = X[i][1] - cc[1], dz = X[i][2] - cc[2];
            part += dx * dx + dy * dy + dz * dz;
        }
        part = km_warp_sum(part);
        if (lane == 0) w.inertia = part;
    }
    __syncthreads();
    if (tid == 0) {
        // best of n_init: lower inertia AND a different clustering (sklearn KMeans.fit)
        int b = 0;
        for (int i = 1; i < n_init; ++i) {
            if (!(W[i].inertia < W[b].inertia)) continue;
            int map[kKmMaxK], inv[kKmMaxK];
            for (int k = 0; k < K; ++k) { map[k] = -1; inv[k] = -1; }
            bool same = true;
            for (int p = 0; p < n_nn && same; ++p) {
                int l1 = W[i].label[p], l2 = W[b].label[p];
                if (map[l1] == -1 && inv[l2] == -1) { map[l1] = l2; inv[l2] = l1; }
                else if (map[l1] != l2) same = false;
            }
            if (!same) b = i;
        }
        best_init = b;
    }
    if (tid < kKmMaxK) { hist[tid] = 0; err_sum[tid] = 0.f; err_cnt[tid] = 0; }
    __syncthreads();
    if (tid < K * 3) {
        int k = tid / 3, a = tid % 3;
        float v = W[best_init].c[k][a] + mean[a];
        fin_c[k][a] = v;
        centers_out[((size_t)s * K + k) * 3 + a] = v;
    }
    __syncthreads();
    // predict over all n points -> label histogram
    int local[kKmMaxK];
    for (int k = 0; k < kKmMaxK; ++k) local[k] = 0;
    for (int i = tid; i < n; i += blockDim.x) {
        float p[3] = {xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2]};
        int l = km_nearest(p, fin_c, K);
        for (int k = 0; k < kKmMaxK; ++k) local[k] += (l == k);
    }
    for (int k = 0; k < K; ++k) {
        int v = local[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && v) atomicAdd(&hist[k], v);
    }
    // mean intra-cluster error over the neighbourhood with predict labels
    if (tid < n_nn) {
        int l = km_nearest(Xraw[tid], fin_c, K);
        float dx = Xraw[tid][0] - fin_c[l][0], dy = Xraw[tid][1] - fin_c[l][1], dz = Xraw[tid][2] - fin_c[l][2];
        atomicAdd(&err_sum[l], sqrtf(dx * dx + dy * dy + dz * dz));
        atomicAdd(&err_cnt[l], 1);
    }
    __syncthreads();
    if (tid == 0) {
        double dmin = 1e30;
        for (int a = 0; a < K; ++a)
            for (int b = a + 1; b < K; ++b) {
                float dx = fin_c[a][0] - fin_c[b][0], dy = fin_c[a][1] - fin_c[b][1], dz = fin_c[a][2] - fin_c[b][2];
                dmin = fmin(dmin, (double)sqrtf(dx * dx + dy * dy + dz * dz));
            }
        double emax = 0.0;
        for (int k = 0; k < K; ++k) {
            double e = err_cnt[k] > 0 ? (double)err_sum[k] / err_cnt[k] : (double)NAN;
            if (!(e <= emax)) emax = e;  // NaN (empty cluster) poisons the test like the reference
        }
        double mu = 0.0, var = 0.0;
        for (int k = 0; k < K; ++k) mu += (double)hist[k] / n;
        mu /= K;
        for (int k = 0; k < K; ++k) { double d = (double)hist[k] / n - mu; var += d * d; }
        stats_out[4 * (size_t)s + 0] = dmin;
        stats_out[4 * (size_t)s + 1] = emax;
        stats_out[4 * (size_t)s + 2] = sqrt(var / K);
        stats_out[4 * (size_t)s + 3] = (double)W[best_init].inertia;
    }
}

__global__ void assign_labels_kernel(const float* __restrict__ xyz, int n, const float* __restrict__ centers, int K,
                                     int32_t* __restrict__ labels, int32_t* __restrict__ counts) {
    __shared__ float c[kKmMaxK][3];
    __shared__ int hist[kKmMaxK];
    if (threadIdx.x < K * 3) c[threadIdx.x / 3][threadIdx.x % 3] = centers[threadIdx.x];
    if (threadIdx.x < kKmMaxK) hist[threadIdx.x] = 0;
    __syncthreads();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        float p[3] = {xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2]};
        int l = km_nearest(p, c, K);
        labels[i] = l;
        if (counts) atomicAdd(&hist[l], 1);
    }
    __syncthreads();
    if (counts && threadIdx.x < K && hist[threadIdx.x]) atomicAdd(&counts[threadIdx.x], hist[threadIdx.x]);
}

// Model selection + labels + segment permutation in one block (no host round trip):
//   thread 0 replays the sequential acceptance rule of utils/symmetry.py:233-235 over the S rows of
//   statistics and, for K = 4, the centre ranking of :239-244 (float32 norms like NumPy's);
//   every thread then labels its contiguous slice of points (KMeans.predict, :246) and the block
//   writes, per segment in the reference's order, the ascending indices of its points.
// out_off [K+2]: segment offsets [K+1], then the winning sample (-1: no model accepted).
__global__ void __launch_bounds__(1024)
segments_kernel(const float* __restrict__ xyz, int n, const float* __restrict__ centers,
                const double* __restrict__ stats, int S, int K, double min_center,
                int64_t* __restrict__ perm, int32_t* __restrict__ labels_out, int32_t* __restrict__ out_off) {
    __shared__ float c[kKmMaxK][3];
    __shared__ int seg_of_label[kKmMaxK];
    __shared__ int s_win;
    __shared__ int cnt[4][1024];
    __shared__ int seg_base[5];
    const int tid = threadIdx.x;
    if (tid == 0) {
        double curr_var = 100.0;
        int win = -1;
        for (int i = 0; i < S; ++i) {
            const double dmin = stats[4 * i], emax = stats[4 * i + 1], sd = stats[4 * i + 2];
            if (dmin > min_center && min_center > emax && curr_var > sd) { curr_var = sd; win = i; }
        }
        s_win = win;
        if (win >= 0) {
            for (int k = 0; k < K; ++k)
                for (int a = 0; a < 3; ++a) c[k][a] = centers[((size_t)win * K + k) * 3 + a];
            int order[4] = {0, 1, 2, 3};
            if (K == 4) {
                float d[3];
                for (int j = 1; j < 4; ++j) {
                    const float dx = c[j][0] - c[0][0], dy = c[j][1] - c[0][1], dz = c[j][2] - c[0][2];
                    d[j - 1] = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
                }
                int r[3] = {0, 1, 2};  // stable argsort of three values
                for (int a = 1; a < 3; ++a)
                    for (int b = a; b > 0 && d[r[b]] < d[r[b - 1]]; --b) { int t = r[b]; r[b] = r[b - 1]; r[b - 1] = t; }
                order[1] = r[0] + 1; order[2] = r[2] + 1; order[3] = r[1] + 1;
            }
            for (int s = 0; s < K; ++s) seg_of_label[order[s]] = s;
        }
    }
    __syncthreads();
    if (s_win < 0) {
        if (tid <= K) out_off[tid] = 0;
        if (tid == 0) out_off[K + 1] = -1;
        return;
    }
    const int per = (n + 1023) / 1024;
    const int i0 = min(n, tid * per), i1 = min(n, i0 + per);
    int my[4] = {0, 0, 0, 0};
    for (int i = i0; i < i1; ++i) {
        const float p[3] = {xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2]};
        const int l = km_nearest(p, c, K);
        if (labels_out) labels_out[i] = l;
        const int s = seg_of_label[l];
#pragma unroll
        for (int q = 0; q < 4; ++q) my[q] += (s == q);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) cnt[q][tid] = my[q];
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    if (warp < 4) {  // exclusive scan of segment `warp` over the 1024 slices
        int v[32], tot = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j) { v[j] = cnt[warp][lane * 32 + j]; tot += v[j]; }
        int inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
        int run = inc - tot;
#pragma unroll
        for (int j = 0; j < 32; ++j) { cnt[warp][lane * 32 + j] = run; run += v[j]; }
        if (lane == 31) seg_base[warp + 1] = inc;
    }
    __syncthreads();
    if (tid == 0) {
        seg_base[0] = 0;
        for (int s = 1; s <= 4; ++s) seg_base[s] += seg_base[s - 1];
        for (int s = 0; s <= K; ++s) out_off[s] = seg_base[s];
        out_off[K + 1] = s_win;
    }
    __syncthreads();
    int pos[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) pos[q] = seg_base[q] + cnt[q][tid];
    for (int i = i0; i < i1; ++i) {
        const float p[3] = {xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2]};
        const int s = seg_of_label[km_nearest(p, c, K)];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (s == q) perm[pos[q]++] = i;
    }
}

extern "C" int cb_symmetry_segments(cb_ctx* ctx, const float* xyz, int n, const float* centers, const double* stats,
                                    int n_samples, int K, double min_center, int64_t* perm, int32_t* labels,
                                    int32_t* seg_off, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (!xyz || !centers || !stats || !perm || !seg_off || n <= 0 || n_samples < 0)
        CB_FAIL(ctx, CB_EINVAL, "cb_symmetry_segments: bad arguments");
    if (K != 2 && K != 4) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_symmetry_segments: K must be 2 or 4 (got %d)", K);
    segments_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(xyz, n, centers, stats, n_samples, K, min_center, perm, labels,
                                                          seg_off);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

static int kmeans_launch(cb_ctx* ctx, const float* xyz, int n, const int32_t* nn_idx, int n_samples, int n_nn, int K,
                         const int32_t* rng_first, const double* rng_unif, int n_init, int n_trials, float* centers,
                         double* stats, cudaStream_t st) {
    if (n_nn > kKmMaxPts || K > kKmMaxK || K > n_nn || n_init < 1 || n_init > kKmMaxInit || n_trials < 1 ||
        n_trials > 4)
        CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_kmeans_batch: n_nn<=64, K<=8, n_init<=16, n_trials<=4 (got %d %d %d %d)",
                n_nn, K, n_init, n_trials);
    size_t smem = sizeof(KmWarp) * n_init;
    cudaError_t e = cudaFuncSetAttribute(kmeans_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "smem opt-in failed: %s", cudaGetErrorString(e));
    cb_timed_scope ts(ctx, CB_T_KMEANS, 0.0, st);
    kmeans_batch_kernel<<<n_samples, 32 * n_init, smem, st>>>(xyz, n, nn_idx, n_nn, K, rng_first, rng_unif, n_init,
                                                             n_trials, centers, stats);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

extern "C" int cb_kmeans_batch(cb_ctx* ctx, const float* xyz, int n, const int32_t* nn_idx, int n_samples,
                               int n_nn, int K, const int32_t* rng_first, const double* rng_unif, int n_init,
                               int n_trials, float* centers, double* stats, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (n_samples == 0) return CB_OK;
    if (!xyz || !nn_idx || !rng_first || !rng_unif || !centers || !stats || n <= 0 || n_nn < 1 || K < 1)
        CB_FAIL(ctx, CB_EINVAL, "cb_kmeans_batch: bad arguments");
    return kmeans_launch(ctx, xyz, n, nn_idx, n_samples, n_nn, K, rng_first, rng_unif, n_init, n_trials, centers, stats,
                         (cudaStream_t)stream);
}

// symmetric_cut4 end to end on one stream (internal; allocates from the arena without rewinding it):
// feature-space neighbourhoods of the sampled rows -> batched k-means -> selection + permutation
size_t cb_segment_workspace(int n, int n_samples, int n_nn, int K) {
    return cb_rows_topk_workspace(n_samples, n, n_nn) + cb_align((size_t)n_samples * n_nn * 4) +
           cb_align((size_t)n_samples * K * 12) + cb_align((size_t)n_samples * 32) + 4096;
}

int cb_segment_impl(cb_ctx* ctx, const float* feat, const float* xyz, int n, int d, int K, const int32_t* d_samples,
                    int n_samples, int n_nn, const int32_t* d_first, const double* d_unif, int n_init, int n_trials,
                    double min_center, int64_t* perm, int32_t* seg_off, cudaStream_t st) {
    if (K != 2 && K != 4) CB_FAIL(ctx, CB_EUNSUPPORTED, "symmetry segmentation: K must be 2 or 4 (got %d)", K);
    if (n_nn < 1 || n_nn > n || n_samples < 1) CB_FAIL(ctx, CB_EINVAL, "symmetry segmentation: bad sample sizes");
    int32_t* nn_idx = cb_alloc_t<int32_t>(ctx, (size_t)n_samples * n_nn);
    float* centers = cb_alloc_t<float>(ctx, (size_t)n_samples * K * 3);
    double* stats = cb_alloc_t<double>(ctx, (size_t)n_samples * 4);
    if (!nn_idx || !centers || !stats) return CB_ENOMEM;
    int rc = cb_rows_topk(ctx, feat, n_samples, feat, n, d, n_nn, 0, 0, 0, 1, nullptr, nn_idx, nullptr, nullptr, 0,
                          d_samples, st);
    if (rc) return rc;
    if ((rc = kmeans_launch(ctx, xyz, n, nn_idx, n_samples, n_nn, K, d_first, d_unif, n_init, n_trials, centers, stats,
                            st)))
        return rc;
    segments_kernel<<<1, 1024, 0, st>>>(xyz, n, centers, stats, n_samples, K, min_center, perm, nullptr, seg_off);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

extern "C" int cb_assign_labels(cb_ctx* ctx, const float* xyz, int n, const float* centers, int K,
                                int32_t* labels, int32_t* counts, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (n == 0) return CB_OK;
    if (!xyz || !centers || !labels || n < 0 || K < 1 || K > kKmMaxK)
        CB_FAIL(ctx, CB_EINVAL, "cb_assign_labels: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    if (counts) CB_CUDA(ctx, cudaMemsetAsync(counts, 0, K * sizeof(int32_t), st));
    assign_labels_kernel<<<cb_div_up(n, 256), 256, 0, st>>>(xyz, n, centers, K, labels, counts);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}
