// Engine E2: exact float64 distance rows + per-row radix select / bitonic sort.
// Serves K2 retrieval (utils/retrieval.py:129-177 of the reference: cdist/pdist + argsort), the
// 50-NN feature rows of symmetric_cut4 (utils/symmetry.py:201-214) and kNN shapes outside E1.
#include "common.cuh"
#include "knn_internal.cuh"

#include <math.h>

// ---------------------------------------------------------------------------- rows
// block (32,8): 32 references x 32 queries per tile, 4 queries per thread.
__global__ void __launch_bounds__(256)
rows_kernel(const float* __restrict__ q, const int32_t* __restrict__ q_gather, int q0, int nqc,
            const float* __restrict__ r, int nr, int d, int metric, int exclude_self, int self_offset,
            double* __restrict__ out) {
    __shared__ float sq[32][33];
    __shared__ float sr[32][33];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int j = blockIdx.x * 32 + tx;
    const int qb = blockIdx.y * 32;
    double acc[4] = {0, 0, 0, 0}, nq2[4] = {0, 0, 0, 0}, nr2 = 0;
    for (int c0 = 0; c0 < d; c0 += 32) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int row = ty + 8 * i;
            int ql = qb + row;
            float v = 0.f, w = 0.f;
            if (c0 + tx < d) {
                if (ql < nqc) {
                    int qrow = q_gather ? q_gather[q0 + ql] : q0 + ql;
                    v = q[(size_t)qrow * d + c0 + tx];
                }
                int rj = blockIdx.x * 32 + row;
                if (rj < nr) w = r[(size_t)rj * d + c0 + tx];
            }
            sq[row][tx] = v;
            sr[row][tx] = w;
        }
        __syncthreads();
        const int cn = min(32, d - c0);
        for (int c = 0; c < cn; ++c) {
            const double rv = (double)sr[tx][c];
            if (metric == 3) nr2 = fma(rv, rv, nr2);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const double qv = (double)sq[ty + 8 * i][c];
                if (metric <= 1) {
                    double df = __dsub_rn(qv, rv);
                    acc[i] = __dadd_rn(acc[i], __dmul_rn(df, df));
                } else if (metric == 2) {
                    acc[i] = __dadd_rn(acc[i], __dmul_rn(qv, rv));
                } else {
                    acc[i] = fma(qv, rv, acc[i]);
                    nq2[i] = fma(qv, qv, nq2[i]);
                }
            }
        }
    }
    if (j >= nr) return;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int ql = qb + ty + 8 * i;
        if (ql >= nqc) continue;
        double v = acc[i];
        if (metric == 1) v = sqrt(v);
        else if (metric == 2) v = -v;
        else if (metric == 3) v = 1.0 - v / (sqrt(nq2[i]) * sqrt(nr2));
        if (exclude_self && j == q0 + ql + self_offset) v = exclude_self == 2 ? 0.0 : INFINITY;
        out[(size_t)ql * nr + j] = v;
    }
}

// -------------------------------------------------------------------------- select
__device__ __forceinline__ uint64_t ord_u64(double v) {
    uint64_t u = (uint64_t)__double_as_longlong(v);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double ord_f64(uint64_t u) {
    u = (u >> 63) ? (u & 0x7fffffffffffffffull) : ~u;
    return __longlong_as_double((long long)u);
}

// one block (256 threads) per row.  dynamic smem: kcap * (8 + 4) bytes.
__global__ void __launch_bounds__(256)
select_kernel(const double* __restrict__ rows, int nr, int k, int kcap, int q0, int sort_by_index,
              int64_t* __restrict__ idx64, int32_t* __restrict__ idx32, double* __restrict__ dist64,
              float* __restrict__ dist32, int dist32_mode) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* skey = reinterpret_cast<uint64_t*>(smem_raw);
    int32_t* sidx = reinterpret_cast<int32_t*>(skey + kcap);
    __shared__ unsigned int hist[256];
    __shared__ uint64_t s_prefix;
    __shared__ int s_need, s_base, s_wsum[8];
    const int tid = threadIdx.x;
    const double* row = rows + (size_t)blockIdx.x * nr;

    // radix select of the k-th smallest key, most significant byte first
    if (tid == 0) { s_prefix = 0; s_need = k; }
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = 56 - 8 * pass;
        hist[tid] = 0;
        __syncthreads();
        const uint64_t prefix = s_prefix;
        const uint64_t mask = pass == 0 ? 0ull : (~0ull << (shift + 8));
        for (int j = tid; j < nr; j += 256) {
            uint64_t u = ord_u64(row[j]);
            if ((u & mask) == prefix) atomicAdd(&hist[(u >> shift) & 255], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            int need = s_need;
            int b = 0;
            for (; b < 256; ++b) {
                int h = (int)hist[b];
                if (need <= h) break;
                need -= h;
            }
            s_need = need;
            s_prefix = prefix | ((uint64_t)b << shift);
        }
        __syncthreads();
    }
    const uint64_t kth = s_prefix;  // key of the k-th smallest element
    const int need_eq = s_need;     // how many elements equal to kth belong to the top-k
    // ordered collect: everything below kth plus the first need_eq equal ones, in index order
    if (tid == 0) s_base = 0;
    __shared__ int s_eq_taken;
    if (tid == 0) s_eq_taken = 0;
    __syncthreads();
    for (int j0 = 0; j0 < nr; j0 += 256) {
        const int j = j0 + tid;
        uint64_t u = 0;
        bool lt = false, eq = false;
        if (j < nr) {
            u = ord_u64(row[j]);
            lt = u < kth;
            eq = u == kth;
        }
        // rank of equal elements within this chunk (index order)
        unsigned beq = __ballot_sync(0xffffffffu, eq);
        int lane = tid & 31, w = tid >> 5;
        if (lane == 0) s_wsum[w] = __popc(beq);
        __syncthreads();
        int eq_before = s_eq_taken;
        for (int x = 0; x < w; ++x) eq_before += s_wsum[x];
        eq_before += __popc(beq & ((1u << lane) - 1));
        int eq_total = 0;
        for (int x = 0; x < 8; ++x) eq_total += s_wsum[x];
        bool take = lt || (eq && eq_before < need_eq);
        __syncthreads();
        unsigned bt = __ballot_sync(0xffffffffu, take);
        if (lane == 0) s_wsum[w] = __popc(bt);
        __syncthreads();
        int pos = s_base;
        for (int x = 0; x < w; ++x) pos += s_wsum[x];
        pos += __popc(bt & ((1u << lane) - 1));
        if (take && pos < kcap) { skey[pos] = u; sidx[pos] = j; }
        int tot = 0;
        for (int x = 0; x < 8; ++x) tot += s_wsum[x];
        __syncthreads();
        if (tid == 0) { s_base += tot; s_eq_taken += eq_total; }
        __syncthreads();
    }
    for (int t = k + tid; t < kcap; t += 256) { skey[t] = ~0ull; sidx[t] = 0x7fffffff; }
    __syncthreads();
    if (!sort_by_index) {
        for (int size = 2; size <= kcap; size <<= 1) {
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                for (int t = tid; t < kcap; t += 256) {
                    int p = t ^ stride;
                    if (p > t) {
                        bool up = (t & size) == 0;
                        uint64_t ka = skey[t], kb = skey[p];
                        int ia = sidx[t], ib = sidx[p];
                        bool gt = ka > kb || (ka == kb && ia > ib);
                        if (gt == up) { skey[t] = kb; skey[p] = ka; sidx[t] = ib; sidx[p] = ia; }
                    }
                }
                __syncthreads();
            }
        }
    }
    const size_t o = (size_t)(q0 + blockIdx.x) * k;
    for (int t = tid; t < k; t += 256) {
        int j = sidx[t];
        double v = ord_f64(skey[t]);
        if (idx64) idx64[o + t] = j;
        if (idx32) idx32[o + t] = j;
        if (dist64) dist64[o + t] = v;
        if (dist32) dist32[o + t] = dist32_mode == 1 ? sqrtf((float)v + 1e-7f) : (dist32_mode == 2 ? (float)(-v) : (float)v);
    }
}

static const size_t kRowsBudget = size_t(256) << 20;

size_t cb_rows_topk_workspace(int nq, int nr, int k) {
    size_t per_row = (size_t)nr * 8;
    size_t rows = kRowsBudget / (per_row ? per_row : 1);
    if (rows < 1) rows = 1;
    if (rows > (size_t)nq) rows = nq;
    return cb_align(rows * per_row) + 4096;
}

int cb_rows_topk(cb_ctx* ctx, const float* q, int nq, const float* r, int nr, int d, int k, int metric,
                 int exclude_self, int self_offset, int sort_by_index, int64_t* idx64, int32_t* idx32,
                 double* dist64, float* dist32, int dist32_mode, const int32_t* q_gather, cudaStream_t st) {
    if (nq == 0) return CB_OK;
    if (k <= 0 || k > nr - (exclude_self == 1 ? 1 : 0))
        CB_FAIL(ctx, CB_EINVAL, "top-k: k=%d outside [1,%d]", k, nr - (exclude_self == 1 ? 1 : 0));
    int kcap = 32;
    while (kcap < k) kcap <<= 1;
    if (kcap > 16384) CB_FAIL(ctx, CB_EUNSUPPORTED, "top-k: k=%d above the 16384 the sort network holds", k);
    size_t smem = (size_t)kcap * 12;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "smem opt-in failed: %s", cudaGetErrorString(e));
    }
    size_t per_row = (size_t)nr * 8;
    int rows_per = (int)(kRowsBudget / per_row);
    if (rows_per < 1) rows_per = 1;
    if (rows_per > nq) rows_per = nq;
    double* rows = cb_alloc_t<double>(ctx, (size_t)rows_per * nr);
    if (!rows) return CB_ENOMEM;
    for (int q0 = 0; q0 < nq; q0 += rows_per) {
        int nqc = min(rows_per, nq - q0);
        cb_timed_scope ts(ctx, CB_T_ROWS_SELECT, (double)nqc * nr * d * 4.0, st);
        dim3 grid(cb_div_up(nr, 32), cb_div_up(nqc, 32));
        rows_kernel<<<grid, dim3(32, 8), 0, st>>>(q, q_gather, q0, nqc, r, nr, d, metric, exclude_self,
                                                  self_offset, rows);
        CB_LAUNCH_CHECK(ctx);
        select_kernel<<<nqc, 256, smem, st>>>(rows, nr, k, kcap, q0, sort_by_index, idx64, idx32, dist64, dist32,
                                              dist32_mode);
        CB_LAUNCH_CHECK(ctx);
    }
    return CB_OK;
}

extern "C" int cb_retrieval_topk(cb_ctx* ctx, const float* q, int nq, const float* db, int ndb, int d, int k,
                                 int metric, int exclude_self, int self_offset, int32_t* idx, double* dist,
                                 void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (nq == 0) return CB_OK;
    if (!q || !db || !idx || nq < 0 || ndb <= 0 || d <= 0 || (metric != 0 && metric != 1))
        CB_FAIL(ctx, CB_EINVAL, "cb_retrieval_topk: bad arguments");
    if ((rc = cb_reserve(ctx, cb_rows_topk_workspace(nq, ndb, k)))) return rc;
    return cb_rows_topk(ctx, q, nq, db, ndb, d, k, metric == 0 ? 1 : 3, exclude_self, self_offset, 0, nullptr,
                        idx, dist, nullptr, 0, nullptr, (cudaStream_t)stream);
}

extern "C" int cb_feature_rows_topk(cb_ctx* ctx, const float* feat, int n, int d, const int32_t* sample_idx,
                                    int n_samples, int n_nn, int32_t* nn_idx, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (n_samples == 0) return CB_OK;
    if (!feat || !sample_idx || !nn_idx || n <= 0 || d <= 0 || n_nn <= 0 || n_nn > n)
        CB_FAIL(ctx, CB_EINVAL, "cb_feature_rows_topk: bad arguments");
    if ((rc = cb_reserve(ctx, cb_rows_topk_workspace(n_samples, n, n_nn)))) return rc;
    return cb_rows_topk(ctx, feat, n_samples, feat, n, d, n_nn, 0, 0, 0, 1, nullptr, nn_idx, nullptr, nullptr, 0,
                        sample_idx, (cudaStream_t)stream);
}
