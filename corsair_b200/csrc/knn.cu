// K1: feature-space kNN  (replaces utils/find_nn.py:34-119 of the reference).
//
// Engine E1 (d <= 32, k <= 12): approximate candidate pass + exact re-rank.
//   1. row norms |r|^2 of the references (fp32) and their maximum
//   2. candidate kernel: GEMM-form score s = |r|^2 - 2 q.r (or -q.r) per (query, ref chunk),
//      keeps the KC best per chunk in registers.  Two variants: SIMT fp32 (this file) and
//      tcgen05 split-bf16 tensor-core tiles (knn_tc.cu).
//   3. re-rank: one warp per query recomputes the S*KC candidates in float64 difference form
//      (unfused, sequential over the feature axis: bit-identical to the oracle), orders them by
//      (distance, index) and PROVES that no dropped reference can belong to the top-k using the
//      score error bound; rows that cannot be proven are queued
//   4. fallback: queued rows are solved by exact float64 brute force.
// Engine E2 (anything else): exact float64 distance rows + per-row radix select (select.cu).
#include "common.cuh"
#include "knn_internal.cuh"

#include <float.h>
#include <math.h>

// ------------------------------------------------------------------------------- norms
__global__ void knn_norms_kernel(const float* __restrict__ r, int nr, int d, float* __restrict__ rn,
                                 unsigned int* __restrict__ rmax_bits) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    float s = 0.f;
    if (j < nr) {
        const float* p = r + (size_t)j * d;
        for (int c = 0; c < d; ++c) s = fmaf(p[c], p[c], s);
        rn[j] = s;
    }
    // block max -> one atomic (non-negative floats order like their bit patterns)
    for (int o = 16; o > 0; o >>= 1) s = fmaxf(s, __shfl_xor_sync(0xffffffffu, s, o));
    if ((threadIdx.x & 31) == 0 && s > 0.f) atomicMax(rmax_bits, __float_as_uint(s));
}

// pad rows of width d to width dp (zero fill) so E1 only sees d in {16, 32}
__global__ void knn_pad_kernel(const float* __restrict__ in, int n, int d, int dp, float* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)n * dp) return;
    int row = (int)(i / dp), c = (int)(i % dp);
    out[i] = c < d ? in[(size_t)row * d + c] : 0.f;
}

// ------------------------------------------------------------- SIMT candidate kernel
// grid (ceil(nq/128), S); block 128: one query per thread, reference chunk tiled through smem.
template <int D, int KC>
__global__ void __launch_bounds__(128)
knn_cand_simt_kernel(const float* __restrict__ q, int nq, const float* __restrict__ r,
                     const float* __restrict__ rn, int nr, int chunk, int maxdot,
                     int32_t* __restrict__ cand_idx, float* __restrict__ cand_s, int S) {
    constexpr int TR = 64;
    __shared__ __align__(16) float sr[TR * D];
    __shared__ float sn[TR];
    const int qi = blockIdx.x * 128 + threadIdx.x;
    const int c = blockIdx.y;
    const int r0 = c * chunk;
    const int r1 = min(nr, r0 + chunk);

    float qv[D];
    if (qi < nq) {
        const float4* qp = reinterpret_cast<const float4*>(q + (size_t)qi * D);
#pragma unroll
        for (int t = 0; t < D / 4; ++t) {
            float4 v = qp[t];
            qv[4 * t] = v.x; qv[4 * t + 1] = v.y; qv[4 * t + 2] = v.z; qv[4 * t + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int t = 0; t < D; ++t) qv[t] = 0.f;
    }
    float bs[KC];
    int bi[KC];
#pragma unroll
    for (int t = 0; t < KC; ++t) { bs[t] = INFINITY; bi[t] = -1; }

    for (int base = r0; base < r1; base += TR) {
        const int nt = min(TR, r1 - base);
        __syncthreads();
        const float4* src = reinterpret_cast<const float4*>(r + (size_t)base * D);
        for (int t = threadIdx.x; t < nt * (D / 4); t += 128) reinterpret_cast<float4*>(sr)[t] = src[t];
        if (threadIdx.x < nt) sn[threadIdx.x] = maxdot ? 0.f : rn[base + threadIdx.x];
        __syncthreads();
        for (int j = 0; j < nt; ++j) {
            const float4* rp = reinterpret_cast<const float4*>(sr + j * D);
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int t = 0; t < D / 4; ++t) {
                float4 v = rp[t];
                a0 = fmaf(qv[4 * t], v.x, a0);
                a1 = fmaf(qv[4 * t + 1], v.y, a1);
                a0 = fmaf(qv[4 * t + 2], v.z, a0);
                a1 = fmaf(qv[4 * t + 3], v.w, a1);
            }
            float dot = a0 + a1;
            float s = maxdot ? -dot : fmaf(-2.f, dot, sn[j]);
            if (s < bs[KC - 1]) {
                bs[KC - 1] = s;
                bi[KC - 1] = base + j;
#pragma unroll
                for (int t = KC - 1; t > 0; --t) {
                    if (bs[t] < bs[t - 1]) {
                        float fs = bs[t]; bs[t] = bs[t - 1]; bs[t - 1] = fs;
                        int fi = bi[t]; bi[t] = bi[t - 1]; bi[t - 1] = fi;
                    }
                }
            }
        }
    }
    if (qi < nq) {
        size_t o = ((size_t)qi * S + c) * KC;
#pragma unroll
        for (int t = 0; t < KC; ++t) { cand_idx[o + t] = bi[t]; cand_s[o + t] = bs[t]; }
    }
}

// --------------------------------------------------------------------- exact re-rank
__device__ __forceinline__ double knn_exact_key(const float* __restrict__ qp, const float* __restrict__ rp,
                                                int d, int maxdot) {
    // unfused, sequential over the feature axis: the same float64 operations as the oracle
    double acc = 0.0;
    if (maxdot) {
        for (int c = 0; c < d; ++c) acc = __dadd_rn(acc, __dmul_rn((double)qp[c], (double)rp[c]));
        return -acc;
    }
    for (int c = 0; c < d; ++c) {
        double df = __dsub_rn((double)qp[c], (double)rp[c]);
        acc = __dadd_rn(acc, __dmul_rn(df, df));
    }
    return acc;
}

// the same operations with the query row already converted (exact) and a compile-time width
template <int D>
__device__ __forceinline__ double knn_exact_key_q64(const double* __restrict__ qd, const float* __restrict__ rp, int d,
                                                    int maxdot) {
    double acc = 0.0;
    if (maxdot) {
#pragma unroll
        for (int c = 0; c < (D ? D : 32); ++c)
            if (D || c < d) acc = __dadd_rn(acc, __dmul_rn(qd[c], (double)rp[c]));
        return -acc;
    }
#pragma unroll
    for (int c = 0; c < (D ? D : 32); ++c)
        if (D || c < d) {
            const double df = __dsub_rn(qd[c], (double)rp[c]);
            acc = __dadd_rn(acc, __dmul_rn(df, df));
        }
    return acc;
}

// ... and the candidate row staged 16-byte aligned in shared memory, read four columns at a time
template <int D>
__device__ __forceinline__ double knn_exact_key_q64_vec(const double* __restrict__ qd, const float* __restrict__ rp, int maxdot) {
    double acc = 0.0;
#pragma unroll
    for (int c = 0; c < D; c += 4) {
        const float4 v = *reinterpret_cast<const float4*>(rp + c);
        const float r4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (maxdot) {
                acc = __dadd_rn(acc, __dmul_rn(qd[c + e], (double)r4[e]));
            } else {
                const double df = __dsub_rn(qd[c + e], (double)r4[e]);
                acc = __dadd_rn(acc, __dmul_rn(df, df));
            }
        }
    }
    return maxdot ? -acc : acc;
}

__device__ __forceinline__ bool knn_less(double a, int ia, double b, int ib) {
    return a < b || (a == b && ia < ib);
}

__device__ __forceinline__ float knn_out_dist(double key, int metric) {
    if (metric == CB_KNN_L2) return sqrtf((float)key + 1e-7f);
    if (metric == CB_KNN_MAXDOT) return (float)(-key);
    return (float)key;
}

// (key, index) of a candidate as integers that order like the pair: the float64 key's bits made
// monotone (negative keys occur for the inner-product metric)
__device__ __forceinline__ unsigned long long knn_sortable(double key) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(key);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double knn_unsortable(unsigned long long s) {
    return __longlong_as_double((long long)((s >> 63) ? (s & 0x7fffffffffffffffull) : ~s));
}

constexpr int kRowPad = 33;  // floats per staged row: lane l reading column c of row l hits bank (l + c) % 32
constexpr int kRowVec = 36;  // ... or 16-byte aligned rows read with LDS.128: conflict-free per quarter-warp (36 l mod 32 = 4 l)

// asynchronous 4-byte copy global -> shared (LDGSTS): the copies of a whole batch are in flight together;
// with a register in between ptxas paired every load with its store and the rows arrived one L2
// latency after the other (SASS, round 2)
__device__ __forceinline__ void knn_cp_async4(float* dst_smem, const float* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void knn_cp_async16(float* dst_smem, const float* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void knn_cp_async_wait() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// one warp per query; ncand = S*KC <= 32*SL candidate slots (SL per lane; unused slots hold -1).
// bound (optional, [nq, S]): per chunk a score no dropped reference of that chunk is below;
// without it the chunk's last (worst kept) candidate score plays that role.
// Candidate rows are gathered COOPERATIVELY: the warp copies each candidate's row with one coalesced
// load into shared memory and every lane then walks its own row there (round 1: every lane walked its
// own global row with scalar loads, 32 cache lines per load instruction — the kernel was bound by L1
// tag traffic, 322 us at 50 000 x 10 000 with ~48 candidates per query).
// The k rounds of "smallest remaining (key, index)" are three warp-wide integer minima (REDUX) each:
// high word of the key, low word among the lanes that match, index among those that still match.
template <int SL>
__global__ void __launch_bounds__(256)
knn_rerank_kernel(const float* __restrict__ q, int nq, const float* __restrict__ r, int nr_total, int d, int dq_stride,
                  int k, int metric, const int32_t* __restrict__ cand_idx, const float* __restrict__ cand_s,
                  int S, int KC, const float* __restrict__ bound, const unsigned int* __restrict__ rmax_bits,
                  double gamma, const int32_t* __restrict__ unscramble, int unscramble_n,
                  int64_t* __restrict__ idx64, int32_t* __restrict__ idx32, float* __restrict__ dist,
                  int32_t* __restrict__ fb_list, int* __restrict__ fb_count) {
    __shared__ __align__(16) float s_rows[8][32][kRowVec];
    __shared__ __align__(16) double s_q[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int qi = blockIdx.x * (blockDim.x >> 5) + warp;
    // rows of 16 or 32 floats at their natural stride: every lane copies ITS candidate's row with 16-byte
    // asynchronous copies and reads it back with LDS.128 (stride 36: conflict-free per quarter-warp)
    const bool vec = (d == 32 || d == 16) && dq_stride == d && ((size_t)r & 15) == 0;
    if (qi >= nq) return;
    const int maxdot = metric == CB_KNN_MAXDOT;
    const int ncand = S * KC;
    const float* qp = q + (size_t)qi * dq_stride;
    float (*rows)[kRowVec] = s_rows[warp];
    s_q[warp][lane] = lane < d ? (double)qp[lane] : 0.0;
    unsigned long long key[SL];
    int id[SL];
#pragma unroll
    for (int t = 0; t < SL; ++t) {
        const int ci = lane + 32 * t;
        int j = ci < ncand ? cand_idx[(size_t)qi * ncand + ci] : -1;
        if (unscramble && j >= 0) {
            CB_DEV_CHECK(j < unscramble_n);
            j = unscramble[j];
        }
        id[t] = 0x7fffffff;
        key[t] = ~0ull;
        if (!__ballot_sync(0xffffffffu, j >= 0)) continue;
        __syncwarp();
        if (vec) {
            if (j >= 0) {
                CB_DEV_CHECK(j < nr_total);
                const float* src = r + (size_t)j * d;
                if (d == 32) {
#pragma unroll
                    for (int c = 0; c < 32; c += 4) knn_cp_async16(&rows[lane][c], src + c);
                } else {
#pragma unroll
                    for (int c = 0; c < 16; c += 4) knn_cp_async16(&rows[lane][c], src + c);
                }
            }
        } else {
#pragma unroll 8
            for (int rr = 0; rr < 32; ++rr) {
                const int jj = __shfl_sync(0xffffffffu, j, rr);
                CB_DEV_CHECK(jj < nr_total);
                if (jj >= 0 && lane < d) knn_cp_async4(&rows[rr][lane], r + (size_t)jj * dq_stride + lane);
            }
        }
        knn_cp_async_wait();
        __syncwarp();
        if (j >= 0) {
            id[t] = j;
            const double kd = vec ? (d == 32 ? knn_exact_key_q64_vec<32>(s_q[warp], rows[lane], maxdot)
                                             : knn_exact_key_q64_vec<16>(s_q[warp], rows[lane], maxdot))
                                  : knn_exact_key_q64<0>(s_q[warp], rows[lane], d, maxdot);
            key[t] = knn_sortable(kd);
        }
    }
    double kth = 0.0;
    int found = 0;
    for (int round = 0; round < k; ++round) {
        unsigned long long bk = ~0ull;
        int bj = 0x7fffffff;
#pragma unroll
        for (int t = 0; t < SL; ++t)
            if (key[t] < bk || (key[t] == bk && id[t] < bj)) { bk = key[t]; bj = id[t]; }
        const unsigned hi = (unsigned)(bk >> 32), lo = (unsigned)bk;
        const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
        const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
        const bool tied = hi == mh && lo == ml;
        const int wj = (int)__reduce_min_sync(0xffffffffu, tied ? (unsigned)bj : 0x7fffffffu);
        if (wj == 0x7fffffff) break;  // fewer than k candidates (cannot happen when k <= nr)
        if (tied && bj == wj) {
#pragma unroll
            for (int t = 0; t < SL; ++t)
                if (id[t] == wj) { id[t] = 0x7fffffff; key[t] = ~0ull; }
        }
        const double wk = knn_unsortable(((unsigned long long)mh << 32) | ml);
        if (lane == 0) {
            size_t o = (size_t)qi * k + round;
            if (idx64) idx64[o] = wj;
            if (idx32) idx32[o] = wj;
            if (dist) dist[o] = knn_out_dist(wk, metric);
        }
        kth = wk;
        ++found;
    }
    // proof obligation: nothing dropped by a chunk can beat the k-th exact key
    double qn = 0.0;
    for (int c = lane; c < d; c += 32) qn += (double)qp[c] * (double)qp[c];
    for (int o = 16; o > 0; o >>= 1) qn += __shfl_xor_sync(0xffffffffu, qn, o);
    const double rmax = (double)__uint_as_float(*rmax_bits);
    double e, shift;
    if (maxdot) {
        e = gamma * sqrt(qn * rmax) + 1e-300;
        shift = 0.0;
    } else {
        e = gamma * (rmax + 2.0 * sqrt(qn * rmax)) + 1e-300;
        shift = qn * (1.0 - 1e-12);
    }
    bool bad = found < k;
    for (int c = lane; c < S; c += 32) {
        float w = bound ? bound[(size_t)qi * S + c] : cand_s[((size_t)qi * S + c) * KC + KC - 1];
        if (!(w == INFINITY) && !((double)w - e + shift > kth)) bad = true;
    }
    bad = __any_sync(0xffffffffu, bad);
    if (bad && lane == 0) fb_list[atomicAdd(fb_count, 1)] = qi;
}

// exact float64 brute force for the queued rows: one block per row.  The block stages 256 references
// at a time in shared memory with coalesced loads; every thread evaluates one of them per tile and
// keeps its own k best in registers (sorted, k <= 12 in engine E1); the block then merges the 256
// sorted lists in k rounds of "smallest head".  (Round 1 rescanned all references once per round, every
// thread walking its own global rows: 900 us for a single queued row at 50 000 x 10 000.)
// With few queued rows (the usual case: 0-2) one block per row leaves the row's 40 x 256-reference steps
// on a single SM, ~50-100 us of pure latency; the references of a row are therefore split over up to
// kFbParts blocks (part p scans the steps t = p mod parts), each writes its k best to a scratch list and
// the last part to finish (atomic ticket) merges them.
constexpr int kFbK = 12;
constexpr int kFbParts = 8;
constexpr int kFbSplitRows = 1024;  // rows the scratch lists hold; more queued rows run one block per row
__global__ void __launch_bounds__(256)
knn_fallback_kernel(const float* __restrict__ q, const float* __restrict__ r, int nr, int d, int stride,
                    int k, int metric, const int32_t* __restrict__ fb_list, const int* __restrict__ fb_count,
                    unsigned long long* __restrict__ part_key, int* __restrict__ part_idx, int* __restrict__ part_done,
                    int64_t* __restrict__ idx64, int32_t* __restrict__ idx32, float* __restrict__ dist) {
    __shared__ float s_rows[256][kRowPad];
    __shared__ float s_q[32];
    __shared__ unsigned long long s_key[8];
    __shared__ int s_id[8];
    __shared__ unsigned long long s_wkey;
    __shared__ int s_wid;
    const int maxdot = metric == CB_KNN_MAXDOT;
    const int n = *fb_count;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ int s_last;
    const int steps = (nr + 255) / 256;
    const int parts = (n > 0 && n <= kFbSplitRows) ? max(1, min(min(kFbParts, steps), (int)gridDim.x / n)) : 1;
    for (int u = blockIdx.x; u < n * parts; u += gridDim.x) {
        const int e = u / parts, part = u % parts;
        const int qi = fb_list[e];
        __syncthreads();
        if (threadIdx.x < 32) s_q[threadIdx.x] = threadIdx.x < d ? q[(size_t)qi * stride + threadIdx.x] : 0.f;
        unsigned long long bk[kFbK];
        int bj[kFbK];
#pragma unroll
        for (int t = 0; t < kFbK; ++t) { bk[t] = ~0ull; bj[t] = 0x7fffffff; }
        for (int base = part * 256; base < nr; base += parts * 256) {
            const int nt = min(256, nr - base);
            __syncthreads();
            // warp w copies rows w, w + 8, ...: one coalesced load per row
            if (lane < d)
                for (int rr = warp; rr < nt; rr += 8) knn_cp_async4(&s_rows[rr][lane], r + (size_t)(base + rr) * stride + lane);
            knn_cp_async_wait();
            __syncthreads();
            if ((int)threadIdx.x < nt) {
                const unsigned long long key = knn_sortable(knn_exact_key(s_q, s_rows[threadIdx.x], d, maxdot));
                const int j = base + threadIdx.x;
                // sorted insertion by (key, index); slots beyond k are never read
                if (key < bk[kFbK - 1] || (key == bk[kFbK - 1] && j < bj[kFbK - 1])) {
                    bk[kFbK - 1] = key;
                    bj[kFbK - 1] = j;
#pragma unroll
                    for (int t = kFbK - 1; t > 0; --t) {
                        const bool sw = bk[t] < bk[t - 1] || (bk[t] == bk[t - 1] && bj[t] < bj[t - 1]);
                        if (sw) {
                            const unsigned long long tk = bk[t]; bk[t] = bk[t - 1]; bk[t - 1] = tk;
                            const int tj = bj[t]; bj[t] = bj[t - 1]; bj[t - 1] = tj;
                        }
                    }
                }
            }
        }
        for (int round = 0; round < k; ++round) {
            // smallest head of the block: warp minima (three REDUX), then the eight warps through shared memory
            const unsigned hi = (unsigned)(bk[0] >> 32), lo = (unsigned)bk[0];
            const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
            const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
            const bool tied = hi == mh && lo == ml;
            const int wj = (int)__reduce_min_sync(0xffffffffu, tied ? (unsigned)bj[0] : 0x7fffffffu);
            if (lane == 0) { s_key[warp] = ((unsigned long long)mh << 32) | ml; s_id[warp] = wj; }
            __syncthreads();
            if (threadIdx.x == 0) {
                unsigned long long wk = s_key[0];
                int wi = s_id[0];
                for (int w = 1; w < 8; ++w)
                    if (s_key[w] < wk || (s_key[w] == wk && s_id[w] < wi)) { wk = s_key[w]; wi = s_id[w]; }
                s_wkey = wk;
                s_wid = wi;
                if (parts > 1) {
                    part_key[((size_t)e * kFbParts + part) * kFbK + round] = wk;
                    part_idx[((size_t)e * kFbParts + part) * kFbK + round] = wi;
                } else if (wi != 0x7fffffff) {
                    const size_t o = (size_t)qi * k + round;
                    if (idx64) idx64[o] = wi;
                    if (idx32) idx32[o] = wi;
                    if (dist) dist[o] = knn_out_dist(knn_unsortable(wk), metric);
                }
            }
            __syncthreads();
            if (bj[0] == s_wid && bk[0] == s_wkey) {
                // the winner pops its head
#pragma unroll
                for (int t = 0; t + 1 < kFbK; ++t) { bk[t] = bk[t + 1]; bj[t] = bj[t + 1]; }
                bk[kFbK - 1] = ~0ull;
                bj[kFbK - 1] = 0x7fffffff;
            }
        }
        if (parts > 1) {
            // the last part of this row to finish merges the parts' sorted lists (warp 0: <= 96 entries, three
            // per lane, k rounds of three REDUX minima)
            if (threadIdx.x == 0) {
                __threadfence();
                s_last = atomicAdd(&part_done[e], 1) == parts - 1;
            }
            __syncthreads();
            if (s_last && warp == 0) {
                __threadfence();
                unsigned long long mk[3];
                int mj[3];
#pragma unroll
                for (int t = 0; t < 3; ++t) {
                    const int i = lane + 32 * t;  // entry (part i / k, round i % k)
                    const bool ok = i < parts * k;
                    const size_t o = ((size_t)e * kFbParts + (ok ? i / k : 0)) * kFbK + (ok ? i % k : 0);
                    mk[t] = ok ? __ldcg(&part_key[o]) : ~0ull;
                    mj[t] = ok ? __ldcg(&part_idx[o]) : 0x7fffffff;
                }
                for (int round = 0; round < k; ++round) {
                    unsigned long long b = ~0ull;
                    int bi = 0x7fffffff;
#pragma unroll
                    for (int t = 0; t < 3; ++t)
                        if (mk[t] < b || (mk[t] == b && mj[t] < bi)) { b = mk[t]; bi = mj[t]; }
                    const unsigned hi = (unsigned)(b >> 32), lo = (unsigned)b;
                    const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
                    const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
                    const bool tied = hi == mh && lo == ml;
                    const int wi = (int)__reduce_min_sync(0xffffffffu, tied ? (unsigned)bi : 0x7fffffffu);
                    if (wi == 0x7fffffff) break;
                    if (tied && bi == wi) {
#pragma unroll
                        for (int t = 0; t < 3; ++t)
                            if (mj[t] == wi) { mk[t] = ~0ull; mj[t] = 0x7fffffff; }
                    }
                    if (lane == 0) {
                        const size_t o = (size_t)qi * k + round;
                        if (idx64) idx64[o] = wi;
                        if (idx32) idx32[o] = wi;
                        if (dist) dist[o] = knn_out_dist(knn_unsortable(((unsigned long long)mh << 32) | ml), metric);
                    }
                }
                if (lane == 0) part_done[e] = 0;
            }
        }
    }
}

// ------------------------------------------------------------------------ host driver
template <int D, int KC>
static void launch_simt(const float* q, int nq, const float* r, const float* rn, int nr, int chunk, int S,
                        int maxdot, int32_t* ci, float* cs, cudaStream_t st) {
    dim3 grid(cb_div_up(nq, 128), S);
    knn_cand_simt_kernel<D, KC><<<grid, 128, 0, st>>>(q, nq, r, rn, nr, chunk, maxdot, ci, cs, S);
}

const char* cb_knn_variant_name(const cb_ctx* ctx);

static size_t knn_workspace(cb_ctx* ctx, int nq, int nr, int d, int k);

int cb_knn_impl(cb_ctx* ctx, const float* q, int nq, const float* r, int nr, int d, int k, int metric,
                int64_t* idx, float* dist, int32_t* idx32, cudaStream_t st) {
    if (nq == 0) return CB_OK;
    if (!q || !r || nq < 0 || nr <= 0 || d <= 0 || d > 256 || k <= 0 || k > 64 || k > nr || metric < 0 ||
        metric > 2)
        CB_FAIL(ctx, CB_EINVAL, "cb_knn: bad arguments (nq=%d nr=%d d=%d k=%d metric=%d)", nq, nr, d, k,
                metric);
    if (d > 32 || k > 12) {
        // Engine E2: exact rows + radix select
        int m2 = metric == CB_KNN_MAXDOT ? 2 : 0;
        int mode = metric == CB_KNN_L2 ? 1 : (metric == CB_KNN_MAXDOT ? 2 : 0);
        return cb_rows_topk(ctx, q, nq, r, nr, d, k, m2, 0, 0, 0, idx, idx32, nullptr, dist, mode, nullptr,
                            st);
    }
    const int dp = d <= 16 ? 16 : 32;
    const int KC = k <= 5 ? 8 : 16;
    // the tcgen05 variant collects every reference at or above a per-(query, chunk) threshold: up to
    // LC = 6 KC list slots per chunk; the SIMT variant keeps exactly the KC best
    const bool use_tc = ctx->knn_variant != 1;
    const int LC = use_tc ? knn_cand_tc_list_slots(KC) : KC;
    const int qblocks = cb_div_up(nq, 128);
    int S, chunk;
    if (use_tc) {
        // reference chunks of whole 256-column tiles, S * LC <= 256 list slots per query: the S with the
        // shortest makespan — rounds of CTAs (one per SM) x tile passes per chunk (+ a per-CTA constant
        // for the prologue); ties go to the smaller S (fewer candidates to re-rank)
        const int tiles = cb_div_up(nr, 256);
        double best_cost = 1e300;
        int best_tps = tiles;
        for (int s_try = 1; s_try <= min(256 / LC, tiles); ++s_try) {
            const int tps = cb_div_up(tiles, s_try);
            const int ss = cb_div_up(tiles, tps);
            const int rounds = cb_div_up(qblocks * ss, ctx->sm_count);
            // every extra chunk adds ~2 KC candidates per query to append and re-rank: 3 % per chunk
            const double cost = rounds * (knn_cand_tc_passes(tps) + 1.5) * (1.0 + 0.03 * ss);
            if (cost < best_cost - 1e-9) { best_cost = cost; best_tps = tps; }
        }
        chunk = best_tps * 256;
        S = cb_div_up(nr, chunk);
    } else {
        S = ctx->sm_count / qblocks;  // about one CTA per SM, rounded down: no second round of CTAs
        S = max(1, min(S, 256 / LC));
        S = min(S, max(1, nr / 256));
        chunk = cb_div_up(nr, S);
        chunk = (chunk + 255) & ~255;
        S = cb_div_up(nr, chunk);
    }

    unsigned int* scal = cb_alloc_t<unsigned int>(ctx, 2);  // [0] rmax bits, [1] fallback count
    int32_t* ci = cb_alloc_t<int32_t>(ctx, (size_t)nq * S * LC);
    float* bound = cb_alloc_t<float>(ctx, (size_t)nq * S);
    int32_t* fb = cb_alloc_t<int32_t>(ctx, nq);
    // scratch of the exact fallback's split rows: per row and part kFbK (key, index) entries, a ticket per row
    const size_t fb_entries = (size_t)kFbSplitRows * kFbParts * kFbK;
    unsigned long long* fb_key = cb_alloc_t<unsigned long long>(ctx, fb_entries);
    int* fb_idx = cb_alloc_t<int>(ctx, fb_entries);
    int* fb_done = cb_alloc_t<int>(ctx, kFbSplitRows);
    if (!scal || !ci || !bound || !fb || !fb_key || !fb_idx || !fb_done) return CB_ENOMEM;
    CB_CUDA(ctx, cudaMemsetAsync(scal, 0, 2 * sizeof(unsigned int), st));
    CB_CUDA(ctx, cudaMemsetAsync(fb_done, 0, kFbSplitRows * sizeof(int), st));
    ctx->knn_fb_count = (const int*)(scal + 1);
    const int maxdot = metric == CB_KNN_MAXDOT;
    double gamma = (dp + 4) * 6.0e-8;
    // rows the exact kernels read: the caller's (stride d) for the tensor-core variant, zero-padded
    // copies (stride dp) for the SIMT one
    const float* qd = q;
    const float* rd = r;
    int stride = d;
    float* cs = nullptr;
    const int32_t* unscramble = nullptr;
    if (use_tc) {
        cb_timed_scope tscope(ctx, CB_T_KNN_CAND, 2.0 * nq * (double)nr * dp, st);
        int rc = knn_cand_tc_launch(ctx, q, nq, r, nr, d, dp, KC, LC, chunk, S, maxdot, ci, bound, scal, &unscramble, st);
        if (rc) return rc;
        gamma = knn_cand_tc_gamma(dp);
    } else {
        // the SIMT kernel reads rows as float4: narrower rows and pointers off a 16-byte boundary go through
        // padded copies in the (aligned) arena
        if (dp != d || (((size_t)q | (size_t)r) & 15) != 0) {
            float* qpad = cb_alloc_t<float>(ctx, (size_t)nq * dp);
            float* rpad = cb_alloc_t<float>(ctx, (size_t)nr * dp);
            if (!qpad || !rpad) return CB_ENOMEM;
            knn_pad_kernel<<<cb_div_up((int)((size_t)nq * dp), 256), 256, 0, st>>>(q, nq, d, dp, qpad);
            CB_LAUNCH_CHECK(ctx);
            knn_pad_kernel<<<cb_div_up((int)((size_t)nr * dp), 256), 256, 0, st>>>(r, nr, d, dp, rpad);
            CB_LAUNCH_CHECK(ctx);
            qd = qpad;
            rd = rpad;
        }
        stride = dp;
        float* rn = cb_alloc_t<float>(ctx, nr);
        cs = cb_alloc_t<float>(ctx, (size_t)nq * S * LC);
        if (!rn || !cs) return CB_ENOMEM;
        knn_norms_kernel<<<cb_div_up(nr, 256), 256, 0, st>>>(rd, nr, dp, rn, scal);
        CB_LAUNCH_CHECK(ctx);
        cb_timed_scope tscope(ctx, CB_T_KNN_CAND, 2.0 * nq * (double)nr * dp, st);
        if (dp == 16 && KC == 8) launch_simt<16, 8>(qd, nq, rd, rn, nr, chunk, S, maxdot, ci, cs, st);
        else if (dp == 16) launch_simt<16, 16>(qd, nq, rd, rn, nr, chunk, S, maxdot, ci, cs, st);
        else if (KC == 8) launch_simt<32, 8>(qd, nq, rd, rn, nr, chunk, S, maxdot, ci, cs, st);
        else launch_simt<32, 16>(qd, nq, rd, rn, nr, chunk, S, maxdot, ci, cs, st);
        CB_LAUNCH_CHECK(ctx);
    }
    cb_timed_scope tscope2(ctx, CB_T_KNN_RERANK, 0.0, st);
    {
        const int SL = cb_div_up(S * LC, 32);
        auto kern = SL <= 1 ? knn_rerank_kernel<1> : SL <= 2 ? knn_rerank_kernel<2> : SL <= 4 ? knn_rerank_kernel<4>
                                                                                               : knn_rerank_kernel<8>;
        kern<<<cb_div_up(nq, 8), 256, 0, st>>>(qd, nq, rd, nr, d, stride, k, metric, ci, cs, S, LC, use_tc ? bound : nullptr,
                                               scal, gamma, unscramble, cb_div_up(nr, 256) * 256, idx, idx32, dist, fb, (int*)(scal + 1));
        CB_LAUNCH_CHECK(ctx);
    }
    knn_fallback_kernel<<<min(nq * kFbParts, 4 * ctx->sm_count), 256, 0, st>>>(qd, rd, nr, d, stride, k, metric, fb,
                                                                                (int*)(scal + 1), fb_key, fb_idx, fb_done,
                                                                                idx, idx32, dist);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

size_t cb_knn_workspace(int nq, int nr, int d, int k) { return knn_workspace(nullptr, nq, nr, d, k); }

static size_t knn_workspace(cb_ctx* ctx, int nq, int nr, int d, int k) {
    if (d > 32 || k > 12) return cb_rows_topk_workspace(nq, nr, k);
    size_t b = 0;
    b += cb_align((size_t)nq * 32 * 4) + cb_align((size_t)nr * 32 * 4);  // pads
    b += cb_align((size_t)nr * 4) + 256;
    b += 2 * cb_align((size_t)nq * 256 * 4) + cb_align((size_t)nq * 32 * 4) + cb_align((size_t)nq * 4);
    b += cb_align((size_t)(nq + 128) * 112 * 2) + cb_align((size_t)(nr + 256) * 112 * 2) + cb_align((size_t)(nr + 256) * 4) + (size_t)kFbSplitRows * (kFbParts * kFbK * 12 + 4) + (1 << 16);  // packed bf16 operand tiles + un-scramble table of the tc variant
    return b;
}

__global__ void knn_mutual_kernel(const int64_t* __restrict__ idx, int k, const int32_t* __restrict__ rev,
                                  int nq, uint8_t* __restrict__ mutual) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nq) mutual[i] = rev[idx[(size_t)i * k]] == i;
}

extern "C" int cb_knn(cb_ctx* ctx, const float* q, int nq, const float* r, int nr, int d, int k, int metric,
                      int64_t* idx, float* dist, uint8_t* mutual, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (nq == 0) return CB_OK;
    if (!idx) CB_FAIL(ctx, CB_EINVAL, "cb_knn: idx is null");
    if (nq < 0 || nr <= 0 || d <= 0 || d > 256 || k <= 0 || k > nr)
        CB_FAIL(ctx, CB_EINVAL, "cb_knn: bad shape (nq=%d nr=%d d=%d k=%d)", nq, nr, d, k);
    size_t need = knn_workspace(ctx, nq, nr, d, k);
    if (mutual) need += knn_workspace(ctx, nr, nq, d, 1) + cb_align((size_t)nr * 4);
    if ((rc = cb_reserve(ctx, need))) return rc;
    if ((rc = cb_knn_impl(ctx, q, nq, r, nr, d, k, metric, idx, dist, nullptr, st))) return rc;
    if (mutual) {
        int32_t* rev = cb_alloc_t<int32_t>(ctx, nr);
        if (!rev) return CB_ENOMEM;
        if ((rc = cb_knn_impl(ctx, r, nr, q, nq, d, 1, metric, nullptr, nullptr, rev, st))) return rc;
        knn_mutual_kernel<<<cb_div_up(nq, 256), 256, 0, st>>>(idx, k, rev, nq, mutual);
        CB_LAUNCH_CHECK(ctx);
    }
    return CB_OK;
}

extern "C" int cb_knn_last_fallback_rows(cb_ctx* ctx, void* stream) {
    if (!ctx || !ctx->knn_fb_count) return -1;
    int n = -1;
    if (cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) return -1;
    if (cudaMemcpy(&n, ctx->knn_fb_count, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return n;
}

extern "C" int cb_knn_ragged(cb_ctx* ctx, const float* q, const float* r, int d, int k, int metric,
                             const int32_t* h_q_off, const int32_t* h_r_off, int n_problems, int64_t* idx,
                             float* dist, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (n_problems < 0 || !h_q_off || !h_r_off) CB_FAIL(ctx, CB_EINVAL, "cb_knn_ragged: bad offsets");
    size_t need = 0;
    for (int p = 0; p < n_problems; ++p) {
        int nq = h_q_off[p + 1] - h_q_off[p], nr = h_r_off[p + 1] - h_r_off[p];
        if (nq < 0 || nr < 0 || (nq > 0 && k > nr))
            CB_FAIL(ctx, CB_EINVAL, "cb_knn_ragged: problem %d has nq=%d nr=%d k=%d", p, nq, nr, k);
        if (nq > 0) need += knn_workspace(ctx, nq, nr, d, k);
    }
    if ((rc = cb_reserve(ctx, need))) return rc;
    for (int p = 0; p < n_problems; ++p) {
        int nq = h_q_off[p + 1] - h_q_off[p], nr = h_r_off[p + 1] - h_r_off[p];
        if (nq == 0) continue;
        rc = cb_knn_impl(ctx, q + (size_t)h_q_off[p] * d, nq, r + (size_t)h_r_off[p] * d, nr, d, k, metric,
                         idx + (size_t)h_q_off[p] * k, dist ? dist + (size_t)h_q_off[p] * k : nullptr,
                         nullptr, st);
        if (rc) return rc;
    }
    return CB_OK;
}
