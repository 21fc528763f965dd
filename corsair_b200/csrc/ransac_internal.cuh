// Internal interfaces between ransac.cu (fit / SIMT scoring / fold) and ransac_tc.cu (tcgen05
// scoring).
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

struct RansacState {
    int budget;       // hypotheses to evaluate in total (shrinks when adaptive)
    int best_h;
    int best_count;
    int n_eval;
    double best_sum;
    double best_rmse;
    double T[16];
    unsigned int smax_bits;  // max |src| (fp32 bits), max |tgt| of the raw correspondences
    unsigned int tmax_bits;
    // tensor-core scoring: correspondences are centred on their means before the features are
    // built (keeps the quadratic form's terms, hence its rounding band, small for clouds far from
    // the origin)
    double csum[6];           // sum of src xyz, sum of tgt xyz over the m correspondences
    unsigned int smaxc_bits;  // max |src - mean|, max |tgt - mean| (fp32 bits, rounded up)
    unsigned int tmaxc_bits;
    int recount_mismatch;     // fold's exact recount of a winning candidate disagreed (must stay 0)
};

// the oracle's float64 evaluation (unfused, same association order) of one pair
__device__ __forceinline__ double ransac_exact_d2_inl(const double* __restrict__ Td, float4 s, float4 q) {
    const double x = s.x, y = s.y, z = s.z;
    double f0 = __dsub_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(Td[0], x), __dmul_rn(Td[1], y)), __dmul_rn(Td[2], z)), Td[3]), (double)q.x);
    double f1 = __dsub_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(Td[4], x), __dmul_rn(Td[5], y)), __dmul_rn(Td[6], z)), Td[7]), (double)q.y);
    double f2 = __dsub_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(Td[8], x), __dmul_rn(Td[9], y)), __dmul_rn(Td[10], z)), Td[11]), (double)q.z);
    return __dadd_rn(__dadd_rn(__dmul_rn(f0, f0), __dmul_rn(f1, f1)), __dmul_rn(f2, f2));
}

// ---- tensor-core scoring: u = d^2 - thr^2 as ONE inner product over a widened K axis ----------
//   d^2 = |R s + t - q|^2 = sum_ij (-2 R_ij) q_i s_j + sum_j 2 (R^T t)_j s_j + sum_i (-2 t_i) q_i
//         + (|s|^2 + |q|^2) + |t|^2                       (R orthonormal; s, q centred, t adjusted)
// 15 bilinear terms w_k * phi_k in split bf16, three pieces each (w = wa + wb + wc, phi = pa + pb +
// pc: 24 significant bits) and the six products of total order <= 2, i.e. fp32-grade accuracy; the
// per-correspondence norm n and the per-hypothesis constant c = |t|^2 - thr^2 carried as three
// bf16 pieces against ones:
//   column        0..14  15..29  30..44  45..59  60..74  75..89  90..92     93..95
//   A (hypoth.)   wa     wa      wb      wa      wc      wb      1 1 1      c0 c1 c2
//   B (corresp.)  pa     pb      pa      pc      pa      pb      n0 n1 n2   1 1 1
constexpr int kRtKP = 96;        // widened K axis (6 MMAs of K = 16)
constexpr int kRtTileM = 128;    // hypotheses per MMA (TMEM lanes)
constexpr int kRtTileN = 128;    // correspondences per MMA / B tile
constexpr int kRtTerms = 15;
// A tiles (of 128 hypotheses) per scoring CTA.  2: one CTA per SM (512 TMEM columns, 221 KB of shared
// memory); 1: two CTAs per SM (256 columns, 106 KB each) whose pipelines interleave.
#ifndef CB_RT_HT
#define CB_RT_HT 2
#endif
constexpr int kRtHT = CB_RT_HT;
constexpr int kRtCtaHyps = kRtHT * kRtTileM;  // hypotheses per scoring CTA

__device__ __forceinline__ __nv_bfloat16 rt_bf16(double x) { return __float2bfloat16_rn((float)x); }
__device__ __forceinline__ void rt_split2(double x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
    hi = rt_bf16(x);
    lo = rt_bf16(x - (double)__bfloat162float(hi));
}
__device__ __forceinline__ void rt_split3(double x, __nv_bfloat16& a, __nv_bfloat16& b, __nv_bfloat16& c) {
    a = rt_bf16(x);
    double r = x - (double)__bfloat162float(a);
    b = rt_bf16(r);
    c = rt_bf16(r - (double)__bfloat162float(b));
}

// One hypothesis' A row (weights) + its rounding band, written by the fit kernel.
// T12 = [R | t] of the RAW coordinates; cs / cq = the correspondence means the features use.
__device__ inline void ransac_tc_write_weights(const double* T12, const double* cs, const double* cq, double thr2,
                                               double smaxc, double tmaxc, int j, __nv_bfloat16* __restrict__ Apack,
                                               float* __restrict__ band) {
    double tp[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        tp[i] = T12[4 * i + 3] + T12[4 * i] * cs[0] + T12[4 * i + 1] * cs[1] + T12[4 * i + 2] * cs[2] - cq[i];
    double w[kRtTerms];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c) w[3 * i + c] = -2.0 * T12[4 * i + c];
#pragma unroll
    for (int c = 0; c < 3; ++c) w[9 + c] = 2.0 * (T12[c] * tp[0] + T12[4 + c] * tp[1] + T12[8 + c] * tp[2]);
#pragma unroll
    for (int i = 0; i < 3; ++i) w[12 + i] = -2.0 * tp[i];
    const double tn2 = tp[0] * tp[0] + tp[1] * tp[1] + tp[2] * tp[2];
    const double cc = tn2 - thr2;
    __nv_bfloat16 row[kRtKP];
#pragma unroll
    for (int k = 0; k < kRtTerms; ++k) {
        __nv_bfloat16 wa, wb, wc;
        rt_split3(w[k], wa, wb, wc);
        row[k] = wa;
        row[15 + k] = wa;
        row[30 + k] = wb;
        row[45 + k] = wa;
        row[60 + k] = wc;
        row[75 + k] = wb;
    }
    row[90] = row[91] = row[92] = __float2bfloat16_rn(1.f);
    rt_split3(cc, row[93], row[94], row[95]);
    const int tile = j / kRtTileM, r = j % kRtTileM;
    char* base = reinterpret_cast<char*>(Apack) + (size_t)tile * kRtTileM * kRtKP * 2;
#pragma unroll
    for (int kc = 0; kc < kRtKP / 8; ++kc)
        *reinterpret_cast<uint4*>(base + tc::tile_offset(kRtTileM, r, kc)) = *reinterpret_cast<const uint4*>(&row[8 * kc]);
    // rounding band of the tensor-core value of u (DESIGN.md "RANSAC exactness"):
    //   split error   <= 1.1 * 2^-25 * sum_k |w_k||phi_k|  (dropped products wb pc, wc pb, wc pc and the
    //                    2^-27 residues of both three-piece splits),
    //                    sum_k <= 2 sqrt(3) |s||q| + 2|t|(|s| + |q|)
    //   accumulation  <= 12 * 2^-23 * (sum_k + |s|^2 + |q|^2 + |c|): the fp32 accumulation of the six
    //                    K=16 MMAs; measured on B200 at <= 2.6 * 2^-23 * sum|terms| (tests/test_gpu_ransac.py
    //                    asserts a 2x margin on every run)
    //   split3 residue of n and c, |R s| vs |s| of the float64 rotation: 2^-26 * (...) + 1e-12 |s|^2
    const double tn = sqrt(tn2);
    const double bmag = 2.0 * 1.7320508075688773 * smaxc * tmaxc + 2.0 * tn * (smaxc + tmaxc);
    const double mag = bmag + smaxc * smaxc + tmaxc * tmaxc + fabs(cc);
    const double b = 1.1 * 2.98023223876953125e-8 * bmag + 12.0 * 1.1920928955078125e-7 * mag + 1.5e-8 * mag + 1e-12 * smaxc * smaxc;
    band[j] = (float)(b * 1.01) + 1e-30f;
}

// features + maxima of the centred correspondences (launch after ransac_pack_kernel)
int ransac_tc_pack_corr_launch(cb_ctx* ctx, const float4* src4, const float4* tgt4, int m, RansacState* state,
                               __nv_bfloat16* Bpack, cudaStream_t st);
// scoring of one wave; atomically ADDS every hypothesis' inlier count (and, when want_sum, its sum of
// squared inlier distances) into pc/ps [w] (zeroed by the fit kernel)
int ransac_tc_score_launch(cb_ctx* ctx, const __nv_bfloat16* Apack, const __nv_bfloat16* Bpack, const float* band,
                           const double* Tw, const float4* src4, const float4* tgt4, int m, int h0, int w,
                           double thr2, const RansacState* state, int32_t* pc, double* ps, bool want_sum, cudaStream_t st);
