// K3: correspondence RANSAC.
// Replaces registration_based_on_corr (utils/eval_pose.py:82-100 of the reference), i.e.
// Open3D's registration_ransac_based_on_correspondence(ransac_n=10, point-to-point without
// scaling, RANSACConvergenceCriteria(100000, 0.999)) on an identity correspondence set
// (SURVEY.md section 8a rows a8-a12).
//
// Stages per wave of hypotheses (all enqueued up front; a wave whose first hypothesis is at or
// beyond the device-side budget exits immediately):
//   fit    one thread per hypothesis: gather the minimal set (explicit, or the counter-based
//          seeded draw), Umeyama without scale in float64 with a one-sided Jacobi 3x3 SVD
//   score  block = 512 hypotheses (four per thread as two register pairs, FFMA2/FADD2 packed
//          fp32) x a slice of the correspondences streamed through shared memory; float64
//          unfused re-evaluation inside the rounding band around max_dist^2 so the inlier COUNT
//          is exactly the float64 one
//   fold   one warp folds the wave in hypothesis order with Open3D's "better" rule
//          (count desc, rmse asc, first wins ties) and its adaptive budget update.
#include "ransac_internal.cuh"

#include <math.h>
#include <stdlib.h>

// -------------------------------------------------------------------------- pack
__global__ void ransac_pack_kernel(const float* __restrict__ src_xyz, const float* __restrict__ tgt_xyz,
                                   const int64_t* __restrict__ src_idx, const int64_t* __restrict__ tgt_idx,
                                   int m, float4* __restrict__ src4, float4* __restrict__ tgt4,
                                   RansacState* __restrict__ state) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    float ns = 0.f, nt = 0.f;
    float4 sv = make_float4(0.f, 0.f, 0.f, 0.f), tv = sv;
    if (i < m) {
        size_t a = src_idx ? (size_t)src_idx[i] : (size_t)i;
        size_t b = tgt_idx ? (size_t)tgt_idx[i] : (size_t)i;
        float4 s = make_float4(src_xyz[3 * a], src_xyz[3 * a + 1], src_xyz[3 * a + 2], 0.f);
        float4 t = make_float4(tgt_xyz[3 * b], tgt_xyz[3 * b + 1], tgt_xyz[3 * b + 2], 0.f);
        src4[i] = s;
        tgt4[i] = t;
        sv = s;
        tv = t;
        ns = sqrtf(s.x * s.x + s.y * s.y + s.z * s.z);
        nt = sqrtf(t.x * t.x + t.y * t.y + t.z * t.z);
    }
    for (int o = 16; o > 0; o >>= 1) {
        ns = fmaxf(ns, __shfl_xor_sync(0xffffffffu, ns, o));
        nt = fmaxf(nt, __shfl_xor_sync(0xffffffffu, nt, o));
    }
    if ((threadIdx.x & 31) == 0) {
        if (ns > 0.f) atomicMax(&state->smax_bits, __float_as_uint(ns));
        if (nt > 0.f) atomicMax(&state->tmax_bits, __float_as_uint(nt));
    }
    // correspondence means for the tensor-core features (any centre is valid: it only tightens
    // the rounding band, the decisions inside the band are re-made exactly)
    double acc[6] = {(double)sv.x, (double)sv.y, (double)sv.z, (double)tv.x, (double)tv.y, (double)tv.z};
    __shared__ double s_acc[8][6];
#pragma unroll
    for (int c = 0; c < 6; ++c) {
        for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
        if ((threadIdx.x & 31) == 0) s_acc[threadIdx.x >> 5][c] = acc[c];
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        double t = 0.0;
        for (int wv = 0; wv < (int)(blockDim.x >> 5); ++wv) t += s_acc[wv][threadIdx.x];
        atomicAdd(&state->csum[threadIdx.x], t);
    }
}

__global__ void ransac_init_kernel(RansacState* state, int h_total) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        state->budget = h_total;
        state->best_h = -1;
        state->best_count = 0;
        state->n_eval = 0;
        state->best_sum = 0.0;
        state->best_rmse = 0.0;
        for (int i = 0; i < 16; ++i) state->T[i] = (i % 5 == 0) ? 1.0 : 0.0;
        state->smax_bits = 0;
        state->tmax_bits = 0;
        for (int i = 0; i < 6; ++i) state->csum[i] = 0.0;
        state->smaxc_bits = 0;
        state->tmaxc_bits = 0;
        state->recount_mismatch = 0;
    }
}

// --------------------------------------------------------------------------- fit
// One-sided (Hestenes) Jacobi SVD of a 3x3 held column-wise in a[col][row]; on exit
// a[:,j] = sigma_j u_j, v[:,j] = v_j.  Same algorithm as the oracle's C restatement.
__device__ void jacobi3(double a[3][3], double v[3][3]) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) v[i][j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 30; ++sweep) {
        int rotated = 0;
#pragma unroll
        for (int p = 0; p < 2; ++p)
#pragma unroll
            for (int q = p + 1; q < 3; ++q) {
                double al = 0, be = 0, ga = 0;
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    al += a[p][r] * a[p][r];
                    be += a[q][r] * a[q][r];
                    ga += a[p][r] * a[q][r];
                }
                if (ga == 0.0 || fabs(ga) <= 1e-300 + 2.3e-16 * sqrt(al * be)) continue;
                rotated = 1;
                double zeta = (be - al) / (2.0 * ga);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    double x = a[p][r], y = a[q][r];
                    a[p][r] = c * x - s * y;
                    a[q][r] = s * x + c * y;
                    x = v[p][r]; y = v[q][r];
                    v[p][r] = c * x - s * y;
                    v[q][r] = s * x + c * y;
                }
            }
        if (!rotated) break;
    }
}

__device__ __forceinline__ void cross3(const double* x, const double* y, double* o) {
    o[0] = x[1] * y[2] - x[2] * y[1];
    o[1] = x[2] * y[0] - x[0] * y[2];
    o[2] = x[0] * y[1] - x[1] * y[0];
}

// Umeyama without scale (Eigen::umeyama as Open3D's TransformationEstimationPointToPoint uses
// it): R = U diag(1,1,det(U)det(V)) V^T of Sigma = 1/n sum (t-mt)(s-ms)^T, t = mt - R ms.
// T12 = [R | t] row-major 3x4.
__device__ void umeyama_fit(const float4* __restrict__ src4, const float4* __restrict__ tgt4, const int* idx,
                            int n, double* T12) {
    double ms[3] = {0, 0, 0}, mt[3] = {0, 0, 0};
    for (int i = 0; i < n; ++i) {
        float4 s = src4[idx[i]], t = tgt4[idx[i]];
        ms[0] += (double)s.x; ms[1] += (double)s.y; ms[2] += (double)s.z;
        mt[0] += (double)t.x; mt[1] += (double)t.y; mt[2] += (double)t.z;
    }
    for (int r = 0; r < 3; ++r) { ms[r] /= n; mt[r] /= n; }
    double a[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < n; ++i) {
        float4 s = src4[idx[i]], t = tgt4[idx[i]];
        double ds[3] = {(double)s.x - ms[0], (double)s.y - ms[1], (double)s.z - ms[2]};
        double dt[3] = {(double)t.x - mt[0], (double)t.y - mt[1], (double)t.z - mt[2]};
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int r = 0; r < 3; ++r) a[c][r] += dt[r] * ds[c];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int r = 0; r < 3; ++r) a[c][r] /= n;
    double v[3][3];
    jacobi3(a, v);
    double nr[3];
    int o[3] = {0, 1, 2};
    for (int c = 0; c < 3; ++c) nr[c] = sqrt(a[c][0] * a[c][0] + a[c][1] * a[c][1] + a[c][2] * a[c][2]);
    for (int i = 0; i < 2; ++i)
        for (int j = i + 1; j < 3; ++j)
            if (nr[o[j]] > nr[o[i]]) { int tmp = o[i]; o[i] = o[j]; o[j] = tmp; }
    double u1[3], u2[3], u3[3], v1[3], v2[3], v3[3], vx[3];
    for (int r = 0; r < 3; ++r) { v1[r] = v[o[0]][r]; v2[r] = v[o[1]][r]; v3[r] = v[o[2]][r]; }
    const double n1 = nr[o[0]], n2 = nr[o[1]];
    if (n1 > 0) { for (int r = 0; r < 3; ++r) u1[r] = a[o[0]][r] / n1; }
    else { u1[0] = 1; u1[1] = 0; u1[2] = 0; }
    if (n2 > 1e-300 && n2 > 1e-14 * n1) {
        for (int r = 0; r < 3; ++r) u2[r] = a[o[1]][r] / n2;
        double d = u1[0] * u2[0] + u1[1] * u2[1] + u1[2] * u2[2];
        for (int r = 0; r < 3; ++r) u2[r] -= d * u1[r];
    } else {
        int k = fabs(u1[0]) < fabs(u1[1]) ? (fabs(u1[0]) < fabs(u1[2]) ? 0 : 2)
                                          : (fabs(u1[1]) < fabs(u1[2]) ? 1 : 2);
        double e[3] = {0, 0, 0};
        e[k] = 1;
        cross3(u1, e, u2);
    }
    const double l2 = sqrt(u2[0] * u2[0] + u2[1] * u2[1] + u2[2] * u2[2]);
    for (int r = 0; r < 3; ++r) u2[r] /= l2;
    cross3(u1, u2, u3);
    cross3(v1, v2, vx);
    const double detv = vx[0] * v3[0] + vx[1] * v3[1] + vx[2] * v3[2];
    const double sg = detv < 0 ? -1.0 : 1.0;
    for (int r = 0; r < 3; ++r) {
        double R0 = u1[r] * v1[0] + u2[r] * v2[0] + sg * u3[r] * v3[0];
        double R1 = u1[r] * v1[1] + u2[r] * v2[1] + sg * u3[r] * v3[1];
        double R2 = u1[r] * v1[2] + u2[r] * v2[2] + sg * u3[r] * v3[2];
        T12[4 * r] = R0; T12[4 * r + 1] = R1; T12[4 * r + 2] = R2;
        T12[4 * r + 3] = mt[r] - (R0 * ms[0] + R1 * ms[1] + R2 * ms[2]);
    }
}

constexpr int kMaxRansacN = 16;

__global__ void __launch_bounds__(128)
ransac_fit_kernel(const float4* __restrict__ src4, const float4* __restrict__ tgt4, int m,
                  const int32_t* __restrict__ minimal, uint64_t seed, int h0, int w, int ransac_n,
                  const RansacState* __restrict__ state, double* __restrict__ Tw, double thr2,
                  __nv_bfloat16* __restrict__ Apack, float* __restrict__ band, int32_t* __restrict__ tot_c,
                  double* __restrict__ tot_s) {
    if (state && h0 >= state->budget) return;
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= w) return;
    const int h = h0 + j;
    if (tot_c) {
        // totals the tensor-core scoring adds into; hypotheses at or beyond the budget are not
        // fitted or scored (the fold never looks at them)
        tot_c[j] = 0;
        tot_s[j] = 0.0;
        if (h >= state->budget) return;
    }
    int idx[kMaxRansacN];
    for (int t = 0; t < ransac_n; ++t)
        idx[t] = minimal ? minimal[(size_t)h * ransac_n + t]
                         : cb_draw_index(seed, (uint64_t)h * (uint64_t)ransac_n + (uint64_t)t, m);
    double T12[12];
    umeyama_fit(src4, tgt4, idx, ransac_n, T12);
    for (int t = 0; t < 12; ++t) Tw[(size_t)j * 12 + t] = T12[t];
    if (Apack) {
        // weight row + rounding band of the tensor-core scoring (ransac_internal.cuh)
        const double inv = 1.0 / (double)m;
        const double cs[3] = {state->csum[0] * inv, state->csum[1] * inv, state->csum[2] * inv};
        const double cq[3] = {state->csum[3] * inv, state->csum[4] * inv, state->csum[5] * inv};
        ransac_tc_write_weights(T12, cs, cq, thr2, (double)__uint_as_float(state->smaxc_bits),
                                (double)__uint_as_float(state->tmaxc_bits), j, Apack, band);
    }
}

__global__ void ransac_draw_kernel(uint64_t seed, int m, size_t total, int32_t* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < total) out[i] = cb_draw_index(seed, (uint64_t)i, m);
}

// ------------------------------------------------------------------------- score
// Packed-fp32 scoring.  Block = 128 threads, each thread owns 2*NP hypotheses held as NP register
// PAIRS (h, h+1) so every arithmetic instruction is an FFMA2/FADD2 over two hypotheses; a slice of
// the correspondences is streamed through shared memory with every coordinate stored twice
// ((x,x),(y,y),...) so one LDS.128 yields two ready-made operand pairs, broadcast to the warp.
// Per (hypothesis, correspondence): e = R s + (t - q), d2 = |e|^2, u = d2 - thr^2;
// inlier mask = (u < 0) as 1.0f/0.0f feeding packed count/sum accumulators.  When |u| falls inside
// the fp32 rounding band the oracle's unfused float64 expression decides instead, so the COUNT is
// exactly the float64 count.  grid (ceil(w / (256*NP)), P).
constexpr int kScoreTile = 256;
typedef unsigned long long f32x2;

__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpk2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

__device__ __noinline__ double ransac_exact_d2(const double* __restrict__ Td, float4 s, float4 q) {
    return ransac_exact_d2_inl(Td, s, q);
}

template <int NP>
__global__ void __launch_bounds__(128)
ransac_score_kernel(const float4* __restrict__ src4, const float4* __restrict__ tgt4, int m, int split,
                    const double* __restrict__ Tw, int h0, int w, double thr2d,
                    const RansacState* __restrict__ state, int32_t* __restrict__ pc, double* __restrict__ ps) {
    if (h0 >= state->budget) return;
    __shared__ float4 sp[kScoreTile * 3];  // per point: (x,x,y,y) (z,z,-qx,-qx) (-qy,-qy,-qz,-qz)
    const int jbase = (blockIdx.x * 128 + threadIdx.x) * 2 * NP;  // first hypothesis of this thread
    f32x2 R[NP][9], Tt[NP][3];
    float tn_max = 0.f;
#pragma unroll
    for (int p = 0; p < NP; ++p) {
        const double* Ta = Tw + (size_t)min(jbase + 2 * p, w - 1) * 12;
        const double* Tb = Tw + (size_t)min(jbase + 2 * p + 1, w - 1) * 12;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
#pragma unroll
            for (int c = 0; c < 3; ++c) R[p][3 * r + c] = pk2((float)Ta[4 * r + c], (float)Tb[4 * r + c]);
            Tt[p][r] = pk2((float)Ta[4 * r + 3], (float)Tb[4 * r + 3]);
        }
        float ta = sqrtf((float)(Ta[3] * Ta[3] + Ta[7] * Ta[7] + Ta[11] * Ta[11]));
        float tb = sqrtf((float)(Tb[3] * Tb[3] + Tb[7] * Tb[7] + Tb[11] * Tb[11]));
        tn_max = fmaxf(tn_max, fmaxf(ta, tb));
    }
    // rounding band of the fp32 evaluation around thr^2 ("RANSAC exactness")
    const float smax = __uint_as_float(state->smax_bits), tmax = __uint_as_float(state->tmax_bits);
    const float errv = 16.f * 5.9604645e-8f * (smax + tmax + tn_max);
    const float thr2 = (float)thr2d;
    const float band = 2.f * sqrtf(thr2) * errv + errv * errv + 8.f * 5.9604645e-8f * thr2;
    const f32x2 nthr2 = pk2(-thr2, -thr2);
    const f32x2 zero2 = pk2(0.f, 0.f);

    const int p0 = blockIdx.y * split, p1 = min(m, p0 + split);
    int cnt[2 * NP];
    double sum[2 * NP];
#pragma unroll
    for (int e = 0; e < 2 * NP; ++e) { cnt[e] = 0; sum[e] = 0.0; }

    for (int base = p0; base < p1; base += kScoreTile) {
        const int nt = min(kScoreTile, p1 - base);
        __syncthreads();
        for (int t = threadIdx.x; t < nt; t += 128) {
            const float4 s = src4[base + t], q = tgt4[base + t];
            sp[3 * t] = make_float4(s.x, s.x, s.y, s.y);
            sp[3 * t + 1] = make_float4(s.z, s.z, -q.x, -q.x);
            sp[3 * t + 2] = make_float4(-q.y, -q.y, -q.z, -q.z);
        }
        __syncthreads();
        f32x2 fc[NP], fs[NP];
#pragma unroll
        for (int p = 0; p < NP; ++p) { fc[p] = zero2; fs[p] = zero2; }
#pragma unroll 2
        for (int t = 0; t < nt; ++t) {
            const ulonglong2 v0 = *reinterpret_cast<const ulonglong2*>(&sp[3 * t]);
            const ulonglong2 v1 = *reinterpret_cast<const ulonglong2*>(&sp[3 * t + 1]);
            const ulonglong2 v2 = *reinterpret_cast<const ulonglong2*>(&sp[3 * t + 2]);
            const f32x2 X = v0.x, Y = v0.y, Z = v1.x, NQX = v1.y, NQY = v2.x, NQZ = v2.y;
            f32x2 d2[NP];
            float mlo[NP], mhi[NP];
            float bmin = INFINITY;
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                f32x2 e0 = fma2(R[p][0], X, fma2(R[p][1], Y, fma2(R[p][2], Z, add2(Tt[p][0], NQX))));
                f32x2 e1 = fma2(R[p][3], X, fma2(R[p][4], Y, fma2(R[p][5], Z, add2(Tt[p][1], NQY))));
                f32x2 e2 = fma2(R[p][6], X, fma2(R[p][7], Y, fma2(R[p][8], Z, add2(Tt[p][2], NQZ))));
                d2[p] = fma2(e2, e2, fma2(e1, e1, mul2(e0, e0)));
                float ulo, uhi;
                unpk2(add2(d2[p], nthr2), ulo, uhi);
                mlo[p] = ulo < 0.f ? 1.f : 0.f;
                mhi[p] = uhi < 0.f ? 1.f : 0.f;
                bmin = fminf(bmin, fminf(fabsf(ulo), fabsf(uhi)));
            }
            if (bmin <= band) {
                // some evaluation of this thread sits inside the band: float64 decides for those
                const float4 s = src4[base + t], q = tgt4[base + t];
#pragma unroll
                for (int p = 0; p < NP; ++p) {
                    float dlo, dhi;
                    unpk2(d2[p], dlo, dhi);
                    if (fabsf(dlo - thr2) <= band) {
                        double dd = ransac_exact_d2(Tw + (size_t)min(jbase + 2 * p, w - 1) * 12, s, q);
                        mlo[p] = dd < thr2d ? 1.f : 0.f;
                        dlo = (float)dd;
                    }
                    if (fabsf(dhi - thr2) <= band) {
                        double dd = ransac_exact_d2(Tw + (size_t)min(jbase + 2 * p + 1, w - 1) * 12, s, q);
                        mhi[p] = dd < thr2d ? 1.f : 0.f;
                        dhi = (float)dd;
                    }
                    d2[p] = pk2(dlo, dhi);
                }
            }
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                const f32x2 mk = pk2(mlo[p], mhi[p]);
                fc[p] = add2(fc[p], mk);
                fs[p] = fma2(mk, d2[p], fs[p]);
            }
        }
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            float a, b2;
            unpk2(fc[p], a, b2);
            cnt[2 * p] += (int)a;
            cnt[2 * p + 1] += (int)b2;
            unpk2(fs[p], a, b2);
            sum[2 * p] += (double)a;
            sum[2 * p + 1] += (double)b2;
        }
    }
#pragma unroll
    for (int e = 0; e < 2 * NP; ++e)
        if (jbase + e < w) {
            pc[(size_t)blockIdx.y * w + jbase + e] = cnt[e];
            ps[(size_t)blockIdx.y * w + jbase + e] = sum[e];
        }
}

// -------------------------------------------------------------------------- fold
__device__ int ransac_budget_from_count(int good, int m, int n, double conf, int cap) {
    if (conf >= 1.0 || good <= 0) return cap;
    double wr = (double)good / (double)m;
    double denom = log(1.0 - pow(wr, (double)n));
    if (denom == 0.0) return cap;
    double k = log(1.0 - conf) / denom;
    return k < (double)cap ? (int)ceil(k) : cap;
}

// key order of Open3D's IsBetterRANSACThan plus "first hypothesis wins exact ties"
__device__ __forceinline__ bool ransac_better(int c, double r, int h, int bc, double br, int bh) {
    return c > bc || (c == bc && (r < br || (r == br && h < bh)));
}

// One block of 1024 threads folds hypotheses h0..h0+w-1 with the semantics of the sequential
// scan: process h in order while h < budget; an improvement (count up, or equal count and lower
// rmse) may shrink the budget, which depends only on the best count so far.  Parallel form:
//   1. totals per hypothesis (sum of the P partials) staged in shared memory
//   2. block max-scan of per-segment maximum counts -> the best count BEFORE each segment
//   3. every thread walks its contiguous segment with the running best count and finds the
//      first h with h >= budget(best count so far); block-min -> where the scan stops
//   4. lexicographic best over [h0, stop) merged with the best of the earlier waves.
constexpr int kFoldThreads = 1024;

__global__ void __launch_bounds__(kFoldThreads)
ransac_fold_kernel(int32_t* __restrict__ pc, double* __restrict__ ps, int P, const double* __restrict__ Tw,
                   int h0, int w, int m, int ransac_n, int adaptive, double conf, RansacState* __restrict__ state,
                   int32_t* __restrict__ counts, double* __restrict__ sums, double* __restrict__ dev_work,
                   const float4* __restrict__ src4, const float4* __restrict__ tgt4, double thr2d) {
    if (h0 >= state->budget) return;
    __shared__ int s_scan[kFoldThreads];
    __shared__ int s_stop;
    __shared__ int s_wc[32], s_wh[32];
    __shared__ double s_wr[32];
    extern __shared__ int s_cnt[];  // [w] total inlier counts of the wave (every later pass reads these)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int budget_in = state->budget;
    const int prev_count = state->best_count;
    if (tid == 0) {
        // evaluations (hypothesis x correspondence) the scoring launch of this wave performed: its CTAs
        // of kRtCtaHyps hypotheses whose first hypothesis lay below the budget it saw (= budget_in: the previous
        // fold finished before that launch started)
        const int scored = min(w, ((min(budget_in, h0 + w) - h0 + kRtCtaHyps - 1) / kRtCtaHyps) * kRtCtaHyps);
        dev_work[CB_T_RANSAC_SCORE] += (double)scored * (double)m;
        dev_work[CB_T_RANSAC_WAVES] += 1.0;
        s_stop = h0 + w;
    }
    // 1. totals into shared memory (independent loads, unrolled so they overlap)
    if (P == 1) {
#pragma unroll 8
        for (int j = tid; j < w; j += kFoldThreads) s_cnt[j] = pc[j];
    } else {
        for (int j = tid; j < w; j += kFoldThreads) {
            int cnt = 0;
            double sm = 0.0;
            for (int p = 0; p < P; ++p) {
                cnt += pc[(size_t)p * w + j];
                sm += ps[(size_t)p * w + j];
            }
            s_cnt[j] = cnt;
            ps[j] = sm;
        }
    }
    __syncthreads();
    // 2. best count before each segment
    const int L = (w + kFoldThreads - 1) / kFoldThreads;
    const int j0 = min(w, tid * L), j1 = min(w, j0 + L);
    int seg_max = 0;
    for (int j = j0; j < j1; ++j) seg_max = max(seg_max, s_cnt[j]);
    s_scan[tid] = seg_max;
    __syncthreads();
    for (int off = 1; off < kFoldThreads; off <<= 1) {
        int v = tid >= off ? s_scan[tid - off] : 0;
        __syncthreads();
        s_scan[tid] = max(s_scan[tid], v);
        __syncthreads();
    }
    int run = max(prev_count, tid > 0 ? s_scan[tid - 1] : 0);
    // 3. first h at or beyond the (shrinking) budget.  The budget is non-increasing in the best
    //    count, so if even the wave's highest count leaves it at or beyond the end of the wave nobody
    //    stops inside (the usual case of a long run): skip the per-thread walks and their pow/log.
    const int wave_max = max(prev_count, s_scan[kFoldThreads - 1]);
    const int b_min = adaptive ? min(budget_in, ransac_budget_from_count(wave_max, m, ransac_n, conf, budget_in)) : budget_in;
    if (b_min < h0 + w && j0 < j1) {
        int b = adaptive ? min(budget_in, ransac_budget_from_count(run, m, ransac_n, conf, budget_in)) : budget_in;
        int stop = -1;
        for (int j = j0; j < j1; ++j) {
            const int h = h0 + j;
            if (h >= b) { stop = h; break; }
            const int c = s_cnt[j];
            if (c > run) {
                run = c;
                if (adaptive) b = min(b, ransac_budget_from_count(run, m, ransac_n, conf, budget_in));
            }
        }
        if (stop < 0 && h0 + j1 >= b && j1 < w) stop = h0 + j1;  // budget reached exactly at the segment end
        if (stop >= 0) atomicMin(&s_stop, stop);
    }
    __syncthreads();
    const int stop = s_stop;
    // 4. the best of [h0, stop): highest count; among hypotheses tied at that count the lowest
    //    rmse, first wins exact ties.  The streamed sums are approximate (fp32 / tensor-core
    //    values), so the sum of every hypothesis tied at the top count is recomputed here with the
    //    oracle's float64 expression: the tie-break, best_sum and best_rmse are exact.
    int cm = -1;
    for (int j = tid; j < w && h0 + j < stop; j += kFoldThreads) {
        const int c = s_cnt[j];
        if (counts) counts[h0 + j] = c;
        if (sums) sums[h0 + j] = ps[j];
        cm = max(cm, c);
    }
    for (int o = 16; o > 0; o >>= 1) cm = max(cm, __shfl_xor_sync(0xffffffffu, cm, o));
    if (lane == 0) s_wc[warp] = cm;
    __syncthreads();
    cm = s_wc[0];
    for (int i = 1; i < 32; ++i) cm = max(cm, s_wc[i]);
    __syncthreads();
    const int prev_best = state->best_count;
    int bj = -1;
    double bsum = 0.0, brm = 0.0;
    if (cm > 0 && cm >= prev_best) {
        const int nrange = min(w, stop - h0);
        int last = -1;
        for (;;) {
            int nxt = 0x7fffffff;
            for (int j = tid; j < nrange; j += kFoldThreads)
                if (j > last && s_cnt[j] == cm) { nxt = j; break; }
            for (int o = 16; o > 0; o >>= 1) nxt = min(nxt, __shfl_xor_sync(0xffffffffu, nxt, o));
            if (lane == 0) s_wh[warp] = nxt;
            __syncthreads();
            nxt = s_wh[0];
            for (int i = 1; i < 32; ++i) nxt = min(nxt, s_wh[i]);
            __syncthreads();
            if (nxt == 0x7fffffff) break;
            last = nxt;
            double Tj[12];
#pragma unroll
            for (int t = 0; t < 12; ++t) Tj[t] = Tw[(size_t)nxt * 12 + t];
            double sm = 0.0;
            int cn = 0;
#pragma unroll 4
            for (int i = tid; i < m; i += kFoldThreads) {
                const double dd = ransac_exact_d2_inl(Tj, src4[i], tgt4[i]);
                if (dd < thr2d) { sm += dd; ++cn; }
            }
            for (int o = 16; o > 0; o >>= 1) {
                sm += __shfl_xor_sync(0xffffffffu, sm, o);
                cn += __shfl_xor_sync(0xffffffffu, cn, o);
            }
            if (lane == 0) { s_wr[warp] = sm; s_wc[warp] = cn; }
            __syncthreads();
            sm = 0.0;
            cn = 0;
            for (int i = 0; i < 32; ++i) { sm += s_wr[i]; cn += s_wc[i]; }
            __syncthreads();
            const double r = sqrt(sm / (double)cm);
            if (tid == 0) {
                if (cn != cm) atomicAdd(&state->recount_mismatch, 1);
                ps[nxt] = sm;
                if (sums) sums[h0 + nxt] = sm;
            }
            if (bj < 0 || r < brm) { bj = nxt; bsum = sm; brm = r; }
        }
    }
    if (tid == 0) {
        // the best of the earlier waves keeps exact ties (lower hypothesis index)
        const bool improve = bj >= 0 && (cm > prev_best || (cm == prev_best && brm < state->best_rmse));
        int best_count = prev_best;
        if (improve) {
            for (int t = 0; t < 12; ++t) state->T[t] = Tw[(size_t)bj * 12 + t];
            state->best_count = cm;
            state->best_rmse = brm;
            state->best_sum = bsum;
            state->best_h = h0 + bj;
            best_count = cm;
        }
        int budget = budget_in;
        if (adaptive) budget = min(budget, ransac_budget_from_count(best_count, m, ransac_n, conf, budget_in));
        state->budget = budget;
        state->n_eval = stop;
    }
}

__global__ void ransac_finish_kernel(const RansacState* __restrict__ state, double* __restrict__ T,
                                     double* __restrict__ stats) {
    int i = threadIdx.x;
    if (i < 16) T[i] = (i >= 12) ? (i == 15 ? 1.0 : 0.0) : state->T[i];
    if (i == 0) {
        stats[0] = state->best_h;
        stats[1] = state->best_count;
        stats[2] = state->best_sum;
        stats[3] = state->n_eval;
    }
}

// -------------------------------------------------------------------------- host
__global__ void ransac_expand_T_kernel(const double* __restrict__ Tw, int h, double* __restrict__ T_all) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= h * 16) return;
    int hh = i / 16, e = i % 16;
    T_all[i] = e < 12 ? Tw[(size_t)hh * 12 + e] : (e == 15 ? 1.0 : 0.0);
}

// Waves of hypotheses: a small first wave (most adaptive runs stop inside it), then waves that give
// every SM one 256-hypothesis CTA of the tensor-core scoring kernel.
static int ransac_wave_first() {
    static const int v = [] {
        const char* e = getenv("CB_RANSAC_WAVE_FIRST");  // tuning knob: hypotheses in the first wave
        int x = e ? atoi(e) : 4096;
        return x < 256 ? 256 : (x > 40960 ? 40960 : x / 256 * 256);
    }();
    return v;
}
static int ransac_wave_max(const cb_ctx* ctx) {
    static const int override_ctas = [] {  // CB_RANSAC_WAVE_CTAS: tuning knob (CTAs of 256 hypotheses per wave)
        const char* e = getenv("CB_RANSAC_WAVE_CTAS");
        return e ? atoi(e) : 0;
    }();
    const int ctas = override_ctas > 0 ? min(override_ctas, 160) : min(ctx->sm_count, 160);
    return max(ransac_wave_first(), ctas * 256);
}
static int ransac_wave_size(const cb_ctx* ctx, int wave_index) {
    return wave_index == 0 ? ransac_wave_first() : ransac_wave_max(ctx);
}

// workspace cb_ransac_impl needs (an upper bound, both scoring variants)
size_t cb_ransac_workspace(const cb_ctx* ctx, int m, int h_total) {
    const size_t wmax = (size_t)ransac_wave_max(ctx);
    const size_t hcap = (size_t)max(h_total, (int)wmax);
    const size_t n_tot = max(hcap, size_t(5) << 18);
    const size_t m_pad = (size_t)cb_div_up(max(m, 1), kRtTileN) * kRtTileN;
    return cb_align(sizeof(RansacState)) + 2 * cb_align((size_t)(m + 1) * 16) + cb_align(hcap * 12 * 8) +
           cb_align(n_tot * 4) + cb_align(n_tot * 8) + cb_align(m_pad * kRtKP * 2) + cb_align((hcap + 256) * kRtKP * 2) +
           cb_align(hcap * 4) + (1 << 16);
}

// cb_ransac without the arena rewind: callable from composite entry points (register.cu)
int cb_ransac_impl(cb_ctx* ctx, const float* src_xyz, const float* tgt_xyz, const int64_t* src_idx,
                   const int64_t* tgt_idx, int m, const int32_t* minimal, uint64_t seed, int h_total, int ransac_n,
                   double max_dist, int adaptive, double confidence, double* T, double* stats, int32_t* counts,
                   double* sums, cudaStream_t st) {
    int rc;
    if (!T || !stats || m < 0 || h_total < 0 || ransac_n > kMaxRansacN)
        CB_FAIL(ctx, CB_EINVAL, "cb_ransac: bad arguments (m=%d h=%d n=%d)", m, h_total, ransac_n);
    const int wmax = ransac_wave_max(ctx);
    const size_t pw_max = size_t(5) << 18;  // entries of the [P, w] partial count / sum arrays
    constexpr int NP = 2;                   // hypothesis pairs per SIMT scoring thread
    const bool use_tc = ctx->ransac_variant != 1 && m < (1 << 24);
    const int m_pad = cb_div_up(max(m, 1), kRtTileN) * kRtTileN;
    // tensor-core path: after the first wave ONE fit launch covers every remaining hypothesis (it
    // fills the machine better than a fit per wave), so the per-hypothesis buffers span the whole run
    const bool merged = use_tc && h_total > ransac_wave_first() && h_total <= (1 << 18);
    const size_t hcap = merged ? (size_t)h_total : (size_t)wmax;
    const size_t n_tot = use_tc ? hcap : pw_max;
    size_t need = cb_align(sizeof(RansacState)) + 2 * cb_align((size_t)(m + 1) * 16) +
                  cb_align(hcap * 12 * 8) + cb_align(n_tot * 4) + cb_align(n_tot * 8);
    if (use_tc)
        need += cb_align((size_t)m_pad * kRtKP * 2) + cb_align((hcap + 256) * kRtKP * 2) + cb_align(hcap * 4);
    if ((rc = cb_reserve(ctx, need))) return rc;
    RansacState* state = cb_alloc_t<RansacState>(ctx, 1);
    if (!state) return CB_ENOMEM;
    ransac_init_kernel<<<1, 32, 0, st>>>(state, h_total);
    CB_LAUNCH_CHECK(ctx);
    // Open3D returns the default (identity, fitness 0) result for degenerate requests
    const bool degenerate = ransac_n < 3 || m < ransac_n || max_dist <= 0.0 || h_total == 0;
    if (!degenerate) {
        if (!src_xyz || !tgt_xyz) CB_FAIL(ctx, CB_EINVAL, "cb_ransac: null point arrays");
        float4* src4 = cb_alloc_t<float4>(ctx, m);
        float4* tgt4 = cb_alloc_t<float4>(ctx, m);
        double* Tw = cb_alloc_t<double>(ctx, hcap * 12);
        int32_t* pc = cb_alloc_t<int32_t>(ctx, n_tot);
        double* ps = cb_alloc_t<double>(ctx, n_tot);
        if (!src4 || !tgt4 || !Tw || !pc || !ps) return CB_ENOMEM;
        __nv_bfloat16 *Apack = nullptr, *Bpack = nullptr;
        float* band = nullptr;
        if (use_tc) {
            Bpack = cb_alloc_t<__nv_bfloat16>(ctx, (size_t)m_pad * kRtKP);
            Apack = cb_alloc_t<__nv_bfloat16>(ctx, (hcap + 256) * kRtKP);
            band = cb_alloc_t<float>(ctx, hcap);
            if (!Apack || !Bpack || !band) return CB_ENOMEM;
        }
        ransac_pack_kernel<<<cb_div_up(m, 256), 256, 0, st>>>(src_xyz, tgt_xyz, src_idx, tgt_idx, m, src4, tgt4,
                                                              state);
        CB_LAUNCH_CHECK(ctx);
        if (use_tc && (rc = ransac_tc_pack_corr_launch(ctx, src4, tgt4, m, state, Bpack, st))) return rc;
        const double thr2 = max_dist * max_dist;
        CB_CUDA(ctx, cudaFuncSetAttribute(ransac_fold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(wmax * sizeof(int))));
        int h0 = 0;
        for (int wave = 0; h0 < h_total; ++wave) {
            const int w = min(ransac_wave_size(ctx, wave), h_total - h0);
            const size_t off = merged ? (size_t)h0 : 0;  // this wave's slice of the per-hypothesis buffers
            double* Tw_w = Tw + off * 12;
            int32_t* pc_w = pc + off;
            double* ps_w = ps + off;
            __nv_bfloat16* Apack_w = Apack ? Apack + off * kRtKP : nullptr;
            float* band_w = band ? band + off : nullptr;
            if (!merged || wave <= 1) {
                const int fit_w = merged && wave == 1 ? h_total - h0 : w;
                cb_timed_scope ts(ctx, CB_T_RANSAC_FIT, 0.0, st);
                ransac_fit_kernel<<<cb_div_up(fit_w, 128), 128, 0, st>>>(src4, tgt4, m, minimal, seed, h0, fit_w, ransac_n,
                                                                         state, Tw_w, thr2, Apack_w, band_w,
                                                                         use_tc ? pc_w : nullptr, use_tc ? ps_w : nullptr);
                CB_LAUNCH_CHECK(ctx);
            }
            int P = 1;
            if (use_tc) {
                cb_timed_scope ts(ctx, CB_T_RANSAC_SCORE, 0.0, st);
                if ((rc = ransac_tc_score_launch(ctx, Apack_w, Bpack, band_w, Tw_w, src4, tgt4, m, h0, w, thr2, state, pc_w, ps_w,
                                                 sums != nullptr, st)))
                    return rc;
            } else {
                const int groups = cb_div_up(w, 128 * 2 * NP);
                P = cb_div_up(8 * ctx->sm_count, groups);  // ~2 waves at 4 resident blocks per SM
                P = max(1, min(min(P, (int)(pw_max / (size_t)w)), cb_div_up(m, kScoreTile)));
                int split = cb_div_up(cb_div_up(m, P), kScoreTile) * kScoreTile;
                P = cb_div_up(m, split);
                cb_timed_scope ts(ctx, CB_T_RANSAC_SCORE, 0.0, st);
                ransac_score_kernel<NP><<<dim3(groups, P), 128, 0, st>>>(src4, tgt4, m, split, Tw_w, h0, w, thr2, state,
                                                                     pc_w, ps_w);
                CB_LAUNCH_CHECK(ctx);
            }
            ransac_fold_kernel<<<1, kFoldThreads, (size_t)w * sizeof(int), st>>>(pc_w, ps_w, P, Tw_w, h0, w, m, ransac_n, adaptive, confidence, state,
                                                 counts, sums, ctx->dev_work, src4, tgt4, thr2);
            CB_LAUNCH_CHECK(ctx);
            h0 += w;
        }
    }
    ransac_finish_kernel<<<1, 32, 0, st>>>(state, T, stats);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

extern "C" int cb_ransac(cb_ctx* ctx, const float* src_xyz, int n_src, const float* tgt_xyz, int n_tgt,
                         const int64_t* src_idx, const int64_t* tgt_idx, int m, const int32_t* minimal,
                         uint64_t seed, int h_total, int ransac_n, double max_dist, int adaptive,
                         double confidence, double* T, double* stats, int32_t* counts, double* sums,
                         void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    (void)n_src;
    (void)n_tgt;
    return cb_ransac_impl(ctx, src_xyz, tgt_xyz, src_idx, tgt_idx, m, minimal, seed, h_total, ransac_n, max_dist,
                          adaptive, confidence, T, stats, counts, sums, (cudaStream_t)stream);
}

extern "C" int cb_set_ransac_variant(cb_ctx* ctx, int variant) {
    if (!ctx || variant < 0 || variant > 2) return CB_EINVAL;
    ctx->ransac_variant = variant;
    return CB_OK;
}

extern "C" const char* cb_ransac_variant(const cb_ctx* ctx) {
    return (ctx && ctx->ransac_variant == 1) ? "simt" : "tcgen05";
}

extern "C" int cb_ransac_gemm_k(void) { return kRtKP; }

extern "C" int cb_ransac_debug_dump(cb_ctx* ctx, float* u_out, int ld) {
    if (!ctx || (u_out && ld <= 0)) return CB_EINVAL;
    ctx->ransac_dump = u_out;
    ctx->ransac_dump_ld = ld;
    return CB_OK;
}

extern "C" int cb_draw_minimal_sets(cb_ctx* ctx, uint64_t seed, int m, int h_total, int ransac_n, int32_t* out,
                                    void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (!out || m <= 0 || h_total < 0 || ransac_n <= 0) CB_FAIL(ctx, CB_EINVAL, "cb_draw_minimal_sets: bad arguments");
    size_t total = (size_t)h_total * ransac_n;
    if (total == 0) return CB_OK;
    ransac_draw_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(seed, m, total, out);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

extern "C" int cb_fit_minimal(cb_ctx* ctx, const float* src_xyz, const float* tgt_xyz, int m,
                              const int32_t* minimal, int h_total, int ransac_n, double* T_all, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (!src_xyz || !tgt_xyz || !minimal || !T_all || m <= 0 || h_total < 0 || ransac_n < 1 || ransac_n > kMaxRansacN)
        CB_FAIL(ctx, CB_EINVAL, "cb_fit_minimal: bad arguments");
    if (h_total == 0) return CB_OK;
    if ((rc = cb_reserve(ctx, 2 * cb_align((size_t)m * 16) + cb_align((size_t)h_total * 96) + 4096))) return rc;
    RansacState* state = cb_alloc_t<RansacState>(ctx, 1);
    float4* src4 = cb_alloc_t<float4>(ctx, m);
    float4* tgt4 = cb_alloc_t<float4>(ctx, m);
    double* Tw = cb_alloc_t<double>(ctx, (size_t)h_total * 12);
    if (!state || !src4 || !tgt4 || !Tw) return CB_ENOMEM;
    ransac_init_kernel<<<1, 32, 0, st>>>(state, h_total);
    CB_LAUNCH_CHECK(ctx);
    ransac_pack_kernel<<<cb_div_up(m, 256), 256, 0, st>>>(src_xyz, tgt_xyz, nullptr, nullptr, m, src4, tgt4, state);
    CB_LAUNCH_CHECK(ctx);
    ransac_fit_kernel<<<cb_div_up(h_total, 128), 128, 0, st>>>(src4, tgt4, m, minimal, 0, 0, h_total, ransac_n,
                                                               nullptr, Tw, 0.0, nullptr, nullptr, nullptr, nullptr);
    CB_LAUNCH_CHECK(ctx);
    ransac_expand_T_kernel<<<cb_div_up(h_total * 16, 256), 256, 0, st>>>(Tw, h_total, T_all);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}
