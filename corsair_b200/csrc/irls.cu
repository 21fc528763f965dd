// K5: est_quad_linear_robust (utils/_transform_estimation.py:110-137 of the reference):
// `iters` IRLS rounds of the small-angle linearised weighted point-to-point solve
//   rows  w [0, z,-y, 1,0,0 | qx-x],  w [-z,0, x, 0,1,0 | qy-y],  w [ y,-x,0, 0,0,1 | qz-z]
//   x = (A^T A)^-1 A^T b;  T_i = [Rz(x2) Ry(x1) Rx(x0) | x3..5];  p <- T_i p;
//   w = par / (|p - q| + par), par = 1 halved at i = 5, 10, 15;  T = T_i T.
// One block does the whole problem: per round a fused float64 reduction of the 21 unique
// A^T A entries and the 6 A^T b entries, a 6x6 solve by one thread, and an in-place update.
#include "common.cuh"

#include <math.h>

constexpr int kIrlsThreads = 1024;
constexpr int kIrlsTerms = 27;  // 21 upper-triangular + 6 rhs

__device__ __forceinline__ void irls_accumulate(double* acc, float w, float x, float y, float z, float bx,
                                                float by, float bz) {
    // A^T A = sum w^2 (r1 r1^T + r2 r2^T + r3 r3^T); index map (i<=j): see irls_unpack
    const double w2 = (double)w * (double)w;
    const double X = x, Y = y, Z = z;
    acc[0] += w2 * (Z * Z + Y * Y);   // (0,0)
    acc[1] += w2 * (-X * Y);          // (0,1)
    acc[2] += w2 * (-X * Z);          // (0,2)
    acc[4] += w2 * (-Z);              // (0,4)
    acc[5] += w2 * (Y);               // (0,5)
    acc[6] += w2 * (Z * Z + X * X);   // (1,1)
    acc[7] += w2 * (-Y * Z);          // (1,2)
    acc[8] += w2 * (Z);               // (1,3)
    acc[10] += w2 * (-X);             // (1,5)
    acc[11] += w2 * (Y * Y + X * X);  // (2,2)
    acc[12] += w2 * (-Y);             // (2,3)
    acc[13] += w2 * (X);              // (2,4)
    acc[15] += w2;                    // (3,3)
    acc[18] += w2;                    // (4,4)
    acc[20] += w2;                    // (5,5)
    const double BX = bx, BY = by, BZ = bz;
    acc[21] += w2 * (-Z * BY + Y * BZ);
    acc[22] += w2 * (Z * BX - X * BZ);
    acc[23] += w2 * (-Y * BX + X * BY);
    acc[24] += w2 * BX;
    acc[25] += w2 * BY;
    acc[26] += w2 * BZ;
}

// Gaussian elimination with partial pivoting on the 6x6 normal equations
__device__ void irls_solve6(const double* acc, double* x) {
    double M[6][7];
    int t = 0;
    for (int i = 0; i < 6; ++i)
        for (int j = i; j < 6; ++j) { M[i][j] = acc[t]; M[j][i] = acc[t]; ++t; }
    for (int i = 0; i < 6; ++i) M[i][6] = acc[21 + i];
    for (int c = 0; c < 6; ++c) {
        int piv = c;
        for (int r = c + 1; r < 6; ++r)
            if (fabs(M[r][c]) > fabs(M[piv][c])) piv = r;
        if (piv != c)
            for (int j = 0; j < 7; ++j) { double tmp = M[c][j]; M[c][j] = M[piv][j]; M[piv][j] = tmp; }
        const double inv = 1.0 / M[c][c];
        for (int r = c + 1; r < 6; ++r) {
            const double f = M[r][c] * inv;
            for (int j = c; j < 7; ++j) M[r][j] -= f * M[c][j];
        }
    }
    for (int i = 5; i >= 0; --i) {
        double s = M[i][6];
        for (int j = i + 1; j < 6; ++j) s -= M[i][j] * x[j];
        x[i] = s / M[i][i];
    }
}

__global__ void __launch_bounds__(kIrlsThreads)
irls_kernel(const float* __restrict__ p0, const float* __restrict__ p1, const float* __restrict__ w_in, int n,
            int iters, float* __restrict__ cur, float* __restrict__ wts, float* __restrict__ T_out) {
    __shared__ double red[32][kIrlsTerms];
    __shared__ float Ti[12];
    __shared__ float Tacc[16];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < n; i += kIrlsThreads) {
        cur[3 * i] = p0[3 * i]; cur[3 * i + 1] = p0[3 * i + 1]; cur[3 * i + 2] = p0[3 * i + 2];
        wts[i] = w_in ? w_in[i] : 1.f;
    }
    if (tid < 16) Tacc[tid] = (tid % 5 == 0) ? 1.f : 0.f;
    __syncthreads();
    float par = 1.f;
    for (int it = 0; it < iters; ++it) {
        if (it > 0 && it % 5 == 0) par *= 0.5f;
        double acc[kIrlsTerms];
#pragma unroll
        for (int t = 0; t < kIrlsTerms; ++t) acc[t] = 0.0;
        for (int i = tid; i < n; i += kIrlsThreads) {
            float x = cur[3 * i], y = cur[3 * i + 1], z = cur[3 * i + 2];
            irls_accumulate(acc, wts[i], x, y, z, p1[3 * i] - x, p1[3 * i + 1] - y, p1[3 * i + 2] - z);
        }
#pragma unroll
        for (int t = 0; t < kIrlsTerms; ++t) {
            double v = acc[t];
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) red[warp][t] = v;
        }
        __syncthreads();
        if (tid == 0) {
            double tot[kIrlsTerms];
            for (int t = 0; t < kIrlsTerms; ++t) {
                double v = 0.0;
                for (int wv = 0; wv < kIrlsThreads / 32; ++wv) v += red[wv][t];
                tot[t] = v;
            }
            double x[6];
            irls_solve6(tot, x);
            // get_trans: R = Rz(x2) Ry(x1) Rx(x0) in float32 like the reference
            const float c0 = cosf((float)x[0]), s0 = sinf((float)x[0]);
            const float c1 = cosf((float)x[1]), s1 = sinf((float)x[1]);
            const float c2 = cosf((float)x[2]), s2 = sinf((float)x[2]);
            Ti[0] = c2 * c1; Ti[1] = c2 * s1 * s0 - s2 * c0; Ti[2] = c2 * s1 * c0 + s2 * s0; Ti[3] = (float)x[3];
            Ti[4] = s2 * c1; Ti[5] = s2 * s1 * s0 + c2 * c0; Ti[6] = s2 * s1 * c0 - c2 * s0; Ti[7] = (float)x[4];
            Ti[8] = -s1;     Ti[9] = c1 * s0;                Ti[10] = c1 * c0;               Ti[11] = (float)x[5];
            // trans = T_i @ trans
            float nt[12];
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 4; ++c) {
                    float v = Ti[4 * r] * Tacc[c] + Ti[4 * r + 1] * Tacc[4 + c] + Ti[4 * r + 2] * Tacc[8 + c];
                    if (c == 3) v += Ti[4 * r + 3];
                    nt[4 * r + c] = v;
                }
            for (int e = 0; e < 12; ++e) Tacc[e] = nt[e];
        }
        __syncthreads();
        for (int i = tid; i < n; i += kIrlsThreads) {
            float x = cur[3 * i], y = cur[3 * i + 1], z = cur[3 * i + 2];
            float nx = Ti[0] * x + Ti[1] * y + Ti[2] * z + Ti[3];
            float ny = Ti[4] * x + Ti[5] * y + Ti[6] * z + Ti[7];
            float nz = Ti[8] * x + Ti[9] * y + Ti[10] * z + Ti[11];
            cur[3 * i] = nx; cur[3 * i + 1] = ny; cur[3 * i + 2] = nz;
            float dx = nx - p1[3 * i], dy = ny - p1[3 * i + 1], dz = nz - p1[3 * i + 2];
            wts[i] = par / (sqrtf(dx * dx + dy * dy + dz * dz) + par);
        }
        __syncthreads();
    }
    if (tid < 16) T_out[tid] = Tacc[tid];
}

extern "C" int cb_irls_pose(cb_ctx* ctx, const float* p0, const float* p1, const float* w, int n, int iters,
                            float* T, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (!p0 || !p1 || !T || n <= 0 || iters < 0) CB_FAIL(ctx, CB_EINVAL, "cb_irls_pose: bad arguments");
    if ((rc = cb_reserve(ctx, (size_t)n * 16 + 4096))) return rc;
    float* cur = cb_alloc_t<float>(ctx, (size_t)n * 3);
    float* wts = cb_alloc_t<float>(ctx, n);
    if (!cur || !wts) return CB_ENOMEM;
    irls_kernel<<<1, kIrlsThreads, 0, (cudaStream_t)stream>>>(p0, p1, w, n, iters, cur, wts, T);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}
