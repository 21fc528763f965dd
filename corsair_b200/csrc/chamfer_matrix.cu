// All-pairs two-direction Chamfer table: compute_dist / chamfer of the reference's
// utils/pc_dist.py:45-99 (every cloud has the same point count P, 2000 there).
// One block per cloud pair: both clouds live in shared memory, each thread owns P/256 points of one
// cloud and scans the other in fp32 difference form; the two directional means are summed.  The
// value for (i,j) and (j,i) comes from the same two sums (a pair whose mirror entry lies in the same
// row block is computed once and written twice), so the table is exactly symmetric.
//
// chamfer_matrix_kernel (P <= 4266): the inputs are raw float32 points, so the fp32 difference-form
// squared distance is already accurate to a few 2^-24 relative — no float64 re-evaluation of the
// chosen neighbour, hence no index tracking: the inner loop is packed-fp32 arithmetic
// (add/mul/fma.f32x2: two query points per instruction, eight per thread against every streamed
// point) and one min per evaluation: ~3.9 instructions per distance instead of ~9.5.
// chamfer_matrix_big_kernel: the earlier index-tracking kernel, kept for 4266 < P <= 6400.
#include "common.cuh"

#include <math.h>

__global__ void __launch_bounds__(256)
chamfer_matrix_big_kernel(const float* __restrict__ pts, int L, int P, int row0, double* __restrict__ out) {
    extern __shared__ float4 cm_smem[];
    float4* A = cm_smem;
    float4* B = cm_smem + P;
    __shared__ double wsum[2][8];
    const int i = row0 + blockIdx.y, j = blockIdx.x;
    double* o = out + (size_t)blockIdx.y * L + j;
    if (i == j) {
        if (threadIdx.x == 0) *o = 200.0;  // eye * 100, then table += table.T (utils/pc_dist.py:66,98)
        return;
    }
    const float* pa = pts + (size_t)i * P * 3;
    const float* pb = pts + (size_t)j * P * 3;
    for (int t = threadIdx.x; t < P; t += 256) {
        A[t] = make_float4(pa[3 * t], pa[3 * t + 1], pa[3 * t + 2], 0.f);
        B[t] = make_float4(pb[3 * t], pb[3 * t + 1], pb[3 * t + 2], 0.f);
    }
    __syncthreads();
    double sums[2] = {0.0, 0.0};
#pragma unroll
    for (int dir = 0; dir < 2; ++dir) {
        const float4* X = dir ? B : A;
        const float4* Y = dir ? A : B;
        for (int t = threadIdx.x; t < P; t += 256) {
            const float4 p = X[t];
            float bd = INFINITY;
            int bj = 0;
#pragma unroll 8
            for (int u = 0; u < P; ++u) {
                const float4 v = Y[u];
                float dx = p.x - v.x, dy = p.y - v.y, dz = p.z - v.z;
                float d = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                if (d < bd) { bd = d; bj = u; }
            }
            const float4 v = Y[bj];
            double dx = (double)p.x - (double)v.x, dy = (double)p.y - (double)v.y, dz = (double)p.z - (double)v.z;
            sums[dir] += sqrt(dx * dx + dy * dy + dz * dz);
        }
    }
#pragma unroll
    for (int dir = 0; dir < 2; ++dir) {
        double s = sums[dir];
        for (int of = 16; of > 0; of >>= 1) s += __shfl_xor_sync(0xffffffffu, s, of);
        if ((threadIdx.x & 31) == 0) wsum[dir][threadIdx.x >> 5] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sa = 0.0, sb = 0.0;
        for (int w = 0; w < 8; ++w) { sa += wsum[0][w]; sb += wsum[1][w]; }
        // order the two directional means by cloud index so (i,j) and (j,i) add identically
        double ma = sa / P, mb = sb / P;
        *o = i < j ? ma + mb : mb + ma;
    }
}


namespace {

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 cm_pk2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void cm_unpk2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 cm_add2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 cm_mul2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 cm_fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

constexpr int kCmThreads = 256;
constexpr int kCmPerThread = 8;  // query points per thread and sweep (four packed pairs)

// mean over the points of X of the distance to the nearest point of Y; Y sits in shared memory as
// three arrays of NEGATED, duplicated coordinates (-v, -v) so that one packed add gives p - v for
// two query points.  Returns this thread's partial sum of distances.
__device__ __forceinline__ double cm_direction(const float* __restrict__ x, int P, const float2* __restrict__ nx,
                                               const float2* __restrict__ ny, const float2* __restrict__ nz) {
    double sum = 0.0;
    for (int base = 0; base < P; base += kCmThreads * kCmPerThread) {
        f32x2 px[kCmPerThread / 2], py[kCmPerThread / 2], pz[kCmPerThread / 2];
        float bd[kCmPerThread];
#pragma unroll
        for (int k = 0; k < kCmPerThread / 2; ++k) {
            const int i0 = min(base + threadIdx.x + (2 * k) * kCmThreads, P - 1);
            const int i1 = min(base + threadIdx.x + (2 * k + 1) * kCmThreads, P - 1);
            px[k] = cm_pk2(x[3 * i0], x[3 * i1]);
            py[k] = cm_pk2(x[3 * i0 + 1], x[3 * i1 + 1]);
            pz[k] = cm_pk2(x[3 * i0 + 2], x[3 * i1 + 2]);
            bd[2 * k] = bd[2 * k + 1] = INFINITY;
        }
#pragma unroll 4
        for (int u = 0; u < P; ++u) {
            const f32x2 vx = *reinterpret_cast<const f32x2*>(&nx[u]);
            const f32x2 vy = *reinterpret_cast<const f32x2*>(&ny[u]);
            const f32x2 vz = *reinterpret_cast<const f32x2*>(&nz[u]);
#pragma unroll
            for (int k = 0; k < kCmPerThread / 2; ++k) {
                const f32x2 dx = cm_add2(px[k], vx), dy = cm_add2(py[k], vy), dz = cm_add2(pz[k], vz);
                const f32x2 d = cm_fma2(dz, dz, cm_fma2(dy, dy, cm_mul2(dx, dx)));
                float d0, d1;
                cm_unpk2(d, d0, d1);
                bd[2 * k] = fminf(bd[2 * k], d0);
                bd[2 * k + 1] = fminf(bd[2 * k + 1], d1);
            }
        }
#pragma unroll
        for (int k = 0; k < kCmPerThread; ++k)
            if (base + threadIdx.x + k * kCmThreads < P) sum += sqrt((double)bd[k]);
    }
    return sum;
}

__global__ void __launch_bounds__(kCmThreads)
chamfer_matrix_kernel(const float* __restrict__ pts, int L, int P, int row0, int row1, int half,
                      double* __restrict__ out) {
    extern __shared__ float2 cm2_smem[];
    __shared__ double wsum[2][8];
    const int i = row0 + blockIdx.y, j = blockIdx.x;
    double* o = out + (size_t)blockIdx.y * L + j;
    if (i == j) {
        if (threadIdx.x == 0) *o = half ? 0.0 : 200.0;  // eye * 100, then table += table.T (utils/pc_dist.py:66,98)
        return;
    }
    bool mirror_here = j >= row0 && j < row1;  // entry (j, i) belongs to this call as well
    if (half) {
        // circulant half: row i owns the pairs (i, i+1 .. i+floor((L-1)/2)) mod L, plus (i, i+L/2) for
        // i < L/2 when L is even: every unordered pair exactly once, the same count in every row
        const int dlt = j - i < 0 ? j - i + L : j - i;
        const bool mine = dlt <= (L - 1) / 2 || (2 * dlt == L && i < L / 2);
        if (!mine) {
            if (threadIdx.x == 0) *o = 0.0;
            return;
        }
        mirror_here = false;
    } else if (mirror_here && j < i) {
        return;  // ... and block (j, i) writes both
    }
    const float* pa = pts + (size_t)i * P * 3;
    const float* pb = pts + (size_t)j * P * 3;
    float2* an = cm2_smem;           // [3][P] negated duplicated coordinates of cloud i
    float2* bn = cm2_smem + 3 * P;   // ... of cloud j
    for (int t = threadIdx.x; t < P; t += kCmThreads) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float va = -pa[3 * t + a], vb = -pb[3 * t + a];
            an[a * P + t] = make_float2(va, va);
            bn[a * P + t] = make_float2(vb, vb);
        }
    }
    __syncthreads();
    double sums[2];
    sums[0] = cm_direction(pa, P, bn, bn + P, bn + 2 * P);
    sums[1] = cm_direction(pb, P, an, an + P, an + 2 * P);
#pragma unroll
    for (int dir = 0; dir < 2; ++dir) {
        double s = sums[dir];
        for (int of = 16; of > 0; of >>= 1) s += __shfl_xor_sync(0xffffffffu, s, of);
        if ((threadIdx.x & 31) == 0) wsum[dir][threadIdx.x >> 5] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sa = 0.0, sb = 0.0;
        for (int w = 0; w < 8; ++w) { sa += wsum[0][w]; sb += wsum[1][w]; }
        // order the two directional means by cloud index so (i,j) and (j,i) add identically
        const double ma = sa / P, mb = sb / P;
        const double v = i < j ? ma + mb : mb + ma;
        *o = v;
        if (mirror_here) out[(size_t)(j - row0) * L + i] = v;
    }
}

}  // namespace

static int chamfer_matrix_run(cb_ctx* ctx, const float* pts, int L, int P, int row0, int row1, int half, double* out,
                              void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (!pts || !out || L <= 0 || P <= 0 || row0 < 0 || row1 > L || row0 > row1)
        CB_FAIL(ctx, CB_EINVAL, "cb_chamfer_matrix: bad arguments");
    if (row1 == row0) return CB_OK;
    const size_t smem2 = (size_t)P * 6 * sizeof(float2);
    if (smem2 <= 200 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(chamfer_matrix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
        if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "smem opt-in failed: %s", cudaGetErrorString(e));
        cb_timed_scope ts(ctx, CB_T_CHAMFER_NN, 8.0 * (double)P * P * 2.0 * (double)(row1 - row0) * L, (cudaStream_t)stream);
        chamfer_matrix_kernel<<<dim3(L, row1 - row0), kCmThreads, smem2, (cudaStream_t)stream>>>(pts, L, P, row0, row1, half,
                                                                                                out);
        CB_LAUNCH_CHECK(ctx);
        return CB_OK;
    }
    if (half) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_chamfer_matrix_half: P=%d above 4266", P);
    size_t smem = (size_t)P * 2 * sizeof(float4);
    if (smem > 200 * 1024) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_chamfer_matrix: P=%d exceeds shared memory (max 6400)", P);
    cudaError_t e = cudaFuncSetAttribute(chamfer_matrix_big_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "smem opt-in failed: %s", cudaGetErrorString(e));
    chamfer_matrix_big_kernel<<<dim3(L, row1 - row0), 256, smem, (cudaStream_t)stream>>>(pts, L, P, row0, out);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

extern "C" int cb_chamfer_matrix(cb_ctx* ctx, const float* pts, int L, int P, int row0, int row1, double* out,
                                 void* stream) {
    return chamfer_matrix_run(ctx, pts, L, P, row0, row1, 0, out, stream);
}

extern "C" int cb_chamfer_matrix_half(cb_ctx* ctx, const float* pts, int L, int P, int row0, int row1, double* out,
                                      void* stream) {
    return chamfer_matrix_run(ctx, pts, L, P, row0, row1, 1, out, stream);
}
