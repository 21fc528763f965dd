// All-pairs two-direction Chamfer table: compute_dist / chamfer of the reference's
// utils/pc_dist.py:45-99 (every cloud has the same point count P, 2000 there).
// One block per (row cloud i, column cloud j != i): both clouds live in shared memory as float4,
// each thread owns P/256 points of one cloud, scans the other cloud (fp32 difference form) and
// re-evaluates the chosen neighbour in float64; the two directional means are summed.  The value
// for (i,j) and (j,i) is computed from the same two sums, so the table is exactly symmetric.
#include "common.cuh"

#include <math.h>

__global__ void __launch_bounds__(256)
chamfer_matrix_kernel(const float* __restrict__ pts, int L, int P, int row0, double* __restrict__ out) {
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

extern "C" int cb_chamfer_matrix(cb_ctx* ctx, const float* pts, int L, int P, int row0, int row1, double* out,
                                 void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (!pts || !out || L <= 0 || P <= 0 || row0 < 0 || row1 > L || row0 > row1)
        CB_FAIL(ctx, CB_EINVAL, "cb_chamfer_matrix: bad arguments");
    if (row1 == row0) return CB_OK;
    size_t smem = (size_t)P * 2 * sizeof(float4);
    if (smem > 200 * 1024) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_chamfer_matrix: P=%d exceeds shared memory (max 6400)", P);
    cudaError_t e = cudaFuncSetAttribute(chamfer_matrix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "smem opt-in failed: %s", cudaGetErrorString(e));
    chamfer_matrix_kernel<<<dim3(L, row1 - row0), 256, smem, (cudaStream_t)stream>>>(pts, L, P, row0, out);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}
