// All-pairs two-direction Chamfer table: compute_dist / chamfer of the reference's
// utils/pc_dist.py:45-99 (every cloud has the same point count P, 2000 there).
// One block per cloud pair: both clouds live in shared memory, each thread owns P/256 points of one
// cloud and scans the other in fp32 difference form; the two directional means are summed.  The
// value for (i,j) and (j,i) comes from the same two sums (a pair whose mirror entry lies in the same
// row block is computed once and written twice), so the table is exactly symmetric.
//
// chamfer_matrix_kernel (P <= 7000): the inputs are raw float32 points, so the fp32 difference-form
// squared distance is already accurate to a few 2^-24 relative — no float64 re-evaluation of the
// chosen neighbour, hence no index tracking: the inner loop is packed-fp32 arithmetic
// (add/mul/fma.f32x2: two query points per instruction, eight per thread against every streamed
// point) and one min per evaluation.
// ONE distance tile feeds both directions, like the reference's single [m, n] matrix with min(0) and
// min(1) (utils/pc_dist.py:45-52): the points of cloud i sit in registers (their minima too), cloud j
// is streamed from shared memory, and the minimum of a streamed point over all register points is a
// per-thread min of its eight distances, one warp REDUX.MIN on the float bits (non-negative floats
// order like unsigned integers) and one shared-memory atomicMin per warp: P^2 evaluations per pair
// instead of the 2 P^2 of rounds 1-2 (two sweeps, kept as chamfer_matrix_two_sweeps_kernel behind the
// environment switch CORSAIR_B200_CHAMFER_TWO_SWEEPS=1: the values and the order of every sum are
// the same, so the two kernels agree bit for bit and a test checks that).
// chamfer_matrix_big_kernel: the earlier index-tracking kernel, kept for 4266 < P <= 6400.
#include "common.cuh"

#include <math.h>
#include <stdlib.h>

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

// build knobs (measured variants: see DESIGN.md K4b)
#ifndef CM_THREADS
#define CM_THREADS 256
#endif
#ifndef CM_PER_THREAD
#define CM_PER_THREAD 8
#endif
#ifndef CM_UNROLL
#define CM_UNROLL 2
#endif
#ifndef CM_MINBLOCKS
#define CM_MINBLOCKS 3
#endif
#define CM_STR2(x) #x
#define CM_STR(x) CM_STR2(x)
constexpr int kCmThreads = CM_THREADS;
constexpr int kCmPerThread = CM_PER_THREAD;  // query points per thread and sweep (packed pairs)
constexpr int kCmWarps = kCmThreads / 32;
constexpr int kCmUnroll = CM_UNROLL;

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
chamfer_matrix_two_sweeps_kernel(const float* __restrict__ pts, int L, int P, int row0, int row1, int half,
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

// which (i, j) entries this call computes: shared by both table kernels.  Returns false when the block
// has nothing to compute (it has then written what the entry needs).
__device__ __forceinline__ bool cm_block_role(int i, int j, int L, int row0, int row1, int half, double* o,
                                              bool& mirror_here) {
    if (i == j) {
        if (threadIdx.x == 0) *o = half ? 0.0 : 200.0;  // eye * 100, then table += table.T (utils/pc_dist.py:66,98)
        return false;
    }
    mirror_here = j >= row0 && j < row1;  // entry (j, i) belongs to this call as well
    if (half) {
        const int dlt = j - i < 0 ? j - i + L : j - i;
        const bool mine = dlt <= (L - 1) / 2 || (2 * dlt == L && i < L / 2);
        if (!mine) {
            if (threadIdx.x == 0) *o = 0.0;
            return false;
        }
        mirror_here = false;
    } else if (mirror_here && j < i) {
        return false;  // ... and block (j, i) writes both
    }
    return true;
}

__global__ void __launch_bounds__(kCmThreads, CM_MINBLOCKS)  // 28 P bytes of shared memory: three blocks per SM at P = 2048
chamfer_matrix_kernel(const float* __restrict__ pts, int L, int P, int row0, int row1, int half,
                      double* __restrict__ out) {
    extern __shared__ float2 cm2_smem[];
    __shared__ double wsum[2][kCmWarps];
    const int i = row0 + blockIdx.y, j = blockIdx.x;
    double* o = out + (size_t)blockIdx.y * L + j;
    bool mirror_here;
    if (!cm_block_role(i, j, L, row0, row1, half, o, mirror_here)) return;
    const float* pa = pts + (size_t)i * P * 3;
    const float* pb = pts + (size_t)j * P * 3;
    // [3][P32] negated duplicated coordinates of cloud j, P rounded up to whole groups of 32 streamed points
    // (the filler repeats the last point: a duplicate changes no minimum of cloud i, and its own minimum is
    // never read)
    const int P32 = (P + 31) & ~31;
    float2* bnx = cm2_smem;
    float2* bny = bnx + P32;
    float2* bnz = bny + P32;
    unsigned* bmin = reinterpret_cast<unsigned*>(bnz + P32);  // [P32] float bits: min over cloud i of d^2 to point u of cloud j
    for (int t = threadIdx.x; t < P32; t += kCmThreads) {
        const int tt = min(t, P - 1);
        const float vx = -pb[3 * tt], vy = -pb[3 * tt + 1], vz = -pb[3 * tt + 2];
        bnx[t] = make_float2(vx, vx);
        bny[t] = make_float2(vy, vy);
        bnz[t] = make_float2(vz, vz);
        bmin[t] = 0x7f800000u;  // +inf
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    double sum_a = 0.0;
    for (int base = 0; base < P; base += kCmThreads * kCmPerThread) {
        f32x2 px[kCmPerThread / 2], py[kCmPerThread / 2], pz[kCmPerThread / 2];
        float bd[kCmPerThread];
#pragma unroll
        for (int k = 0; k < kCmPerThread / 2; ++k) {
            // rows beyond P repeat the last point as well
            const int i0 = min(base + threadIdx.x + (2 * k) * kCmThreads, P - 1);
            const int i1 = min(base + threadIdx.x + (2 * k + 1) * kCmThreads, P - 1);
            px[k] = cm_pk2(pa[3 * i0], pa[3 * i1]);
            py[k] = cm_pk2(pa[3 * i0 + 1], pa[3 * i1 + 1]);
            pz[k] = cm_pk2(pa[3 * i0 + 2], pa[3 * i1 + 2]);
            bd[2 * k] = bd[2 * k + 1] = INFINITY;
        }
        for (int ub = 0; ub < P32; ub += 32) {
            // lane l keeps the warp's minimum for streamed point ub + l; one conflict-free atomicMin per 32 points
            unsigned mine = 0x7f800000u;
#pragma unroll kCmUnroll
            for (int t = 0; t < 32; t += 2) {
                // two streamed points per step: one 16-byte shared-memory read per coordinate, and the register
                // points' minima take both distances in one three-input minimum
                const float4 vx4 = *reinterpret_cast<const float4*>(&bnx[ub + t]);
                const float4 vy4 = *reinterpret_cast<const float4*>(&bny[ub + t]);
                const float4 vz4 = *reinterpret_cast<const float4*>(&bnz[ub + t]);
                const f32x2 vx0 = cm_pk2(vx4.x, vx4.y), vx1 = cm_pk2(vx4.z, vx4.w);
                const f32x2 vy0 = cm_pk2(vy4.x, vy4.y), vy1 = cm_pk2(vy4.z, vy4.w);
                const f32x2 vz0 = cm_pk2(vz4.x, vz4.y), vz1 = cm_pk2(vz4.z, vz4.w);
                float d0[kCmPerThread], d1[kCmPerThread];
#pragma unroll
                for (int k = 0; k < kCmPerThread / 2; ++k) {
                    const f32x2 ax = cm_add2(px[k], vx0), ay = cm_add2(py[k], vy0), az = cm_add2(pz[k], vz0);
                    const f32x2 bx = cm_add2(px[k], vx1), by = cm_add2(py[k], vy1), bz = cm_add2(pz[k], vz1);
                    cm_unpk2(cm_fma2(az, az, cm_fma2(ay, ay, cm_mul2(ax, ax))), d0[2 * k], d0[2 * k + 1]);
                    cm_unpk2(cm_fma2(bz, bz, cm_fma2(by, by, cm_mul2(bx, bx))), d1[2 * k], d1[2 * k + 1]);
                    bd[2 * k] = fminf(fminf(bd[2 * k], d0[2 * k]), d1[2 * k]);
                    bd[2 * k + 1] = fminf(fminf(bd[2 * k + 1], d0[2 * k + 1]), d1[2 * k + 1]);
                }
                // each streamed point's minimum over this warp's 256 register points
                float m0 = fminf(fminf(fminf(d0[0], d0[1]), d0[2]),
                                 fminf(fminf(fminf(d0[3], d0[4]), d0[5]), fminf(d0[6], d0[7])));
                float m1 = fminf(fminf(fminf(d1[0], d1[1]), d1[2]),
                                 fminf(fminf(fminf(d1[3], d1[4]), d1[5]), fminf(d1[6], d1[7])));
#pragma unroll
                for (int k = 8; k + 1 < kCmPerThread; k += 2) {
                    m0 = fminf(fminf(m0, d0[k]), d0[k + 1]);
                    m1 = fminf(fminf(m1, d1[k]), d1[k + 1]);
                }
                const unsigned w0 = __reduce_min_sync(0xffffffffu, __float_as_uint(m0));
                const unsigned w1 = __reduce_min_sync(0xffffffffu, __float_as_uint(m1));
                mine = lane == t ? w0 : mine;
                mine = lane == t + 1 ? w1 : mine;
            }
            atomicMin(&bmin[ub + lane], mine);
        }
#pragma unroll
        for (int k = 0; k < kCmPerThread; ++k)
            if (base + threadIdx.x + k * kCmThreads < P) sum_a += sqrt((double)bd[k]);
    }
    __syncthreads();
    // direction j -> i: the same points per thread and the same order of additions as a sweep of its own
    double sum_b = 0.0;
    for (int base = 0; base < P; base += kCmThreads * kCmPerThread)
#pragma unroll
        for (int k = 0; k < kCmPerThread; ++k) {
            const int u = base + threadIdx.x + k * kCmThreads;
            if (u < P) sum_b += sqrt((double)__uint_as_float(bmin[u]));
        }
    double sums[2] = {sum_a, sum_b};
#pragma unroll
    for (int dir = 0; dir < 2; ++dir) {
        double s = sums[dir];
        for (int of = 16; of > 0; of >>= 1) s += __shfl_xor_sync(0xffffffffu, s, of);
        if ((threadIdx.x & 31) == 0) wsum[dir][threadIdx.x >> 5] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sa = 0.0, sb = 0.0;
        for (int w = 0; w < kCmWarps; ++w) { sa += wsum[0][w]; sb += wsum[1][w]; }
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
    // test hook: the two-sweep kernel of rounds 1-2 (same values, twice the evaluations)
    const char* env = getenv("CORSAIR_B200_CHAMFER_TWO_SWEEPS");
    const bool two_sweeps = env && env[0] == '1';
    const size_t smem2 = two_sweeps ? (size_t)P * 6 * sizeof(float2) : (size_t)((P + 31) & ~31) * (3 * sizeof(float2) + sizeof(unsigned));
    if (smem2 <= 200 * 1024) {
        auto kern = two_sweeps ? chamfer_matrix_two_sweeps_kernel : chamfer_matrix_kernel;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
        if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "smem opt-in failed: %s", cudaGetErrorString(e));
        // work = 8 flop per point pair of every computed entry, each distance evaluated once
        cb_timed_scope ts(ctx, CB_T_CHAMFER_NN, 8.0 * (double)P * P * (double)(row1 - row0) * L, (cudaStream_t)stream);
        kern<<<dim3(L, row1 - row0), kCmThreads, smem2, (cudaStream_t)stream>>>(pts, L, P, row0, row1, half, out);
        CB_LAUNCH_CHECK(ctx);
        return CB_OK;
    }
    if (half) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_chamfer_matrix_half: P=%d above 7000", P);
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
