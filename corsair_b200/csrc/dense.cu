// Dense [n,m] matrices kept for API parity with the reference: pdist (utils/find_nn.py:24-31,
// fp32 difference form) and F0 F1^T (utils/_transform_estimation.py:149, return_corr=True).
// The fused kNN never materialises these; they exist for callers that ask for the matrix.
#include "common.cuh"

// block (32,8); tile 32 x 32, 4 rows per thread.  mode 0 SquareL2, 1 L2 (+1e-7), 2 dot.
__global__ void __launch_bounds__(256)
dense_kernel(const float* __restrict__ a, int n, const float* __restrict__ b, int m, int d, int mode,
             float* __restrict__ out) {
    __shared__ float sa[32][33];
    __shared__ float sb[32][33];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int j = blockIdx.x * 32 + tx;
    const int ib = blockIdx.y * 32;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c0 = 0; c0 < d; c0 += 32) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int row = ty + 8 * i;
            float v = 0.f, w = 0.f;
            if (c0 + tx < d) {
                if (ib + row < n) v = a[(size_t)(ib + row) * d + c0 + tx];
                if (blockIdx.x * 32 + row < m) w = b[(size_t)(blockIdx.x * 32 + row) * d + c0 + tx];
            }
            sa[row][tx] = v;
            sb[row][tx] = w;
        }
        __syncthreads();
        const int cn = min(32, d - c0);
        for (int c = 0; c < cn; ++c) {
            const float bv = sb[tx][c];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float av = sa[ty + 8 * i][c];
                if (mode == 2) acc[i] = fmaf(av, bv, acc[i]);
                else { float df = av - bv; acc[i] = fmaf(df, df, acc[i]); }
            }
        }
    }
    if (j >= m) return;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int row = ib + ty + 8 * i;
        if (row < n) out[(size_t)row * m + j] = mode == 1 ? sqrtf(acc[i] + 1e-7f) : acc[i];
    }
}

static int dense_launch(cb_ctx* ctx, const float* a, int n, const float* b, int m, int d, int mode, float* out,
                        void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (n == 0 || m == 0) return CB_OK;
    if (!a || !b || !out || n < 0 || m < 0 || d <= 0) CB_FAIL(ctx, CB_EINVAL, "dense matrix: bad arguments");
    dim3 grid(cb_div_up(m, 32), cb_div_up(n, 32));
    dense_kernel<<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(a, n, b, m, d, mode, out);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

extern "C" int cb_pdist(cb_ctx* ctx, const float* a, int n, const float* b, int m, int d, int dist_type,
                        float* out, void* stream) {
    if (ctx && dist_type != 0 && dist_type != 1) CB_FAIL(ctx, CB_EUNSUPPORTED, "Not implemented");
    return dense_launch(ctx, a, n, b, m, d, dist_type, out, stream);
}

extern "C" int cb_dot_matrix(cb_ctx* ctx, const float* a, int n, const float* b, int m, int d, float* out,
                             void* stream) {
    return dense_launch(ctx, a, n, b, m, d, 2, out, stream);
}
