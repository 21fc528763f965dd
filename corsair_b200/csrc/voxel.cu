// Voxel de-duplication of a point cloud on the device: the
// `ME.utils.sparse_quantize(np.floor(pc / voxel_size), return_index=True)` step of quantize_pc
// (evaluation-shapenet.py:97-107; utils/Info/CADLib.py:106-121 does the same for the CAD library).
// Keeps, for every occupied voxel, the point with the LOWEST index, and returns those indices in
// ascending order — numpy.unique(..., axis=0, return_index=True) followed by a sort of the indices,
// the rule corsair_b200/synth.py and SURVEY.md §8(d) use.  MinkowskiEngine itself is not installed, so
// which duplicate ME keeps cannot be checked here (parity unpinned for this call).
//
// Three small kernels: (1) voxel key = three 21-bit biased integer coordinates, inserted into an
// open-addressing hash table (atomicCAS on the key, atomicMin on the point index); (2)+(3) one block
// flags the points that own their voxel, scans the flags and writes indices + integer grid coordinates.
#include "common.cuh"

#include <math.h>

namespace {

constexpr unsigned long long kEmpty = ~0ull;
constexpr int kBias = 1 << 20;  // |floor(p / voxel)| < 2^20

__device__ __forceinline__ bool voxel_key(const float* __restrict__ xyz, int i, float voxel, unsigned long long& key,
                                          int (&g)[3]) {
    bool ok = true;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float q = floorf(__fdiv_rn(xyz[3 * (size_t)i + a], voxel));  // float32 division like NumPy's
        ok = ok && fabsf(q) < (float)kBias;
        g[a] = (int)q;
    }
    key = ((unsigned long long)(g[0] + kBias) << 42) | ((unsigned long long)(g[1] + kBias) << 21) |
          (unsigned long long)(g[2] + kBias);
    return ok;
}

__global__ void voxel_insert_kernel(const float* __restrict__ xyz, int n, float voxel, unsigned mask,
                                    unsigned long long* __restrict__ tkey, int* __restrict__ tmin,
                                    int* __restrict__ slot_of, int* __restrict__ bad) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long key;
    int g[3];
    if (!voxel_key(xyz, i, voxel, key, g)) { atomicExch(bad, 1); slot_of[i] = 0; return; }
    unsigned slot = (unsigned)(cb_splitmix64(key) & mask);
    for (;;) {
        const unsigned long long old = atomicCAS(&tkey[slot], kEmpty, key);
        if (old == kEmpty || old == key) break;
        slot = (slot + 1) & mask;
    }
    atomicMin(&tmin[slot], i);
    slot_of[i] = (int)slot;
}

// one block: flags -> exclusive scan over contiguous slices -> ascending indices + grid coordinates
__global__ void __launch_bounds__(1024)
voxel_compact_kernel(const float* __restrict__ xyz, int n, float voxel, const int* __restrict__ tmin,
                     const int* __restrict__ slot_of, const int* __restrict__ bad, int64_t* __restrict__ idx_out,
                     int32_t* __restrict__ grid_out, int32_t* __restrict__ count) {
    __shared__ int part[1024];
    const int tid = threadIdx.x;
    if (*bad) {  // a coordinate outside +-2^20 voxels (or not finite): no result
        if (tid == 0) *count = -1;
        return;
    }
    const int per = (n + 1023) / 1024;
    const int i0 = min(n, tid * per), i1 = min(n, i0 + per);
    int c = 0;
    for (int i = i0; i < i1; ++i) c += tmin[slot_of[i]] == i;
    part[tid] = c;
    __syncthreads();
    if (tid < 32) {  // exclusive scan of the 1024 slice counts by one warp
        int v[32], tot = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j) { v[j] = part[tid * 32 + j]; tot += v[j]; }
        int inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, inc, o); if (tid >= o) inc += y; }
        int run = inc - tot;
#pragma unroll
        for (int j = 0; j < 32; ++j) { part[tid * 32 + j] = run; run += v[j]; }
        if (tid == 31) *count = inc;
    }
    __syncthreads();
    int pos = part[tid];
    for (int i = i0; i < i1; ++i) {
        if (tmin[slot_of[i]] != i) continue;
        idx_out[pos] = i;
        if (grid_out) {
            unsigned long long key;
            int g[3];
            voxel_key(xyz, i, voxel, key, g);
            grid_out[3 * (size_t)pos] = g[0];
            grid_out[3 * (size_t)pos + 1] = g[1];
            grid_out[3 * (size_t)pos + 2] = g[2];
        }
        ++pos;
    }
}

}  // namespace

extern "C" int cb_voxel_unique(cb_ctx* ctx, const float* xyz, int n, float voxel, int64_t* idx_out, int32_t* grid_out,
                               int32_t* count, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (!count || n < 0 || !(voxel > 0.f)) CB_FAIL(ctx, CB_EINVAL, "cb_voxel_unique: bad arguments");
    if (n == 0) {
        CB_CUDA(ctx, cudaMemsetAsync(count, 0, 4, st));
        return CB_OK;
    }
    if (!xyz || !idx_out) CB_FAIL(ctx, CB_EINVAL, "cb_voxel_unique: null pointer");
    unsigned cap = 1024;
    while (cap < 2u * (unsigned)n) cap <<= 1;
    if ((rc = cb_reserve(ctx, (size_t)cap * 12 + (size_t)n * 4 + 4096))) return rc;
    unsigned long long* tkey = cb_alloc_t<unsigned long long>(ctx, cap);
    int* tmin = cb_alloc_t<int>(ctx, cap);
    int* slot_of = cb_alloc_t<int>(ctx, n);
    int* bad = cb_alloc_t<int>(ctx, 1);
    if (!tkey || !tmin || !slot_of || !bad) return CB_ENOMEM;
    CB_CUDA(ctx, cudaMemsetAsync(tkey, 0xff, (size_t)cap * 8, st));
    CB_CUDA(ctx, cudaMemsetAsync(tmin, 0x7f, (size_t)cap * 4, st));
    CB_CUDA(ctx, cudaMemsetAsync(bad, 0, 4, st));
    voxel_insert_kernel<<<cb_div_up(n, 256), 256, 0, st>>>(xyz, n, voxel, cap - 1, tkey, tmin, slot_of, bad);
    CB_LAUNCH_CHECK(ctx);
    voxel_compact_kernel<<<1, 1024, 0, st>>>(xyz, n, voxel, tmin, slot_of, bad, idx_out, grid_out, count);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}
