// K4: Chamfer / nearest-neighbour distances in 3-D.
// Replaces chamfer_kdtree_1direction(apply_transform(..)) (utils/preprocess.py:39-48,67-70),
// chamfer_gpu / chamfer_gpu_1direction (:51-64) and pc_dist.chamfer / compute_dist
// (utils/pc_dist.py:45-99) of the reference.
//
// The search runs in fp32 difference form (brute force, references tiled through shared
// memory, packed (distance bits, index) atomicMin across reference chunks); the distance that
// enters the mean is then recomputed in float64 for the chosen neighbour, so the result is
// within ~1e-7 relative of the float64 KD-tree answer (contract: 1e-5).
#include "common.cuh"

#include <math.h>
#include <stdarg.h>

#include <mutex>

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 ch_pk2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void ch_unpk2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 ch_add2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 ch_mul2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 ch_fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

constexpr int kNnThreads = 256;
constexpr int kNnPerThread = 4;  // x points per thread (two packed pairs)

// brute-force NN: four x points per thread as two packed fp32 pairs (add/mul/fma.f32x2: two distances
// per instruction), the y chunk tiled through shared memory as negated duplicated coordinates so that
// p - v is one packed add; poses stay on the device.  The arithmetic per distance is the scalar
// kernel's (dx*dx, then two fused multiply-adds), first minimum wins inside a chunk, (distance,
// index) atomicMin across chunks.
__global__ void __launch_bounds__(kNnThreads)
nn3_kernel(const float4* __restrict__ x, int nx, const float4* __restrict__ y, int ny, int chunk,
           unsigned long long* __restrict__ best) {
    constexpr int TY = 256;
    __shared__ float2 snx[TY], sny[TY], snz[TY];
    const int i0 = blockIdx.x * (kNnThreads * kNnPerThread) + threadIdx.x;
    const int y0 = blockIdx.y * chunk, y1 = min(ny, y0 + chunk);
    f32x2 px[kNnPerThread / 2], py[kNnPerThread / 2], pz[kNnPerThread / 2];
    float bd[kNnPerThread];
    int bj[kNnPerThread];
#pragma unroll
    for (int k = 0; k < kNnPerThread / 2; ++k) {
        const float4 a = x[min(i0 + (2 * k) * kNnThreads, nx - 1)];
        const float4 b = x[min(i0 + (2 * k + 1) * kNnThreads, nx - 1)];
        px[k] = ch_pk2(a.x, b.x);
        py[k] = ch_pk2(a.y, b.y);
        pz[k] = ch_pk2(a.z, b.z);
        bd[2 * k] = bd[2 * k + 1] = INFINITY;
        bj[2 * k] = bj[2 * k + 1] = 0;
    }
    for (int base = y0; base < y1; base += TY) {
        const int nt = min(TY, y1 - base);
        __syncthreads();
        if (threadIdx.x < nt) {
            const float4 v = y[base + threadIdx.x];
            snx[threadIdx.x] = make_float2(-v.x, -v.x);
            sny[threadIdx.x] = make_float2(-v.y, -v.y);
            snz[threadIdx.x] = make_float2(-v.z, -v.z);
        }
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < nt; ++j) {
            const f32x2 vx = *reinterpret_cast<const f32x2*>(&snx[j]);
            const f32x2 vy = *reinterpret_cast<const f32x2*>(&sny[j]);
            const f32x2 vz = *reinterpret_cast<const f32x2*>(&snz[j]);
            const int jj = base + j;
#pragma unroll
            for (int k = 0; k < kNnPerThread / 2; ++k) {
                const f32x2 dx = ch_add2(px[k], vx), dy = ch_add2(py[k], vy), dz = ch_add2(pz[k], vz);
                const f32x2 d = ch_fma2(dz, dz, ch_fma2(dy, dy, ch_mul2(dx, dx)));
                float d0, d1;
                ch_unpk2(d, d0, d1);
                if (d0 < bd[2 * k]) { bd[2 * k] = d0; bj[2 * k] = jj; }
                if (d1 < bd[2 * k + 1]) { bd[2 * k + 1] = d1; bj[2 * k + 1] = jj; }
            }
        }
    }
    if (y1 > y0) {
#pragma unroll
        for (int k = 0; k < kNnPerThread; ++k) {
            const int i = i0 + k * kNnThreads;
            if (i < nx) {
                unsigned long long key = ((unsigned long long)__float_as_uint(bd[k]) << 32) | (unsigned int)bj[k];
                atomicMin(&best[i], key);
            }
        }
    }
}

__device__ __forceinline__ void apply_pose(const double* T, double& x, double& y, double& z) {
    double tx = T[0] * x + T[1] * y + T[2] * z + T[3];
    double ty = T[4] * x + T[5] * y + T[6] * z + T[7];
    double tz = T[8] * x + T[9] * y + T[10] * z + T[11];
    x = tx; y = ty; z = tz;
}

// exact float64 distance of every p to its chosen neighbour; per-block partial sum / max
__global__ void __launch_bounds__(256)
nn3_finalize_kernel(const float* __restrict__ p, int np, const double* __restrict__ Tp,
                    const float* __restrict__ q, const double* __restrict__ Tq,
                    const unsigned long long* __restrict__ best, double* __restrict__ partial_sum,
                    double* __restrict__ partial_max, int32_t* __restrict__ nn_idx) {
    __shared__ double ws[8], wm[8];
    const int i = blockIdx.x * 256 + threadIdx.x;
    double d = 0.0;
    if (i < np) {
        int j = (int)(best[i] & 0xffffffffull);
        double x = p[3 * (size_t)i], y = p[3 * (size_t)i + 1], z = p[3 * (size_t)i + 2];
        double u = q[3 * (size_t)j], v = q[3 * (size_t)j + 1], w = q[3 * (size_t)j + 2];
        if (Tp) apply_pose(Tp, x, y, z);
        if (Tq) apply_pose(Tq, u, v, w);
        double dx = x - u, dy = y - v, dz = z - w;
        d = sqrt(dx * dx + dy * dy + dz * dz);
        if (nn_idx) nn_idx[i] = j;
    }
    double s = d, m = d;
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    }
    if ((threadIdx.x & 31) == 0) { ws[threadIdx.x >> 5] = s; wm[threadIdx.x >> 5] = m; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { s += ws[w]; m = fmax(m, wm[w]); }
        partial_sum[blockIdx.x] = s;
        partial_max[blockIdx.x] = m;
    }
}

// out[0] (+)= sum(partials)/n   or   max(partials)
__global__ void nn3_reduce_kernel(const double* __restrict__ partial_sum, const double* __restrict__ partial_max,
                                  int nblocks, int n, int mode /*0 mean set,1 mean add,2 max set,3 max of both*/,
                                  double* __restrict__ out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double s = 0.0, m = 0.0;
    for (int b = 0; b < nblocks; ++b) { s += partial_sum[b]; m = fmax(m, partial_max[b]); }
    if (mode == 0) out[0] = s / (double)n;
    else if (mode == 1) out[0] += s / (double)n;
    else if (mode == 2) out[0] = m;
    else out[0] = fmax(out[0], m);
}

__global__ void pack_points_dev_pose_kernel(const float* __restrict__ xyz, int n, const double* __restrict__ T,
                                            float4* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double x = xyz[3 * (size_t)i], y = xyz[3 * (size_t)i + 1], z = xyz[3 * (size_t)i + 2];
    if (T) apply_pose(T, x, y, z);
    out[i] = make_float4((float)x, (float)y, (float)z, 0.f);
}

static int nn3_direction(cb_ctx* ctx, const float* p, int np, const double* Tp, const float4* p4,
                         const float* q, int nq, const double* Tq, const float4* q4, int mode, double* out,
                         int32_t* nn_idx, cudaStream_t st) {
    unsigned long long* best = cb_alloc_t<unsigned long long>(ctx, np);
    int nblk = cb_div_up(np, 256);
    double* ps = cb_alloc_t<double>(ctx, nblk);
    double* pm = cb_alloc_t<double>(ctx, nblk);
    if (!best || !ps || !pm) return CB_ENOMEM;
    CB_CUDA(ctx, cudaMemsetAsync(best, 0xff, (size_t)np * 8, st));
    const int nblk_nn = cb_div_up(np, kNnThreads * kNnPerThread);
    int S = cb_div_up(4 * ctx->sm_count, nblk_nn);  // ~4 resident CTAs per SM
    S = max(1, min(S, cb_div_up(nq, 256)));
    int chunk = cb_div_up(cb_div_up(nq, S), 256) * 256;
    S = cb_div_up(nq, chunk);
    {
        cb_timed_scope ts(ctx, CB_T_CHAMFER_NN, 8.0 * np * (double)nq, st);
        nn3_kernel<<<dim3(nblk_nn, S), kNnThreads, 0, st>>>(p4, np, q4, nq, chunk, best);
        CB_LAUNCH_CHECK(ctx);
    }
    nn3_finalize_kernel<<<nblk, 256, 0, st>>>(p, np, Tp, q, Tq, best, ps, pm, nn_idx);
    CB_LAUNCH_CHECK(ctx);
    nn3_reduce_kernel<<<1, 32, 0, st>>>(ps, pm, nblk, np, mode, out);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

static int chamfer_impl(cb_ctx* ctx, const float* a, int na, const float* b, int nb, const double* T,
                        int two_dir, int use_max, double* out, int32_t* nn_idx, cudaStream_t st) {
    if (!a || !b || !out || na <= 0 || nb <= 0) CB_FAIL(ctx, CB_EINVAL, "chamfer: empty cloud or null pointer");
    int rc = cb_reserve(ctx, (size_t)(na + nb) * 40 + (1 << 16));
    if (rc) return rc;
    float4* a4 = cb_alloc_t<float4>(ctx, na);
    float4* b4 = cb_alloc_t<float4>(ctx, nb);
    if (!a4 || !b4) return CB_ENOMEM;
    pack_points_dev_pose_kernel<<<cb_div_up(na, 256), 256, 0, st>>>(a, na, T, a4);
    CB_LAUNCH_CHECK(ctx);
    pack_points_dev_pose_kernel<<<cb_div_up(nb, 256), 256, 0, st>>>(b, nb, nullptr, b4);
    CB_LAUNCH_CHECK(ctx);
    rc = nn3_direction(ctx, a, na, T, a4, b, nb, nullptr, b4, use_max ? 2 : 0, out, nn_idx, st);
    if (rc) return rc;
    if (two_dir) rc = nn3_direction(ctx, b, nb, nullptr, b4, a, na, T, a4, use_max ? 3 : 1, out, nullptr, st);
    return rc;
}

// one-direction Chamfer without the arena rewind (composite entry points, register.cu)
int cb_chamfer_impl(cb_ctx* ctx, const float* a, int na, const float* b, int nb, const double* T, double* out,
                    cudaStream_t st) {
    return chamfer_impl(ctx, a, na, b, nb, T, 0, 0, out, nullptr, st);
}

extern "C" int cb_chamfer(cb_ctx* ctx, const float* a, int na, const float* b, int nb, const double* T,
                          int two_dir, double* out, int32_t* nn_idx, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    return chamfer_impl(ctx, a, na, b, nb, T, two_dir, 0, out, nn_idx, (cudaStream_t)stream);
}

extern "C" int cb_nn_max(cb_ctx* ctx, const float* a, int na, const float* b, int nb, const double* T,
                         double* out, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    return chamfer_impl(ctx, a, na, b, nb, T, 0, 1, out, nullptr, (cudaStream_t)stream);
}

// chamfer_max of a cloud against rigidly moved copies of itself (evaluation-shapenet.py:122-146):
// out[r] = max over BOTH directions of the nearest-neighbour distance between xyz and T_r(xyz).
extern "C" int cb_self_hausdorff(cb_ctx* ctx, const float* xyz, int n, const double* T, int n_poses, double* out,
                                 void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (n_poses == 0) return CB_OK;
    if (!xyz || !T || !out || n <= 0 || n_poses < 0) CB_FAIL(ctx, CB_EINVAL, "cb_self_hausdorff: bad arguments");
    for (int r = 0; r < n_poses; ++r) {
        ctx->arena_off = 0;  // the stream orders the poses: each reuses the scratch of the one before
        if ((rc = chamfer_impl(ctx, xyz, n, xyz, n, T + 16 * (size_t)r, 1, 1, out + r, nullptr, (cudaStream_t)stream)))
            return rc;
    }
    return CB_OK;
}

// Cross table of two-direction Chamfer distances between two stacks of equal-sized clouds — the
// GT-CAD x reconstruction table of utils/precompute_chamfer_splats.py:113-149 (15 000 points each
// there): out[i, j] = chamfer_kdtree_1direction(A_i, B_j) + chamfer_kdtree_1direction(B_j, A_i).
// One pair Chamfer per entry on the stream; at that size each is ~200 us of search work, so the
// launches are not the limit.
extern "C" int cb_chamfer_cross(cb_ctx* ctx, const float* a, int la, int pa, const float* b, int lb, int pb,
                                double* out, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    if (la == 0 || lb == 0) return CB_OK;
    if (!a || !b || !out || la < 0 || lb < 0 || pa <= 0 || pb <= 0) CB_FAIL(ctx, CB_EINVAL, "cb_chamfer_cross: bad arguments");
    for (int i = 0; i < la; ++i)
        for (int j = 0; j < lb; ++j) {
            ctx->arena_off = 0;  // stream order makes the scratch reusable pair after pair
            if ((rc = chamfer_impl(ctx, a + (size_t)i * pa * 3, pa, b + (size_t)j * pb * 3, pb, nullptr, 1, 0,
                                   out + (size_t)i * lb + j, nullptr, (cudaStream_t)stream)))
                return rc;
        }
    return CB_OK;
}

// ---------------------------------------------------------------------------------------------------
// cb_chamfer_static: cb_chamfer for callers that are not allowed to allocate device memory.
//
// Everything above takes its scratch (packed float4 copies of both clouds, the per-point
// (distance, index) keys, the per-block partial sums) from a cb_ctx arena, and cb_create / cb_reserve
// obtain that arena with cudaMalloc.  Some hosts of this library cannot do that: a plugin running
// inside an application that has handed all of its device memory to its own pool up front (a
// PyTorch / TensorRT-style caching allocator set to grab the card, an inference server with a fixed
// memory budget per model), a real-time loop where cudaMalloc's implicit device synchronisation and
// unbounded latency are forbidden, or a process whose memory accounting forbids allocations it did
// not make itself.  For those callers the scratch lives in the module image instead: one file-scope
// __device__ array of fixed size that every GPU holds from the moment it loads this library until the
// process exits.  The price is the hard size limit below and that all calls share the one array.
//
// The array's name below is only a handle on the host side; the real per-device address comes from
// cudaGetSymbolAddress, asked once per device with that device current (static_scratch_for).
constexpr int kStaticMaxPoints = CB_CHAMFER_STATIC_MAX_POINTS;           // per cloud
constexpr size_t kStaticPartials = (kStaticMaxPoints + 255) / 256;       // blocks of nn3_finalize_kernel
constexpr size_t kStaticOffA4 = 0;                                        // float4 [max]
constexpr size_t kStaticOffB4 = kStaticOffA4 + sizeof(float4) * kStaticMaxPoints;
constexpr size_t kStaticOffBestA = kStaticOffB4 + sizeof(float4) * kStaticMaxPoints;   // u64 [max]
constexpr size_t kStaticOffBestB = kStaticOffBestA + 8 * (size_t)kStaticMaxPoints;
constexpr size_t kStaticOffPart = kStaticOffBestB + 8 * (size_t)kStaticMaxPoints;      // 4 x double [partials]
constexpr size_t kStaticScratchBytes = kStaticOffPart + 4 * 8 * kStaticPartials;
static_assert(kStaticOffB4 % 256 == 0 && kStaticOffBestA % 256 == 0 && kStaticOffBestB % 256 == 0 &&
              kStaticOffPart % 256 == 0 && (8 * kStaticPartials) % 256 == 0, "scratch carve-up must stay 256-byte aligned");
static_assert(kStaticScratchBytes == CB_CHAMFER_STATIC_SCRATCH_BYTES, "header and kernel file disagree on the scratch size");

static __device__ __align__(256) unsigned char g_chamfer_static_scratch[kStaticScratchBytes];

namespace {
constexpr int kStaticMaxDevices = 64;
struct StaticDevice {
    unsigned char* base = nullptr;  // device address of g_chamfer_static_scratch on this GPU
    int sm_count = 0;
};
StaticDevice g_static_dev[kStaticMaxDevices];
std::mutex g_static_mutex;

void static_err(char* h_err, size_t h_err_len, const char* fmt, ...) __attribute__((format(printf, 3, 4)));
void static_err(char* h_err, size_t h_err_len, const char* fmt, ...) {
    if (!h_err || h_err_len == 0) return;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(h_err, h_err_len, fmt, ap);
    va_end(ap);
}

// First call on a device: resolve the array's address there and read the SM count.  Neither call
// allocates on behalf of the caller, but resolving a symbol may make the driver load this module's
// image, which is why the first call per device belongs outside any stream capture.
int static_scratch_for(int device, StaticDevice* out, char* h_err, size_t h_err_len) {
    if (device < 0 || device >= kStaticMaxDevices) {
        static_err(h_err, h_err_len, "cb_chamfer_static: device ordinal %d outside [0, %d)", device, kStaticMaxDevices);
        return CB_EUNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_static_mutex);
    StaticDevice& sd = g_static_dev[device];
    if (!sd.base) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_chamfer_static_scratch);
        if (e != cudaSuccess || !p) {
            static_err(h_err, h_err_len, "cb_chamfer_static: cudaGetSymbolAddress on device %d failed: %s", device,
                       cudaGetErrorString(e));
            return CB_ECUDA;
        }
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || sms <= 0) {
            static_err(h_err, h_err_len, "cb_chamfer_static: SM count of device %d unavailable: %s", device,
                       cudaGetErrorString(e));
            return CB_ECUDA;
        }
        sd.sm_count = sms;
        sd.base = static_cast<unsigned char*>(p);
    }
    *out = sd;
    return CB_OK;
}

#define CB_STATIC_CUDA(expr)                                                                          \
    do {                                                                                              \
        cudaError_t e_ = (expr);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            static_err(h_err, h_err_len, "cb_chamfer_static: %s failed: %s", #expr, cudaGetErrorString(e_)); \
            return CB_ECUDA;                                                                          \
        }                                                                                             \
    } while (0)

// one search direction with caller-provided scratch; the launch geometry is nn3_direction's, so the
// (distance, index) keys and the order of every float64 sum are the arena path's
int static_direction(const StaticDevice& sd, const float* p, int np, const double* Tp, const float4* p4, const float* q,
                     int nq, const double* Tq, const float4* q4, unsigned long long* best, double* ps, double* pm,
                     int mode, double* out, int32_t* nn_idx, cudaStream_t st, char* h_err, size_t h_err_len) {
    // the keys are folded with atomicMin and the array still holds the previous call's: reset them
    CB_STATIC_CUDA(cudaMemsetAsync(best, 0xff, (size_t)np * 8, st));
    const int nblk = cb_div_up(np, 256);
    const int nblk_nn = cb_div_up(np, kNnThreads * kNnPerThread);
    int S = cb_div_up(4 * sd.sm_count, nblk_nn);
    S = max(1, min(S, cb_div_up(nq, 256)));
    const int chunk = cb_div_up(cb_div_up(nq, S), 256) * 256;
    S = cb_div_up(nq, chunk);
    nn3_kernel<<<dim3(nblk_nn, S), kNnThreads, 0, st>>>(p4, np, q4, nq, chunk, best);
    CB_STATIC_CUDA(cudaGetLastError());
    nn3_finalize_kernel<<<nblk, 256, 0, st>>>(p, np, Tp, q, Tq, best, ps, pm, nn_idx);
    CB_STATIC_CUDA(cudaGetLastError());
    nn3_reduce_kernel<<<1, 32, 0, st>>>(ps, pm, nblk, np, mode, out);
    CB_STATIC_CUDA(cudaGetLastError());
    return CB_OK;
}
}  // namespace

extern "C" int cb_chamfer_static(const float* a, int na, const float* b, int nb, const double* T, int two_dir,
                                 double* out, int32_t* nn_idx, char* h_err, size_t h_err_len, void* stream) {
    if (h_err && h_err_len) h_err[0] = 0;
    if (!a || !b || !out || na <= 0 || nb <= 0) {
        static_err(h_err, h_err_len, "cb_chamfer_static: empty cloud or null pointer");
        return CB_EINVAL;
    }
    if (na > kStaticMaxPoints || nb > kStaticMaxPoints) {
        static_err(h_err, h_err_len,
                   "cb_chamfer_static: clouds of %d and %d points exceed the built-in scratch (at most %d points per "
                   "cloud); use cb_chamfer with a cb_ctx for larger clouds",
                   na, nb, kStaticMaxPoints);
        return CB_EUNSUPPORTED;
    }
    int device = -1;
    CB_STATIC_CUDA(cudaGetDevice(&device));
    StaticDevice sd;
    int rc = static_scratch_for(device, &sd, h_err, h_err_len);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    float4* a4 = reinterpret_cast<float4*>(sd.base + kStaticOffA4);
    float4* b4 = reinterpret_cast<float4*>(sd.base + kStaticOffB4);
    unsigned long long* best_a = reinterpret_cast<unsigned long long*>(sd.base + kStaticOffBestA);
    unsigned long long* best_b = reinterpret_cast<unsigned long long*>(sd.base + kStaticOffBestB);
    double* part = reinterpret_cast<double*>(sd.base + kStaticOffPart);
    pack_points_dev_pose_kernel<<<cb_div_up(na, 256), 256, 0, st>>>(a, na, T, a4);
    CB_STATIC_CUDA(cudaGetLastError());
    pack_points_dev_pose_kernel<<<cb_div_up(nb, 256), 256, 0, st>>>(b, nb, nullptr, b4);
    CB_STATIC_CUDA(cudaGetLastError());
    rc = static_direction(sd, a, na, T, a4, b, nb, nullptr, b4, best_a, part, part + kStaticPartials, 0, out, nn_idx, st,
                          h_err, h_err_len);
    if (rc) return rc;
    if (two_dir)
        rc = static_direction(sd, b, nb, nullptr, b4, a, na, T, a4, best_b, part + 2 * kStaticPartials,
                              part + 3 * kStaticPartials, 1, out, nullptr, st, h_err, h_err_len);
    return rc;
}
