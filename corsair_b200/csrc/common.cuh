// Shared context / workspace / error plumbing for the corsair_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "../../include/corsair_b200.h"

struct cb_ctx {
    int device = 0;
    int sm_count = 148;
    char* arena = nullptr;      // device workspace
    size_t arena_bytes = 0;
    size_t arena_off = 0;       // bump pointer, reset at the start of every public call
    int64_t launches = 0;
    int64_t h2d_bytes = 0, d2h_bytes = 0;  // host<->device copies the library itself issued (cb_copy_bytes)
    int knn_variant = 0;        // 0 auto, 1 simt, 2 tcgen05
    int ransac_variant = 0;     // 0 auto (tcgen05), 1 simt fp32, 2 tcgen05
    float* ransac_dump = nullptr;  // test hook: the tensor-core values U[h, c] of the next cb_ransac
    int ransac_dump_ld = 0;
    const int* knn_fb_count = nullptr;  // device counter of the last E1 kNN problem (diagnostic)
    char err[512] = {0};
    // bench instrumentation (cb_timing_*): CUDA-event brackets around the tracked kernel classes
    bool timing = false;
    struct TimedLaunch { int cls; double work; cudaEvent_t e0, e1; };
    std::vector<TimedLaunch> timed;
    std::vector<cudaEvent_t> event_pool;
    double cls_ms[CB_TIMING_CLASSES] = {0};
    double cls_work[CB_TIMING_CLASSES] = {0};
    int64_t cls_launches[CB_TIMING_CLASSES] = {0};
    double* dev_work = nullptr;  // [CB_TIMING_CLASSES] device-side work counters (RANSAC waves that ran)
    // cb_sym_pose: pinned staging for its small host<->device traffic, and the sub-contexts (own
    // arena + stream each) its independent registrations run on
    char* pinned = nullptr;
    std::vector<cb_ctx*> sub;
    std::vector<cudaStream_t> sub_stream;
    std::vector<cudaEvent_t> sub_event;
    cudaEvent_t fork_event = nullptr;
    // cb_sym_pose_many: contexts (own stream, arena, staging, internal streams) that hold the pairs in flight
    std::vector<cb_ctx*> slots;
    std::vector<cudaStream_t> slot_stream;
    std::vector<cudaEvent_t> slot_event;
    cudaEvent_t sync_event = nullptr;  // blocking-sync event: the host thread sleeps instead of spinning
};

// Bracket one tracked launch with events when timing is on.
struct cb_timed_scope {
    cb_ctx* ctx;
    cudaStream_t st;
    cb_ctx::TimedLaunch t;
    bool on;
    cb_timed_scope(cb_ctx* c, int cls, double work, cudaStream_t s);
    ~cb_timed_scope();
};

#define CB_FAIL(ctx, code, ...)                                  \
    do {                                                         \
        snprintf((ctx)->err, sizeof((ctx)->err), __VA_ARGS__);   \
        return (code);                                           \
    } while (0)

#define CB_CUDA(ctx, expr)                                                                   \
    do {                                                                                     \
        cudaError_t e_ = (expr);                                                             \
        if (e_ != cudaSuccess)                                                               \
            CB_FAIL(ctx, CB_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_),   \
                    __FILE__, __LINE__);                                                     \
    } while (0)

#define CB_LAUNCH_CHECK(ctx)                                                                 \
    do {                                                                                     \
        (ctx)->launches++;                                                                   \
        cudaError_t e_ = cudaGetLastError();                                                 \
        if (e_ != cudaSuccess)                                                               \
            CB_FAIL(ctx, CB_ECUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), \
                    __FILE__, __LINE__);                                                     \
    } while (0)

// Start of a public call: select the device and rewind the arena.
int cb_begin(cb_ctx* ctx);
// Bump-allocate `bytes` (256-byte aligned) from the arena; grows the arena (synchronising the
// device) when it is too small.  Returns nullptr and sets ctx->err on failure.
void* cb_alloc(cb_ctx* ctx, size_t bytes);
// Make sure the arena holds at least `bytes` in total BEFORE any cb_alloc of this call (growing
// in the middle of a call would invalidate earlier sub-allocations).
int cb_reserve(cb_ctx* ctx, size_t bytes);

static inline size_t cb_align(size_t b) { return (b + 255) & ~size_t(255); }

template <typename T>
static inline T* cb_alloc_t(cb_ctx* ctx, size_t n) {
    return reinterpret_cast<T*>(cb_alloc(ctx, n * sizeof(T)));
}

static inline int cb_div_up(int a, int b) { return (a + b - 1) / b; }

// ---- debug build (make debug -> lib/libcorsair_b200_dbg.so): device-side range checks -----------
// compute-sanitizer is closed on the GPU pool this was developed on; the debug library traps (the
// launch then fails loudly) when an index the multi-role kernels compute leaves its range.
#ifdef CB_DEBUG_CHECKS
#define CB_DEV_CHECK(cond)                                                        \
    do {                                                                          \
        if (!(cond)) {                                                            \
            printf("corsair_b200 device check failed: %s (%s:%d)\n", #cond, __FILE__, __LINE__); \
            __trap();                                                             \
        }                                                                         \
    } while (0)
#else
#define CB_DEV_CHECK(cond) ((void)0)
#endif

// ---- device helpers ---------------------------------------------------------------------
__device__ __forceinline__ uint64_t cb_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// minimal-set draw shared by the fit kernel and cb_draw_minimal_sets (see the header)
__device__ __forceinline__ int cb_draw_index(uint64_t seed, uint64_t counter, int m) {
    uint64_t x = cb_splitmix64(seed * 0x9E3779B97F4A7C15ull + counter);
    return (int)(((x >> 32) * (uint64_t)(uint32_t)m) >> 32);
}

// internal entry points shared between translation units (none of them rewinds the arena)
size_t cb_knn_workspace(int nq, int nr, int d, int k);
size_t cb_ransac_workspace(const cb_ctx* ctx, int m, int h_total);
int cb_ransac_impl(cb_ctx* ctx, const float* src_xyz, const float* tgt_xyz, const int64_t* src_idx,
                   const int64_t* tgt_idx, int m, const int32_t* minimal, uint64_t seed, int h_total, int ransac_n,
                   double max_dist, int adaptive, double confidence, double* T, double* stats, int32_t* counts,
                   double* sums, cudaStream_t st);
int cb_chamfer_impl(cb_ctx* ctx, const float* a, int na, const float* b, int nb, const double* T, double* out,
                    cudaStream_t st);
int cb_knn_impl(cb_ctx* ctx, const float* q, int nq, const float* r, int nr, int d, int k,
                int metric, int64_t* idx, float* dist, int32_t* idx32, cudaStream_t st);
size_t cb_segment_workspace(int n, int n_samples, int n_nn, int K);
int cb_segment_impl(cb_ctx* ctx, const float* feat, const float* xyz, int n, int d, int K, const int32_t* d_samples,
                    int n_samples, int n_nn, const int32_t* d_first, const double* d_unif, int n_init, int n_trials,
                    double min_center, int64_t* perm, int32_t* seg_off, cudaStream_t st);
