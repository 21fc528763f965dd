// Context, workspace arena and error reporting of the C ABI (include/corsair_b200.h).
#include "common.cuh"

#include <new>

extern "C" int cb_version(void) { return 100; }

extern "C" int cb_create(cb_ctx** out, int device, size_t workspace_bytes) {
    if (!out) return CB_EINVAL;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
        cudaGetLastError();
        return CB_ECUDA;  // no CPU fallback: without a CUDA device there is no context
    }
    cb_ctx* ctx = new (std::nothrow) cb_ctx();
    if (!ctx) return CB_ENOMEM;
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess) {
        delete ctx;
        return CB_ECUDA;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
    if (workspace_bytes < (size_t(64) << 20)) workspace_bytes = size_t(64) << 20;
    if (cudaMalloc(&ctx->arena, workspace_bytes) != cudaSuccess) {
        cudaGetLastError();
        delete ctx;
        return CB_ENOMEM;
    }
    ctx->arena_bytes = workspace_bytes;
    if (cudaMalloc(&ctx->dev_work, CB_TIMING_CLASSES * sizeof(double)) != cudaSuccess ||
        cudaMemset(ctx->dev_work, 0, CB_TIMING_CLASSES * sizeof(double)) != cudaSuccess) {
        cudaGetLastError();
        cudaFree(ctx->arena);
        delete ctx;
        return CB_ENOMEM;
    }
    *out = ctx;
    return CB_OK;
}

extern "C" int cb_destroy(cb_ctx* ctx) {
    if (!ctx) return CB_EINVAL;
    cudaSetDevice(ctx->device);
    if (ctx->arena) cudaFree(ctx->arena);
    if (ctx->dev_work) cudaFree(ctx->dev_work);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    for (auto e : ctx->sub_event) cudaEventDestroy(e);
    for (auto s : ctx->sub_stream) cudaStreamDestroy(s);
    for (auto c : ctx->sub) cb_destroy(c);
    for (auto e : ctx->slot_event) cudaEventDestroy(e);
    for (auto s : ctx->slot_stream) cudaStreamDestroy(s);
    for (auto c : ctx->slots) cb_destroy(c);
    if (ctx->fork_event) cudaEventDestroy(ctx->fork_event);
    if (ctx->sync_event) cudaEventDestroy(ctx->sync_event);
    for (auto& t : ctx->timed) { cudaEventDestroy(t.e0); cudaEventDestroy(t.e1); }
    for (auto e : ctx->event_pool) cudaEventDestroy(e);
    delete ctx;
    return CB_OK;
}

extern "C" const char* cb_last_error(const cb_ctx* ctx) { return ctx ? ctx->err : "null context"; }

extern "C" int64_t cb_launch_count(const cb_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" int cb_copy_bytes(const cb_ctx* ctx, int64_t* h2d, int64_t* d2h) {
    if (!ctx) return CB_EINVAL;
    if (h2d) *h2d = ctx->h2d_bytes;
    if (d2h) *d2h = ctx->d2h_bytes;
    return CB_OK;
}

extern "C" int cb_set_knn_variant(cb_ctx* ctx, int variant) {
    if (!ctx || variant < 0 || variant > 2) return CB_EINVAL;
    ctx->knn_variant = variant;
    return CB_OK;
}

int cb_begin(cb_ctx* ctx) {
    if (!ctx) return CB_EINVAL;
    ctx->err[0] = 0;
    ctx->arena_off = 0;
    CB_CUDA(ctx, cudaSetDevice(ctx->device));
    return CB_OK;
}

int cb_reserve(cb_ctx* ctx, size_t bytes) {
    bytes = cb_align(bytes) + (size_t(1) << 20);
    if (bytes <= ctx->arena_bytes) return CB_OK;
    if (ctx->arena_off != 0) CB_FAIL(ctx, CB_ENOMEM, "cb_reserve after cb_alloc");
    // earlier calls on the stream may still read the old arena
    CB_CUDA(ctx, cudaDeviceSynchronize());
    if (ctx->arena) cudaFree(ctx->arena);
    ctx->arena = nullptr;
    ctx->knn_fb_count = nullptr;
    ctx->arena_bytes = 0;
    size_t want = bytes + bytes / 4;
    if (cudaMalloc(&ctx->arena, want) != cudaSuccess) {
        cudaGetLastError();
        CB_FAIL(ctx, CB_ENOMEM, "workspace of %zu bytes could not be allocated", want);
    }
    ctx->arena_bytes = want;
    return CB_OK;
}

void* cb_alloc(cb_ctx* ctx, size_t bytes) {
    bytes = cb_align(bytes ? bytes : 1);
    if (ctx->arena_off + bytes > ctx->arena_bytes) {
        snprintf(ctx->err, sizeof(ctx->err), "workspace exhausted (%zu + %zu > %zu): missing cb_reserve",
                 ctx->arena_off, bytes, ctx->arena_bytes);
        return nullptr;
    }
    void* p = ctx->arena + ctx->arena_off;
    ctx->arena_off += bytes;
    return p;
}

// ------------------------------------------------------------------ bench instrumentation
static cudaEvent_t take_event(cb_ctx* ctx) {
    if (!ctx->event_pool.empty()) {
        cudaEvent_t e = ctx->event_pool.back();
        ctx->event_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

cb_timed_scope::cb_timed_scope(cb_ctx* c, int cls, double work, cudaStream_t s) : ctx(c), st(s), on(c->timing) {
    if (!on) return;
    t.cls = cls;
    t.work = work;
    t.e0 = take_event(ctx);
    t.e1 = take_event(ctx);
    cudaEventRecord(t.e0, st);
}

cb_timed_scope::~cb_timed_scope() {
    if (!on) return;
    cudaEventRecord(t.e1, st);
    ctx->timed.push_back(t);
}

static int timing_drain(cb_ctx* ctx) {
    CB_CUDA(ctx, cudaSetDevice(ctx->device));
    CB_CUDA(ctx, cudaDeviceSynchronize());
    for (auto& t : ctx->timed) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.e0, t.e1) == cudaSuccess) ctx->cls_ms[t.cls] += ms;
        ctx->cls_work[t.cls] += t.work;
        ctx->cls_launches[t.cls] += 1;
        ctx->event_pool.push_back(t.e0);
        ctx->event_pool.push_back(t.e1);
    }
    ctx->timed.clear();
    return CB_OK;
}

extern "C" int cb_timing_enable(cb_ctx* ctx, int on) {
    if (!ctx) return CB_EINVAL;
    ctx->timing = on != 0;
    return CB_OK;
}

extern "C" int cb_timing_read(cb_ctx* ctx, int cls, double* ms, double* work, int64_t* launches) {
    if (!ctx || cls < 0 || cls >= CB_TIMING_CLASSES) return CB_EINVAL;
    int rc = timing_drain(ctx);
    if (rc) return rc;
    double dev[CB_TIMING_CLASSES];
    CB_CUDA(ctx, cudaMemcpy(dev, ctx->dev_work, sizeof(dev), cudaMemcpyDeviceToHost));
    if (ms) *ms = ctx->cls_ms[cls];
    if (work) *work = ctx->cls_work[cls] + dev[cls];
    if (launches) *launches = ctx->cls_launches[cls];
    return CB_OK;
}

extern "C" int cb_timing_reset(cb_ctx* ctx) {
    if (!ctx) return CB_EINVAL;
    int rc = timing_drain(ctx);
    if (rc) return rc;
    for (int i = 0; i < CB_TIMING_CLASSES; ++i) { ctx->cls_ms[i] = 0; ctx->cls_work[i] = 0; ctx->cls_launches[i] = 0; }
    CB_CUDA(ctx, cudaMemset(ctx->dev_work, 0, CB_TIMING_CLASSES * sizeof(double)));
    return CB_OK;
}
