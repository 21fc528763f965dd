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
    *out = ctx;
    return CB_OK;
}

extern "C" int cb_destroy(cb_ctx* ctx) {
    if (!ctx) return CB_EINVAL;
    cudaSetDevice(ctx->device);
    if (ctx->arena) cudaFree(ctx->arena);
    delete ctx;
    return CB_OK;
}

extern "C" const char* cb_last_error(const cb_ctx* ctx) { return ctx ? ctx->err : "null context"; }

extern "C" int64_t cb_launch_count(const cb_ctx* ctx) { return ctx ? ctx->launches : 0; }

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
