// FP64 / FP32 FMA issue rate per SM on the box's GPU:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/fp64_rate scripts/microbench/fp64_rate.cu && /tmp/fp64_rate
#include <cstdio>
#include <cuda_runtime.h>
template <typename T>
__global__ void fma_chain(T* out, int iters, T a, T b) {
    T x[8];
    for (int i = 0; i < 8; ++i) x[i] = (T)(threadIdx.x + i);
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = x[i] * a + b;
    T s = 0;
    for (int i = 0; i < 8; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename T>
double run(const char* name, int warps_per_sm) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int threads = 256, blocks = sms * warps_per_sm / 8, iters = 4096;
    T* out;
    cudaMalloc(&out, sizeof(T) * threads * blocks);
    fma_chain<T><<<blocks, threads>>>(out, 16, (T)1.0001, (T)0.5);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    fma_chain<T><<<blocks, threads>>>(out, iters, (T)1.0001, (T)0.5);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fmas = (double)blocks * threads * iters * 8;
    const double per_clk_sm = fmas / (ms * 1e-3) / (clk_khz * 1e3) / sms;
    printf("%s: %d warps/SM: %.3f ms, %.1f FMA lanes per (max-)clock per SM, %.2f TFLOP/s\n", name, warps_per_sm, ms, per_clk_sm,
           2 * fmas / (ms * 1e-3) / 1e12);
    cudaFree(out);
    return per_clk_sm;
}
int main() {
    run<float>("fp32", 32);
    run<double>("fp64", 8);
    run<double>("fp64", 32);
    return 0;
}
