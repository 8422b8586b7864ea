// Issue-rate microbenchmark: scalar FFMA against packed FFMA2 (fma.rn.f32x2), plain and with a scalar-broadcast
// operand, 8 independent accumulator chains per thread, 8 warps per SM sub-partition.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_rate ffma2_rate.cu && ./ffma2_rate
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t pk(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float2 upk(uint64_t r) { float2 d; asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(r)); return d; }

template <int MODE>
__global__ void __launch_bounds__(1024) k(float* out, float s, int iters) {
    float2 acc[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[q] = make_float2(threadIdx.x * 1e-3f + q, q * 0.5f);
    const float2 b = make_float2(s, s * 0.5f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                if (MODE == 0) { acc[q].x = fmaf(acc[q].x, b.x, b.y); acc[q].y = fmaf(acc[q].y, b.y, b.x); }
                if (MODE == 1) { uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(pk(acc[q].x, acc[q].y)), "l"(pk(b.x, b.y)), "l"(pk(b.y, b.x))); acc[q] = upk(d); }
                if (MODE == 2) { uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(pk(s, s)), "l"(pk(acc[q].x, acc[q].y)), "l"(pk(b.y, b.x))); acc[q] = upk(d); }
            }
    }
    float r = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) r += acc[q].x + acc[q].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
static void run(const char* name, float* out, int sms) {
    const int iters = 4096;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<sms, 1024>>>(out, 0.999f, 16);
    cudaEventRecord(e0);
    k<MODE><<<sms, 1024>>>(out, 0.999f, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double fma_per_thread = (double)iters * 64 * 2;   // scalar-equivalent FMAs
    const double tf = fma_per_thread * 1024 * sms * 2 / (ms * 1e-3) / 1e12;
    printf("%-28s %8.3f ms  %6.1f TFLOP/s fp32 (scalar-equivalent)\n", name, ms, tf);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    float* out; cudaMalloc(&out, (size_t)p.multiProcessorCount * 1024 * 4);
    run<0>("FFMA  (2 per pair)", out, p.multiProcessorCount);
    run<1>("FFMA2 pair operands", out, p.multiProcessorCount);
    run<2>("FFMA2 scalar broadcast", out, p.multiProcessorCount);
    return 0;
}
