// layout_bw.cu -- how much HBM bandwidth a g2p-shaped access pattern gets from the layout alone.
// Every warp handles batches of 32 consecutive particles: reads NR 16-byte records per particle, writes NW, from /
// to (a) NR + NW separate float4 arrays (the library's SoA: a warp touches 512 contiguous bytes in each of 16 arrays),
// (b) an AoSoA of 32-particle blocks (the same 512-byte pieces, adjacent: one contiguous (NR|NW) x 512 B run per batch).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o layout_bw layout_bw.cu && ./layout_bw
#include <cstdio>
#include <cuda_runtime.h>

template <int NR, int NW, bool AOSOA>
__global__ void __launch_bounds__(128, 5) k(const float4* __restrict__ in, float4* __restrict__ out, long n, unsigned* cursor) {
    const int lane = threadIdx.x & 31;
    const long nb = n / 32;
    for (;;) {
        unsigned b = 0;
        if (lane == 0) b = atomicAdd(cursor, 8u);
        b = __shfl_sync(0xffffffffu, b, 0);
        if (b >= nb) return;
        for (unsigned bb = b; bb < b + 8 && bb < nb; ++bb) {
            float4 acc = make_float4(0, 0, 0, 0);
            float4 v[NR];
#pragma unroll
            for (int a = 0; a < NR; ++a) {
                const long idx = AOSOA ? ((long)bb * NR + a) * 32 + lane : (long)a * n + (long)bb * 32 + lane;
                v[a] = in[idx];
            }
#pragma unroll
            for (int a = 0; a < NR; ++a) { acc.x += v[a].x; acc.y += v[a].y; acc.z += v[a].z; acc.w += v[a].w; }
#pragma unroll
            for (int a = 0; a < NW; ++a) {
                const long idx = AOSOA ? ((long)bb * NW + a) * 32 + lane : (long)a * n + (long)bb * 32 + lane;
                out[idx] = make_float4(acc.x + a, acc.y, acc.z, acc.w);
            }
        }
    }
}

template <int NR, int NW, bool AOSOA>
static void run(const char* name, float4* in, float4* out, long n, unsigned* cursor) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f;
    for (int it = 0; it < 6; ++it) {
        cudaMemset(cursor, 0, 4);
        cudaEventRecord(e0);
        k<NR, NW, AOSOA><<<148 * 5, 128>>>(in, out, n, cursor);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (it && ms < best) best = ms;
    }
    printf("%-28s %7.3f ms  %7.1f GB/s\n", name, best, (double)n * 16 * (NR + NW) / best / 1e6);
}

int main() {
    const long n = 4096000;
    float4 *in, *out; unsigned* cursor;
    cudaMalloc(&in, n * 16 * 6); cudaMalloc(&out, n * 16 * 11); cudaMalloc(&cursor, 4);
    cudaMemset(in, 0, n * 16 * 6); cudaMemset(out, 0, n * 16 * 11);
    run<5, 10, false>("SoA   5 reads 10 writes", in, out, n, cursor);
    run<5, 10, true>("AoSoA 5 reads 10 writes", in, out, n, cursor);
    run<6, 0, false>("SoA   6 reads", in, out, n, cursor);
    run<6, 0, true>("AoSoA 6 reads", in, out, n, cursor);
    run<1, 10, false>("SoA   1 read 10 writes", in, out, n, cursor);
    run<1, 10, true>("AoSoA 1 read 10 writes", in, out, n, cursor);
    run<5, 5, false>("SoA   5 reads 5 writes", in, out, n, cursor);
    run<5, 5, true>("AoSoA 5 reads 5 writes", in, out, n, cursor);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
