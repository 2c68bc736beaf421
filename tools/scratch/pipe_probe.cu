// pipe_probe.cu -- which issue pipe do the packed min/max flavours use on sm_100a?  Times dependent-free chains of
// VIMNMX3.S16x2, HMNMX2 (half2 min/max) and both interleaved: if the mixed kernel takes max(a, b) rather than a + b the two
// run on different pipes.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_probe tools/scratch/pipe_probe.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdio.h>

__device__ __forceinline__ unsigned hmax2u(unsigned a, unsigned b)
{
    unsigned r;
    asm volatile("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ unsigned hmin2u(unsigned a, unsigned b)
{
    unsigned r;
    asm volatile("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ unsigned imax3(unsigned a, unsigned b, unsigned c)
{
    return __vimax3_s16x2(a, b, c);
}
__device__ __forceinline__ unsigned imin3(unsigned a, unsigned b, unsigned c)
{
    return __vimin3_s16x2(a, b, c);
}
__device__ __forceinline__ unsigned imax2(unsigned a, unsigned b)
{
    return __vmaxs2(a, b);
}

template <int MODE> __global__ void k(unsigned *out, int iters)
{
    unsigned a[8], h[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        a[i] = (threadIdx.x * 2654435761u + i * 40503u) & 0x00FF00FFu | 0x64006400u;
        h[i] = (threadIdx.x * 1664525u + i * 1013904223u) & 0x00FF00FFu | 0x64006400u;
    }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0 || MODE == 2) { a[i] = imax3(a[i], a[(i + 1) & 7], a[(i + 3) & 7]); a[i] = imin3(a[i], a[(i + 2) & 7], a[(i + 5) & 7]); }
            if (MODE == 1 || MODE == 2) { h[i] = hmax2u(h[i], h[(i + 1) & 7]); h[i] = hmin2u(h[i], h[(i + 3) & 7]); }
            if (MODE == 3) { a[i] = imax2(a[i], a[(i + 1) & 7]); a[i] = imax2(a[i], a[(i + 3) & 7]); }
            if (MODE == 4) { // 2-input int + half mixed
                a[i] = imax2(a[i], a[(i + 1) & 7]); a[i] = imax2(a[i], a[(i + 3) & 7]);
                h[i] = hmax2u(h[i], h[(i + 1) & 7]); h[i] = hmin2u(h[i], h[(i + 3) & 7]);
            }
        }
    }
    unsigned acc = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) acc += a[i] ^ h[i];
    if (acc == 0x12345u) out[0] = acc;
}

template <int MODE> static float run(unsigned *d, int iters)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e9f;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0);
        k<MODE><<<148 * 8, 256>>>(d, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r > 0 && ms < best) best = ms;
    }
    return best;
}

int main()
{
    unsigned *d;
    cudaMalloc(&d, 256);
    const int iters = 4096;
    const double warpInstr = 148.0 * 8 * 8 * iters * 16; // 16 instructions of each flavour per iteration per warp
    const float t0 = run<0>(d, iters), t1 = run<1>(d, iters), t2 = run<2>(d, iters), t3 = run<3>(d, iters), t4 = run<4>(d, iters);
    printf("VIMNMX3.S16x2 only : %.3f ms  (%.2f warp-instr/clk/SM at 1.965 GHz)\n", t0, warpInstr / (t0 * 1e-3) / 148 / 1.965e9);
    printf("HMNMX2 only        : %.3f ms  (%.2f)\n", t1, warpInstr / (t1 * 1e-3) / 148 / 1.965e9);
    printf("both interleaved   : %.3f ms  (sum %.3f, max %.3f)\n", t2, t0 + t1, t0 > t1 ? t0 : t1);
    printf("VIMNMX.S16x2 2-in  : %.3f ms  (%.2f)\n", t3, warpInstr / (t3 * 1e-3) / 148 / 1.965e9);
    printf("2-in int + half    : %.3f ms  (sum %.3f)\n", t4, t3 + t1);
    return 0;
}
