// pipes.cu -- which pipe do VIMNMX3.S16x2 / HMNMX2 / HFMA2 issue on?  nvcc -arch=sm_100a -O3 -o pipes pipes.cu
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
template <int MODE> __global__ void k(unsigned int *out, int iters, unsigned int seed)
{
    unsigned int a[8];
    __half2 h[8];
    for (int i = 0; i < 8; i++) {
        a[i] = seed * (threadIdx.x + 1 + i);
        h[i] = __floats2half2_rn((float)((threadIdx.x + i) & 255), (float)((threadIdx.x * 3 + i) & 255));
    }
    const __half2 c = __floats2half2_rn(3.f, 5.f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0 || MODE == 2) a[i] = __vimax3_s16x2(a[i], a[(i + 1) & 7] ^ 0x10001u, seed);
            if (MODE == 1 || MODE == 2) h[i] = __hmax2(h[i], __hmin2(h[(i + 1) & 7], c));
            if (MODE == 3) h[i] = __hfma2(h[i], c, h[(i + 1) & 7]);
            if (MODE == 4) { a[i] = __vimax3_s16x2(a[i], a[(i + 1) & 7] ^ 0x10001u, seed); h[i] = __hfma2(h[i], c, h[(i + 1) & 7]); }
        }
    }
    unsigned int r = 0;
    for (int i = 0; i < 8; i++) r ^= a[i] ^ *reinterpret_cast<unsigned int *>(&h[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char *name, int opsPerIter)
{
    unsigned int *d;
    cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    k<MODE><<<148 * 8, 256>>>(d, 100, 7);
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(d, iters, 7);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double warpInstr = 148.0 * 8 * 8 * (double)iters * opsPerIter;
    printf("%-28s %.3f ms  -> %.2f warp-instr/clk/SM (at 1.965 GHz)\n", name, ms, warpInstr / (ms * 1e-3) / 148 / 1.965e9);
    cudaFree(d);
}
int main()
{
    run<0>("VIMNMX3.S16x2 (+LOP3)", 16);
    run<1>("HMNMX2 x2", 16);
    run<2>("VIMNMX3+LOP3 and HMNMX2 x2", 32);
    run<3>("HFMA2", 8);
    run<4>("VIMNMX3+LOP3 and HFMA2", 24);
    return 0;
}
