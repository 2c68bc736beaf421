// common.cuh -- shared host/device helpers of liborbb200 (sm_100a only; no CPU fallback).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/orbb200.h"

namespace orbb200 {

void set_error(const char *fmt, ...);

#define ORB_CUDA_OK(expr)                                                                        \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess) {                                                                 \
            orbb200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return ORBB200_ERR_CUDA;                                                             \
        }                                                                                        \
    } while (0)

#define ORB_ARG_OK(cond, msg)                                    \
    do {                                                         \
        if (!(cond)) {                                           \
            orbb200::set_error("%s: %s", __func__, msg);         \
            return ORBB200_ERR_ARG;                              \
        }                                                        \
    } while (0)

static inline int div_up(int a, int b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#ifdef __CUDACC__
// block-wide exclusive scan of one int per thread; returns the exclusive prefix, *total = block sum.
// tmp: >= 33 ints of shared memory.  All threads of the block must call it.
__device__ __forceinline__ int block_excl_scan(int v, int *tmp, int *total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads(); // protect tmp from a previous call
    if (lane == 31) tmp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = lane < nw ? tmp[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        if (lane < nw) tmp[lane] = winc - w;
        if (lane == 31) tmp[32] = winc;
    }
    __syncthreads();
    *total = tmp[32];
    return tmp[wid] + inc - v;
}

__device__ __forceinline__ int warp_sum(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
#endif

} // namespace orbb200
