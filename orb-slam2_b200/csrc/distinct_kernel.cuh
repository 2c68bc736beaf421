// distinct_kernel.cuh -- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:266-340), batched over map points.
//
// Reference, per map point with N observed descriptors: Distances[N][N] by ORBmatcher::DescriptorDistance (:306-316,
// diagonal 0), every row sorted, median = vDists[0.5 * (N - 1)] (:327-331, index truncated = (N - 1) >> 1), the FIRST row
// with the strictly smallest median wins (:333-337).
//
// sm_100a formulation: one warp per map point (groups are tiny: N is the number of keyframes that observe the point).
//   * N <= 32: lane j keeps descriptor j in registers; row i's words arrive by shuffle; the (N-1)/2-th order statistic of
//     the 32 lane values is a 9-step binary search over the value range [0, 256] with one ballot per step.
//   * N > 32: distances of a row go into a per-warp 288-bin shared-memory histogram (bin = distance), the order statistic
//     is the first bin whose inclusive prefix reaches (N-1)/2 + 1 (9 bins per lane + one warp scan).
// Both are exact integer work: results are bit-identical to the sort.
#pragma once
#include "common.cuh"

namespace orbb200 {

constexpr int DD_WARPS = 8;
constexpr int DD_BINS = 288; // 257 distance values rounded up to 9 per lane

__device__ __forceinline__ int dd_popc8(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1)
{
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__device__ __forceinline__ uint4 dd_shfl4(const uint4 v, int src)
{
    return make_uint4(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src),
                      __shfl_sync(0xffffffffu, v.z, src), __shfl_sync(0xffffffffu, v.w, src));
}

// desc: rows of 32 bytes, 16-byte aligned; group g owns rows [offsets[g], offsets[g + 1])
__global__ void __launch_bounds__(DD_WARPS * 32) k_distinctive(const uint4 *__restrict__ desc, const int *__restrict__ offsets,
                                                               int ngroups, int *__restrict__ bestIdx,
                                                               int *__restrict__ bestMedian)
{
    __shared__ int sHist[DD_WARPS][DD_BINS];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = blockIdx.x * DD_WARPS + wid;
    if (g >= ngroups) return;
    const int o = offsets[g], N = offsets[g + 1] - o;
    if (N <= 0) { // observations.empty() / vDescriptors.empty(): the reference returns without touching mDescriptor
        if (lane == 0) { bestIdx[g] = -1; bestMedian[g] = 0; }
        return;
    }
    const int k1 = ((N - 1) >> 1) + 1; // the median is the k1-th smallest (1-based)
    int best = 0x7fffffff, bestI = 0;
    if (N <= 32) {
        uint4 m0 = make_uint4(0, 0, 0, 0), m1 = m0;
        if (lane < N) { m0 = desc[2 * (size_t)(o + lane)]; m1 = desc[2 * (size_t)(o + lane) + 1]; }
        for (int i = 0; i < N; i++) {
            const uint4 r0 = dd_shfl4(m0, i), r1 = dd_shfl4(m1, i);
            const int d = lane < N ? dd_popc8(m0, m1, r0, r1) : 1024; // idle lanes never count
            int lo = 0, hi = 256; // smallest v with #{d <= v} >= k1
#pragma unroll
            for (int s = 0; s < 9; s++) {
                const int mid = (lo + hi) >> 1;
                const int c = __popc(__ballot_sync(0xffffffffu, d <= mid));
                if (c >= k1) hi = mid; else lo = mid + 1;
            }
            if (lo < best) { best = lo; bestI = i; }
        }
    } else {
        int *h = sHist[wid];
        for (int i = 0; i < N; i++) {
#pragma unroll
            for (int b = 0; b < DD_BINS / 32; b++) h[b * 32 + lane] = 0;
            __syncwarp();
            const uint4 r0 = desc[2 * (size_t)(o + i)], r1 = desc[2 * (size_t)(o + i) + 1];
            for (int j = lane; j < N; j += 32) {
                const uint4 c0 = desc[2 * (size_t)(o + j)], c1 = desc[2 * (size_t)(o + j) + 1];
                atomicAdd(&h[dd_popc8(r0, r1, c0, c1)], 1);
            }
            __syncwarp();
            int c[DD_BINS / 32], sum = 0;
#pragma unroll
            for (int b = 0; b < DD_BINS / 32; b++) { c[b] = h[lane * (DD_BINS / 32) + b]; sum += c[b]; }
            int inc = sum;
#pragma unroll
            for (int s = 1; s < 32; s <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, s);
                if (lane >= s) inc += t;
            }
            const unsigned int reach = __ballot_sync(0xffffffffu, inc >= k1);
            const int src = __ffs(reach) - 1; // the lane whose bins hold the k1-th value
            int med = 0;
            if (lane == src) {
                int run = inc - sum;
#pragma unroll
                for (int b = 0; b < DD_BINS / 32; b++) {
                    if (run < k1 && run + c[b] >= k1) med = lane * (DD_BINS / 32) + b;
                    run += c[b];
                }
            }
            med = __shfl_sync(0xffffffffu, med, src);
            if (med < best) { best = med; bestI = i; }
            __syncwarp();
        }
    }
    if (lane == 0) { bestIdx[g] = bestI; bestMedian[g] = best; }
}

} // namespace orbb200
