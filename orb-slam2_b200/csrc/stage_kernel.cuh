// stage_kernel.cuh -- level-0 staging for the device API: a caller's image whose base or row pitch is not 16-byte
// aligned (cv::Mat rows of 1241 bytes, ROIs) is copied once into the level-0 slot of the handle's pyramid buffer
// (pitch % 16 == 0), so that every later kernel takes its 16-byte vector path.  Pure HBM copy: 2 * P0 bytes.
// Reads are aligned 16-byte loads of the two chunks that straddle each output chunk, recombined with funnel shifts.
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

#ifdef ORBB200_EXTRACT_IMPL

constexpr int ST_THREADS = 256;

__global__ void __launch_bounds__(ST_THREADS) k_stage_level0(const uint8_t *__restrict__ src, size_t srcImgStride, int srcPitch,
                                                             uint8_t *__restrict__ dst, size_t dstImgStride, int dstPitch,
                                                             int w, int h)
{
    const int chunksPerRow = dstPitch >> 4;
    const int b = blockIdx.y;
    const uint8_t *simg = src + (size_t)b * srcImgStride;
    const uint8_t *send = simg + (size_t)(h - 1) * srcPitch + w; // one past the last valid byte of this image
    uint8_t *dimg = dst + (size_t)b * dstImgStride;
    for (int i = blockIdx.x * ST_THREADS + threadIdx.x; i < chunksPerRow * h; i += gridDim.x * ST_THREADS) {
        const int row = i / chunksPerRow, c = i - row * chunksPerRow, x = c << 4;
        uint4 out = make_uint4(0, 0, 0, 0);
        if (x < w) {
            const uint8_t *a = simg + (size_t)row * srcPitch + x;
            const uint8_t *last = a + min(15, w - 1 - x); // last needed byte
            const size_t m = (size_t)a & 15;
            const uint8_t *a0 = a - m;
            unsigned int W[9];
            const uint4 lo = ldg_nc_v4(a0);
            W[0] = lo.x; W[1] = lo.y; W[2] = lo.z; W[3] = lo.w;
            uint4 hi = make_uint4(0, 0, 0, 0);
            if (m != 0 && a0 + 16 <= last && a0 + 16 < send) hi = ldg_nc_v4(a0 + 16);
            W[4] = hi.x; W[5] = hi.y; W[6] = hi.z; W[7] = hi.w; W[8] = 0;
            const int ws = (int)(m >> 2), bs = (int)(m & 3) * 8;
            unsigned int o[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                unsigned int l0 = W[j], l1 = W[j + 1];
#pragma unroll
                for (int k = 1; k < 4; k++)
                    if (ws == k) { l0 = W[j + k]; l1 = W[j + k + 1]; }
                o[j] = __funnelshift_r(l0, l1, bs);
            }
            out = make_uint4(o[0], o[1], o[2], o[3]);
        }
        *reinterpret_cast<uint4 *>(dimg + (size_t)row * dstPitch + x) = out;
    }
}

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
