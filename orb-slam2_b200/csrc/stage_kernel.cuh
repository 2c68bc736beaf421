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

// Colour input: the cv::cvtColor(im, gray, CV_RGB2GRAY / CV_BGR2GRAY / CV_RGBA2GRAY / CV_BGRA2GRAY) prologue of
// Tracking::GrabImage* (src/Tracking.cc:177-202, 218-231, 255-266) fused into the level-0 staging: interleaved 3- or
// 4-channel rows in, the aligned gray level-0 slot out.  OpenCV's 8-bit path (pinned against cv2 4.13):
// gray = (R * 9798 + G * 19235 + B * 3735 + 16384) >> 15.  One thread per 16 output pixels: CH + 1 aligned 16-byte loads,
// funnel-shift realignment, 16 x (3 byte extracts, 3 IMAD, shift).  HBM traffic (CH + 1) * P0 bytes.
template <int CH>
__global__ void __launch_bounds__(ST_THREADS) k_stage_gray(const uint8_t *__restrict__ src, size_t srcImgStride, int srcPitch,
                                                           uint8_t *__restrict__ dst, size_t dstImgStride, int dstPitch, int w, int h,
                                                           int rFirst)
{
    const int chunksPerRow = dstPitch >> 4;
    const int b = blockIdx.y;
    const uint8_t *simg = src + (size_t)b * srcImgStride;
    const uint8_t *send = simg + (size_t)(h - 1) * srcPitch + (size_t)w * CH; // one past the last valid byte of this image
    uint8_t *dimg = dst + (size_t)b * dstImgStride;
    const unsigned int c0 = rFirst ? 9798u : 3735u, c2 = rFirst ? 3735u : 9798u;
    for (int i = blockIdx.x * ST_THREADS + threadIdx.x; i < chunksPerRow * h; i += gridDim.x * ST_THREADS) {
        const int row = i / chunksPerRow, c = i - row * chunksPerRow, x = c << 4;
        uint4 out = make_uint4(0, 0, 0, 0);
        if (x < w) {
            const uint8_t *a = simg + (size_t)row * srcPitch + (size_t)x * CH;
            const uint8_t *last = a + (size_t)min(16, w - x) * CH - 1; // last needed byte
            const size_t m = (size_t)a & 15;
            const uint8_t *a0 = a - m;
            unsigned int W[4 * (CH + 1) + 1];
#pragma unroll
            for (int j = 0; j <= CH; j++) {
                uint4 v = make_uint4(0, 0, 0, 0);
                if (a0 + 16 * j <= last && a0 + 16 * j < send) v = ldg_nc_v4(a0 + 16 * j);
                W[4 * j] = v.x; W[4 * j + 1] = v.y; W[4 * j + 2] = v.z; W[4 * j + 3] = v.w;
            }
            W[4 * (CH + 1)] = 0;
            const int ws = (int)(m >> 2), bs = (int)(m & 3) * 8;
            unsigned int o[4 * CH];
#pragma unroll
            for (int j = 0; j < 4 * CH; j++) {
                unsigned int l0 = W[j], l1 = W[j + 1];
#pragma unroll
                for (int k = 1; k < 4; k++)
                    if (ws == k) { l0 = W[j + k]; l1 = W[j + k + 1]; }
                o[j] = __funnelshift_r(l0, l1, bs);
            }
            unsigned int g[4] = {0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const int i0 = CH * k, i1 = CH * k + 1, i2 = CH * k + 2;
                const unsigned int p0 = (o[i0 >> 2] >> ((i0 & 3) * 8)) & 0xFFu, p1 = (o[i1 >> 2] >> ((i1 & 3) * 8)) & 0xFFu,
                                   p2 = (o[i2 >> 2] >> ((i2 & 3) * 8)) & 0xFFu;
                const unsigned int y = (p0 * c0 + p1 * 19235u + p2 * c2 + 16384u) >> 15;
                g[k >> 2] |= y << ((k & 3) * 8);
            }
            out = make_uint4(g[0], g[1], g[2], g[3]);
        }
        *reinterpret_cast<uint4 *>(dimg + (size_t)row * dstPitch + x) = out;
    }
}

// Frame::UndistortKeyPoints (src/Frame.cc:470-515): cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) on the N x 2
// float matrix of keypoint coordinates.  OpenCV's arithmetic (pinned against cv2 4.13, bit-exact): everything in double,
// x = (u - cx) * (1 / fx); 5 fixed-point iterations x = (x0 - deltaX) * icdist with
// icdist = (1 + ((k6 r2 + k5) r2 + k4) r2) / (1 + ((k3 r2 + k2) r2 + k1) r2)  (k4..k6 = 0 for the reference's 5 coefficients),
// deltaX = 2 p1 x y + p2 (r2 + 2 x^2); an iteration with icdist < 0 falls back to the undistorted normalised point; then
// the new projection P = K, one narrowing to float.  No FMA contraction (-fmad=false), IEEE div.rn.f64.
// cam = {fx, fy, cx, cy} and dist = {k1, k2, p1, p2, k3} are the FLOAT values of mK / mDistCoef, widened here.
struct UndistortParams {
    float fx, fy, cx, cy;
    float k1, k2, p1, p2, k3;
};

__global__ void k_undistort_keypoints(const orbb200_keypoint *__restrict__ in, orbb200_keypoint *__restrict__ out, int n,
                                      const UndistortParams U)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    orbb200_keypoint kp = in[i];
    const double fx = U.fx, fy = U.fy, cx = U.cx, cy = U.cy;
    const double ifx = 1. / fx, ify = 1. / fy;
    const double k0 = U.k1, k1 = U.k2, k2 = U.p1, k3 = U.p2, k4 = U.k3;
    const double u = kp.x, v = kp.y;
    double x = (u - cx) * ifx, y = (v - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
        const double r2 = x * x + y * y;
        const double icdist = (1 + ((0. * r2 + 0.) * r2 + 0.) * r2) / (1 + ((k4 * r2 + k1) * r2 + k0) * r2);
        if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
        const double deltaX = 2 * k2 * x * y + k3 * (r2 + 2 * x * x) + 0. * r2 + 0. * r2 * r2;
        const double deltaY = k2 * (r2 + 2 * y * y) + 2 * k3 * x * y + 0. * r2 + 0. * r2 * r2;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
    }
    const double xx = fx * x + 0. * y + cx, yy = 0. * x + fy * y + cy, ww = 1. / (0. * x + 0. * y + 1.);
    kp.x = (float)(xx * ww);
    kp.y = (float)(yy * ww);
    out[i] = kp;
}

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
