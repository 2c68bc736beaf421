// describe_kernel.cuh -- E6 + E8 + E9: IC_Angle orientation (src/ORBextractor.cc:83-111), 256-bit rBRIEF
// (:116-157) and output assembly (:1288-1338) in one kernel, one warp per selected keypoint.
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

constexpr int DS_WARPS = 8;

__constant__ signed char c_pattern[256][4] = {
#include "orb_pattern.inc"
};

// cv::fastAtan2 (OpenCV 4.x scalar path), evaluated with explicit round-to-nearest ops: no FMA contraction.
__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    const float eps = (float)2.2204460492503131e-16;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

__global__ void __launch_bounds__(DS_WARPS * 32) k_orient_describe(const __grid_constant__ ExtractParams P)
{
    __shared__ signed char sPat[1024];
    for (int i = threadIdx.x; i < 256; i += blockDim.x)
        reinterpret_cast<int *>(sPat)[i] = reinterpret_cast<const int *>(c_pattern)[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int slot = blockIdx.x * DS_WARPS + wid, b = blockIdx.y;
    // per-level counts of this image -> output offset of each level (levels concatenated 0..L-1, :1302-1337)
    const int cnt = lane < P.nlevels ? P.lvlCount[(size_t)b * P.nlevels + lane] : 0;
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    const int total = __shfl_sync(0xffffffffu, inc, 31);
    if (blockIdx.x == 0 && wid == 0 && lane == 0) P.nOut[b] = total;
    if (slot >= P.nodeTotal) return;
    int l = 0;
    while (l + 1 < P.nlevels && slot >= P.geom[l + 1].nodeOff) l++;
    const LevelGeom g = P.geom[l];
    const int pos = slot - g.nodeOff;
    const int cntL = __shfl_sync(0xffffffffu, cnt, l), offL = __shfl_sync(0xffffffffu, inc - cnt, l);
    if (pos >= cntL) return;
    const int outIdx = offL + pos;
    if (outIdx >= P.cap) return;
    const uint32_t packed = P.lvlKp[(size_t)b * P.nodeTotal + slot];
    const int x = (int)(packed & 0xFFF) + 16, y = (int)((packed >> 12) & 0xFFF) + 16;
    const int resp = (int)(packed >> 24);

    // E6: intensity-centroid moments over the radius-15 disc of the UN-blurred level; lane <-> column u
    int pitch, valid;
    const uint8_t *img = level_ptr(P, g, l, b, pitch, valid);
    int m10 = 0, m01 = 0;
    {
        const int u = lane - 15;
        const int au = abs(u);
        const uint8_t *c = img + (size_t)y * pitch + x + u;
        if (lane < 31) {
#pragma unroll 1
            for (int v = -15; v <= 15; v++) {
                if (au <= __ldg(P.umax + abs(v))) {
                    const int val = __ldg(c + v * pitch);
                    m10 += u * val;
                    m01 += v * val;
                }
            }
        }
    }
    m10 = warp_sum(m10);
    m01 = warp_sum(m01);
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // E8: steered BRIEF on the blurred level; lane <-> descriptor byte
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    const float ang = __fmul_rn(angle, factorPI);
    // libm cosf/sinf on the CPU are (almost always) the correctly rounded values: take them through double
    const float a = (float)cos((double)ang), bsn = (float)sin((double)ang);
    const uint8_t *bl = P.blur + (size_t)b * P.pyrStride + g.imgOff + (size_t)y * g.pitch + x;
    int val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const signed char *p = sPat + (lane * 8 + k) * 4;
        const float x0 = (float)p[0], y0 = (float)p[1], x1 = (float)p[2], y1 = (float)p[3];
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, bsn), __fmul_rn(y0, a)));
        const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, bsn)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, bsn), __fmul_rn(y1, a)));
        const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, bsn)));
        const int t0 = __ldg(bl + r0 * g.pitch + c0), t1 = __ldg(bl + r1 * g.pitch + c1);
        val |= (t0 < t1) << k;
    }
    P.desc[((size_t)b * P.cap + outIdx) * 32 + lane] = (uint8_t)val;

    // E9: keypoint record (== cv::KeyPoint); pt scaled to level-0 coordinates for level > 0 (:1329-1333)
    if (lane < 7) {
        float f;
        const float fx = l ? __fmul_rn((float)x, g.scale) : (float)x;
        const float fy = l ? __fmul_rn((float)y, g.scale) : (float)y;
        switch (lane) {
        case 0: f = fx; break;
        case 1: f = fy; break;
        case 2: f = g.patch; break;
        case 3: f = angle; break;
        case 4: f = (float)resp; break;
        case 5: f = __int_as_float(l); break;
        default: f = __int_as_float(-1); break;
        }
        reinterpret_cast<float *>(P.kps + (size_t)b * P.cap + outIdx)[lane] = f;
    }
}

} // namespace orbb200
