// describe_kernel.cuh -- E6 + E8 + E9: IC_Angle orientation (src/ORBextractor.cc:83-111), 256-bit rBRIEF
// (:116-157) and output assembly (:1288-1338) in one kernel, one warp per selected keypoint.
//
// Round-1 ncu of the first version: 1430 warp instructions per keypoint, long-scoreboard bound (16 scattered byte
// gathers per lane = 512 L1 sectors per keypoint), XU pipe busy with I2F/F2I.  This version:
//   * the 37-row blurred patch of the keypoint is staged in shared memory with 16-byte row loads (74 sectors), and
//     the 512 samples become LDS.U8;
//   * the pattern lives in shared memory as floats, transposed so that the lanes' float4 reads are conflict free;
//   * cvRound = add 1.5*2^23 and take the mantissa (round-half-even, exact for |x| < 2^22) instead of F2I;
//   * the orientation disc is fully unrolled with the (fixed) umax table as immediates, all 31 loads in flight;
//   * cos/sin of the angle come from one double-precision fdlibm-style sincos on [0, 2*pi], rounded to float
//     (glibc's cosf/sinf are correctly rounded).
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

constexpr int DS_WARPS = 8;
constexpr int DS_PROWS = 37, DS_PPITCH = 64; // staged blurred patch: rows y-18..y+18, 64 bytes from (x-18) & ~15

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

// sin and cos of a in [0, 2*pi] in double (fdlibm kernels after a two-term Cody-Waite reduction; error < 1e-16), to be
// rounded to float: the CPU's cosf/sinf are correctly rounded, and so is this except for astronomically rare ties.
// Cheaper than the generic sincos(double): no slow-path range reduction, no stack frame.
__device__ __forceinline__ void sincos_0_2pi(double x, double *sn, double *cs)
{
    const int k = __double2int_rn(x * 6.36619772367581382433e-01);
    double r = __fma_rn(-(double)k, 1.57079632673412561417e+00, x);
    r = __fma_rn(-(double)k, 6.07710050650619224932e-11, r);
    const double z = r * r;
    double ps = __fma_rn(z, 1.58969099521155010221e-10, -2.50507602534068634195e-08);
    ps = __fma_rn(z, ps, 2.75573137070700676789e-06);
    ps = __fma_rn(z, ps, -1.98412698298579493134e-04);
    ps = __fma_rn(z, ps, 8.33333333332248946124e-03);
    ps = __fma_rn(z, ps, -1.66666666666666324348e-01);
    const double s = __fma_rn(z * r, ps, r);
    double pc = __fma_rn(z, -1.13596475577881948265e-11, 2.08757232129817482790e-09);
    pc = __fma_rn(z, pc, -2.75573143513906633035e-07);
    pc = __fma_rn(z, pc, 2.48015872894767294178e-05);
    pc = __fma_rn(z, pc, -1.38888888888741095749e-03);
    pc = __fma_rn(z, pc, 4.16666666666666019037e-02);
    const double c = __fma_rn(z * z, pc, __fma_rn(z, -0.5, 1.0));
    const int q = k & 3;
    *sn = (q == 0) ? s : (q == 1) ? c : (q == 2) ? -s : -c;
    *cs = (q == 0) ? c : (q == 1) ? -s : (q == 2) ? -c : s;
}

// cvRound for |v| < 2^22: round-half-to-even through the float adder
__device__ __forceinline__ int cv_round_magic(float v)
{
    return __float_as_int(__fadd_rn(v, 12582912.0f)) - 0x4B400000;
}

__global__ void __launch_bounds__(DS_WARPS * 32) k_orient_describe(const __grid_constant__ ExtractParams P)
{
    __shared__ __align__(16) float4 sPat[8 * 32];                               // [bit k of the byte][lane]
    __shared__ __align__(16) uint8_t sPatch[DS_WARPS][DS_PROWS * DS_PPITCH];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        // pattern pair i belongs to descriptor byte i / 8 (= lane), bit i % 8
        sPat[(i & 7) * 32 + (i >> 3)] = make_float4((float)c_pattern[i][0], (float)c_pattern[i][1], (float)c_pattern[i][2], (float)c_pattern[i][3]);
    }
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

    // ---- stage the blurred 37 x 37 patch (rows y-18..y+18, 16-byte chunks from (x-18) & ~15): issued first so that
    //      its latency overlaps the orientation below
    const uint8_t *bl = P.blur + (size_t)b * P.pyrStride + g.imgOff;
    const int px0 = (x - 18) & ~15;
    uint8_t *patch = sPatch[wid];
    {
        uint4 v[5];
#pragma unroll
        for (int i = 0; i < 5; i++) {
            const int idx = lane + 32 * i; // chunk index: row = idx / 4, chunk = idx % 4
            const int r = idx >> 2, ch = idx & 3;
            const int gx = px0 + (ch << 4);
            v[i] = make_uint4(0, 0, 0, 0);
            if (idx < DS_PROWS * 4 && gx < g.pitch) v[i] = ldg_nc_v4(bl + (size_t)(y - 18 + r) * g.pitch + gx);
        }
#pragma unroll
        for (int i = 0; i < 5; i++) {
            const int idx = lane + 32 * i;
            if (idx < DS_PROWS * 4) *reinterpret_cast<uint4 *>(patch + (idx >> 2) * DS_PPITCH + ((idx & 3) << 4)) = v[i];
        }
    }

    // ---- E6: intensity-centroid moments over the radius-15 disc of the UN-blurred level; lane <-> column u
    int pitch, valid;
    const uint8_t *img = level_ptr(P, g, l, b, pitch, valid);
    int m10 = 0, m01 = 0;
    {
        // umax of src/ORBextractor.cc:510-533 for HALF_PATCH_SIZE = 15 (orbb200_compute_level_tables reproduces it)
        constexpr int UM[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
        const int u = lane - 15;
        const int au = abs(u);
        const uint8_t *c = img + (size_t)y * pitch + x + u;
        int vals[31];
#pragma unroll
        for (int v = -15; v <= 15; v++) {
            const int av = v < 0 ? -v : v;
            vals[v + 15] = (lane < 31 && au <= UM[av]) ? (int)__ldg(c + v * pitch) : 0;
        }
        int s = 0;
#pragma unroll
        for (int v = -15; v <= 15; v++) {
            s += vals[v + 15];
            m01 += v * vals[v + 15];
        }
        m10 = u * s;
    }
    m10 = warp_sum(m10);
    m01 = warp_sum(m01);
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- E8: steered BRIEF on the staged blurred patch; lane <-> descriptor byte
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    const float ang = __fmul_rn(angle, factorPI);
    // libm cosf/sinf on the CPU are (almost always) the correctly rounded values: take them through double
    double sd, cd;
    sincos_0_2pi((double)ang, &sd, &cd);
    const float a = (float)cd, bsn = (float)sd;
    __syncwarp();
    const uint8_t *pc = patch + 18 * DS_PPITCH + (x - px0); // patch centre
    int val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float4 p = sPat[k * 32 + lane];
        const int r0 = cv_round_magic(__fadd_rn(__fmul_rn(p.x, bsn), __fmul_rn(p.y, a)));
        const int c0 = cv_round_magic(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, bsn)));
        const int r1 = cv_round_magic(__fadd_rn(__fmul_rn(p.z, bsn), __fmul_rn(p.w, a)));
        const int c1 = cv_round_magic(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, bsn)));
        const int t0 = pc[r0 * DS_PPITCH + c0], t1 = pc[r1 * DS_PPITCH + c1];
        val |= (t0 < t1) << k;
    }
    P.desc[((size_t)b * P.cap + outIdx) * 32 + lane] = (uint8_t)val;

    // ---- E9: keypoint record (== cv::KeyPoint); pt scaled to level-0 coordinates for level > 0 (:1329-1333)
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
