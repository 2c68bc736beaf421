// describe_kernel.cuh -- E6 + E8 + E9: IC_Angle orientation (src/ORBextractor.cc:83-111), 256-bit rBRIEF
// (:116-157) and output assembly (:1288-1338) in one kernel; a warp walks kpw selected keypoints.
//
// Round-1 ncu of the first version: 1430 warp instructions per keypoint, long-scoreboard bound (16 scattered byte
// gathers per lane = 512 L1 sectors per keypoint), XU pipe busy with I2F/F2I; second version (smem-staged blurred
// patch): 1000 instructions, half of them the 31 predicated byte loads per lane of the orientation disc and a
// per-keypoint prologue (pattern fill, level search, geometry load).  This version:
//   * block prologue once per 32 keypoints: pattern to smem, per-level tables held one level per lane (level of a slot
//     = one ballot + popc, its geometry = shuffles), disc masks built in registers;
//   * BOTH patches of a keypoint are staged in shared memory with 16-byte row loads (blurred 37 x 64, un-blurred
//     31 x 48): 8 loads per lane in flight instead of 31 + 16 dependent byte gathers;
//   * orientation: lane <-> disc ROW; the 31 bytes of the row are 3 conflict-free LDS.128, aligned by one funnel
//     shift per word, masked to the disc, and reduced with DP4A (row sum, and column-weighted sum for m10) --
//     16 DP4A replace 31 x (load, predicate, 2 multiply-adds);
//   * the pattern lives in shared memory as floats, transposed so that the lanes' float4 reads are conflict free;
//   * cvRound = add 1.5*2^23 and take the mantissa (round-half-even, exact for |x| < 2^22) instead of F2I;
//   * cos/sin of the angle come from one double-precision fdlibm-style sincos on [0, 2*pi], rounded to float
//     (glibc's cosf/sinf are correctly rounded).
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

constexpr int DS_WARPS = 8;                   // warps per block
constexpr int DS_KPW_MAX = 16, DS_KPW_MIN = 2; // keypoint slots per warp: 16 for batches (block prologue amortised, the next keypoint's patches always in
                                              // flight: 0.70 ms at 4 -> 0.58 ms at 16 per 320 KITTI images), fewer when that would leave SMs without a block
constexpr int DS_PROWS = 37, DS_PPITCH = 64; // staged blurred patch: rows y-18..y+18, 64 bytes from (x-18) & ~15
constexpr int DS_RROWS = 32, DS_RPITCH = 48; // staged un-blurred patch: rows y-15..y+15 (+1 pad), 48 bytes from (x-15) & ~15

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

// 16-byte asynchronous copy global -> shared, bypassing L1 (each patch is read once); !in: 16 zero bytes, nothing is read
__device__ __forceinline__ void cp_async16_zfill(void *smem, const void *gmem, bool in)
{
    const unsigned int sa = (unsigned int)__cvta_generic_to_shared(smem);
    const int n = in ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sa), "l"(gmem), "r"(n));
}

__device__ __forceinline__ void ds_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void ds_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
constexpr size_t DS_SMEM = 8 * 32 * 16 + (size_t)DS_WARPS * (2 * DS_PROWS * DS_PPITCH + DS_RROWS * DS_RPITCH);

__device__ __forceinline__ int dp4a_us(unsigned int a, unsigned int b, int c)
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// m10 / m01 of the radius-15 disc from the staged un-blurred patch; lane <-> row v = lane - 15.
// KW = word offset of the row's first disc byte inside the lane's 12 staged words (warp-uniform, 0..3).
template <int KW> __device__ __forceinline__ void disc_row_sums(const uint4 &q0, const uint4 &q1, const uint4 &q2, unsigned int sb,
                                                                const unsigned int (&mk)[8], int &s, int &t)
{
    const unsigned int w[12] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w};
    s = 0; t = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const unsigned int lo = w[KW + j], hi = (KW + j + 1 < 12) ? w[KW + j + 1] : 0u;
        const unsigned int px = __funnelshift_r(lo, hi, sb) & mk[j];
        // byte k of the row is column u = k - 15: weights (4j-15, 4j-14, 4j-13, 4j-12) as signed bytes
        const unsigned int wt = (unsigned int)((4 * j - 15) & 0xFF) | ((unsigned int)((4 * j - 14) & 0xFF) << 8) |
                                ((unsigned int)((4 * j - 13) & 0xFF) << 16) | ((unsigned int)((4 * j - 12) & 0xFF) << 24);
        s = (int)__dp4a(px, 0x01010101u, (unsigned int)s);
        t = dp4a_us(px, wt, t);
    }
}

__global__ void __launch_bounds__(DS_WARPS * 32, 4) k_orient_describe(const __grid_constant__ ExtractParams P, int kpw)
{
    extern __shared__ __align__(16) unsigned char ds_smem[];
    float4 *sPat = reinterpret_cast<float4 *>(ds_smem);                                // [bit k of the byte][lane]
    // per warp: blurred patch double buffered (this keypoint / the next one in flight); ONE un-blurred patch -- it is only read at the
    // start of a keypoint (orientation), so the next keypoint's copy is issued right after that, under the descriptor computation
    uint8_t *sPatchW = ds_smem + 8 * 32 * sizeof(float4) + (size_t)(threadIdx.x >> 5) * (2 * DS_PROWS * DS_PPITCH + DS_RROWS * DS_RPITCH);
    uint8_t *sRawW = sPatchW + 2 * DS_PROWS * DS_PPITCH;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        // pattern pair i belongs to descriptor byte i / 8 (= lane), bit i % 8
        sPat[(i & 7) * 32 + (i >> 3)] = make_float4((float)c_pattern[i][0], (float)c_pattern[i][1], (float)c_pattern[i][2], (float)c_pattern[i][3]);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.y;
    // ---- per-level tables, one level per lane: node slot range, keypoint count, output offset, image slot, scale
    int tNodeOff = 0x7fffffff, tCnt = 0, tPitch = 0;
    unsigned int tImgOff = 0;
    float tScale = 1.f, tPatch = 31.f;
    if (lane < P.nlevels) {
        const LevelGeom *g = P.geom + lane;
        tNodeOff = g->nodeOff; tPitch = g->pitch; tImgOff = g->imgOff; tScale = g->scale; tPatch = g->patch;
        tCnt = P.lvlCount[(size_t)b * P.nlevels + lane];
    }
    int inc = tCnt; // levels concatenated 0..L-1 in the output (:1302-1337)
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    const int tOutOff = inc - tCnt;
    if (blockIdx.x == 0 && wid == 0 && lane == 31) P.nOut[b] = inc;
    // ---- disc mask of this lane's row v = lane - 15: bytes k with |k - 15| <= umax[|v|]
    //      (umax of src/ORBextractor.cc:510-533 for HALF_PATCH_SIZE = 15; orbb200_compute_level_tables reproduces it)
    unsigned int mk[8];
    {
        constexpr unsigned long long UMP = 0x3689ABCDDEEEFFFFull; // umax[0..15] = 15 15 15 15 14 14 14 13 13 12 11 10 9 8 6 3, 4 bits each
        const int av = abs(lane - 15);
        const int um = av < 16 ? (int)((UMP >> (4 * av)) & 15) : -1;
        const unsigned int bits = um >= 0 ? (((1u << (2 * um + 1)) - 1u) << (15 - um)) : 0u;
#pragma unroll
        for (int j = 0; j < 8; j++) mk[j] = ((((bits >> (4 * j)) & 15u) * 0x00204081u) & 0x01010101u) * 0xFFu;
    }
    __syncthreads();

    // ---- a warp walks its DS_KPW keypoint slots with the patches of the NEXT keypoint in flight (cp.async, 16 bytes per lane and
    //      request, straight into the other half of the warp's double-buffered shared-memory patches) while it works on this one:
    //      the kernel used to spend a third of its issue slots waiting for the eight staging loads of each keypoint
    struct Kp { int valid, l, outIdx, x, y, resp; };
    auto decode = [&](int slot) {
        Kp k;
        k.valid = 0; k.l = 0; k.outIdx = 0; k.x = 0; k.y = 0; k.resp = 0;
        if (slot >= P.nodeTotal) return k;
        const int l = __popc(__ballot_sync(0xffffffffu, slot >= tNodeOff)) - 1; // nodeOff ascends with the level
        const int pos = slot - __shfl_sync(0xffffffffu, tNodeOff, l);
        const int cnt = __shfl_sync(0xffffffffu, tCnt, l), oo = __shfl_sync(0xffffffffu, tOutOff, l);
        if (pos >= cnt || oo + pos >= P.cap) return k;
        const uint32_t packed = P.lvlKp[(size_t)b * P.nodeTotal + slot];
        k.valid = 1; k.l = l; k.outIdx = oo + pos;
        k.x = (int)(packed & 0xFFF) + 16; k.y = (int)((packed >> 12) & 0xFFF) + 16; k.resp = (int)(packed >> 24);
        return k;
    };
    // blurred 37 x 37 patch (rows y-18..y+18, 16-byte chunks from (x-18) & ~15) and un-blurred 31 x 31 one (rows y-15..y+15 from
    // (x-15) & ~15, row 31 = zeros); chunks beyond the row pitch are zero-filled (src-size 0)
    auto prefetch_blur = [&](const Kp &k, uint8_t *patch) {
        const int gpitch = __shfl_sync(0xffffffffu, tPitch, k.l);
        const uint8_t *bl = P.blur + (size_t)b * P.pyrStride + __shfl_sync(0xffffffffu, tImgOff, k.l);
        const int px0 = (k.x - 18) & ~15;
#pragma unroll
        for (int i = 0; i < 5; i++) {
            const int idx = lane + 32 * i; // chunk index: row = idx / 4, chunk = idx % 4
            const int r = idx >> 2, gx = px0 + ((idx & 3) << 4);
            if (idx < DS_PROWS * 4) {
                const bool in = gx < gpitch;
                cp_async16_zfill(patch + r * DS_PPITCH + ((idx & 3) << 4), bl + (size_t)(k.y - 18 + r) * gpitch + (in ? gx : 0), in);
            }
        }
    };
    auto prefetch_raw = [&](const Kp &k, uint8_t *raw) {
        const int gpitch = __shfl_sync(0xffffffffu, tPitch, k.l);
        const unsigned int imgOff = __shfl_sync(0xffffffffu, tImgOff, k.l);
        const uint8_t *img = k.l ? P.pyr + (size_t)b * P.pyrStride + imgOff : P.l0 + (size_t)b * P.l0ImgStride;
        const int ipitch = k.l ? gpitch : P.l0Pitch;
        const int px1 = (k.x - 15) & ~15;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const int idx = lane + 32 * i; // row = idx / 3, chunk = idx % 3
            const int r = (idx * 171) >> 9, gx = px1 + ((idx - 3 * r) << 4);
            const bool in = idx < 31 * 3 && gx < ipitch;
            cp_async16_zfill(raw + idx * 16, img + (size_t)(k.y - 15 + (in ? r : 0)) * ipitch + (in ? gx : 0), in);
        }
    };
    const int slot0 = (blockIdx.x * DS_WARPS + wid) * kpw;
    Kp cur = decode(slot0);
    if (cur.valid) { prefetch_blur(cur, sPatchW); prefetch_raw(cur, sRawW); ds_cp_commit(); }
    for (int kpi = 0; kpi < kpw; kpi++) {
        Kp nxt;
        nxt.valid = 0;
        if (kpi + 1 < kpw) nxt = decode(slot0 + kpi + 1);
        // every lane is done with the blurred buffer the next keypoint goes to (it held keypoint kpi - 1)
        __syncwarp();
        if (nxt.valid) { prefetch_blur(nxt, sPatchW + ((kpi + 1) & 1) * DS_PROWS * DS_PPITCH); ds_cp_commit(); }
        if (!cur.valid) {
            if (nxt.valid) { prefetch_raw(nxt, sRawW); ds_cp_commit(); }
            cur = nxt;
            continue;
        }
        // outstanding groups, oldest first: blurred(cur), raw(cur) [, blurred(nxt)]
        if (nxt.valid) ds_cp_wait<1>(); else ds_cp_wait<0>();
        __syncwarp();
        const uint8_t *patch = sPatchW + (kpi & 1) * DS_PROWS * DS_PPITCH, *raw = sRawW;
        const int l = cur.l, outIdx = cur.outIdx, x = cur.x, y = cur.y, resp = cur.resp;
        const int px0 = (x - 18) & ~15, px1 = (x - 15) & ~15;

        // ---- E6: intensity-centroid moments over the radius-15 disc of the UN-blurred level; lane <-> row v
        int m10, m01;
        {
            const uint4 *rowp = reinterpret_cast<const uint4 *>(raw + lane * DS_RPITCH);
            const uint4 q0 = rowp[0], q1 = rowp[1], q2 = rowp[2];
            __syncwarp(); // every lane holds its row: the un-blurred buffer is free for the next keypoint
            if (nxt.valid) { prefetch_raw(nxt, sRawW); ds_cp_commit(); }
            const int o = (x - 15) - px1; // first disc column inside the staged row (warp-uniform)
            const unsigned int sb = (unsigned int)(o & 3) * 8u;
            int s, t;
            switch (o >> 2) {
            case 0: disc_row_sums<0>(q0, q1, q2, sb, mk, s, t); break;
            case 1: disc_row_sums<1>(q0, q1, q2, sb, mk, s, t); break;
            case 2: disc_row_sums<2>(q0, q1, q2, sb, mk, s, t); break;
            default: disc_row_sums<3>(q0, q1, q2, sb, mk, s, t); break;
            }
            m10 = warp_sum(t);
            m01 = warp_sum((lane - 15) * s);
        }
        const float angle = fast_atan2_deg((float)m01, (float)m10);

        // ---- E8: steered BRIEF on the staged blurred patch; lane <-> descriptor byte
        const float factorPI = (float)(3.14159265358979323846 / 180.f);
        const float ang = __fmul_rn(angle, factorPI);
        // libm cosf/sinf on the CPU are (almost always) the correctly rounded values: take them through double
        double sd, cd;
        sincos_0_2pi((double)ang, &sd, &cd);
        const float a = (float)cd, bsn = (float)sd;
        // cvRound by the magic-number add leaves K + n in the float's bits (K = 0x4B400000); instead of subtracting K from every row
        // and column, the patch-centre address carries -(DS_PPITCH + 1) * K once (32-bit shared-window arithmetic, wraps exactly)
        static_assert(DS_PPITCH == 64, "the row stride is applied as a shift");
        const unsigned int pcK = (unsigned int)__cvta_generic_to_shared(patch + 18 * DS_PPITCH + (x - px0)) - 65u * 0x4B400000u;
        auto round_bits = [](float v) { return (unsigned int)__float_as_int(__fadd_rn(v, 12582912.0f)); }; // |v| < 2^22
        auto lds_u8 = [](unsigned int addr) { unsigned int r; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r) : "r"(addr)); return r; };
        int val = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const float4 p = sPat[k * 32 + lane];
            const unsigned int r0 = round_bits(__fadd_rn(__fmul_rn(p.x, bsn), __fmul_rn(p.y, a)));
            const unsigned int c0 = round_bits(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, bsn)));
            const unsigned int r1 = round_bits(__fadd_rn(__fmul_rn(p.z, bsn), __fmul_rn(p.w, a)));
            const unsigned int c1 = round_bits(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, bsn)));
            const unsigned int t0 = lds_u8(pcK + (r0 << 6) + c0), t1 = lds_u8(pcK + (r1 << 6) + c1);
            val |= (t0 < t1) << k;
        }
        P.desc[((size_t)b * P.cap + outIdx) * 32 + lane] = (uint8_t)val;

        // ---- E9: keypoint record (== cv::KeyPoint); pt scaled to level-0 coordinates for level > 0 (:1329-1333)
        const float scale = __shfl_sync(0xffffffffu, tScale, l), psize = __shfl_sync(0xffffffffu, tPatch, l);
        if (lane < 7) {
            float f;
            const float fx = l ? __fmul_rn((float)x, scale) : (float)x;
            const float fy = l ? __fmul_rn((float)y, scale) : (float)y;
            switch (lane) {
            case 0: f = fx; break;
            case 1: f = fy; break;
            case 2: f = psize; break;
            case 3: f = angle; break;
            case 4: f = (float)resp; break;
            case 5: f = __int_as_float(l); break;
            default: f = __int_as_float(-1); break;
            }
            reinterpret_cast<float *>(P.kps + (size_t)b * P.cap + outIdx)[lane] = f;
        }
        cur = nxt;
    }
}

} // namespace orbb200
