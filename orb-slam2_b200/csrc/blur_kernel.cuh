// blur_kernel.cuh -- E7  blur7: cv::GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) on every level
// (src/ORBextractor.cc:1312-1314), 8U fixed point exactly as OpenCV 4.x: Q8 taps 18 34 48 56 48 34 18 (sum 256),
// out = (sum_ij k_i k_j p_ij + 32768) >> 16.  The double sum is exact integer arithmetic, so the separable passes may
// run in either order; here vertical first (fits 16 bits: <= 256*255), then horizontal in 32 bits.
//
// sm_100a formulation (HBM-streaming stage, no shared memory):
//   * a warp owns a strip of 128 columns (120 outputs + 4-column halo lanes on each side) and slides down the rows;
//     every lane owns 4 adjacent columns = one aligned 32-bit word per row, so a row of the strip is one coalesced
//     128-byte request;
//   * vertical pass on two packed 16x2 words per lane: a 32-bit IMAD by a tap multiplies both lanes at once because
//     no lane can exceed 65 280; the 7-row window lives in registers (loop unrolled by 7: static ring indices);
//   * the 6 halo values of the horizontal pass come from the neighbouring lanes by 4 warp shuffles;
//   * horizontal pass with DP2A: one instruction multiplies a packed (V_x, V_x+1) pair by two 8-bit taps and
//     accumulates in 32 bits -- 4 DP2A per output pixel.
//   * REFLECT_101 columns: only lanes that touch x < 0 or x >= W gather bytes; rows are reflected by index.
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

constexpr int BL_TW = 120, BL_TH = 64; // outputs per warp strip (columns) and rows per work item
constexpr int BL_WARPS = 4, BL_THREADS = BL_WARPS * 32;
#ifndef ORBB200_BL_MINBLOCKS
#define ORBB200_BL_MINBLOCKS 12
#endif
constexpr int BL_MINBLOCKS = ORBB200_BL_MINBLOCKS;

#ifdef ORBB200_EXTRACT_IMPL

__device__ __forceinline__ int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p;
    return p;
}

// REFLECT_101 row index without the loop when one reflection is enough (|overshoot| <= 6 < H - 1)
__device__ __forceinline__ int reflect_row(int r, int H)
{
    if (H < 8) return reflect101(r, H);
    r = r < 0 ? -r : r;
    return r >= H ? 2 * (H - 1) - r : r;
}

// 4 pixels of a row: one 32-bit load for lanes inside the image, 4 byte loads at the lane's (row-independent,
// precomputed) reflected columns for the few lanes at the left / right edge
__device__ __forceinline__ unsigned int blur_load4(const uint8_t *row, int c, const int (&cx)[4], bool interior)
{
    if (interior) return __ldg(reinterpret_cast<const unsigned int *>(row + c));
    return (unsigned int)__ldg(row + cx[0]) | ((unsigned int)__ldg(row + cx[1]) << 8) | ((unsigned int)__ldg(row + cx[2]) << 16) |
           ((unsigned int)__ldg(row + cx[3]) << 24);
}

#define BL_TAPS(a, b) ((unsigned int)(a) | ((unsigned int)(b) << 8))

// rows [y0, yEnd) of one 120-column strip.  FASTROWS: rows up to yEnd+2 exist (no reflection inside the loop, the source row is a
// running pointer); ALLINT: every lane loads an aligned word (no per-lane branch around the load)
template <bool FASTROWS, bool ALLINT>
__device__ __forceinline__ void blur_strip(const uint8_t *src, uint8_t *dst, int pitch, int dpitch, int y0, int yEnd, int H, int c,
                                           const int (&cx)[4], bool interior, bool store)
{
    auto load_row = [&](int r) -> unsigned int { // the seven rows loaded before the loop may lie above the image also when FASTROWS
        const uint8_t *row = src + (size_t)reflect_row(r, H) * pitch;
        if (ALLINT) return __ldg(reinterpret_cast<const unsigned int *>(row + c));
        return blur_load4(row, c, cx, interior);
    };
    unsigned int ra[7], rb[7]; // ring of 7 rows, widened: ra = (p0, p1), rb = (p2, p3) in 16-bit lanes
#pragma unroll
    for (int k = 0; k < 6; k++) {
        const unsigned int w = load_row(y0 - 3 + k);
        ra[k] = __byte_perm(w, 0u, 0x4140);
        rb[k] = __byte_perm(w, 0u, 0x4342);
    }
    unsigned int nxt = load_row(y0 + 3);
    const uint8_t *nrow = src + (size_t)(y0 + 4) * pitch + c; // FASTROWS && ALLINT: the next source word, one pitch further per row
    uint8_t *drow = dst + (size_t)y0 * dpitch + c;
    for (int base = y0; base < yEnd; base += 7) {
#pragma unroll
        for (int k = 0; k < 7; k++) {
            const int y = base + k;
            if (y < yEnd) { // warp-uniform
                ra[(k + 6) % 7] = __byte_perm(nxt, 0u, 0x4140);
                rb[(k + 6) % 7] = __byte_perm(nxt, 0u, 0x4342);
                if (y + 1 < yEnd) {
                    if (FASTROWS && ALLINT) nxt = __ldg(reinterpret_cast<const unsigned int *>(nrow));
                    else nxt = load_row(y + 4);
                }
                nrow += pitch;
                // vertical pass: rows y-3 .. y+3 are ring slots k .. k+6 (mod 7)
                const unsigned int Va = 18u * (ra[k % 7] + ra[(k + 6) % 7]) + 34u * (ra[(k + 1) % 7] + ra[(k + 5) % 7]) +
                                        48u * (ra[(k + 2) % 7] + ra[(k + 4) % 7]) + 56u * ra[(k + 3) % 7];
                const unsigned int Vb = 18u * (rb[k % 7] + rb[(k + 6) % 7]) + 34u * (rb[(k + 1) % 7] + rb[(k + 5) % 7]) +
                                        48u * (rb[(k + 2) % 7] + rb[(k + 4) % 7]) + 56u * rb[(k + 3) % 7];
                // halo from the neighbouring lanes: La = (V-4, V-3), Lb = (V-2, V-1), Ra = (V4, V5), Rb = (V6, V7)
                const unsigned int La = __shfl_up_sync(0xffffffffu, Va, 1), Lb = __shfl_up_sync(0xffffffffu, Vb, 1);
                const unsigned int Ra = __shfl_down_sync(0xffffffffu, Va, 1), Rb = __shfl_down_sync(0xffffffffu, Vb, 1);
                // horizontal pass: 7 taps over packed pairs, 32-bit accumulators seeded with the rounding constant
                unsigned int o0 = __dp2a_lo(La, BL_TAPS(0, 18), 32768u);
                o0 = __dp2a_lo(Lb, BL_TAPS(34, 48), o0);
                o0 = __dp2a_lo(Va, BL_TAPS(56, 48), o0);
                o0 = __dp2a_lo(Vb, BL_TAPS(34, 18), o0);
                unsigned int o1 = __dp2a_lo(Lb, BL_TAPS(18, 34), 32768u);
                o1 = __dp2a_lo(Va, BL_TAPS(48, 56), o1);
                o1 = __dp2a_lo(Vb, BL_TAPS(48, 34), o1);
                o1 = __dp2a_lo(Ra, BL_TAPS(18, 0), o1);
                unsigned int o2 = __dp2a_lo(Lb, BL_TAPS(0, 18), 32768u);
                o2 = __dp2a_lo(Va, BL_TAPS(34, 48), o2);
                o2 = __dp2a_lo(Vb, BL_TAPS(56, 48), o2);
                o2 = __dp2a_lo(Ra, BL_TAPS(34, 18), o2);
                unsigned int o3 = __dp2a_lo(Va, BL_TAPS(18, 34), 32768u);
                o3 = __dp2a_lo(Vb, BL_TAPS(48, 56), o3);
                o3 = __dp2a_lo(Ra, BL_TAPS(48, 34), o3);
                o3 = __dp2a_lo(Rb, BL_TAPS(18, 0), o3);
                if (store) {
                    const unsigned int lo = __byte_perm(o0, o1, 0x0062); // (o0 >> 16) | (o1 >> 16) << 8
                    const unsigned int hi = __byte_perm(o2, o3, 0x0062);
                    *reinterpret_cast<unsigned int *>(drow) = __byte_perm(lo, hi, 0x5410);
                }
                drow += dpitch;
            }
        }
    }
}

__global__ void __launch_bounds__(BL_THREADS, BL_MINBLOCKS) k_blur7(const __grid_constant__ ExtractParams P)
{
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * BL_WARPS + (threadIdx.x >> 5);
    if (item >= P.totalBlurTiles) return;
    int l = 0;
    while (l + 1 < P.nlevels && item >= P.geom[l + 1].blurTileBase) l++;
    const LevelGeom g = P.geom[l];
    const int b = blockIdx.y;
    const int t = item - g.blurTileBase;
    const int sx = t % g.blurTilesX, sy = t / g.blurTilesX;
    const int W = g.w, H = g.h;
    const int y0 = sy * BL_TH, yEnd = min(y0 + BL_TH, H);
    const int c = sx * BL_TW - 4 + 4 * lane; // first of the lane's 4 virtual columns
    int pitch, valid;
    const uint8_t *src = level_ptr(P, g, l, b, pitch, valid);
    uint8_t *dst = P.blur + (size_t)b * P.pyrStride + g.imgOff;
    // lanes entirely beyond the reflected range never contribute (their neighbours' halo stays within +-3 of [0, W))
    const bool interior = c >= 0 && c + 3 < W && ((((size_t)src) | (size_t)pitch) & 3) == 0;
    const bool store = lane >= 1 && lane <= 30 && c < W;
    int cx[4];
#pragma unroll
    for (int k = 0; k < 4; k++) cx[k] = (c + k < W + 3) ? reflect101(c + k, W) : 0; // columns past W + 2 feed no output

    // two warp-uniform properties pick the loop variant: no row loaded inside the loop needs reflection (all but the last strip of a
    // level) and no lane needs reflected columns (all but the strips at the left / right edge).  The generic loop spent 40 % of its
    // instructions finding the word to load and the word to store.
    const bool fastRows = yEnd + 3 <= H;
    const bool allInt = __all_sync(0xffffffffu, interior);
    if (fastRows && allInt) blur_strip<true, true>(src, dst, pitch, g.pitch, y0, yEnd, H, c, cx, interior, store);
    else blur_strip<false, false>(src, dst, pitch, g.pitch, y0, yEnd, H, c, cx, interior, store);
}

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
