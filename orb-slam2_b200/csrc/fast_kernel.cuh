// fast_kernel.cuh -- E3+E4  fast_cells: one thread block per 30-px grid cell of any level
// (src/ORBextractor.cc:953-1006: cell loop, cv::FAST(roi, iniThFAST, nms=true), retry with minThFAST when empty).
//
// cv::FAST semantics used (derived and pinned against cv2 in tests/test_oracle_pin.py):
//   score S(p) = max over the 16 contiguous 9-arcs of min|v - ring| - 1 (threshold independent);
//   corner at th  <=>  S >= th;   NMS keeps p iff S(p) > S(q) for the 8 neighbours q, where q outside the ROI
//   interior or not a corner at th counts 0.  For S(p) >= th the masked comparison equals the raw one, so the
//   survivors at iniTh are the local maxima with S >= iniTh and the minTh retry (taken only when that set is
//   empty) yields the local maxima with S >= minTh -- one scoring pass serves both thresholds.
//   ROI interiors tile [19, W-19) x [19, H-19) without overlap: every pixel is scored exactly once.
//
// sm_100a formulation (integer-ALU bound stage; ncu round 1: 245 thread instructions per pixel in the scalar form):
//   * two pixels per thread in packed 16-bit lanes: VIMNMX3.S16x2 is a native instruction on sm_100a.
//   * the arc minimum is translation invariant: max_arcs min9(v - ring) = v - min_arcs max9(ring), so the network
//     runs on raw ring values (no 16 subtractions):  A = max_arcs min9(ring), B = min_arcs max9(ring),
//     strength M = max(A - v, v - B); corner at t <=> M > t; score = M - 1.
//     (The negation is applied once, outside the min/max chains: ptxas 12.9 drops a negation folded into
//     VIMNMX3 -- see DESIGN.md.)
//   * the ROI is staged in shared memory widened to 16 bits, twice: copy A holds pixel pairs starting at even
//     positions, copy B the pairs starting at odd positions, so all 17 packed reads of a pixel pair are aligned
//     conflict-free LDS.32 with no byte shuffling.
//   * a warp-uniform quick test on the 4 compass ring pixels skips flat regions.
//   * NMS flags become warp ballots; each warp owns a contiguous band of rows, so the reference's row-major order
//     falls out of lane order + a 4-entry prefix -- no atomics, no block scan.
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

#ifdef ORBB200_EXTRACT_IMPL

constexpr int FC_THREADS = 128, FC_WARPS = 4;
constexpr int FC_SPW = 48;     // staged row pitch in 32-bit words per copy (96 halfwords)
constexpr int FC_SCP = 64;     // score row pitch in bytes: interior <= 60 wide at column x + 2, zero border
constexpr int FC_MAXROWS = 66; // hCell + 6 <= 66 (hCell < 60 because nRows = floor(height / 30))
constexpr int FC_MAXG = 64;    // row groups per cell

__host__ __device__ inline size_t fast_cells_smem(int rows)
{
    return (size_t)rows * FC_SPW * 4 * 2 + (size_t)(rows - 4) * FC_SCP + (size_t)FC_MAXG * 16 + 64;
}

__device__ __forceinline__ unsigned int widen_lo(unsigned int w) { return __byte_perm(w, 0u, 0x4140); }
__device__ __forceinline__ unsigned int widen_hi(unsigned int w) { return __byte_perm(w, 0u, 0x4342); }

__global__ void __launch_bounds__(FC_THREADS) k_fast_cells(const __grid_constant__ ExtractParams P, int rowsCap)
{
    extern __shared__ __align__(16) unsigned char fsm[];
    unsigned int *sA = reinterpret_cast<unsigned int *>(fsm);
    unsigned int *sB = sA + rowsCap * FC_SPW;
    uint8_t *sSc = reinterpret_cast<uint8_t *>(sB + rowsCap * FC_SPW);
    uint4 *sBal = reinterpret_cast<uint4 *>(sSc + (rowsCap - 4) * FC_SCP);
    int *sCnt = reinterpret_cast<int *>(sBal + FC_MAXG); // [FC_WARPS][2]

    const int c = blockIdx.x, b = blockIdx.y;
    int l = 0;
    while (l + 1 < P.nlevels && c >= P.geom[l + 1].cellBase) l++;
    const LevelGeom g = P.geom[l];
    const int cc = c - g.cellBase, ci = cc / g.nCols, cj = cc - ci * g.nCols;
    const int maxBX = g.w - 16, maxBY = g.h - 16;
    const int iniY = 16 + ci * g.hCell, iniX = 16 + cj * g.wCell;
    const int maxY = min(iniY + g.hCell + 6, maxBY), maxX = min(iniX + g.wCell + 6, maxBX);
    const int rw = maxX - iniX, rh = maxY - iniY;
    int *cellCount = P.cellCount + (size_t)b * P.totalCells + c;
    if (iniY >= maxBY - 3 || iniX >= maxBX - 6 || rw < 7 || rh < 7) { // skipped cell (:957-976) or empty interior
        if (threadIdx.x == 0) *cellCount = 0;
        return;
    }
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int pitch, valid;
    const uint8_t *img = level_ptr(P, g, l, b, pitch, valid);
    const int xb = iniX & ~15;
    const int nbytes = ((maxX - xb) + 15) & ~15; // <= 96
    const int o = (iniX + 3 - xb) & 1;           // staged index of image column gx: gx - xb + o
    const int a0 = iniX + 3 - xb + o;            // staged index of the first interior column (even)
    // ---- stage the ROI, widened to 16 bit, as copy A (pairs at even indices) and copy B (pairs at odd indices)
    {
        const bool vec = ((((size_t)img) | (size_t)pitch) & 15) == 0;
        const int chunks = nbytes >> 4;
        for (int i = tid; i < rh * chunks; i += FC_THREADS) {
            const int r = i / chunks, ch = i - r * chunks, x = xb + (ch << 4);
            const uint8_t *src = img + (size_t)(iniY + r) * pitch + x;
            unsigned int wd[5];
            if (vec && x + 16 <= valid) {
                const uint4 v = ldg_nc_v4(src);
                wd[0] = v.x; wd[1] = v.y; wd[2] = v.z; wd[3] = v.w;
            } else {
                wd[0] = wd[1] = wd[2] = wd[3] = 0;
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if (x + k < g.w) wd[k >> 2] |= (unsigned int)__ldg(src + k) << ((k & 3) * 8);
            }
            wd[4] = (x + 16 < g.w) ? (unsigned int)__ldg(src + 16) : 0u;
            unsigned int U[8], S[8]; // U[j] = pixels (2j, 2j+1), S[j] = pixels (2j+1, 2j+2) of the chunk
#pragma unroll
            for (int j = 0; j < 4; j++) {
                U[2 * j] = widen_lo(wd[j]);
                U[2 * j + 1] = widen_hi(wd[j]);
                const unsigned int sh = __funnelshift_r(wd[j], wd[j + 1], 8);
                S[2 * j] = widen_lo(sh);
                S[2 * j + 1] = widen_hi(sh);
            }
            unsigned int *rowA = sA + r * FC_SPW + 8 * ch, *rowB = sB + r * FC_SPW + 8 * ch;
            const int room = FC_SPW - 8 * ch - 1; // words 8ch+1 .. 8ch+8 must stay inside the row
            if (o == 0) {
                *reinterpret_cast<uint4 *>(rowA) = make_uint4(U[0], U[1], U[2], U[3]);
                *reinterpret_cast<uint4 *>(rowA + 4) = make_uint4(U[4], U[5], U[6], U[7]);
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if (j < room) rowB[1 + j] = S[j];
            } else {
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if (j < room) { rowA[1 + j] = S[j]; rowB[1 + j] = U[j]; }
            }
        }
    }
    const int iw = rw - 6, ih = rh - 6;
    for (int i = tid; i < ((ih + 2) * FC_SCP) >> 2; i += FC_THREADS) reinterpret_cast<int *>(sSc)[i] = 0;
    __syncthreads();

    // ---- thread <-> pixel pair; a warp covers 1, 2 or 4 interior rows; each warp owns a contiguous band of groups
    const int ppr = (iw + 1) >> 1;
    const int lprShift = ppr <= 8 ? 3 : (ppr <= 16 ? 4 : 5);
    const int rpw = 32 >> lprShift;
    const int ngroups = (ih + rpw - 1) / rpw;
    const int gpw = (ngroups + FC_WARPS - 1) / FC_WARPS;
    const int g0 = wid * gpw, g1 = min(g0 + gpw, ngroups);
    const int xl = lane & ((1 << lprShift) - 1), yl = lane >> lprShift;
    const int x = 2 * xl;
    const int minTh = P.minTh, iniTh = P.iniTh;
    const unsigned int thp1 = (unsigned int)(minTh + 1) * 0x10001u;

    for (int gi = g0; gi < g1; gi++) {
        const int y = gi * rpw + yl;
        const bool act = (y < ih) && (xl < ppr);
        const int wofs = (act ? (y + 3) * FC_SPW : 3 * FC_SPW) + ((a0 + (act ? x : 0)) >> 1);
        const unsigned int *pa = sA + wofs, *pb = sB + wofs;
        const unsigned int v = pa[0];
        unsigned int r[16];
        r[0] = pa[3 * FC_SPW];        // ( 0, 3)
        r[8] = pa[-3 * FC_SPW];       // ( 0,-3)
        r[4] = pb[2];                 // ( 3, 0)
        r[12] = pb[-1];               // (-3, 0)
        // quick test: every 9-arc holds ring pixel k or k+8 for each k
        {
            const unsigned int tb = __vmins2(__vmaxs2(r[0], r[8]), __vmaxs2(r[4], r[12]));
            const unsigned int td = __vmaxs2(__vmins2(r[0], r[8]), __vmins2(r[4], r[12]));
            const unsigned int q = __vmaxs2(__vsub2(tb, v), __vsub2(v, td)); // > minTh in a lane <=> a corner is possible
            const bool possible = act && ((__vsub2(q, thp1) & 0x80008000u) != 0x80008000u);
            if (!__any_sync(0xffffffffu, possible)) continue;
        }
        r[1] = pb[3 * FC_SPW + 1];    // ( 1, 3)
        r[2] = pa[2 * FC_SPW + 1];    // ( 2, 2)
        r[3] = pb[FC_SPW + 2];        // ( 3, 1)
        r[5] = pb[-FC_SPW + 2];       // ( 3,-1)
        r[6] = pa[-2 * FC_SPW + 1];   // ( 2,-2)
        r[7] = pb[-3 * FC_SPW + 1];   // ( 1,-3)
        r[9] = pb[-3 * FC_SPW];       // (-1,-3)
        r[10] = pa[-2 * FC_SPW - 1];  // (-2,-2)
        r[11] = pb[-FC_SPW - 1];      // (-3,-1)
        r[13] = pb[FC_SPW - 1];       // (-3, 1)
        r[14] = pa[2 * FC_SPW - 1];   // (-2, 2)
        r[15] = pb[3 * FC_SPW];       // (-1, 3)
        unsigned int mn3[16], mx3[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            mn3[k] = __vimin3_s16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
            mx3[k] = __vimax3_s16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
        }
        unsigned int mn9[16], mx9[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            mn9[k] = __vimin3_s16x2(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]);
            mx9[k] = __vimax3_s16x2(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]);
        }
        unsigned int A = mn9[0], Bm = mx9[0];
#pragma unroll
        for (int k = 1; k < 15; k += 2) {
            A = __vimax3_s16x2(A, mn9[k], mn9[k + 1]);
            Bm = __vimin3_s16x2(Bm, mx9[k], mx9[k + 1]);
        }
        A = __vmaxs2(A, mn9[15]);
        Bm = __vmins2(Bm, mx9[15]);
        // strength per lane: darker ring v - B, brighter ring A - v
        const unsigned int M = __vmaxs2(__vsub2(A, v), __vsub2(v, Bm));
        const int mlo = (int)(short)(M & 0xFFFFu), mhi = (int)M >> 16;
        const int slo = mlo > minTh ? mlo - 1 : 0;
        const int shi = (mhi > minTh && x + 1 < iw) ? mhi - 1 : 0;
        if (act) *reinterpret_cast<unsigned short *>(sSc + (y + 1) * FC_SCP + x + 2) = (unsigned short)(slo | (shi << 8));
    }
    __syncthreads();

    // ---- local maxima as ballots, per group; counts of the iniTh set and of the minTh set
    int cntLm = 0, cntIni = 0;
    for (int gi = g0; gi < g1; gi++) {
        const int y = gi * rpw + yl;
        const bool act = (y < ih) && (xl < ppr);
        const uint8_t *s = sSc + ((act ? y : 0) + 1) * FC_SCP + (act ? x : 0) + 2;
        const unsigned int mid = act ? *reinterpret_cast<const unsigned short *>(s) : 0u;
        bool lmLo = false, lmHi = false, inLo = false, inHi = false;
        if (__any_sync(0xffffffffu, mid != 0u)) {
            if (mid != 0u) {
                const int v0 = mid & 0xFF, v1 = mid >> 8;
                const int t0 = s[-FC_SCP - 1], t12 = *reinterpret_cast<const unsigned short *>(s - FC_SCP), t3 = s[-FC_SCP + 2];
                const int b0 = s[FC_SCP - 1], b12 = *reinterpret_cast<const unsigned short *>(s + FC_SCP), b3 = s[FC_SCP + 2];
                const int m0 = s[-1], m3 = s[2];
                const int tm = max(t12 & 0xFF, t12 >> 8), bm = max(b12 & 0xFF, b12 >> 8);
                const int n0 = __vimax3_s32(__vimax3_s32(t0, tm, b0), bm, max(m0, v1));
                const int n1 = __vimax3_s32(__vimax3_s32(t3, tm, b3), bm, max(m3, v0));
                lmLo = v0 > n0;
                lmHi = v1 > n1;
                inLo = lmLo && v0 >= iniTh;
                inHi = lmHi && v1 >= iniTh;
            }
        }
        const unsigned int bl = __ballot_sync(0xffffffffu, lmLo), bh = __ballot_sync(0xffffffffu, lmHi);
        const unsigned int il = __ballot_sync(0xffffffffu, inLo), ihb = __ballot_sync(0xffffffffu, inHi);
        if (lane == 0) sBal[gi] = make_uint4(bl, bh, il, ihb);
        cntLm += __popc(bl) + __popc(bh);
        cntIni += __popc(il) + __popc(ihb);
    }
    if (lane == 0) { sCnt[2 * wid] = cntLm; sCnt[2 * wid + 1] = cntIni; }
    __syncthreads();
    int anyIni = 0, off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < FC_WARPS; w++) anyIni |= sCnt[2 * w + 1];
    const int sel = anyIni ? 1 : 0; // the iniTh set when it is not empty, else the minTh retry (:991-995)
#pragma unroll
    for (int w = 0; w < FC_WARPS; w++) {
        if (w < wid) off += sCnt[2 * w + sel];
        total += sCnt[2 * w + sel];
    }
    // ---- ordered write: groups in row order, lanes in column order, low pixel before high pixel
    uint32_t *out = P.cand + (size_t)b * P.candStride + g.candOff + (size_t)cc * g.cellCap;
    const unsigned int lt = (1u << lane) - 1u;
    for (int gi = g0; gi < g1; gi++) {
        const uint4 bal = sBal[gi];
        const unsigned int ml = sel ? bal.z : bal.x, mh = sel ? bal.w : bal.y;
        if ((ml | mh) == 0u) continue;
        const int y = gi * rpw + yl;
        const int mine = ((ml >> lane) & 1) + ((mh >> lane) & 1);
        if (mine) {
            int pos = off + __popc(ml & lt) + __popc(mh & lt);
            const uint8_t *s = sSc + (y + 1) * FC_SCP + x + 2;
            const uint32_t base = (uint32_t)(cj * g.wCell + x + 3) | ((uint32_t)(ci * g.hCell + y + 3) << 12);
            if ((ml >> lane) & 1) out[pos++] = base | ((uint32_t)s[0] << 24);
            if ((mh >> lane) & 1) out[pos] = (base + 1u) | ((uint32_t)s[1] << 24);
        }
        off += __popc(ml) + __popc(mh);
    }
    if (tid == 0) *cellCount = total;
}

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
