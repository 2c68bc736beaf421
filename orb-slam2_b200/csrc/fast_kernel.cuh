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
// sm_100a formulation (integer-ALU / issue bound stage):
//   * two pixels per thread in packed 16-bit lanes: VIMNMX3.S16x2 is a native instruction on sm_100a.
//   * the arc minimum is translation invariant: max_arcs min9(v - ring) = v - min_arcs max9(ring), so the network
//     runs on raw ring values (no 16 subtractions):  A = max_arcs min9(ring), B = min_arcs max9(ring),
//     strength M = max(A - v, v - B); corner at t <=> M > t; score = M - 1.
//     (The negation is applied once, outside the min/max chains: ptxas 12.9 drops a negation folded into
//     VIMNMX3 -- see DESIGN.md.)
//   * the ROI is staged in shared memory widened to 16 bits, twice: copy U holds the pixel pairs that start at even
//     chunk columns, copy S the pairs that start at odd columns (both written with 128-bit stores), so all 17
//     packed reads of a pixel pair are aligned conflict-free LDS.32 with no byte shuffling; which copy serves the
//     even ring offsets depends on the parity of the cell's first interior column.
//   * a warp-uniform quick test on the 4 compass ring pixels skips flat regions.
//   * NMS is separable and stays packed: the scoring pass exchanges scores with the two neighbouring lanes
//     (2 shuffles), keeps s' = s where s beats its horizontal neighbours and the 3-wide row maximum m3, and stores
//     one word {s'0, s'1, m3_0, m3_1} per pixel pair; the second pass needs the words above and below only.
//   * each warp owns a contiguous band of rows and appends its survivors to its own shared-memory list in lane
//     order (= the reference's row-major order); after the block-wide iniTh/minTh vote the lists are copied to the
//     cell's fixed slot -- no atomics, no block scan.
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

#ifdef ORBB200_EXTRACT_IMPL

constexpr int FC_THREADS = 128, FC_WARPS = 4;
constexpr int FC_SPW = 48;      // staged row pitch in 32-bit words per copy (96 pixels = 6 chunks of 16)
constexpr int FC_MAXROWS = 66;  // hCell + 6 <= 66 (hCell < 60 because nRows = floor(height / 30))
constexpr int FC_LISTCAP = 256; // survivors per warp: <= ceil(15 / 2) * 30 strict local maxima in its <= 15 rows

__host__ __device__ inline size_t fast_cells_smem(int rows)
{
    // two widened copies + one score word per pixel pair (rows - 6 interior rows rounded up to a group, 2 border rows,
    // 32 words per row at most) + the per-warp survivor lists
    return (size_t)rows * FC_SPW * 4 * 2 + (size_t)(rows + 2) * 32 * 4 + (size_t)FC_WARPS * FC_LISTCAP * 4;
}

__device__ __forceinline__ unsigned int widen_lo(unsigned int w) { return __byte_perm(w, 0u, 0x4140); }
__device__ __forceinline__ unsigned int widen_hi(unsigned int w) { return __byte_perm(w, 0u, 0x4342); }
// a - b per 16-bit half for halves in [0, 0x7FFF], BIASED by 0x8000: one plain 32-bit subtraction, no borrow can cross
// the halves, bit 15 of a half is clear <=> a < b there; biased values order like unsigned halves (VIMNMX.U16x2).
// (Deliberately not __vsub2: nvcc 12.9 splits its 16x2 negation constant wrongly in a peeled loop iteration --
// the low half lost its +1 -- see DESIGN.md.)
__device__ __forceinline__ unsigned int bsub2(unsigned int a, unsigned int b) { return (a | 0x80008000u) - b; }
// 0xFFFF in each 16-bit half whose bit 15 is set (prmt replicates the msb of bytes 1 and 3)
__device__ __forceinline__ unsigned int neg_mask2(unsigned int w)
{
    unsigned int r;
    asm("prmt.b32 %0, %1, 0, 0xBB99;" : "=r"(r) : "r"(w));
    return r;
}

__global__ void __launch_bounds__(FC_THREADS, 8) k_fast_cells(const __grid_constant__ ExtractParams P, int rowsCap)
{
    extern __shared__ __align__(16) unsigned char fsm[];
    unsigned int *sU = reinterpret_cast<unsigned int *>(fsm);
    unsigned int *sS = sU + rowsCap * FC_SPW;
    unsigned int *sW = sS + rowsCap * FC_SPW;                 // [(rows + 2)][lanes per row], dense
    unsigned int *sList = sW + (rowsCap + 2) * 32;            // [FC_WARPS][FC_LISTCAP]
    __shared__ int sCnt[2 * FC_WARPS];

    const int c = blockIdx.x, b = blockIdx.y;
    // per-cell record (built once per shape on the host): ROI, level, image slot, output slot
    const uint4 r0 = ldg_nc_v4(P.fastCells + 2 * c), r1 = ldg_nc_v4(P.fastCells + 2 * c + 1);
    const int rw = (int)(r0.y & 0xFFFFu), rh = (int)(r0.y >> 16);
    int *cellCount = P.cellCount + (size_t)b * P.totalCells + c;
    if (rw == 0) { // skipped cell (:957-976) or empty interior
        if (threadIdx.x == 0) *cellCount = 0;
        return;
    }
    const int iniX = (int)(r0.x & 0xFFFFu), iniY = (int)(r0.x >> 16), maxX = iniX + rw;
    const int gw = (int)(r0.w >> 16);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int pitch, valid;
    const uint8_t *img;
    if (r1.z == 0u) { pitch = P.l0Pitch; valid = P.l0Valid; img = P.l0 + (size_t)b * P.l0ImgStride; }
    else { pitch = (int)(r0.w & 0xFFFFu); valid = pitch; img = P.pyr + (size_t)b * P.pyrStride + r0.z; }
    const int xb = iniX & ~15;
    const int chunks = (((maxX - xb) + 15) & ~15) >> 4; // 3 .. 6
    const int a = iniX + 3 - xb;                        // chunk column of the first interior pixel
    // ---- stage the ROI, widened to 16 bit: U[k] = pixels (2k, 2k+1), S[k] = pixels (2k+1, 2k+2) of the row
    {
        const bool vec = ((((size_t)img) | (size_t)pitch) & 15) == 0;
        const unsigned int magic = r1.w; // ceil(65536 / chunks): i / chunks for i < 400
        for (int i = tid; i < rh * chunks; i += FC_THREADS) {
            const int r = (int)(((unsigned int)i * magic) >> 16), ch = i - r * chunks, x = xb + (ch << 4);
            const uint8_t *src = img + (size_t)(iniY + r) * pitch + x;
            unsigned int wd[5];
            if (vec && x + 16 <= valid) {
                const uint4 v = ldg_nc_v4(src);
                wd[0] = v.x; wd[1] = v.y; wd[2] = v.z; wd[3] = v.w;
            } else {
                wd[0] = wd[1] = wd[2] = wd[3] = 0;
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if (x + k < gw) wd[k >> 2] |= (unsigned int)__ldg(src + k) << ((k & 3) * 8);
            }
            wd[4] = (x + 16 < gw) ? (unsigned int)__ldg(src + 16) : 0u;
            unsigned int U[8], S[8];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                U[2 * j] = widen_lo(wd[j]);
                U[2 * j + 1] = widen_hi(wd[j]);
                const unsigned int sh = __funnelshift_r(wd[j], wd[j + 1], 8);
                S[2 * j] = widen_lo(sh);
                S[2 * j + 1] = widen_hi(sh);
            }
            uint4 *rowU = reinterpret_cast<uint4 *>(sU + r * FC_SPW + 8 * ch), *rowS = reinterpret_cast<uint4 *>(sS + r * FC_SPW + 8 * ch);
            rowU[0] = make_uint4(U[0], U[1], U[2], U[3]);
            rowU[1] = make_uint4(U[4], U[5], U[6], U[7]);
            rowS[0] = make_uint4(S[0], S[1], S[2], S[3]);
            rowS[1] = make_uint4(S[4], S[5], S[6], S[7]);
        }
    }
    const int iw = rw - 6, ih = rh - 6;

    // ---- thread <-> pixel pair; a warp covers 1, 2 or 4 interior rows; each warp owns a contiguous band of groups
    const int ppr = (iw + 1) >> 1;
    const int lprShift = ppr <= 8 ? 3 : (ppr <= 16 ? 4 : 5);
    const int lpr = 1 << lprShift, rpw = 32 >> lprShift;
    const int ngroups = (ih + rpw - 1) / rpw;
    const int gpw = (ngroups + FC_WARPS - 1) / FC_WARPS;
    const int g0 = wid * gpw, g1 = min(g0 + gpw, ngroups);
    const int xl = lane & (lpr - 1), yl = lane >> lprShift;
    const int x = 2 * xl;
    // score words: row y of the interior is word row y + 1; word rows 0 and (ngroups * rpw) + 1 are the zero border
    for (int i = tid; i < lpr; i += FC_THREADS) {
        sW[i] = 0u;
        sW[(ngroups * rpw + 1) * lpr + i] = 0u;
    }
    __syncthreads();

    const int minTh = P.minTh, iniTh = P.iniTh;
    const unsigned int thp1 = (unsigned int)(minTh + 1) * 0x10001u, thm = (unsigned int)minTh * 0x10001u;
    const unsigned int xmask = (xl < ppr) ? ((x + 1 < iw) ? 0xFFFFFFFFu : 0x0000FFFFu) : 0u;
    // which copy serves the even ring offsets depends on the parity of the first interior column
    const unsigned int *evenCopy = (a & 1) ? sS : sU, *oddCopy = (a & 1) ? sU : sS;
    const int cofs = (a & 1) ? ((a - 1) >> 1) : (a >> 1);
    const int oofs = (a & 1) ? cofs : cofs - 1;

    // running pointers: inactive lanes (columns past the interior, rows past it in the last group) read staged or
    // stale shared memory that stays inside the block's allocation; their results are masked out below
    const unsigned int *pa = evenCopy + (g0 * rpw + yl + 3) * FC_SPW + xl + cofs;
    const unsigned int *pb = oddCopy + (g0 * rpw + yl + 3) * FC_SPW + xl + oofs;
    unsigned int *wout = sW + g0 * 32 + lane + lpr;
    for (int gi = g0; gi < g1; gi++, pa += rpw * FC_SPW, pb += rpw * FC_SPW, wout += 32) {
        const unsigned int vmask = (gi * rpw + yl < ih) ? xmask : 0u;
        const bool act = vmask != 0u;
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
            const unsigned int q = __vmaxu2(bsub2(tb, v), bsub2(v, td)); // biased; > minTh in a half <=> a corner is possible
            const bool possible = act && (((q - thp1) & 0x80008000u) != 0u);
            if (!__any_sync(0xffffffffu, possible)) {
                *wout = 0u;
                continue;
            }
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
        // strength per half: darker ring v - B, brighter ring A - v (biased by 0x8000, see bsub2);
        // score = M - 1 where M > minTh, else 0
        const unsigned int Mb = __vmaxu2(bsub2(A, v), bsub2(v, Bm));
        const unsigned int cth = Mb - thp1;                                    // biased M - (minTh + 1): bit 15 set <=> M > minTh
        const unsigned int S = (cth + thm) & neg_mask2(cth) & (vmask & 0x00FF00FFu); // biased M - 1, bias and losers masked off
        // horizontal part of the NMS: scores of the neighbouring pairs come from the neighbouring lanes
        unsigned int SL = __shfl_up_sync(0xffffffffu, S, 1), SR = __shfl_down_sync(0xffffffffu, S, 1);
        if (xl == 0) SL = 0u;
        if (xl == lpr - 1) SR = 0u;
        const unsigned int Sl = __byte_perm(SL, S, 0x5432); // (s(x-1), s(x))
        const unsigned int Sr = __byte_perm(S, SR, 0x5432); // (s(x+1), s(x+2))
        const unsigned int h = __vmaxs2(Sl, Sr);
        const unsigned int m3 = __vmaxs2(h, S);
        const unsigned int Sp = S & ~neg_mask2(bsub2(h, S));  // s where s > both horizontal neighbours
        *wout = __byte_perm(Sp, m3, 0x6420);                  // {s'0, s'1, m3_0, m3_1}
    }
    __syncthreads();

    // ---- vertical part of the NMS; survivors appended to the warp's list in lane order = row-major order
    unsigned int *list = sList + wid * FC_LISTCAP;
    const unsigned int lt = (1u << lane) - 1u;
    const uint32_t base0 = r1.y + (uint32_t)x + ((uint32_t)yl << 12);
    int wcnt = 0;
    const unsigned int *w = sW + g0 * 32 + lane + lpr;
    uint32_t base = base0 + ((uint32_t)(g0 * rpw) << 12);
    for (int gi = g0; gi < g1; gi++, w += 32) {
        const unsigned int W = w[0];
        if (!__any_sync(0xffffffffu, (W & 0xFFFFu) != 0u)) { base += (uint32_t)rpw << 12; continue; }
        const unsigned int mx = __vmaxs2(widen_hi(w[-lpr]), widen_hi(w[lpr]));
        const unsigned int s = widen_lo(W);
        const unsigned int sk = s & ~neg_mask2(bsub2(mx, s)); // s' where it also beats the rows above and below
        // the two pixels of a pair are neighbours, so at most one of them is a strict local maximum: one ballot,
        // one entry per lane
        const unsigned int bs = __ballot_sync(0xffffffffu, sk != 0u);
        if (sk != 0u) {
            const int pos = wcnt + __popc(bs & lt);
            const unsigned int hi = sk >> 16; // 0 when the survivor is the low pixel
            if (pos < FC_LISTCAP)
                list[pos] = (base + (hi ? 1u : 0u)) | ((sk | hi) << 24); // score << 24: bits above 31 fall off
        }
        wcnt += __popc(bs);
        base += (uint32_t)rpw << 12;
    }
    wcnt = min(wcnt, FC_LISTCAP);
    __syncwarp();
    int nIni = 0;
    for (int i0 = 0; i0 < wcnt; i0 += 32) {
        const int i = i0 + lane;
        nIni += __popc(__ballot_sync(0xffffffffu, i < wcnt && (int)(list[i] >> 24) >= iniTh));
    }
    if (lane == 0) { sCnt[2 * wid] = wcnt; sCnt[2 * wid + 1] = nIni; }
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
    // ---- ordered copy of the selected set into the cell's fixed slot
    uint32_t *out = P.cand + (size_t)b * P.candStride + r1.x + off;
    if (!sel) {
        for (int i = lane; i < wcnt; i += 32) out[i] = list[i];
    } else {
        int run = 0;
        for (int i0 = 0; i0 < wcnt; i0 += 32) {
            const int i = i0 + lane;
            const uint32_t e = i < wcnt ? list[i] : 0u;
            const bool ok = (int)(e >> 24) >= iniTh;
            const unsigned int m = __ballot_sync(0xffffffffu, ok);
            if (ok) out[run + __popc(m & lt)] = e;
            run += __popc(m);
        }
    }
    if (tid == 0) *cellCount = total;
}

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
