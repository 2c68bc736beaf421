// fast2_kernel.cuh -- E3+E4  fast_cells, second formulation: bound test on every pixel, full score only where it can matter.
// (src/ORBextractor.cc:953-1006: cell loop, cv::FAST(roi, iniThFAST, nms=true), retry with minThFAST when empty.)
//
// Same cv::FAST semantics as fast_kernel.cuh (score S = max over the 16 contiguous 9-arcs of min|v - ring| - 1, corner at
// th <=> S >= th, 3x3 strict NMS against neighbours that count 0 outside the ROI interior / below th).  What changes is
// where the 80-instruction min/max network runs.  k_fast_cells ran it on every pixel and was bound by the ALU pipe (packed
// min/max issues at half rate: 2 warp instructions / clk / SM, and half2 min/max shares that pipe -- tools/scratch/pipe_probe.cu).
//
//   Phase A (every pixel, two per thread in packed 16-bit lanes): every 9-arc of the ring holds pixel k or k + 8 for each
//     k, so  A = max_arcs min9 <= ub = min_k max(r[k], r[k+8])  and  B = min_arcs max9 >= lb = max_k min(r[k], r[k+8]):
//     24 packed min/max give the bound q = max(ub - v, v - lb) >= M = S + 1.  Pixels with q > th (about 18 % at th = 20,
//     36 % at th = 7 on textured frames) are appended to the warp's queue in row-major order (2 ballots).
//   Phase B (queued pixels only, two per lane): ring values by 16-bit shared-memory loads, the exact network, score byte
//     written to a score map of the cell (zero elsewhere: a pixel that is not queued has S < th).
//   NMS (queued pixels with S >= th): 8 neighbour bytes from the score map; survivors appended to the warp's list in
//     queue order = row-major order; per-warp lists are copied to the cell's fixed slot as before.
//   The block first runs with th = iniThFAST; only when that leaves the cell empty (:991-995) it runs again with
//     th = minThFAST (flat cells: few pixels pass the bound).
#pragma once
#include "fast_kernel.cuh"

namespace orbb200 {

#ifdef ORBB200_EXTRACT_IMPL

constexpr int F2_MAPPITCH = 64; // score map row pitch in bytes (interior width < 60, + 1 border column each side)

__host__ __device__ inline int fast2_qcap(int rows) { return (((rows - 6) + FC_WARPS - 1) / FC_WARPS + 1) * 64; } // entries per warp
__host__ __device__ inline size_t fast2_smem(int rows)
{
    // two widened copies + score map ((rows - 6) interior rows + 2 border rows) + per-warp queues (u16) + per-warp survivor lists
    return (size_t)rows * FC_SPW * 4 * 2 + (size_t)(rows - 4) * F2_MAPPITCH + (size_t)FC_WARPS * fast2_qcap(rows) * 2 +
           (size_t)FC_WARPS * FC_LISTCAP * 4;
}

__global__ void __launch_bounds__(FC_THREADS, 6) k_fast_cells2(const __grid_constant__ ExtractParams P, int rowsCap)
{
    extern __shared__ __align__(16) unsigned char fsm[];
    unsigned int *sU = reinterpret_cast<unsigned int *>(fsm);
    unsigned int *sS = sU + rowsCap * FC_SPW;
    unsigned char *sM = reinterpret_cast<unsigned char *>(sS + rowsCap * FC_SPW);        // [(rowsCap - 4)][F2_MAPPITCH]
    unsigned short *sQ = reinterpret_cast<unsigned short *>(sM + (rowsCap - 4) * F2_MAPPITCH); // [FC_WARPS][qcap]
    const int qcap = fast2_qcap(rowsCap);
    unsigned int *sList = reinterpret_cast<unsigned int *>(sQ + FC_WARPS * qcap);       // [FC_WARPS][FC_LISTCAP]
    __shared__ int sCnt[FC_WARPS];

    const int c = blockIdx.x, b = blockIdx.y;
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
    // zero score map: rows 0 .. ih + 1 (border rows included)
    for (int i = tid; i < ((ih + 2) * F2_MAPPITCH) >> 4; i += FC_THREADS) reinterpret_cast<uint4 *>(sM)[i] = make_uint4(0, 0, 0, 0);

    // ---- thread <-> pixel pair; a warp covers 1, 2 or 4 interior rows per trip; each warp owns a contiguous band of trips
    const int ppr = (iw + 1) >> 1;
    const int lprShift = ppr <= 8 ? 3 : (ppr <= 16 ? 4 : 5);
    const int lpr = 1 << lprShift, rpw = 32 >> lprShift;
    const int ngroups = (int)((unsigned int)(ih + rpw - 1) >> (5 - lprShift)); // / rpw
    const int gpw = (int)((unsigned int)(ngroups + FC_WARPS - 1) / (unsigned int)FC_WARPS);
    const int g0 = wid * gpw, g1 = min(g0 + gpw, ngroups);
    const int xl = lane & (lpr - 1), yl = lane >> lprShift;
    const int x = 2 * xl;
    const unsigned int xmask = (xl < ppr) ? ((x + 1 < iw) ? 0xFFFFFFFFu : 0x0000FFFFu) : 0u;
    const unsigned int *evenCopy = (a & 1) ? sS : sU, *oddCopy = (a & 1) ? sU : sS;
    const int cofs = (a & 1) ? ((a - 1) >> 1) : (a >> 1);
    const int oofs = (a & 1) ? cofs : cofs - 1;
    const unsigned int lt = (1u << lane) - 1u;
    unsigned short *queue = sQ + wid * qcap;
    unsigned int *list = sList + wid * FC_LISTCAP;
    const unsigned short *hU = reinterpret_cast<const unsigned short *>(sU); // staged row as 16-bit pixels, pitch 2 * FC_SPW
    __syncthreads();

    int total = 0, off = 0, wcnt = 0;
    for (int pass = 0; pass < 2; pass++) {
        const int th = pass ? P.minTh : P.iniTh;
        const unsigned int thp1 = (unsigned int)(th + 1) * 0x10001u;
        // ---- Phase A: bound test, queue the pixels that can reach th
        int qn = 0;
        {
            const unsigned int *pa = evenCopy + (g0 * rpw + yl + 3) * FC_SPW + xl + cofs;
            const unsigned int *pb = oddCopy + (g0 * rpw + yl + 3) * FC_SPW + xl + oofs;
            for (int gi = g0; gi < g1; gi++, pa += rpw * FC_SPW, pb += rpw * FC_SPW) {
                const int y = gi * rpw + yl;
                const unsigned int vmask = (y < ih) ? xmask : 0u;
                const unsigned int v = pa[0];
                unsigned int r[16];
                r[0] = pa[3 * FC_SPW];        // ( 0, 3)
                r[8] = pa[-3 * FC_SPW];       // ( 0,-3)
                r[4] = pb[2];                 // ( 3, 0)
                r[12] = pb[-1];               // (-3, 0)
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
                unsigned int mx[8], mn[8];
#pragma unroll
                for (int k = 0; k < 8; k++) { mx[k] = __vmaxs2(r[k], r[k + 8]); mn[k] = __vmins2(r[k], r[k + 8]); }
                const unsigned int ub = __vimin3_s16x2(__vimin3_s16x2(mx[0], mx[1], mx[2]), __vimin3_s16x2(mx[3], mx[4], mx[5]), __vmins2(mx[6], mx[7]));
                const unsigned int lb = __vimax3_s16x2(__vimax3_s16x2(mn[0], mn[1], mn[2]), __vimax3_s16x2(mn[3], mn[4], mn[5]), __vmaxs2(mn[6], mn[7]));
                const unsigned int q = __vmaxu2(bsub2(ub, v), bsub2(v, lb)); // biased bound on the strength M
                const unsigned int f = (q - thp1) & vmask;                    // bit 15 / 31 set <=> the pixel can be a corner at th
                const bool flo = (f & 0x8000u) != 0u, fhi = (f & 0x80000000u) != 0u;
                const unsigned int bl = __ballot_sync(0xffffffffu, flo), bh = __ballot_sync(0xffffffffu, fhi);
                const int pos = qn + __popc(bl & lt) + __popc(bh & lt);
                const unsigned int e = (unsigned int)x | ((unsigned int)y << 6);
                if (flo) queue[pos] = (unsigned short)e;
                if (fhi) queue[pos + (flo ? 1 : 0)] = (unsigned short)(e + 1u);
                qn += __popc(bl) + __popc(bh);
            }
        }
        __syncwarp();
        // ---- Phase B: exact score of the queued pixels, two per lane
        for (int i0 = 0; i0 < qn; i0 += 64) {
            const int ia = i0 + lane, ib = i0 + 32 + lane;
            const bool va = ia < qn, vb = ib < qn;
            const unsigned int ea = queue[va ? ia : i0], eb = queue[vb ? ib : i0];
            const int xa = (int)(ea & 63u), ya = (int)(ea >> 6), xbq = (int)(eb & 63u), ybq = (int)(eb >> 6);
            const unsigned short *ca = hU + (ya + 3) * (2 * FC_SPW) + xa + a;
            const unsigned short *cb = hU + (ybq + 3) * (2 * FC_SPW) + xbq + a;
            constexpr int RP = 2 * FC_SPW;
#define F2_LD(dx, dy) ((unsigned int)ca[(dy) * RP + (dx)] | ((unsigned int)cb[(dy) * RP + (dx)] << 16))
            const unsigned int v = F2_LD(0, 0);
            unsigned int r[16];
            r[0] = F2_LD(0, 3);   r[1] = F2_LD(1, 3);    r[2] = F2_LD(2, 2);    r[3] = F2_LD(3, 1);
            r[4] = F2_LD(3, 0);   r[5] = F2_LD(3, -1);   r[6] = F2_LD(2, -2);   r[7] = F2_LD(1, -3);
            r[8] = F2_LD(0, -3);  r[9] = F2_LD(-1, -3);  r[10] = F2_LD(-2, -2); r[11] = F2_LD(-3, -1);
            r[12] = F2_LD(-3, 0); r[13] = F2_LD(-3, 1);  r[14] = F2_LD(-2, 2);  r[15] = F2_LD(-1, 3);
#undef F2_LD
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
            const unsigned int Mb = __vmaxu2(bsub2(A, v), bsub2(v, Bm)); // biased strength M per half
            const unsigned int c1 = Mb - 0x00010001u;                    // half = 0x7FFF + M: bit 15 set <=> M >= 1
            const unsigned int S = c1 & neg_mask2(c1) & 0x00FF00FFu;      // score M - 1 where M >= 1, else 0
            if (va) sM[(ya + 1) * F2_MAPPITCH + xa + 1] = (unsigned char)(S & 0xFFu);
            if (vb) sM[(ybq + 1) * F2_MAPPITCH + xbq + 1] = (unsigned char)(S >> 16);
        }
        __syncthreads(); // every warp's scores are in the map
        // ---- NMS over the queued pixels; survivors appended in queue order = row-major order
        wcnt = 0;
        for (int i0 = 0; i0 < qn; i0 += 32) {
            const int i = i0 + lane;
            const bool vq = i < qn;
            const unsigned int e = queue[vq ? i : i0];
            const int xq = (int)(e & 63u), yq = (int)(e >> 6);
            const unsigned char *m = sM + (yq + 1) * F2_MAPPITCH + xq + 1;
            const int s = vq ? (int)m[0] : 0;
            bool keep = s >= th;
            if (!__any_sync(0xffffffffu, keep)) continue;
            if (keep) {
                const int n0 = max(max((int)m[-F2_MAPPITCH - 1], (int)m[-F2_MAPPITCH]), max((int)m[-F2_MAPPITCH + 1], (int)m[-1]));
                const int n1 = max(max((int)m[1], (int)m[F2_MAPPITCH - 1]), max((int)m[F2_MAPPITCH], (int)m[F2_MAPPITCH + 1]));
                keep = s > max(n0, n1);
            }
            const unsigned int bs = __ballot_sync(0xffffffffu, keep);
            if (keep) {
                const int pos = wcnt + __popc(bs & lt);
                if (pos < FC_LISTCAP) list[pos] = (r1.y + (uint32_t)xq + ((uint32_t)yq << 12)) | ((uint32_t)s << 24);
            }
            wcnt += __popc(bs);
        }
        wcnt = min(wcnt, FC_LISTCAP);
        if (lane == 0) sCnt[wid] = wcnt;
        __syncthreads();
        total = 0; off = 0;
#pragma unroll
        for (int w = 0; w < FC_WARPS; w++) {
            if (w < wid) off += sCnt[w];
            total += sCnt[w];
        }
        if (total > 0) break; // the iniTh set when it is not empty, else the minTh retry (:991-995)
        __syncthreads();      // sCnt is rewritten by the next pass
    }
    // ---- ordered copy of the warp's survivors into the cell's fixed slot
    uint32_t *out = P.cand + (size_t)b * P.candStride + r1.x + off;
    for (int i = lane; i < wcnt; i += 32) out[i] = list[i];
    if (tid == 0) *cellCount = total;
}

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
