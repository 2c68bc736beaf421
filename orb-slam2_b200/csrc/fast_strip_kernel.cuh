// fast_strip_kernel.cuh -- E3+E4  k_fast_strips: the cell loop + cv::FAST(roi, th, nms = true) + the iniTh / minTh retry
// (src/ORBextractor.cc:925-1009), one thread block per STRIP of grid cells.
//
// cv::FAST semantics used (derived and pinned against cv2 in tests/test_oracle_pin.py, against the reference in
// tests/test_ref_pin.py):
//   score S(p) = max over the 16 contiguous 9-arcs of min|v - ring| - 1 (threshold independent); corner at th <=> S >= th;
//   NMS keeps p iff S(p) > S(q) for its 8 neighbours q, where q outside the ROI interior or not a corner at th counts 0.
//   For S(p) >= th the masked comparison equals the raw one, so the survivors at th are the strict local maxima with
//   S >= th of a map that holds exact scores wherever S >= th (anything below th elsewhere).
//   The interiors of the cell ROIs tile the bordered level without overlap: every pixel is scored once, and "outside the
//   ROI interior" becomes "in another cell" -- a seam mask on the horizontal neighbours plus zero rows above / below.
//
// Why strips (ncu of the per-cell kernel it replaces, profiles/r1_fast_cells2_full_details.txt): one 128-thread block
// per ~35 x 35 cell spent a third of its 118 thread-instructions per pixel in block prologue / staging / set-up, ran its
// bound pass with 18 of 32 lanes busy (a cell row is ~18 pixel pairs) and staged 1.37 x the pixels it scored.  A strip is
// one cell row x up to 160 interior pixels (2..5 cells): the fixed costs are spread over 4 x the pixels, the bound pass
// walks the strip as ONE linear index space (all lanes busy whatever the cell width) and the halo shrinks to 1.2 x.
//
//   stage   ROI rows widened to 16 bit in two parities (U: pairs at even tile columns, S: odd), 128-bit loads / stores
//   scan    Phase A, every pixel, two per thread in packed 16-bit lanes: every 9-arc holds ring pixel k or k + 8, so
//           q = max(min_k max(r_k, r_k+8) - v, v - max_k min(r_k, r_k+8)) >= M = S + 1 over any subset of k; with the four
//           even k (8 loads) 24 % of the pixels at th = 20 have q > th and enter the warp's ring queue (2 ballots; all eight
//           k: 19 %, at twice the loads and min/max).  Phase B whenever 64 entries wait: ring by 16-bit LDS,
//           the exact network (80 VIMNMX3.S16x2), score byte into the strip's score map where S >= th.
//   nms     over the non-zero bytes of the map (a word test skips 4 empty pixels at once) -> one bit per survivor
//   cells   a warp per cell: per-row popcounts of the cell's bit columns -> scan -> the cell's survivor count; a cell that
//           is empty at iniTh is scanned again by that warp alone with minTh (:991-995) and its bits recounted; the bits
//           are enumerated row by row into the cell's fixed slot = the reference's cell-major / row-major order.
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

#ifdef ORBB200_EXTRACT_IMPL

#ifndef ORBB200_FS_THREADS
#define ORBB200_FS_THREADS 128
#endif
constexpr int FS_THREADS = ORBB200_FS_THREADS, FS_WARPS = FS_THREADS / 32;
#ifndef ORBB200_FS_MINBLOCKS
#define ORBB200_FS_MINBLOCKS 8
#endif
constexpr int FS_MINBLOCKS = ORBB200_FS_MINBLOCKS; // blocks per SM the register allocation aims at
#ifndef ORBB200_FS_MAXIW
#define ORBB200_FS_MAXIW 128
#endif
constexpr int FS_MAXIW = ORBB200_FS_MAXIW;                        // interior pixels per strip row
constexpr int FS_SPW = ((FS_MAXIW + 15 + 3 + 3 + 1 + 15) / 16) * 8; // staged row pitch in 32-bit words per copy (16-pixel chunks)
constexpr int FS_MAXROWS = 66;                                      // hCell + 6 <= 66 (hCell < 60 because nRows = floor(height / 30))
constexpr int FS_MAPPITCH = ((FS_MAXIW + 5 + 15) / 16) * 16;      // score map row pitch in bytes; interior x lives at column x + 4
constexpr int FS_BMW = (FS_MAXIW + 31) / 32;                        // bitmap words per interior row
constexpr int FS_QCAP = 128;      // ring queue entries per warp (at most 63 waiting + 64 appended per trip)
constexpr int FS_REC = 3;         // uint4 per strip record
constexpr int FS_LISTCAP = 512;   // corner list entries per warp; a warp that finds more sends the block's NMS down the map-scan path

__host__ __device__ inline size_t fast_strip_smem(int rows)
{
    return (size_t)rows * FS_SPW * 4 + (size_t)(rows - 4) * FS_MAPPITCH;
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

struct FsCtx {
    const unsigned int *pairs;              // the widened copy: word k of a row = tile pixels (e0 + 2k, e0 + 2k + 1), e0 = parity of interior x = 0
    const unsigned short *hU;               // the same copy as 16-bit pixels, pitch 2 * FS_SPW
    unsigned char *sM;                      // score map
    unsigned short *queue;                  // this warp's ring queue
    unsigned short *corners;                // this warp's list of the pixels it scored >= th (input of the NMS)
    int cofs, a;                            // word of interior pair 0; 16-bit index of interior x = 0
};

// Phase B on up to 64 queue entries starting at the ring head: exact score, stored where it reaches th
__device__ __forceinline__ void fs_phase_b(const FsCtx &c, int qh, int n, int th, int lane, int &ncorn)
{
    const int ia = lane, ib = lane + 32;
    const bool va = ia < n, vb = ib < n;
    const unsigned int ea = c.queue[(qh + (va ? ia : 0)) & (FS_QCAP - 1)], eb = c.queue[(qh + (vb ? ib : 0)) & (FS_QCAP - 1)];
    const int xa = (int)(ea & 255u), ya = (int)(ea >> 8), xb = (int)(eb & 255u), yb = (int)(eb >> 8);
    constexpr int RP = 2 * FS_SPW;
    const unsigned short *ca = c.hU + (ya + 3) * RP + xa + c.a;
    const unsigned short *cb = c.hU + (yb + 3) * RP + xb + c.a;
#define FS_LD(dx, dy) ((unsigned int)ca[(dy) * RP + (dx)] | ((unsigned int)cb[(dy) * RP + (dx)] << 16))
    const unsigned int v = FS_LD(0, 0);
    unsigned int r[16];
    r[0] = FS_LD(0, 3);   r[1] = FS_LD(1, 3);    r[2] = FS_LD(2, 2);    r[3] = FS_LD(3, 1);
    r[4] = FS_LD(3, 0);   r[5] = FS_LD(3, -1);   r[6] = FS_LD(2, -2);   r[7] = FS_LD(1, -3);
    r[8] = FS_LD(0, -3);  r[9] = FS_LD(-1, -3);  r[10] = FS_LD(-2, -2); r[11] = FS_LD(-3, -1);
    r[12] = FS_LD(-3, 0); r[13] = FS_LD(-3, 1);  r[14] = FS_LD(-2, 2);  r[15] = FS_LD(-1, 3);
#undef FS_LD
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
    // A = max_arcs min9, B = min_arcs max9.  The two reductions stay separate and the negation is applied once below:
    // ptxas 12.9 drops a negation folded into VIMNMX3.
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
    const int sa = (int)(S & 0xFFu), sb = (int)(S >> 16);
    const bool ka = va && sa >= th, kb = vb && sb >= th;
    const unsigned int ba = __ballot_sync(0xffffffffu, ka), bb = __ballot_sync(0xffffffffu, kb);
    if ((ba | bb) == 0u) return;
    const unsigned int lt = (1u << lane) - 1u;
    if (ka) {
        c.sM[(ya + 1) * FS_MAPPITCH + xa + 4] = (unsigned char)sa;
        const int pos = ncorn + __popc(ba & lt);
        if (pos < FS_LISTCAP) c.corners[pos] = (unsigned short)ea;
    }
    if (kb) {
        c.sM[(yb + 1) * FS_MAPPITCH + xb + 4] = (unsigned char)sb;
        const int pos = ncorn + __popc(ba) + __popc(bb & lt);
        if (pos < FS_LISTCAP) c.corners[pos] = (unsigned short)eb;
    }
    ncorn += __popc(ba) + __popc(bb);
}

// Phases A + B over interior columns [x0, x1) x rows row0, row0 + rowStep, ... < ih of the strip.  A trip is one row x 32 pixel
// pairs (lane <-> pair: consecutive words of one staged row, conflict-free; a trip that wrapped into the next row would put two
// lanes on one bank for all 17 loads).  The pair grid is the strip's: pairs start at even interior x, pixels outside [x0, x1)
// are masked.
__device__ __forceinline__ void fs_scan(const FsCtx &c, int x0, int x1, int ih, int th, int lane, int row0, int rowStep, int &ncorn)
{
    const int p0 = x0 >> 1, ppr = ((x1 + 1) >> 1) - p0;
    if (ppr <= 0) return;
    const unsigned int thp1 = (unsigned int)(th + 1) * 0x10001u;
    const unsigned int lt = (1u << lane) - 1u;
    int qh = 0, qn = 0;
    for (int seg = 0; seg < ppr; seg += 32) {
        const int xr = seg + lane; // pair within the region's row
        const bool inr = xr < ppr;
        const int xl = p0 + (inr ? xr : 0);
        const int x = 2 * xl;
        unsigned int vmask = 0u;
        if (inr) vmask = ((x >= x0) ? 0x0000FFFFu : 0u) | ((x + 1 < x1) ? 0xFFFF0000u : 0u);
        const unsigned int *pa = c.pairs + (row0 + 3) * FS_SPW + xl + c.cofs;
        unsigned int e = (unsigned int)x | ((unsigned int)row0 << 8);
        for (int y = row0; y < ih; y += rowStep, pa += rowStep * FS_SPW, e += (unsigned int)rowStep << 8) {
            const unsigned int v = pa[0];
            // every 9-arc holds ring pixel k or k + 8 for EVERY k; the four even pairs (compass + diagonals: 8 loads instead of 16)
            // already reject ~76 % of the pixels at th = 20 (all eight: 81 %), and the exact network of Phase B costs less than
            // the eight loads and sixteen min/max saved on every pixel (measured: 1.78 -> 1.71 ms; the two compass pairs alone: 1.79 ms)
            const unsigned int r0 = pa[3 * FS_SPW];        // ( 0, 3)
            const unsigned int r8 = pa[-3 * FS_SPW];       // ( 0,-3)
            const unsigned int r4 = __byte_perm(pa[1], pa[2], 0x5432);    // ( 3, 0): odd offset = second pixel of word + 1, first of word + 2
            const unsigned int r12 = __byte_perm(pa[-2], pa[-1], 0x5432); // (-3, 0)
            const unsigned int r2 = pa[2 * FS_SPW + 1];    // ( 2, 2)
            const unsigned int r6 = pa[-2 * FS_SPW + 1];   // ( 2,-2)
            const unsigned int r10 = pa[-2 * FS_SPW - 1];  // (-2,-2)
            const unsigned int r14 = pa[2 * FS_SPW - 1];   // (-2, 2)
            const unsigned int ub = __vimin3_s16x2(__vmaxs2(r0, r8), __vmaxs2(r4, r12), __vmins2(__vmaxs2(r2, r10), __vmaxs2(r6, r14)));
            const unsigned int lb = __vimax3_s16x2(__vmins2(r0, r8), __vmins2(r4, r12), __vmaxs2(__vmins2(r2, r10), __vmins2(r6, r14)));
            const unsigned int q = __vmaxu2(bsub2(ub, v), bsub2(v, lb)); // biased bound on the strength M
            const unsigned int f = (q - thp1) & vmask;                    // bit 15 / 31 set <=> the pixel can be a corner at th
            const bool flo = (f & 0x8000u) != 0u, fhi = (f & 0x80000000u) != 0u;
            const unsigned int bl = __ballot_sync(0xffffffffu, flo), bh = __ballot_sync(0xffffffffu, fhi);
            if ((bl | bh) != 0u) {
                const int pos = qh + qn + __popc(bl & lt) + __popc(bh & lt);
                if (flo) c.queue[pos & (FS_QCAP - 1)] = (unsigned short)e;
                if (fhi) c.queue[(pos + (flo ? 1 : 0)) & (FS_QCAP - 1)] = (unsigned short)(e + 1u);
                qn += __popc(bl) + __popc(bh);
                if (qn >= 64) {
                    __syncwarp();
                    fs_phase_b(c, qh, 64, th, lane, ncorn);
                    __syncwarp();
                    qh = (qh + 64) & (FS_QCAP - 1);
                    qn -= 64;
                }
            }
        }
    }
    if (qn > 0) {
        __syncwarp();
        fs_phase_b(c, qh, qn, th, lane, ncorn);
    }
    __syncwarp();
}

// 3x3 strict NMS of one scored pixel against the map: one bit per survivor
__device__ __forceinline__ void fs_nms_pixel(const unsigned char *sM, const unsigned char *seam, unsigned int *bitmap, int x, int y)
{
    const unsigned char *m = sM + (y + 1) * FS_MAPPITCH + 4 + x;
    const int s = (int)m[0], sm = seam[x];
    int n = max((int)m[-FS_MAPPITCH], (int)m[FS_MAPPITCH]);
    if (!(sm & 1)) n = max(n, max(max((int)m[-FS_MAPPITCH - 1], (int)m[-1]), (int)m[FS_MAPPITCH - 1]));
    if (!(sm & 2)) n = max(n, max(max((int)m[-FS_MAPPITCH + 1], (int)m[1]), (int)m[FS_MAPPITCH + 1]));
    if (s > n) atomicOr(bitmap + y * FS_BMW + (x >> 5), 1u << (x & 31));
}
// ... over a warp's corner list (the normal path: all lanes busy, only scored pixels are visited)
__device__ __forceinline__ void fs_nms_list(const unsigned char *sM, const unsigned char *seam, unsigned int *bitmap, const unsigned short *list,
                                            int n, int lane)
{
    for (int i = lane; i < n; i += 32) {
        const unsigned int e = list[i];
        fs_nms_pixel(sM, seam, bitmap, (int)(e & 255u), (int)(e >> 8));
    }
}

// ... over the non-zero score bytes of interior columns [x0, x1) (fallback when a corner list overflowed: a strip where more
// than ~40 % of the pixels are corners).  Work items (row, 4-pixel word) item0, item0 + itemStep, ...
__device__ __forceinline__ void fs_nms(const unsigned char *sM, const unsigned char *seam, unsigned int *bitmap, int x0, int x1, int ih,
                                       int item0, int itemStep)
{
    const int w0 = x0 >> 2, nw = ((x1 + 3) >> 2) - w0, total = ih * nw;
    if (total <= 0) return;
    const unsigned int magic = 0xFFFFFFFFu / (unsigned int)nw + 1u; // ceil(2^32 / nw); wraps to 0 for nw == 1 (handled below)
    for (int i = item0; i < total; i += itemStep) {
        const int y = nw == 1 ? i : (int)__umulhi((unsigned int)i, magic);
        const int xw = w0 + (i - y * nw);
        const unsigned char *row = sM + (y + 1) * FS_MAPPITCH + 4;
        const unsigned int word = *reinterpret_cast<const unsigned int *>(row + 4 * xw);
        if (word == 0u) continue;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int x = 4 * xw + k;
            if (((word >> (8 * k)) & 0xFFu) == 0u || x < x0 || x >= x1) continue;
            fs_nms_pixel(sM, seam, bitmap, x, y);
        }
    }
}

// the survivor bits of row y inside interior columns [x0, x1) of word wi, as a mask
__device__ __forceinline__ unsigned int fs_row_bits(const unsigned int *bitmap, int y, int wi, int x0, int x1)
{
    unsigned int w = bitmap[y * FS_BMW + wi];
    const int lo = x0 - 32 * wi, hi = x1 - 32 * wi; // keep bits [lo, hi)
    if (lo > 0) w &= ~((1u << lo) - 1u);
    if (hi < 32) w &= (hi <= 0) ? 0u : ((1u << hi) - 1u);
    return w;
}

__global__ void __launch_bounds__(FS_THREADS, FS_MINBLOCKS) k_fast_strips(const __grid_constant__ ExtractParams P, int rowsCap)
{
    extern __shared__ __align__(16) unsigned char fsm[];
    unsigned int *sU = reinterpret_cast<unsigned int *>(fsm);                     // [rowsCap][FS_SPW] the widened ROI
    unsigned char *sM = reinterpret_cast<unsigned char *>(sU + rowsCap * FS_SPW); // [(rowsCap - 4)][FS_MAPPITCH]
    __shared__ unsigned short sQ[FS_WARPS][FS_QCAP];
    __shared__ unsigned int sBits[(FS_MAXROWS - 4) * FS_BMW];
    __shared__ unsigned char sSeam[FS_MAXIW + 16];
    __shared__ unsigned short sCorn[FS_WARPS][FS_LISTCAP];
    __shared__ int sOverflow;

    const int sidx = blockIdx.x, b = blockIdx.y;
    const uint4 r0 = ldg_nc_v4(P.fastStrips + FS_REC * sidx), r1 = ldg_nc_v4(P.fastStrips + FS_REC * sidx + 1),
                r2 = ldg_nc_v4(P.fastStrips + FS_REC * sidx + 2);
    const int rw = (int)(r0.y & 0xFFFFu), rh = (int)(r0.y >> 16);
    const int nAct = (int)(r1.z >> 4) & 15, nAll = (int)(r1.z >> 8) & 255, wCell = (int)(r1.z >> 16) & 255;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int *cellCount = P.cellCount + (size_t)b * P.totalCells + r2.x;
    if (nAct == 0 || rw == 0) { // skipped cell row / cells (:957-976)
        for (int c = tid; c < nAll; c += FS_THREADS) cellCount[c] = 0;
        return;
    }
    const int iniX = (int)(r0.x & 0xFFFFu), iniY = (int)(r0.x >> 16), maxX = iniX + rw;
    const int gw = (int)(r0.w >> 16);
    int pitch, valid;
    const uint8_t *img;
    if ((r1.z & 15u) == 0u) { pitch = P.l0Pitch; valid = P.l0Valid; img = P.l0 + (size_t)b * P.l0ImgStride; }
    else { pitch = (int)(r0.w & 0xFFFFu); valid = pitch; img = P.pyr + (size_t)b * P.pyrStride + r0.z; }
    // tile origin: 16-byte aligned, and such that the ring of interior x = 0 (tile column a - 3) stays inside the widened copy,
    // whose first pixel is tile column e0 = a & 1
    int xb = iniX & ~15;
    if (iniX - xb == 0) xb -= 16;                       // a would be 3 (odd): column 0 would be missing; iniX >= 16, so xb stays >= 0
    const int chunks = (((maxX - xb) + 15) & ~15) >> 4; // <= FS_SPW / 8
    const int a = iniX + 3 - xb;                        // tile column of interior x = 0
    const int e0 = a & 1;
    const int iw = rw - 6, ih = rh - 6;
    // ---- stage the ROI widened to 16 bit, ONE copy: word k of a row = tile pixels (e0 + 2k, e0 + 2k + 1), so that the interior
    //      pixel pairs (2j, 2j + 1) are whole words.  (A second copy of the other parity served the odd ring offsets with plain
    //      loads; it cost a third of the block's shared memory -- 5 blocks per SM instead of 7 -- and half of the staging work.)
    {
        const bool vec = ((((size_t)img) | (size_t)pitch) & 15) == 0;
        const unsigned int magic = r2.z; // ceil(65536 / chunks), chunks as computed above (orbb200_extract.cu): i / chunks for i < 66 * 12
        for (int i = tid; i < rh * chunks; i += FS_THREADS) {
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
            unsigned int E[8];
            if (e0) {
                wd[4] = (x + 16 < gw) ? (unsigned int)__ldg(src + 16) : 0u;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const unsigned int sh = __funnelshift_r(wd[j], wd[j + 1], 8);
                    E[2 * j] = widen_lo(sh);
                    E[2 * j + 1] = widen_hi(sh);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    E[2 * j] = widen_lo(wd[j]);
                    E[2 * j + 1] = widen_hi(wd[j]);
                }
            }
            uint4 *rowE = reinterpret_cast<uint4 *>(sU + r * FS_SPW + 8 * ch);
            rowE[0] = make_uint4(E[0], E[1], E[2], E[3]);
            rowE[1] = make_uint4(E[4], E[5], E[6], E[7]);
        }
    }
    // zero the score map (rows 0 .. ih + 1: the border rows included), the survivor bits; seam flags of every interior column
    for (int i = tid; i < ((ih + 2) * FS_MAPPITCH) >> 4; i += FS_THREADS) reinterpret_cast<uint4 *>(sM)[i] = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < ih * FS_BMW; i += FS_THREADS) sBits[i] = 0u;
    for (int x = tid; x < iw; x += FS_THREADS) {
        const int xc = x % wCell;
        sSeam[x] = (unsigned char)((xc == 0 ? 1 : 0) | ((xc == wCell - 1 || x == iw - 1) ? 2 : 0)); // bit 0 / 1: left / right neighbour is another cell
    }
    FsCtx c;
    c.pairs = sU;
    c.cofs = (a - e0) >> 1;
    c.a = a - e0;
    c.hU = reinterpret_cast<const unsigned short *>(sU);
    c.sM = sM;
    c.queue = sQ[wid];
    c.corners = sCorn[wid];
    if (tid == 0) sOverflow = 0;
    __syncthreads();

    // ---- the whole strip at iniThFAST: bound pass + exact scores, then NMS -> survivor bits
    int ncorn = 0;
    fs_scan(c, 0, iw, ih, P.iniTh, lane, wid, FS_WARPS, ncorn);
    if (ncorn > FS_LISTCAP && lane == 0) sOverflow = 1;
    __syncthreads();
    if (!sOverflow) fs_nms_list(sM, sSeam, sBits, c.corners, ncorn, lane);
    else fs_nms(sM, sSeam, sBits, 0, iw, ih, tid, FS_THREADS);
    __syncthreads();

    // ---- a warp per cell: count its survivor bits, write them out in order; cells that are empty at iniThFAST are collected
    //      and scanned again with minThFAST (:991-995) by the WHOLE block, then counted and written the same way
    const int cellCap = (int)r2.y;
    __shared__ int sRetry[16];
    __shared__ int sNRetry;
    if (tid == 0) sNRetry = 0;
    __syncthreads();
    for (int round = 0; round < 2; round++) {
        const int ncand = round ? sNRetry : nAll;
        for (int k = wid; k < ncand; k += FS_WARPS) {
            const int cc = round ? sRetry[k] : k;
            if (cc >= nAct) {
                if (lane == 0) cellCount[cc] = 0;
                continue;
            }
            const int cx0 = cc * wCell, cx1 = min(cx0 + wCell, iw);
            const int wi0 = cx0 >> 5, wi1 = (cx1 - 1) >> 5;
            const int ya = lane, yb = lane + 32;
            int ca = 0, cb = 0;
            for (int wi = wi0; wi <= wi1; wi++) {
                if (ya < ih) ca += __popc(fs_row_bits(sBits, ya, wi, cx0, cx1));
                if (yb < ih) cb += __popc(fs_row_bits(sBits, yb, wi, cx0, cx1));
            }
            int sa = ca, sb = cb;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int ta = __shfl_up_sync(0xffffffffu, sa, d), tb = __shfl_up_sync(0xffffffffu, sb, d);
                if (lane >= d) { sa += ta; sb += tb; }
            }
            const int totA = __shfl_sync(0xffffffffu, sa, 31);
            const int tot = totA + __shfl_sync(0xffffffffu, sb, 31);
            if (tot == 0 && round == 0) {
                if (lane == 0) sRetry[atomicAdd(&sNRetry, 1)] = cc;
                continue;
            }
            uint32_t *out = P.cand + (size_t)b * P.candStride + r1.x + (size_t)cc * cellCap;
#pragma unroll
            for (int half = 0; half < 2; half++) {
                const int y = half ? yb : ya;
                int o = half ? (totA + sb - cb) : (sa - ca);
                if (y >= ih) continue;
                for (int wi = wi0; wi <= wi1; wi++) {
                    unsigned int bits = fs_row_bits(sBits, y, wi, cx0, cx1);
                    while (bits) {
                        const int x = 32 * wi + (__ffs(bits) - 1);
                        bits &= bits - 1u;
                        const uint32_t sc = sM[(y + 1) * FS_MAPPITCH + x + 4];
                        out[o++] = (r1.y + (uint32_t)x + ((uint32_t)y << 12)) | (sc << 24);
                    }
                }
            }
            if (lane == 0) cellCount[cc] = tot;
        }
        if (round == 1) break;
        __syncthreads();
        const int nRetry = sNRetry;
        if (nRetry == 0) break;
        // cv::FAST again with minThFAST on the empty cells: all warps share every cell's trips / items
        if (tid == 0) sOverflow = 0; // (every thread has read it before the barrier above)
        ncorn = 0;
        for (int k = 0; k < nRetry; k++) {
            const int cc = sRetry[k], cx0 = cc * wCell;
            fs_scan(c, cx0, min(cx0 + wCell, iw), ih, P.minTh, lane, wid, FS_WARPS, ncorn);
        }
        __syncthreads();
        if (ncorn > FS_LISTCAP && lane == 0) sOverflow = 1;
        __syncthreads();
        if (!sOverflow) fs_nms_list(sM, sSeam, sBits, c.corners, ncorn, lane);
        else
            for (int k = 0; k < nRetry; k++) {
                const int cc = sRetry[k], cx0 = cc * wCell;
                fs_nms(sM, sSeam, sBits, cx0, min(cx0 + wCell, iw), ih, tid, FS_THREADS);
            }
        __syncthreads();
    }
}

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
