// extract_kernels.cuh -- sm_100a kernels of the ORB extraction path.
//
// Data layout in HBM (per extractor handle, batch of B same-shape images):
//   pyr   [B][pyrStride]  un-blurred levels, level l at +imgOff_l, rows `pitch_l` apart (pitch % 16 == 0)
//   blur  [B][pyrStride]  7x7 Gaussian of every level, same offsets
//   cand  [B][candStride] u32 candidates (x | y<<12 | score<<24, relative to the 16-px border),
//                         level l at +candOff_l, cell c owns the fixed slot [c*cellCap_l, (c+1)*cellCap_l)
//   ord / knode           same shape as cand: candidates gathered in the reference's order, node id per key
//   lvlKp [B][nodeTotal]  selected keypoints per level (packed like cand), lvlCount [B][nlevels]
// Integer stages are exact; float expressions use explicit _rn intrinsics (and -fmad=false) so that no FMA
// contraction can change a rounding.
#pragma once
#include "common.cuh"

namespace orbb200 {

struct LevelGeom {
    int w, h, pitch;
    unsigned int imgOff;
    int nCols, nRows, wCell, hCell;
    int cellBase, nCells, cellCap;
    unsigned int candOff;
    int candCap;
    int nfeat, nIni;
    float hX;
    int nodeCap, nodeOff, sortLen;
    unsigned int nodeWsOff;
    float scale, patch;
    int blurTileBase, blurTilesX, blurTilesY;
    int rzTileSP, rzTileRows; // shared-memory source tile of the resize INTO this level
    unsigned int coefOff;     // offset of this level's resize tables
};

struct ExtractParams {
    int nlevels, batch, iniTh, minTh;
    const LevelGeom *geom;
    const uint8_t *l0; // level 0: caller's image (device API) or the level-0 slot of pyr (host API)
    size_t l0ImgStride;
    int l0Pitch, l0Valid;
    uint8_t *pyr, *blur;
    size_t pyrStride;
    uint32_t *cand, *ord;
    int *knode;
    size_t candStride;
    int *cellCount, *cellScan;
    int totalCells;
    int *nodeWs;
    size_t nodeWsStride;
    uint32_t *lvlKp;
    int nodeTotal;
    int *lvlCount;
    const int *rzXofs, *rzYofs; // resize tables, all levels concatenated (coefOff)
    const short2 *rzXa, *rzYb;
    int totalBlurTiles;
    const int *umax;
    orbb200_keypoint *kps;
    uint8_t *desc;
    int cap;
    int *nOut;
};

#ifdef ORBB200_EXTRACT_IMPL // kernel definitions: only orbb200_extract.cu instantiates them

__device__ __forceinline__ const uint8_t *level_ptr(const ExtractParams &P, const LevelGeom &g, int l, int b,
                                                    int &pitch, int &valid)
{
    if (l == 0) {
        pitch = P.l0Pitch;
        valid = P.l0Valid;
        return P.l0 + (size_t)b * P.l0ImgStride;
    }
    pitch = g.pitch;
    valid = g.pitch;
    return P.pyr + (size_t)b * P.pyrStride + g.imgOff;
}

__device__ __forceinline__ uint4 ldg_nc_v4(const void *p)
{
    uint4 r;
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// Stage rows [y0, y0+rows) x bytes [xb, xb+nbytes) of an image into shared memory (s[r*sp + c]).
// xb, nbytes, sp are multiples of 16.  16-byte loads where the source allows, byte loads otherwise; bytes at
// x >= w are zero-filled (never consumed).
__device__ __forceinline__ void load_tile(uint8_t *s, int sp, const uint8_t *img, int pitch, int valid, int w, int xb,
                                          int nbytes, int y0, int rows)
{
    const bool vec = ((((size_t)img) | (size_t)pitch) & 15) == 0;
    const int chunks = nbytes >> 4;
    for (int i = threadIdx.x; i < rows * chunks; i += blockDim.x) {
        const int r = i / chunks, c = i - r * chunks, x = xb + (c << 4);
        const uint8_t *src = img + (size_t)(y0 + r) * pitch + x;
        uint4 v;
        if (vec && x + 16 <= valid) {
            v = ldg_nc_v4(src);
        } else {
            unsigned int wd[4] = {0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < 16; k++)
                if (x + k < w) wd[k >> 2] |= (unsigned int)__ldg(src + k) << ((k & 3) * 8);
            v = make_uint4(wd[0], wd[1], wd[2], wd[3]);
        }
        *reinterpret_cast<uint4 *>(s + r * sp + (c << 4)) = v;
    }
}

// ---------------------------------------------------------------------------------------------------
// E2  pyramid_resize: level l from level l-1, cv::resize INTER_LINEAR 8U fixed-point arithmetic
//     (reference call site src/ORBextractor.cc:1366).  Tile = 128 x 16 outputs, 256 threads, each thread
//     4 adjacent pixels on 2 rows; the source footprint is staged in shared memory with 16-byte loads.
// ---------------------------------------------------------------------------------------------------
constexpr int RZ_TW = 128, RZ_TH = 16, RZ_THREADS = 256;

__global__ void __launch_bounds__(RZ_THREADS) k_pyramid_resize(const __grid_constant__ ExtractParams P, int l)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const LevelGeom g = P.geom[l], gs = P.geom[l - 1];
    const int b = blockIdx.z;
    int spitch, svalid;
    const uint8_t *src = level_ptr(P, gs, l - 1, b, spitch, svalid);
    uint8_t *dst = P.pyr + (size_t)b * P.pyrStride + g.imgOff;
    const int *xofs = P.rzXofs + g.coefOff, *yofs = P.rzYofs + g.coefOff;
    const short2 *xa = P.rzXa + g.coefOff, *yb = P.rzYb + g.coefOff;
    const int x0 = blockIdx.x * RZ_TW, y0 = blockIdx.y * RZ_TH;
    const int x1 = min(x0 + RZ_TW, g.w) - 1, y1 = min(y0 + RZ_TH, g.h) - 1;
    const int sw = gs.w, sh = gs.h;
    const int xb = xofs[x0] & ~15;
    const int sxLast = min(xofs[x1] + 1, sw - 1);
    const int nbytes = ((sxLast - xb + 1) + 15) & ~15;
    const int rFirst = min(max(yofs[y0], 0), sh - 1), rLast = min(max(yofs[y1] + 1, 0), sh - 1);
    const int sp = g.rzTileSP;
    load_tile(smem, sp, src, spitch, svalid, sw, xb, nbytes, rFirst, rLast - rFirst + 1);
    __syncthreads();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int dx0 = x0 + tx * 4;
    if (dx0 >= g.w) return;
    int sx[4], sx1[4];
    short2 a[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int dx = min(dx0 + k, g.w - 1);
        sx[k] = xofs[dx];
        sx1[k] = min(sx[k] + 1, sw - 1) - xb;
        sx[k] -= xb;
        a[k] = xa[dx];
    }
#pragma unroll
    for (int rr = 0; rr < 2; rr++) {
        const int dy = y0 + ty + rr * 8;
        if (dy > y1) break;
        const int sy = yofs[dy];
        const short2 bb = yb[dy];
        const uint8_t *S0 = smem + (min(max(sy, 0), sh - 1) - rFirst) * sp;
        const uint8_t *S1 = smem + (min(max(sy + 1, 0), sh - 1) - rFirst) * sp;
        unsigned int packed = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int H0 = S0[sx[k]] * a[k].x + S0[sx1[k]] * a[k].y;
            const int H1 = S1[sx[k]] * a[k].x + S1[sx1[k]] * a[k].y;
            int v = (((bb.x * (H0 >> 4)) >> 16) + ((bb.y * (H1 >> 4)) >> 16) + 2) >> 2;
            v = min(max(v, 0), 255);
            packed |= (unsigned int)v << (8 * k);
        }
        *reinterpret_cast<unsigned int *>(dst + (size_t)dy * g.pitch + dx0) = packed;
    }
}

// ---------------------------------------------------------------------------------------------------
// E7  blur7: cv::GaussianBlur 7x7 sigma 2, 8U fixed point (Q8 taps 18 34 48 56 48 34 18), REFLECT_101 at the
//     level's own edges (src/ORBextractor.cc:1312-1314).  All levels in one launch through a tile table.
//     Tile 128 x 32, 256 threads: horizontal pass u8 -> u16 in shared memory, vertical pass -> u8.
// ---------------------------------------------------------------------------------------------------
constexpr int BL_TW = 128, BL_TH = 32, BL_THREADS = 256;
constexpr int BL_SP = BL_TW + 32; // staged row: 16-aligned start, 3-px halo each side

__device__ __forceinline__ int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p;
    return p;
}

__global__ void __launch_bounds__(BL_THREADS) k_blur7(const __grid_constant__ ExtractParams P)
{
    __shared__ __align__(16) uint8_t sIn[(BL_TH + 6) * BL_SP];
    __shared__ __align__(16) unsigned short sH[(BL_TH + 6) * BL_TW];
    int l = 0;
    while (l + 1 < P.nlevels && (int)blockIdx.x >= P.geom[l + 1].blurTileBase) l++;
    const LevelGeom g = P.geom[l];
    const int b = blockIdx.y;
    const int t = blockIdx.x - g.blurTileBase;
    const int x0 = (t % g.blurTilesX) * BL_TW, y0 = (t / g.blurTilesX) * BL_TH;
    int pitch, valid;
    const uint8_t *src = level_ptr(P, g, l, b, pitch, valid);
    uint8_t *dst = P.blur + (size_t)b * P.pyrStride + g.imgOff;
    const int W = g.w, H = g.h;
    // stage rows y0-3 .. y0+TH+2 (reflected), columns [x0-16, x0+TW+16) clipped to the row (reflected later)
    const int xb = max(x0 - 16, 0);
    const int xe = min(x0 + BL_TW + 16, (W + 15) & ~15);
    const int nbytes = xe - xb;
    const int rows = min(BL_TH, H - y0) + 6;
    {
        const bool vec = ((((size_t)src) | (size_t)pitch) & 15) == 0;
        const int chunks = nbytes >> 4;
        for (int i = threadIdx.x; i < rows * chunks; i += blockDim.x) {
            const int r = i / chunks, c = i - r * chunks, x = xb + (c << 4);
            const int sy = reflect101(y0 - 3 + r, H);
            const uint8_t *p = src + (size_t)sy * pitch + x;
            uint4 v;
            if (vec && x + 16 <= valid) {
                v = ldg_nc_v4(p);
            } else {
                unsigned int wd[4] = {0, 0, 0, 0};
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if (x + k < W) wd[k >> 2] |= (unsigned int)__ldg(p + k) << ((k & 3) * 8);
                v = make_uint4(wd[0], wd[1], wd[2], wd[3]);
            }
            *reinterpret_cast<uint4 *>(sIn + r * BL_SP + (c << 4)) = v;
        }
    }
    __syncthreads();
    const int tw = min(BL_TW, W - x0);
    // horizontal: every staged row, tw columns
    for (int i = threadIdx.x; i < rows * BL_TW; i += blockDim.x) {
        const int r = i / BL_TW, x = i - r * BL_TW;
        if (x >= tw) continue;
        const uint8_t *row = sIn + r * BL_SP - xb;
        const int gx = x0 + x;
        int acc;
        if (gx >= 3 && gx < W - 3) {
            acc = 18 * (row[gx - 3] + row[gx + 3]) + 34 * (row[gx - 2] + row[gx + 2]) +
                  48 * (row[gx - 1] + row[gx + 1]) + 56 * row[gx];
        } else {
            acc = 18 * (row[reflect101(gx - 3, W)] + row[reflect101(gx + 3, W)]) +
                  34 * (row[reflect101(gx - 2, W)] + row[reflect101(gx + 2, W)]) +
                  48 * (row[reflect101(gx - 1, W)] + row[reflect101(gx + 1, W)]) + 56 * row[gx];
        }
        sH[r * BL_TW + x] = (unsigned short)acc;
    }
    __syncthreads();
    // vertical: thread -> 4 adjacent columns, rows strided by 8
    const int tx = (threadIdx.x & 31) * 4, ty = threadIdx.x >> 5;
    if (tx < tw) {
        for (int y = ty; y < rows - 6; y += 8) {
            unsigned int packed = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const unsigned short *c = sH + y * BL_TW + tx + k;
                unsigned int acc = 18u * (c[0] + c[6 * BL_TW]) + 34u * (c[BL_TW] + c[5 * BL_TW]) +
                                   48u * (c[2 * BL_TW] + c[4 * BL_TW]) + 56u * c[3 * BL_TW];
                packed |= ((acc + 32768u) >> 16) << (8 * k);
            }
            *reinterpret_cast<unsigned int *>(dst + (size_t)(y0 + y) * g.pitch + x0 + tx) = packed;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// E3+E4  fast_cells: one thread block per 30-px grid cell of any level (src/ORBextractor.cc:953-1006).
//   cv::FAST(roi, th, nms=true) semantics, derived in DESIGN.md:
//     score S(p) = max over the 16 contiguous 9-arcs of min|v - ring| - 1 (threshold independent);
//     corner at th  <=>  S >= th;   NMS keeps p iff S(p) > S(q) for the 8 neighbours q, where q outside the
//     ROI interior or not a corner at th counts 0.  For S(p) >= th the masked comparison equals the raw
//     one, so the survivors at iniTh are the local maxima with S >= iniTh and the minTh retry (taken only
//     when that set is empty) yields the local maxima with S >= minTh -- one scoring pass serves both.
//   The ROI interiors tile [19, W-19) x [19, H-19) without overlap, so every pixel is scored exactly once.
//   Output: the cell's candidates in row-major order in its fixed slot + the count.
// ---------------------------------------------------------------------------------------------------
constexpr int FC_THREADS = 128;
constexpr int FC_SP = 96;   // staged ROI row: <= 15 (alignment) + 66 (ROI) bytes
constexpr int FC_ROWS = 66; // hCell + 6 <= 66 (hCell < 60 because nRows = floor(height/30))
constexpr int FC_SCP = 64;  // score row pitch: interior <= 60 wide + zero border

// strength M of the FAST-9 test at a pixel: corner at threshold t  <=>  M > t.  c points into shared memory.
__device__ __forceinline__ int fast_arc_strength(const uint8_t *c, int th)
{
    const int v = c[0];
    int d[16];
    d[0] = v - c[3 * FC_SP];       d[8] = v - c[-3 * FC_SP];
    d[4] = v - c[3];               d[12] = v - c[-3];
    // quick reject: every 9-arc contains k or k+8 for each k
    bool br = ((d[0] > th) | (d[8] > th)) & ((d[4] > th) | (d[12] > th));
    bool dk = ((d[0] < -th) | (d[8] < -th)) & ((d[4] < -th) | (d[12] < -th));
    if (!(br | dk)) return 0;
    d[2] = v - c[2 * FC_SP + 2];   d[10] = v - c[-2 * FC_SP - 2];
    d[6] = v - c[-2 * FC_SP + 2];  d[14] = v - c[2 * FC_SP - 2];
    br = br & ((d[2] > th) | (d[10] > th)) & ((d[6] > th) | (d[14] > th));
    dk = dk & ((d[2] < -th) | (d[10] < -th)) & ((d[6] < -th) | (d[14] < -th));
    if (!(br | dk)) return 0;
    d[1] = v - c[3 * FC_SP + 1];   d[9] = v - c[-3 * FC_SP - 1];
    d[3] = v - c[FC_SP + 3];       d[11] = v - c[-FC_SP - 3];
    d[5] = v - c[-FC_SP + 3];      d[13] = v - c[FC_SP - 3];
    d[7] = v - c[-3 * FC_SP + 1];  d[15] = v - c[3 * FC_SP - 1];
    // sliding 9-window min and max over the circular 16-array with 3-input min/max (VIMNMX3)
    int mn3[16], mx3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        mn3[k] = __vimin3_s32(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
        mx3[k] = __vimax3_s32(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
    }
    // NOTE: keep the two reductions separate and negate once at the end.  ptxas 12.9 (sm_100a) miscompiles
    // max(best, mn9, -mx9) -- it folds the negation into VIMNMX3 and drops it (verified on a B200 with
    // tools/scratch/vimnmx.cu: even the plain-loop form returns max(d) instead of the arc strength).
    int bmn = 0, bmx = 0x7fffffff;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const int mn9 = __vimin3_s32(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]);
        const int mx9 = __vimax3_s32(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]);
        bmn = max(bmn, mn9);
        bmx = min(bmx, mx9);
    }
    return max(bmn, -bmx);
}

__global__ void __launch_bounds__(FC_THREADS) k_fast_cells(const __grid_constant__ ExtractParams P)
{
    __shared__ __align__(16) uint8_t sImg[FC_ROWS * FC_SP];
    __shared__ uint8_t sSc[(FC_ROWS - 6 + 2) * FC_SCP];
    __shared__ int scanTmp[34];
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
    int pitch, valid;
    const uint8_t *img = level_ptr(P, g, l, b, pitch, valid);
    const int xb = iniX & ~15;
    const int nbytes = ((maxX - xb) + 15) & ~15;
    load_tile(sImg, FC_SP, img, pitch, valid, g.w, xb, nbytes, iniY, rh);
    for (int i = threadIdx.x; i < (FC_ROWS - 4) * FC_SCP / 4; i += FC_THREADS) reinterpret_cast<int *>(sSc)[i] = 0;
    __syncthreads();
    const int iw = rw - 6, ih = rh - 6, npx = iw * ih;
    const int minTh = P.minTh, iniTh = P.iniTh;
    const uint8_t *org = sImg + 3 * FC_SP + (iniX - xb) + 3;
    for (int p = threadIdx.x; p < npx; p += FC_THREADS) {
        const int y = p / iw, x = p - y * iw;
        const int M = fast_arc_strength(org + y * FC_SP + x, minTh);
        if (M > minTh) sSc[(y + 1) * FC_SCP + x + 1] = (uint8_t)(M - 1);
    }
    __syncthreads();
    // local maxima, emitted in row-major order: thread t owns pixels [t*chunk, (t+1)*chunk)
    const int chunk = (npx + FC_THREADS - 1) / FC_THREADS; // <= 29
    const int p0 = threadIdx.x * chunk, p1 = min(p0 + chunk, npx);
    unsigned int lmMask = 0, iniMask = 0;
    for (int p = p0; p < p1; p++) {
        const int y = p / iw, x = p - y * iw;
        const uint8_t *s = sSc + (y + 1) * FC_SCP + x + 1;
        const int v = s[0];
        if (v == 0) continue;
        if (v > s[-1] && v > s[1] && v > s[-FC_SCP - 1] && v > s[-FC_SCP] && v > s[-FC_SCP + 1] && v > s[FC_SCP - 1] &&
            v > s[FC_SCP] && v > s[FC_SCP + 1]) {
            lmMask |= 1u << (p - p0);
            if (v >= iniTh) iniMask |= 1u << (p - p0);
        }
    }
    const int anyIni = __syncthreads_or(iniMask != 0);
    const unsigned int mask = anyIni ? iniMask : lmMask;
    int total;
    int off = block_excl_scan(__popc(mask), scanTmp, &total);
    uint32_t *out = P.cand + (size_t)b * P.candStride + g.candOff + (size_t)cc * g.cellCap;
    unsigned int m = mask;
    while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        const int p = p0 + k, y = p / iw, x = p - y * iw;
        const int s = sSc[(y + 1) * FC_SCP + x + 1];
        out[off++] = (uint32_t)(cj * g.wCell + x + 3) | ((uint32_t)(ci * g.hCell + y + 3) << 12) | ((uint32_t)s << 24);
    }
    if (threadIdx.x == 0) *cellCount = total;
}

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
