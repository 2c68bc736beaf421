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
    int octDepth;             // depth of the quadtree kernel's histogram (0: iterative formulation only), octree_kernel.cuh
    unsigned int nodeWsOff;
    float scale, patch;
    int blurTileBase, blurTilesX, blurTilesY;
    int rzTileSP, rzTileRows; // shared-memory source tile of the resize INTO this level
    unsigned int coefOff;     // offset of this level's resize tables
};

struct ExtractParams {
    int nlevels, batch, iniTh, minTh;
    const LevelGeom *geom;
    const uint4 *fastStrips; // FS_REC uint4 per FAST strip (all levels): ROI, image slot, cells, output slot (orbb200_extract.cu)
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
    int octDepthCap;          // orbb200_extractor_set_octree_depth: caps LevelGeom::octDepth (tests force the fallback with it)
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
//     (reference call site src/ORBextractor.cc:1366).  Tile = 128 x 64 outputs, 256 threads, each thread 4 adjacent
//     pixels on 8 rows; the source footprint is staged in shared memory with 16-byte loads.
//     ncu of the first version: issue bound (IPC 3.4) at 60 thread instructions per pixel.  Now per source row a thread
//     reads 4 aligned words and two PRMTs with per-thread selectors turn them into the four (p[sx], p[sx+1]) pairs;
//     each horizontal interpolation p0*a0 + p1*a1 is one DP2A (16-bit coefficients x 8-bit pixels), each vertical
//     term (b * (H >> 4)) >> 16 one shift + one IMAD.HI against b << 16.  The result never leaves [0, 255]
//     (a0 + a1 = b0 + b1 = 2048), so cv's saturate_cast needs no code.
// ---------------------------------------------------------------------------------------------------
constexpr int RZ_TW = 128, RZ_TH = 64, RZ_THREADS = 256, RZ_SMEM_PAD = 16;

__device__ __forceinline__ unsigned int prmt_dyn(unsigned int a, unsigned int b, unsigned int sel)
{
    unsigned int r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}

__global__ void __launch_bounds__(RZ_THREADS, 8) k_pyramid_resize(const __grid_constant__ ExtractParams P, int l)
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
    // per-thread column set-up: staged offsets of the 4 source columns, packed 16-bit coefficient pairs, and the byte
    // selectors that pull (p[sx_k], p[sx_k + 1]) for k = 0,1 out of the two words at base01 and for k = 2,3 at base23
    int so[4];
    unsigned int coef[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int dx = min(dx0 + k, g.w - 1);
        so[k] = xofs[dx] - xb;
        const short2 a = xa[dx];
        coef[k] = (unsigned int)(unsigned short)a.x | ((unsigned int)(unsigned short)a.y << 16);
    }
    const int base01 = so[0] & ~3, base23 = so[2] & ~3;
    const int o0 = so[0] - base01, o1 = so[1] - base01, o2 = so[2] - base23, o3 = so[3] - base23;
    const bool fast = (o1 >= 0 && o1 + 1 < 8 && o3 >= 0 && o3 + 1 < 8); // always true for scale factors up to 2
    const unsigned int sel01 = (unsigned int)o0 | ((unsigned int)(o0 + 1) << 4) | ((unsigned int)o1 << 8) | ((unsigned int)(o1 + 1) << 12);
    const unsigned int sel23 = (unsigned int)o2 | ((unsigned int)(o2 + 1) << 4) | ((unsigned int)o3 << 8) | ((unsigned int)(o3 + 1) << 12);
#pragma unroll
    for (int rr = 0; rr < RZ_TH / 8; rr++) {
        const int dy = y0 + ty + rr * 8;
        if (dy > y1) break;
        const int sy = yofs[dy];
        const short2 bb = yb[dy];
        const unsigned int bx16 = (unsigned int)(unsigned short)bb.x << 16, by16 = (unsigned int)(unsigned short)bb.y << 16;
        const uint8_t *S0 = smem + (min(max(sy, 0), sh - 1) - rFirst) * sp;
        const uint8_t *S1 = smem + (min(max(sy + 1, 0), sh - 1) - rFirst) * sp;
        unsigned int p0a, p0b, p1a, p1b; // row 0 / row 1, pixel pairs (k = 0,1) / (k = 2,3)
        if (fast) {
            const unsigned int *q0 = reinterpret_cast<const unsigned int *>(S0 + base01), *q1 = reinterpret_cast<const unsigned int *>(S0 + base23);
            const unsigned int *r0 = reinterpret_cast<const unsigned int *>(S1 + base01), *r1 = reinterpret_cast<const unsigned int *>(S1 + base23);
            p0a = prmt_dyn(q0[0], q0[1], sel01);
            p0b = prmt_dyn(q1[0], q1[1], sel23);
            p1a = prmt_dyn(r0[0], r0[1], sel01);
            p1b = prmt_dyn(r1[0], r1[1], sel23);
        } else {
            p0a = S0[so[0]] | (S0[so[0] + 1] << 8) | (S0[so[1]] << 16) | (S0[so[1] + 1] << 24);
            p0b = S0[so[2]] | (S0[so[2] + 1] << 8) | (S0[so[3]] << 16) | (S0[so[3] + 1] << 24);
            p1a = S1[so[0]] | (S1[so[0] + 1] << 8) | (S1[so[1]] << 16) | (S1[so[1] + 1] << 24);
            p1b = S1[so[2]] | (S1[so[2] + 1] << 8) | (S1[so[3]] << 16) | (S1[so[3] + 1] << 24);
        }
        // horizontal: H = p[sx] * a0 + p[sx + 1] * a1 (the column past the right edge carries a1 = 0)
        const unsigned int H00 = __dp2a_lo(coef[0], p0a, 0u), H01 = __dp2a_hi(coef[1], p0a, 0u);
        const unsigned int H02 = __dp2a_lo(coef[2], p0b, 0u), H03 = __dp2a_hi(coef[3], p0b, 0u);
        const unsigned int H10 = __dp2a_lo(coef[0], p1a, 0u), H11 = __dp2a_hi(coef[1], p1a, 0u);
        const unsigned int H12 = __dp2a_lo(coef[2], p1b, 0u), H13 = __dp2a_hi(coef[3], p1b, 0u);
        // vertical: (((b0 * (H0 >> 4)) >> 16) + ((b1 * (H1 >> 4)) >> 16) + 2) >> 2
        const unsigned int v0 = (__umulhi(H00 >> 4, bx16) + __umulhi(H10 >> 4, by16) + 2u) >> 2;
        const unsigned int v1 = (__umulhi(H01 >> 4, bx16) + __umulhi(H11 >> 4, by16) + 2u) >> 2;
        const unsigned int v2 = (__umulhi(H02 >> 4, bx16) + __umulhi(H12 >> 4, by16) + 2u) >> 2;
        const unsigned int v3 = (__umulhi(H03 >> 4, bx16) + __umulhi(H13 >> 4, by16) + 2u) >> 2;
        *reinterpret_cast<unsigned int *>(dst + (size_t)dy * g.pitch + dx0) = v0 | (v1 << 8) | (v2 << 16) | (v3 << 24);
    }
}

// E7 blur7: see blur_kernel.cuh
// E3+E4 FAST per grid cell: see fast_strip_kernel.cuh

#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
