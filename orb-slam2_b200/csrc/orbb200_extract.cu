// orbb200_extract.cu -- C-ABI of the extractor (include/orbb200.h) over the sm_100a kernels.
// Mirrors ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:53-135, src/ORBextractor.cc).
#include <math.h>
#include <stdarg.h>

#include <algorithm>
#include <mutex>
#include <sched.h>
#include <new>
#include <stdlib.h>

#define ORBB200_EXTRACT_IMPL
#include "blur_kernel.cuh"
#include "describe_kernel.cuh"
#include "extractor_host.cuh"
#include "fast_strip_kernel.cuh"
#include "stage_kernel.cuh"
#include "octree_kernel.cuh"

namespace orbb200 {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static inline int cv_round_f(float v) { return (int)lrintf(v); }

template <typename T> static int dev_alloc(T **p, size_t count)
{
    if (*p) { cudaFree(*p); *p = nullptr; }
    ORB_CUDA_OK(cudaMalloc((void **)p, std::max<size_t>(count, 1) * sizeof(T)));
    return ORBB200_OK;
}

static short sat_short(int v) { return (short)std::min(32767, std::max(-32768, v)); }

// Per-shape geometry: pyramid sizes (:1351-1353), 30-px cell grid (:934-951), quadtree roots (:627-629),
// cv::resize coefficient tables, workspace offsets.
// MaxDynamicSharedMemorySize is an attribute of the FUNCTION (per device), not of a handle: the reference runs its left and right
// extractor on two host threads, so every kernel with dynamic shared memory is opted in ONCE per device to the 200 KB the size
// checks of this file allow, instead of being re-set to a per-launch size before every launch.
static int opt_in_shared_memory(int device)
{
    static std::mutex mu;
    static bool done[64] = {};
    std::lock_guard<std::mutex> lk(mu);
    if (device < 0 || device >= 64 || done[device]) return ORBB200_OK;
    const int kMax = 200 * 1024;
    ORB_CUDA_OK(cudaFuncSetAttribute(k_pyramid_resize, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
    ORB_CUDA_OK(cudaFuncSetAttribute(k_fast_strips, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
    ORB_CUDA_OK(cudaFuncSetAttribute(k_orient_describe, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
    ORB_CUDA_OK(cudaFuncSetAttribute(k_octree_select<OCT_THREADS, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
    ORB_CUDA_OK(cudaFuncSetAttribute(k_octree_select<OCT_THREADS_BIG, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
    done[device] = true;
    return ORBB200_OK;
}

int extractor_ensure_shape(orbb200_extractor *e, int w, int h, int batch)
{
    const bool sameShape = (w == e->w && h == e->h);
    if (sameShape && batch <= e->batchCap) return ORBB200_OK;
    ORB_CUDA_OK(cudaStreamSynchronize(e->stream));
    {
        const int rc = opt_in_shared_memory(e->device);
        if (rc) return rc;
    }
    const int L = e->nlevels;
    if (!sameShape) {
        size_t imgOff = 0, candOff = 0, wsOff = 0;
        int cellBase = 0, nodeOff = 0, tileBase = 0, coefOff = 0;
        size_t octSmem = 0;
        const bool bigImage = (size_t)w * h > ((size_t)1 << 20); // the quadtree kernel's 1024-thread variant (extractor_api.inc)
        for (int l = 0; l < L; l++) {
            LevelGeom &g = e->geom[l];
            memset(&g, 0, sizeof(g));
            g.w = cv_round_f((float)w * e->isf[l]);
            g.h = cv_round_f((float)h * e->isf[l]);
            if (g.w < 1 || g.h < 1) { set_error("level %d of a %dx%d image is empty", l, w, h); return ORBB200_ERR_ARG; }
            if (g.w > 4127 || g.h > 4127) { set_error("level %d (%dx%d) exceeds the 12-bit candidate packing", l, g.w, g.h); return ORBB200_ERR_ARG; }
            g.pitch = (int)align_up(g.w, 16);
            g.imgOff = (unsigned int)imgOff;
            imgOff += align_up((size_t)g.pitch * g.h, 256);
            const float width = (float)(g.w - 32), height = (float)(g.h - 32);
            g.nCols = width > 0 ? (int)(width / 30.f) : 0;
            g.nRows = height > 0 ? (int)(height / 30.f) : 0;
            if (g.nCols < 1 || g.nRows < 1) { g.nCols = g.nRows = 0; g.wCell = g.hCell = 1; }
            else { g.wCell = (int)ceilf(width / g.nCols); g.hCell = (int)ceilf(height / g.nRows); }
            g.cellBase = cellBase;
            g.nCells = g.nCols * g.nRows;
            cellBase += g.nCells;
            g.cellCap = ((g.wCell + 1) / 2) * ((g.hCell + 1) / 2);
            g.candOff = (unsigned int)candOff;
            g.candCap = g.nCells * g.cellCap;
            if (g.candCap > (1 << 24) - 1) { set_error("level %d: candidate capacity %d exceeds 2^24", l, g.candCap); return ORBB200_ERR_ARG; }
            candOff += align_up((size_t)g.candCap, 64);
            g.nfeat = e->nfeat[l];
            int nIni = 1;
            if (width > 0 && height > 0) nIni = (int)roundf(width / (g.h - 32));
            if (nIni < 1) nIni = 1; // the reference divides by zero for portrait images (:627-628)
            g.nIni = nIni;
            g.hX = width > 0 ? width / nIni : 1.f;
            g.nodeCap = std::max(g.nfeat + 3, 4 * nIni) + 8;
            g.sortLen = 1;
            while (g.sortLen < g.nodeCap) g.sortLen <<= 1;
            g.nodeOff = nodeOff;
            nodeOff += g.nodeCap;
            g.nodeWsOff = (unsigned int)wsOff;
            wsOff += align_up((size_t)21 * g.nodeCap, 32);
            octSmem = std::max(octSmem, align_up((size_t)44 * g.nodeCap, 16) + (size_t)8 * g.sortLen);
            // histogram formulation: the deepest table that fits the block's shared-memory budget (6 blocks of 256 threads per SM --
            // the node passes are barrier / latency bound, so blocks per SM count for more than table depth: 0.31 ms at 3 blocks and
            // depth 6, 0.24 ms at 6 blocks and depth 5 per 320 KITTI images; one block of 1024 threads per SM for images above 1 Mpixel)
            g.octDepth = 0;
            if (width > 0 && height > 0 && g.nodeCap < 65535 && nIni <= 255) {
                const size_t budget = bigImage ? (size_t)196 * 1024 : (size_t)36 * 1024;
                for (int D = OCT_MAX_DEPTH; D >= 3; D--)
                    if (oct_fast_smem(nIni, D, g.nodeCap, g.sortLen, (int)width, (int)height) <= budget) { g.octDepth = D; break; }
                if (g.octDepth) octSmem = std::max(octSmem, oct_fast_smem(nIni, g.octDepth, g.nodeCap, g.sortLen, (int)width, (int)height));
            }
            g.scale = e->sf[l];
            g.patch = (float)(int)(31 * e->sf[l]); // scaledPatchSize (:1023)
            g.blurTileBase = tileBase;
            g.blurTilesX = div_up(g.w, BL_TW);
            g.blurTilesY = div_up(g.h, BL_TH);
            tileBase += g.blurTilesX * g.blurTilesY;
            g.coefOff = (unsigned int)coefOff;
            coefOff += (int)align_up(std::max(g.w, g.h), 4);
        }
        e->pyrStride = imgOff;
        e->candStride = candOff;
        e->nodeWsStride = wsOff;
        e->totalCells = cellBase;
        e->nodeTotal = nodeOff;
        e->totalBlurTiles = tileBase;
        e->coefTotal = coefOff;
        e->octreeSmem = octSmem;
        e->fastRows = 7;
        for (int l = 0; l < L; l++)
            if (e->geom[l].nCells > 0) e->fastRows = std::min(std::max(e->fastRows, e->geom[l].hCell + 6), (int)FS_MAXROWS);
        if (octSmem > 200 * 1024) { set_error("nfeatures too large for the quadtree kernel's shared memory (%zu B)", octSmem); return ORBB200_ERR_ARG; }
        // cv::resize INTER_LINEAR tables (scale = 1/((double)dst/src), 11-bit coefficients)
        std::vector<int> xofs(coefOff, 0), yofs(coefOff, 0);
        std::vector<short2> xa(coefOff, make_short2(0, 0)), yb(coefOff, make_short2(0, 0));
        for (int l = 1; l < L; l++) {
            LevelGeom &g = e->geom[l];
            const LevelGeom &s = e->geom[l - 1];
            const double scale_x = 1. / ((double)g.w / s.w), scale_y = 1. / ((double)g.h / s.h);
            for (int dx = 0; dx < g.w; dx++) {
                float fx = (float)((dx + 0.5) * scale_x - 0.5);
                int sx = (int)floorf(fx);
                fx -= sx;
                if (sx < 0) { fx = 0; sx = 0; }
                if (sx >= s.w - 1) { fx = 0; sx = s.w - 1; }
                xofs[g.coefOff + dx] = sx;
                xa[g.coefOff + dx] = make_short2(sat_short(cv_round_f((1.f - fx) * 2048.f)), sat_short(cv_round_f(fx * 2048.f)));
            }
            for (int dy = 0; dy < g.h; dy++) {
                float fy = (float)((dy + 0.5) * scale_y - 0.5);
                int sy = (int)floorf(fy);
                fy -= sy;
                yofs[g.coefOff + dy] = sy;
                yb[g.coefOff + dy] = make_short2(sat_short(cv_round_f((1.f - fy) * 2048.f)), sat_short(cv_round_f(fy * 2048.f)));
            }
            int maxBytes = 16, maxRows = 1;
            for (int x0 = 0; x0 < g.w; x0 += RZ_TW) {
                const int x1 = std::min(x0 + RZ_TW, g.w) - 1;
                const int xb = xofs[g.coefOff + x0] & ~15;
                const int last = std::min(xofs[g.coefOff + x1] + 1, s.w - 1);
                maxBytes = std::max(maxBytes, (int)align_up(last - xb + 1, 16));
            }
            for (int y0 = 0; y0 < g.h; y0 += RZ_TH) {
                const int y1 = std::min(y0 + RZ_TH, g.h) - 1;
                const int r0 = std::min(std::max(yofs[g.coefOff + y0], 0), s.h - 1);
                const int r1 = std::min(std::max(yofs[g.coefOff + y1] + 1, 0), s.h - 1);
                maxRows = std::max(maxRows, r1 - r0 + 1);
            }
            g.rzTileSP = maxBytes;
            g.rzTileRows = maxRows;
            if ((size_t)maxBytes * maxRows > 160 * 1024) { set_error("scale factor too large for the resize tile"); return ORBB200_ERR_ARG; }
        }
        if (dev_alloc(&e->d_geom, L)) return ORBB200_ERR_CUDA;
        {
            // strip records of k_fast_strips (fast_strip_kernel.cuh): the cell ROIs of src/ORBextractor.cc:953-985, worked out once
            // per shape.  A strip = one cell row x a run of cells whose interiors add up to <= FS_MAXIW pixels; the cells the
            // reference skips (:957-976) trail a row (or fill a skipped row) and only get their counts zeroed.
            std::vector<uint4> strips;
            for (int l = 0; l < L; l++) {
                const LevelGeom &g = e->geom[l];
                if (g.nCells == 0) continue;
                const int maxBX = g.w - 16, maxBY = g.h - 16;
                int nActCols = 0;
                for (int cj = 0; cj < g.nCols; cj++) {
                    const int iniX = 16 + cj * g.wCell;
                    if (iniX >= maxBX - 6 || std::min(iniX + g.wCell + 6, maxBX) - iniX < 7) break;
                    nActCols++;
                }
                // cells per strip: a bound-pass trip is one row x 32 pixel pairs, so the strip width with the fewest idle lanes in its
                // last trip wins (ties: the wider strip, whose fixed costs are spread over more pixels)
                const int perStripMax = std::max(1, FS_MAXIW / g.wCell);
                int perStripBest = 1;
                double bestEff = -1;
                for (int C = 1; C <= perStripMax; C++) {
                    const int ppr = (C * g.wCell + 1) / 2;
                    const double eff = (double)ppr / (32.0 * div_up(ppr, 32));
                    if (eff >= bestEff - 0.02) { bestEff = std::max(bestEff, eff); perStripBest = C; }
                }
                const int perStrip = perStripBest;
                for (int ci = 0; ci < g.nRows; ci++) {
                    const int iniY = 16 + ci * g.hCell, maxY = std::min(iniY + g.hCell + 6, maxBY);
                    const bool rowActive = !(iniY >= maxBY - 3) && (maxY - iniY) >= 7 && nActCols > 0;
                    for (int j0 = 0; j0 < g.nCols;) {
                        int nAct = 0, nAll;
                        if (rowActive && j0 < nActCols) {
                            nAct = std::min(perStrip, nActCols - j0);
                            nAll = nAct;
                            if (j0 + nAct == nActCols) nAll = std::min(g.nCols - j0, 255); // the skipped cells that trail the row
                        } else {
                            nAll = std::min(g.nCols - j0, 255);
                        }
                        const int iniX = 16 + j0 * g.wCell;
                        const int maxX = nAct ? std::min(iniX + nAct * g.wCell + 6, maxBX) : iniX;
                        const int rw = nAct ? maxX - iniX : 0, rh = nAct ? maxY - iniY : 0;
                        const int xbTile = (iniX & 15) ? (iniX & ~15) : iniX - 16; // tile origin of k_fast_strips
                        const int chunks = rw ? ((((maxX - xbTile) + 15) & ~15) >> 4) : 1;
                        if (rw - 6 > FS_MAXIW || chunks > FS_SPW / 8) { set_error("internal: FAST strip too wide (%d px)", rw); return ORBB200_ERR_ARG; }
                        strips.push_back(make_uint4((unsigned)iniX | ((unsigned)iniY << 16), (unsigned)rw | ((unsigned)rh << 16), g.imgOff,
                                                    (unsigned)g.pitch | ((unsigned)g.w << 16)));
                        strips.push_back(make_uint4(g.candOff + (unsigned)(ci * g.nCols + j0) * g.cellCap,
                                                    (unsigned)(j0 * g.wCell + 3) | ((unsigned)(ci * g.hCell + 3) << 12),
                                                    (unsigned)l | ((unsigned)nAct << 4) | ((unsigned)nAll << 8) | ((unsigned)g.wCell << 16), 0u));
                        strips.push_back(make_uint4((unsigned)(g.cellBase + ci * g.nCols + j0), (unsigned)g.cellCap, (65536u + chunks - 1) / chunks, 0u));
                        j0 += nAll;
                    }
                }
            }
            e->totalStrips = (int)(strips.size() / FS_REC);
            if (strips.empty()) strips.assign(FS_REC, make_uint4(0, 0, 0, 0));
            if (dev_alloc(&e->d_fastStrips, strips.size())) return ORBB200_ERR_CUDA;
            ORB_CUDA_OK(cudaMemcpy(e->d_fastStrips, strips.data(), sizeof(uint4) * strips.size(), cudaMemcpyHostToDevice));
        }
        if (dev_alloc(&e->d_rzXofs, coefOff) || dev_alloc(&e->d_rzYofs, coefOff) || dev_alloc(&e->d_rzXa, coefOff) ||
            dev_alloc(&e->d_rzYb, coefOff))
            return ORBB200_ERR_CUDA;
        ORB_CUDA_OK(cudaMemcpy(e->d_geom, e->geom, sizeof(LevelGeom) * L, cudaMemcpyHostToDevice));
        if (coefOff > 0) {
            ORB_CUDA_OK(cudaMemcpy(e->d_rzXofs, xofs.data(), sizeof(int) * coefOff, cudaMemcpyHostToDevice));
            ORB_CUDA_OK(cudaMemcpy(e->d_rzYofs, yofs.data(), sizeof(int) * coefOff, cudaMemcpyHostToDevice));
            ORB_CUDA_OK(cudaMemcpy(e->d_rzXa, xa.data(), sizeof(short2) * coefOff, cudaMemcpyHostToDevice));
            ORB_CUDA_OK(cudaMemcpy(e->d_rzYb, yb.data(), sizeof(short2) * coefOff, cudaMemcpyHostToDevice));
        }
        e->w = w;
        e->h = h;
        e->batchCap = 0;
    }
    const size_t B = (size_t)batch;
    if (dev_alloc(&e->d_pyr, e->pyrStride * B + 256) || dev_alloc(&e->d_blur, e->pyrStride * B + 256) ||
        dev_alloc(&e->d_cand, e->candStride * B) || dev_alloc(&e->d_ord, e->candStride * B) ||
        dev_alloc(&e->d_knode, e->candStride * B) || dev_alloc(&e->d_cellCount, (size_t)e->totalCells * B) ||
        dev_alloc(&e->d_cellScan, (size_t)e->totalCells * B) || dev_alloc(&e->d_nodeWs, e->nodeWsStride * B) ||
        dev_alloc(&e->d_lvlKp, (size_t)e->nodeTotal * B) || dev_alloc(&e->d_lvlCount, (size_t)L * B))
        return ORBB200_ERR_CUDA;
    e->batchCap = batch;
    e->allocEpoch++;
    return ORBB200_OK;
}

int extractor_ensure_outputs(orbb200_extractor *e, int batch, int cap)
{
    if (batch <= e->outBatch && cap <= e->outCap) return ORBB200_OK;
    ORB_CUDA_OK(cudaStreamSynchronize(e->stream));
    const int nb = std::max(batch, e->outBatch), nc = std::max(cap, e->outCap);
    if (dev_alloc(&e->d_kps, (size_t)nb * nc) || dev_alloc(&e->d_desc, (size_t)nb * nc * 32) || dev_alloc(&e->d_nOut, nb))
        return ORBB200_ERR_CUDA;
    e->outBatch = nb;
    e->outCap = nc;
    return ORBB200_OK;
}

int extractor_ensure_pipeline(orbb200_extractor *e, size_t stageBytes)
{
    if (!e->copyStream) ORB_CUDA_OK(cudaStreamCreateWithFlags(&e->copyStream, cudaStreamNonBlocking));
    if (!e->d2hStream) ORB_CUDA_OK(cudaStreamCreateWithFlags(&e->d2hStream, cudaStreamNonBlocking));
    for (int i = 0; i < orbb200_extractor::PIPE_STAGES; i++) {
        if (!e->evH2D[i]) ORB_CUDA_OK(cudaEventCreateWithFlags(&e->evH2D[i], cudaEventDisableTiming));
        if (!e->evDone[i]) ORB_CUDA_OK(cudaEventCreateWithFlags(&e->evDone[i], cudaEventDisableTiming));
        if (!e->evStaged[i]) ORB_CUDA_OK(cudaEventCreateWithFlags(&e->evStaged[i], cudaEventDisableTiming));
    }
    if (!e->evBegin) ORB_CUDA_OK(cudaEventCreateWithFlags(&e->evBegin, cudaEventDisableTiming));
    for (int i = 0; i < orbb200_extractor::PIPE_STREAMS; i++) {
        if (!e->workStream[i]) ORB_CUDA_OK(cudaStreamCreateWithFlags(&e->workStream[i], cudaStreamNonBlocking));
        if (!e->evWork[i]) ORB_CUDA_OK(cudaEventCreateWithFlags(&e->evWork[i], cudaEventDisableTiming));
    }
    if (e->stageBytes < stageBytes) {
        ORB_CUDA_OK(cudaStreamSynchronize(e->stream));
        ORB_CUDA_OK(cudaStreamSynchronize(e->copyStream));
        for (int i = 0; i < orbb200_extractor::PIPE_STREAMS; i++) ORB_CUDA_OK(cudaStreamSynchronize(e->workStream[i]));
        for (int i = 0; i < orbb200_extractor::PIPE_STAGES; i++) {
            if (e->d_stage[i]) { cudaFree(e->d_stage[i]); e->d_stage[i] = nullptr; }
            ORB_CUDA_OK(cudaMalloc((void **)&e->d_stage[i], stageBytes + 256));
        }
        e->stageBytes = stageBytes;
    }
    return ORBB200_OK;
}

} // namespace orbb200

#include "extractor_api.inc"
