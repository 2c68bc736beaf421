// octree_kernel.cuh -- E5: deterministic GPU DistributeOctTree (src/ORBextractor.cc:551-609, 617-915).
//
// One thread block per (image, level).  The reference's std::list of nodes is an array indexed by list
// position; keys never move: each key carries the list position of its node (knode), and because
// DivideNode is a stable partition, "first key with the maximum response" (:902-908) is the smallest
// original index among the node's maximal keys.  One reference pass becomes
//   (1) visit list        mode 0 (:733-799): every node with > 1 key, in list order
//                         mode 1 (:814-884): nodes with > 1 key sorted by (nKeys, creation seq) and
//                                            taken from the back until size >= N (:879)
//   (2) per-key quadrant count (warp-aggregated shared-memory atomics)
//   (3) scan over the visit list -> creation index of every non-empty child
//   (4) new list = reverse(children in creation order) ++ unsplit nodes in old order  (push_front, :748...)
//   (5) per-key node id update
// The pointer tie-break of std::sort on pair<int,ExtractorNode*> (:832) is replaced by the creation sequence,
// exactly like oracle/orb_ref.c.
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

constexpr int OCT_THREADS = 256; // threads per block
constexpr int OCT_THREADS_BIG = 1024;      // large images (level 0 above 1 Mpixel, e.g. 4K: ~250 k candidates in one level): the per-key
                                           // loops of a pass are L2-latency bound, four times the threads keep four times the keys in flight

struct NodeBuf {
    int *x0, *x1, *y0, *y1, *cnt, *seq;
};

__device__ __forceinline__ NodeBuf node_buf(int *base, int C)
{
    NodeBuf n;
    n.x0 = base; n.x1 = base + C; n.y0 = base + 2 * C; n.y1 = base + 3 * C; n.cnt = base + 4 * C; n.seq = base + 5 * C;
    return n;
}

// atomicAdd(base[key], 1) for every active lane, aggregated per distinct key inside the warp.
// Must be called by all 32 lanes.
__device__ __forceinline__ void warp_agg_inc(int *base, int key, bool active)
{
    const unsigned int act = __ballot_sync(0xffffffffu, active);
    if (active) {
        const unsigned int peers = __match_any_sync(act, key);
        if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(base + key, __popc(peers));
    }
}

// quadrant of a key inside a node split at (xs, ys) (DivideNode, :551-609)
__device__ __forceinline__ int key_quadrant(uint32_t k, int xs, int ys)
{
    const int x = k & 0xFFF, y = (k >> 12) & 0xFFF;
    return (x < xs ? 0 : 1) + (y < ys ? 0 : 2);
}

// =====================================================================================================================
// Fast path: the tree is built from a HISTOGRAM of the keys, the keys themselves are read twice.
//
// DivideNode's split points depend on the node's bounds only (x0 + ceil((x1 - x0) / 2), :553-554), never on its keys, so the
// path of a key through the tree -- root, then one x-bit and one y-bit per depth -- is a function of its coordinates: two small
// tables (x -> root and D x-bits, y -> D y-bits) built per block.  One pass over the keys counts them in the 2^D x 2^D deepest
// cells of their root (16-bit shared-memory counters), a pyramid of sums gives the key count of every possible node at every
// depth, and the reference's passes (:733-884) then run on the NODE LIST alone: "how many keys does child q get" is a table
// lookup instead of a pass over the keys, and no key ever has to be told which node it moved to.  When the list is final, every
// deepest cell is painted with the list position of the node that covers it and a second pass over the keys takes, per node,
// the maximum response with the reference's first-in-list-order tie-break (:895-912): the order key is the key's position in the
// cell-major candidate array (cell-row-major, row-major inside a cell = the order the reference appends them in, :925-1009).
// A node that must be split below depth D (keys clustered in one deepest cell) makes the block fall back to the iterative
// formulation below, which has no depth limit.  4K level 0 (328 k keys): 12 passes over the keys -> 2.
constexpr int OCT_MAX_DEPTH = 7;

__host__ __device__ inline size_t oct_fast_up_total(int nIni, int D) { return (size_t)nIni * (size_t)(((1 << (2 * D)) - 1) / 3); }
// shared memory of the fast path: pyramid levels 0..D-1 (int), deepest level (u16), 10 node arrays of C ints, sort keys, x / y tables
__host__ __device__ inline size_t oct_fast_smem(int nIni, int D, int C, int sortLen, int width, int height)
{
    size_t b = 4 * oct_fast_up_total(nIni, D) + 2 * ((size_t)nIni << (2 * D)) + 4 * (size_t)10 * C;
    b = (b + 7) & ~(size_t)7;
    b += 8 * (size_t)sortLen + 2 * (size_t)((width + 1) & ~1) + (size_t)height;
    return (b + 15) & ~(size_t)15;
}

struct OctCtl { int size, finish, mode, seq, nexp, rstar, ovf; };

__device__ __forceinline__ int oct_pack(int r, int d, int xp, int yp) { return xp | (yp << 8) | (d << 16) | (r << 20); }

// key count of the node (root r, depth d, x prefix xp, y prefix yp)
__device__ __forceinline__ int oct_count(const int *up, const unsigned short *deep, int nIni, int D, int r, int d, int xp, int yp)
{
    const int i = (r << (2 * d)) + (yp << d) + xp;
    return d == D ? (int)deep[i] : up[nIni * (((1 << (2 * d)) - 1) / 3) + i];
}

// f(key, pos) for every candidate of the level; pos = its position in the cell-major candidate array (ascends in the reference's
// order).  A warp takes 32 consecutive cells, scans their counts and walks the concatenated keys with all lanes busy, OCT_FKU
// independent loads per lane in flight.
constexpr int OCT_FKU = 4;
template <class F> __device__ __forceinline__ void oct_for_each_key(const uint32_t *cand, const int *cellCount, int nCells, int cellCap, F f)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int c0 = wid * 32; c0 < nCells; c0 += nw * 32) {
        const int cnt = c0 + lane < nCells ? min(cellCount[c0 + lane], cellCap) : 0;
        int inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        const int excl = inc - cnt, T = __shfl_sync(0xffffffffu, inc, 31);
        for (int t0 = 0; t0 < T; t0 += 32 * OCT_FKU) {
            uint32_t key[OCT_FKU];
            int pos[OCT_FKU];
#pragma unroll
            for (int u = 0; u < OCT_FKU; u++) {
                const int t = t0 + u * 32 + lane;
                int j = 0; // number of cells whose inclusive count is <= t = the cell that holds key t
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const int v = __shfl_sync(0xffffffffu, inc, j + step - 1);
                    if (v <= t) j += step;
                }
                const int e = __shfl_sync(0xffffffffu, excl, j & 31);
                pos[u] = -1;
                key[u] = 0u;
                if (t < T) {
                    pos[u] = (c0 + j) * cellCap + (t - e);
                    key[u] = __ldg(cand + pos[u]);
                }
            }
#pragma unroll
            for (int u = 0; u < OCT_FKU; u++)
                if (pos[u] >= 0) f(key[u], pos[u]);
        }
    }
}

// returns false (block-uniform) when a node at the depth limit has to be split; nothing has been written then
template <int MAXT> __device__ bool octree_fast(const ExtractParams &P, const LevelGeom &g, int l, int b, int D, unsigned char *dsm,
                                                int *scanTmp, OctCtl *ctl)
{
    const int tid = threadIdx.x, NT = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = NT >> 5;
    const int C = g.nodeCap, N = g.nfeat, nIni = g.nIni;
    const int width = max(g.w - 32, 1), height = max(g.h - 32, 1);
    const int S2 = 1 << (2 * D);
    int *up = reinterpret_cast<int *>(dsm);
    const int upTotal = (int)oct_fast_up_total(nIni, D);
    unsigned short *deep = reinterpret_cast<unsigned short *>(up + upTotal);
    int *nodeBase = reinterpret_cast<int *>(deep + (size_t)nIni * S2);
    int *curCode = nodeBase, *curCnt = nodeBase + C, *curSeq = nodeBase + 2 * C;
    int *nxtCode = nodeBase + 3 * C, *nxtCnt = nodeBase + 4 * C, *nxtSeq = nodeBase + 5 * C;
    int *vrank = nodeBase + 6 * C, *vis = nodeBase + 7 * C, *basea = nodeBase + 8 * C, *cum = nodeBase + 9 * C;
    size_t off = 4 * (size_t)upTotal + 2 * (size_t)nIni * S2 + 4 * (size_t)10 * C;
    off = (off + 7) & ~(size_t)7;
    unsigned long long *skeys = reinterpret_cast<unsigned long long *>(dsm + off);
    unsigned short *xtab = reinterpret_cast<unsigned short *>(skeys + g.sortLen);
    unsigned char *ytab = reinterpret_cast<unsigned char *>(xtab + ((width + 1) & ~1));
    const int *cellCount = P.cellCount + (size_t)b * P.totalCells + g.cellBase;
    const uint32_t *cand = P.cand + (size_t)b * P.candStride + g.candOff;

    // ---- coordinate tables: root (:627-704) and the D split decisions along each axis (DivideNode :551-609)
    for (int x = tid; x < width; x += NT) {
        int r = (int)__fdiv_rn((float)x, g.hX);
        r = min(r, nIni - 1);
        int x0 = (int)__fmul_rn(g.hX, (float)r), x1 = (int)__fmul_rn(g.hX, (float)(r + 1)), xb = 0;
        for (int d = 0; d < D; d++) {
            const int xs = x0 + ((x1 - x0 + 1) >> 1);
            const int bit = x < xs ? 0 : 1;
            if (bit) x0 = xs; else x1 = xs;
            xb = 2 * xb + bit;
        }
        xtab[x] = (unsigned short)(xb | (r << 8));
    }
    for (int y = tid; y < height; y += NT) {
        int y0 = 0, y1 = g.h - 32, yb = 0;
        for (int d = 0; d < D; d++) {
            const int ys = y0 + ((y1 - y0 + 1) >> 1);
            const int bit = y < ys ? 0 : 1;
            if (bit) y0 = ys; else y1 = ys;
            yb = 2 * yb + bit;
        }
        ytab[y] = (unsigned char)yb;
    }
    {
        unsigned int *z = reinterpret_cast<unsigned int *>(deep);
        for (int i = tid; i < (nIni * S2) >> 1; i += NT) z[i] = 0u;
    }
    __syncthreads();
    // ---- pass 1 over the keys: deepest-cell histogram (two 16-bit counters per word; a cell holds far fewer than 65 536 pixels)
    {
        unsigned int *h = reinterpret_cast<unsigned int *>(deep);
        oct_for_each_key(cand, cellCount, g.nCells, g.cellCap, [&](uint32_t key, int) {
            const int xt = xtab[min((int)(key & 0xFFF), width - 1)], yb = ytab[min((int)((key >> 12) & 0xFFF), height - 1)];
            const int idx = ((xt >> 8) << (2 * D)) + (yb << D) + (xt & 0xFF);
            atomicAdd(h + (idx >> 1), 1u << (16 * (idx & 1)));
        });
    }
    __syncthreads();
    // ---- pyramid of sums: level d cell (r, yp, xp) = its four children at level d + 1
    for (int d = D - 1; d >= 0; d--) {
        int *dst = up + nIni * (((1 << (2 * d)) - 1) / 3);
        const int cells = nIni << (2 * d), m = (1 << d) - 1;
        for (int i = tid; i < cells; i += NT) {
            const int r = i >> (2 * d), yp = (i >> d) & m, xp = i & m;
            dst[i] = oct_count(up, deep, nIni, D, r, d + 1, 2 * xp, 2 * yp) + oct_count(up, deep, nIni, D, r, d + 1, 2 * xp + 1, 2 * yp) +
                     oct_count(up, deep, nIni, D, r, d + 1, 2 * xp, 2 * yp + 1) + oct_count(up, deep, nIni, D, r, d + 1, 2 * xp + 1, 2 * yp + 1);
        }
        __syncthreads();
    }
    // ---- roots
    if (tid == 0) {
        int pos = 0;
        for (int r = 0; r < nIni; r++) {
            const int c = up[r];
            if (c > 0) { curCode[pos] = oct_pack(r, 0, 0, 0); curCnt[pos] = c; curSeq[pos] = r; pos++; }
        }
        ctl->size = pos; ctl->finish = 0; ctl->mode = 0; ctl->seq = nIni; ctl->ovf = 0x7fffffff;
    }

    // ---- the reference's passes on the node list
    int *cur[3] = {curCode, curCnt, curSeq}, *nxt[3] = {nxtCode, nxtCnt, nxtSeq};
    while (true) {
        __syncthreads();
        if (ctl->finish) break;
        const int size = ctl->size, mode = ctl->mode, seq0 = ctl->seq;
        const int *cCode = cur[0], *cCnt = cur[1], *cSeq = cur[2];
        int *nCode = nxt[0], *nCnt = nxt[1], *nSeq = nxt[2];
        int nvis = 0;
        // (1) visit list: mode 0 (:733-799) every node with more than one key in list order; mode 1 (:814-884) those nodes sorted
        //     by (nKeys, creation sequence), largest first
        if (mode == 0) {
            for (int base = 0; base < size; base += NT) {
                const int p = base + tid;
                const int f = (p < size && cCnt[p] > 1) ? 1 : 0;
                int tot;
                const int ex = block_excl_scan(f, scanTmp, &tot);
                if (p < size) vrank[p] = f ? nvis + ex : -1;
                if (f) vis[nvis + ex] = p;
                nvis += tot;
            }
        } else {
            for (int base = 0; base < size; base += NT) {
                const int p = base + tid;
                const int f = (p < size && cCnt[p] > 1) ? 1 : 0;
                int tot;
                const int ex = block_excl_scan(f, scanTmp, &tot);
                if (p < size) vrank[p] = -1;
                if (f)
                    skeys[nvis + ex] = ((unsigned long long)cCnt[p] << 40) | ((unsigned long long)(cSeq[p] & 0xFFFFF) << 20) | (unsigned long long)p;
                nvis += tot;
            }
            __syncthreads();
            // order by rank counting: keys are unique (they end in the list position), so the rank of a key = the number of larger
            // keys; every thread reads the whole key array (broadcast reads, no barrier inside) -- for the few hundred nodes of a
            // level this beats a bitonic network's log^2 barrier rounds
            for (int r = tid; r < nvis; r += NT) {
                const unsigned long long key = skeys[r];
                int rank = 0;
                for (int j = 0; j < nvis; j++) rank += skeys[j] > key ? 1 : 0;
                const int p = (int)(key & 0xFFFFFull);
                vis[rank] = p;
                vrank[p] = rank;
            }
        }
        if (tid == 0) { ctl->nexp = 0; ctl->rstar = nvis - 1; ctl->ovf = 0x7fffffff; }
        __syncthreads();
        // (3) creation index of the children (non-empty quadrants, :567-609); a visited node at the depth limit has no child table
        int T = 0;
        for (int base = 0; base < nvis; base += NT) {
            const int r = base + tid;
            int nch = 0;
            if (r < nvis) {
                const int code = cCode[vis[r]];
                const int d = (code >> 16) & 15, xp = code & 255, yp = (code >> 8) & 255, rt = code >> 20;
                if (d >= D) atomicMin(&ctl->ovf, r);
                else
                    nch = (oct_count(up, deep, nIni, D, rt, d + 1, 2 * xp, 2 * yp) > 0) + (oct_count(up, deep, nIni, D, rt, d + 1, 2 * xp + 1, 2 * yp) > 0) +
                          (oct_count(up, deep, nIni, D, rt, d + 1, 2 * xp, 2 * yp + 1) > 0) + (oct_count(up, deep, nIni, D, rt, d + 1, 2 * xp + 1, 2 * yp + 1) > 0);
            }
            int tot;
            const int ex = block_excl_scan(nch, scanTmp, &tot);
            if (r < nvis) {
                basea[r] = T + ex;
                cum[r] = T + ex + nch - (r + 1); // list growth after splitting visit ranks 0..r
            }
            T += tot;
        }
        int nsplit = nvis;
        if (mode == 1) {
            for (int r = tid; r < nvis; r += NT)
                if (size + cum[r] >= N) atomicMin(&ctl->rstar, r);
            __syncthreads();
            nsplit = ctl->rstar + 1;
        }
        __syncthreads();
        if (ctl->ovf < nsplit) return false; // ranks below the first depth-limited node were exact, so this decision is too
        if (mode == 1) {
            for (int r = nsplit + tid; r < nvis; r += NT) vrank[vis[r]] = -1;
            T = 0;
            if (nsplit > 0) {
                const int code = cCode[vis[nsplit - 1]];
                const int d = (code >> 16) & 15, xp = code & 255, yp = (code >> 8) & 255, rt = code >> 20;
                T = basea[nsplit - 1] + (oct_count(up, deep, nIni, D, rt, d + 1, 2 * xp, 2 * yp) > 0) + (oct_count(up, deep, nIni, D, rt, d + 1, 2 * xp + 1, 2 * yp) > 0) +
                    (oct_count(up, deep, nIni, D, rt, d + 1, 2 * xp, 2 * yp + 1) > 0) + (oct_count(up, deep, nIni, D, rt, d + 1, 2 * xp + 1, 2 * yp + 1) > 0);
            }
            __syncthreads();
        }
        // (4) children at the front in reverse creation order (push_front, :748...) ...
        int nexp = 0;
        for (int r = tid; r < nsplit; r += NT) {
            const int code = cCode[vis[r]];
            const int d = (code >> 16) & 15, xp = code & 255, yp = (code >> 8) & 255, rt = code >> 20;
            int ci = basea[r];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int cx = 2 * xp + (q & 1), cy = 2 * yp + (q >> 1);
                const int c = oct_count(up, deep, nIni, D, rt, d + 1, cx, cy);
                if (c > 0) {
                    const int np = T - 1 - ci;
                    nCode[np] = oct_pack(rt, d + 1, cx, cy);
                    nCnt[np] = c;
                    nSeq[np] = seq0 + ci;
                    nexp += c > 1;
                    ci++;
                }
            }
        }
        if (nexp) atomicAdd(&ctl->nexp, nexp);
        // ... then the unsplit nodes in their old order
        int nsurv = 0;
        for (int base = 0; base < size; base += NT) {
            const int p = base + tid;
            const int f = (p < size && vrank[p] < 0) ? 1 : 0;
            int tot;
            const int ex = block_excl_scan(f, scanTmp, &tot);
            if (f) {
                const int np = T + nsurv + ex;
                nCode[np] = cCode[p]; nCnt[np] = cCnt[p]; nSeq[np] = cSeq[p];
            }
            nsurv += tot;
        }
        __syncthreads();
        const int newSize = T + nsurv;
        if (tid == 0) {
            ctl->size = newSize;
            ctl->seq = seq0 + T;
            if (newSize >= N || newSize == size) ctl->finish = 1;
            else if (mode == 0 && newSize + ctl->nexp * 3 > N) ctl->mode = 1;
        }
#pragma unroll
        for (int k = 0; k < 3; k++) { int *t = cur[k]; cur[k] = nxt[k]; nxt[k] = t; }
    }
    // ---- paint every deepest cell with the list position of the final node that covers it
    const int size = ctl->size;
    const int *fCode = cur[0];
    for (int p = wid; p < size; p += nw) {
        const int code = fCode[p];
        const int d = (code >> 16) & 15, xp = code & 255, yp = (code >> 8) & 255, rt = code >> 20;
        const int sh = D - d, side = 1 << sh;
        unsigned short *base = deep + ((size_t)rt << (2 * D)) + ((size_t)(yp << sh) << D) + (xp << sh);
        for (int i = lane; i < side * side; i += 32) base[((i >> sh) << D) + (i & (side - 1))] = (unsigned short)p;
    }
    unsigned int *best = reinterpret_cast<unsigned int *>(vrank);
    for (int p = tid; p < size; p += NT) best[p] = 0u;
    __syncthreads();
    // ---- pass 2 over the keys: one keypoint per node, maximum response, first in the reference's order on ties (:895-912)
    oct_for_each_key(cand, cellCount, g.nCells, g.cellCap, [&](uint32_t key, int pos) {
        const int xt = xtab[min((int)(key & 0xFFF), width - 1)], yb = ytab[min((int)((key >> 12) & 0xFFF), height - 1)];
        const int node = deep[((xt >> 8) << (2 * D)) + (yb << D) + (xt & 0xFF)];
        atomicMax(best + node, ((key >> 24) << 24) | (0xFFFFFFu - (unsigned int)pos));
    });
    __syncthreads();
    uint32_t *out = P.lvlKp + (size_t)b * P.nodeTotal + g.nodeOff;
    for (int p = tid; p < size; p += NT) out[p] = cand[0xFFFFFFu - (best[p] & 0xFFFFFFu)];
    if (tid == 0) P.lvlCount[(size_t)b * P.nlevels + l] = size;
    return true;
}

// =====================================================================================================================
// Iterative formulation (no depth limit): the fallback of the fast path, and the only path when the fast path's tables do not
// fit in shared memory (extreme aspect ratios: many roots).
// workspace (ints) per (image, level): 2 node buffers (6C each) + vrank C + vis C + base C + cum C (21C reserved);
// shared memory: everything the two per-key loops of a pass touch per node -- quadrant counts cc 4C, child positions
// 4C, new positions C, split coordinates recX / recY C each (recX < 0: node not split in this pass) -- + sort keys
// (sortLen u64).  With the node arrays in global memory a key cost a chain of three dependent L2 loads per loop.
template <int MAXT, int MINB> __global__ void __launch_bounds__(MAXT, MINB) k_octree_select(const __grid_constant__ ExtractParams P)
{
    constexpr int OCT_KU = 4; // keys per thread per trip of the per-key loops, all their loads in flight together (8 measured slower on 4K)
    extern __shared__ __align__(16) unsigned char dsm[];
    __shared__ int scanTmp[34];
    __shared__ int sh_size, sh_finish, sh_mode, sh_seq, sh_nexp, sh_rstar;
    const int l = blockIdx.y, b = blockIdx.x; // level-major launch order: the heavy level-0 blocks start first, the light ones fill the tail
    const LevelGeom g = P.geom[l];
    const int tid = threadIdx.x, NT = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = NT >> 5;
    const int C = g.nodeCap, N = g.nfeat;
    int *cc = reinterpret_cast<int *>(dsm);                                   // [4C]
    int *childpos = cc + 4 * C, *newpos = cc + 8 * C, *recX = cc + 9 * C, *recY = cc + 10 * C;
    unsigned long long *skeys = reinterpret_cast<unsigned long long *>(dsm + (((size_t)11 * C * 4 + 15) & ~(size_t)15)); // [sortLen]
    const int *cellCount = P.cellCount + (size_t)b * P.totalCells + g.cellBase;
    int *cellScan = P.cellScan + (size_t)b * P.totalCells + g.cellBase;
    const uint32_t *cand = P.cand + (size_t)b * P.candStride + g.candOff;
    uint32_t *ord = P.ord + (size_t)b * P.candStride + g.candOff;
    int *knode = P.knode + (size_t)b * P.candStride + g.candOff;
    int *ws = P.nodeWs + (size_t)b * P.nodeWsStride + g.nodeWsOff;
    NodeBuf cur = node_buf(ws, C), nxt = node_buf(ws + 6 * C, C);
    int *vrank = ws + 16 * C, *vis = ws + 18 * C, *basea = ws + 19 * C, *cum = ws + 20 * C;

    {
        __shared__ OctCtl fctl;
        const int D = min(g.octDepth, P.octDepthCap);
        if (D > 0) {
            if (octree_fast<MAXT>(P, g, l, b, D, dsm, scanTmp, &fctl)) return;
            __syncthreads();
        }
    }
    // ---- gather the cells' candidates in the reference's order (cell-row-major, row-major inside a cell)
    int n = 0;
    for (int base = 0; base < g.nCells; base += NT) {
        const int c = base + tid < g.nCells ? cellCount[base + tid] : 0;
        int tot;
        const int ex = block_excl_scan(c, scanTmp, &tot);
        if (base + tid < g.nCells) cellScan[base + tid] = n + ex;
        n += tot;
    }
    __syncthreads();
    for (int cell = wid; cell < g.nCells; cell += nw) {
        const int cnt = cellCount[cell], dst = cellScan[cell];
        const uint32_t *src = cand + (size_t)cell * g.cellCap;
        for (int k = lane; k < cnt; k += 32) ord[dst + k] = src[k];
    }
    // ---- roots (:627-704)
    const int nIni = g.nIni;
    for (int i = tid; i < 4 * C; i += NT) cc[i] = 0;
    __syncthreads();
    const int nPad = (n + 31) & ~31;
    for (int i = tid; i < nPad; i += NT) {
        int r = 0;
        if (i < n) {
            r = (int)__fdiv_rn((float)(ord[i] & 0xFFF), g.hX);
            r = min(r, nIni - 1);
            knode[i] = r;
        }
        warp_agg_inc(cc, r, i < n);
    }
    __syncthreads();
    if (tid == 0) {
        int pos = 0;
        for (int r = 0; r < nIni; r++) {
            const int c = cc[r];
            childpos[r] = c > 0 ? pos : -1;
            if (c > 0) {
                cur.x0[pos] = (int)__fmul_rn(g.hX, (float)r);
                cur.x1[pos] = (int)__fmul_rn(g.hX, (float)(r + 1));
                cur.y0[pos] = 0;
                cur.y1[pos] = g.h - 32;
                cur.cnt[pos] = c;
                cur.seq[pos] = r;
                pos++;
            }
        }
        sh_size = pos; sh_finish = 0; sh_mode = 0; sh_seq = nIni;
    }
    __syncthreads();
    for (int i = tid; i < n; i += NT) knode[i] = childpos[knode[i]];

    // ---- passes
    while (true) {
        __syncthreads();
        if (sh_finish) break;
        const int size = sh_size, mode = sh_mode, seq0 = sh_seq;
        int nvis = 0;
        // (1) visit list
        if (mode == 0) {
            for (int base = 0; base < size; base += NT) {
                const int p = base + tid;
                const int f = (p < size && cur.cnt[p] > 1) ? 1 : 0;
                int tot;
                const int ex = block_excl_scan(f, scanTmp, &tot);
                if (p < size) { vrank[p] = f ? nvis + ex : -1; recX[p] = -1; }
                if (f) vis[nvis + ex] = p;
                nvis += tot;
            }
        } else {
            for (int base = 0; base < size; base += NT) {
                const int p = base + tid;
                const int f = (p < size && cur.cnt[p] > 1) ? 1 : 0;
                int tot;
                const int ex = block_excl_scan(f, scanTmp, &tot);
                if (p < size) { vrank[p] = -1; recX[p] = -1; }
                if (f)
                    skeys[nvis + ex] = ((unsigned long long)cur.cnt[p] << 40) | ((unsigned long long)(cur.seq[p] & 0xFFFFF) << 20) |
                                       (unsigned long long)p;
                nvis += tot;
            }
            int S = 1;
            while (S < nvis) S <<= 1;
            for (int i = nvis + tid; i < S; i += NT) skeys[i] = 0ull;
            __syncthreads();
            for (int k = 2; k <= S; k <<= 1)
                for (int j = k >> 1; j > 0; j >>= 1) {
                    for (int i = tid; i < S; i += NT) {
                        const int ixj = i ^ j;
                        if (ixj > i) {
                            const unsigned long long a = skeys[i], c = skeys[ixj];
                            const bool desc = (i & k) == 0;
                            if (desc ? a < c : a > c) { skeys[i] = c; skeys[ixj] = a; }
                        }
                    }
                    __syncthreads();
                }
            for (int r = tid; r < nvis; r += NT) {
                const int p = (int)(skeys[r] & 0xFFFFFull);
                vis[r] = p;
                vrank[p] = r;
            }
        }
        __syncthreads();
        // (2) quadrant counts of the visited nodes
        for (int r = tid; r < nvis; r += NT) {
            const int p = vis[r];
            cc[4 * p] = 0; cc[4 * p + 1] = 0; cc[4 * p + 2] = 0; cc[4 * p + 3] = 0;
            const int x0 = cur.x0[p], y0 = cur.y0[p];
            recX[p] = x0 + ((cur.x1[p] - x0 + 1) >> 1); // x0 + ceil(d / 2.f) for d >= 0 (:553-554)
            recY[p] = y0 + ((cur.y1[p] - y0 + 1) >> 1);
        }
        if (tid == 0) { sh_nexp = 0; sh_rstar = nvis - 1; }
        __syncthreads();
        // four keys per thread per trip: the eight loads are issued before the first dependent shared-memory read (one
        // key at a time the loop ran at L2 latency, ~500 cycles per trip)
        for (int i0 = tid; i0 < nPad; i0 += OCT_KU * NT) {
            int pk[OCT_KU];
            uint32_t ok[OCT_KU];
#pragma unroll
            for (int u = 0; u < OCT_KU; u++) {
                const int i = i0 + u * NT;
                pk[u] = i < n ? knode[i] : -1;
                ok[u] = i < n ? ord[i] : 0u;
            }
#pragma unroll
            for (int u = 0; u < OCT_KU; u++) {
                if (i0 - lane + u * NT >= nPad) break; // warp-uniform: nPad and NT are multiples of 32
                int key = 0;
                bool act = false;
                if (pk[u] >= 0) {
                    const int xs = recX[pk[u]];
                    if (xs >= 0) {
                        key = 4 * pk[u] + key_quadrant(ok[u], xs, recY[pk[u]]);
                        act = true;
                    }
                }
                warp_agg_inc(cc, key, act);
            }
        }
        __syncthreads();
        // (3) creation index of the children; in mode 1 also the cut (:879)
        int T = 0;
        for (int base = 0; base < nvis; base += NT) {
            const int r = base + tid;
            int nch = 0;
            if (r < nvis) {
                const int p = vis[r];
                nch = (cc[4 * p] > 0) + (cc[4 * p + 1] > 0) + (cc[4 * p + 2] > 0) + (cc[4 * p + 3] > 0);
            }
            int tot;
            const int ex = block_excl_scan(nch, scanTmp, &tot);
            if (r < nvis) {
                basea[r] = T + ex;
                cum[r] = T + ex + nch - (r + 1); // list growth after splitting visit ranks 0..r
            }
            T += tot;
        }
        int nsplit = nvis;
        if (mode == 1) {
            for (int r = tid; r < nvis; r += NT)
                if (size + cum[r] >= N) atomicMin(&sh_rstar, r);
            __syncthreads();
            nsplit = sh_rstar + 1;
            for (int r = nsplit + tid; r < nvis; r += NT) { vrank[vis[r]] = -1; recX[vis[r]] = -1; }
            T = nsplit > 0 ? basea[nsplit - 1] : 0;
            if (nsplit > 0) {
                const int p = vis[nsplit - 1];
                T += (cc[4 * p] > 0) + (cc[4 * p + 1] > 0) + (cc[4 * p + 2] > 0) + (cc[4 * p + 3] > 0);
            }
        }
        __syncthreads();
        // (4) children at the front in reverse creation order ...
        int nexp = 0;
        for (int r = tid; r < nsplit; r += NT) {
            const int p = vis[r];
            const int x0 = cur.x0[p], x1 = cur.x1[p], y0 = cur.y0[p], y1 = cur.y1[p];
            const int hx = (x1 - x0 + 1) >> 1, hy = (y1 - y0 + 1) >> 1;
            int ci = basea[r];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int c = cc[4 * p + q];
                if (c > 0) {
                    const int np = T - 1 - ci;
                    nxt.x0[np] = (q & 1) ? x0 + hx : x0;
                    nxt.x1[np] = (q & 1) ? x1 : x0 + hx;
                    nxt.y0[np] = (q & 2) ? y0 + hy : y0;
                    nxt.y1[np] = (q & 2) ? y1 : y0 + hy;
                    nxt.cnt[np] = c;
                    nxt.seq[np] = seq0 + ci;
                    childpos[4 * p + q] = np;
                    nexp += c > 1;
                    ci++;
                }
            }
        }
        if (nexp) atomicAdd(&sh_nexp, nexp);
        // ... then the unsplit nodes in their old order
        int nsurv = 0;
        for (int base = 0; base < size; base += NT) {
            const int p = base + tid;
            const int f = (p < size && vrank[p] < 0) ? 1 : 0;
            int tot;
            const int ex = block_excl_scan(f, scanTmp, &tot);
            if (f) {
                const int np = T + nsurv + ex;
                nxt.x0[np] = cur.x0[p]; nxt.x1[np] = cur.x1[p]; nxt.y0[np] = cur.y0[p]; nxt.y1[np] = cur.y1[p];
                nxt.cnt[np] = cur.cnt[p]; nxt.seq[np] = cur.seq[p];
                newpos[p] = np;
            }
            nsurv += tot;
        }
        __syncthreads();
        // (5) keys follow their node
        for (int i0 = tid; i0 < n; i0 += OCT_KU * NT) {
            int pk[OCT_KU];
            uint32_t ok[OCT_KU];
#pragma unroll
            for (int u = 0; u < OCT_KU; u++) {
                const int i = i0 + u * NT;
                pk[u] = i < n ? knode[i] : -1;
                ok[u] = i < n ? ord[i] : 0u;
            }
#pragma unroll
            for (int u = 0; u < OCT_KU; u++) {
                if (pk[u] >= 0) {
                    const int xs = recX[pk[u]];
                    knode[i0 + u * NT] = xs >= 0 ? childpos[4 * pk[u] + key_quadrant(ok[u], xs, recY[pk[u]])] : newpos[pk[u]];
                }
            }
        }
        __syncthreads();
        const int newSize = T + nsurv;
        if (tid == 0) {
            sh_size = newSize;
            sh_seq = seq0 + T;
            if (newSize >= N || newSize == size) sh_finish = 1;
            else if (mode == 0 && newSize + sh_nexp * 3 > N) sh_mode = 1;
        }
        NodeBuf t = cur; cur = nxt; nxt = t;
    }
    // ---- one keypoint per node: max response, first in original order on ties (:895-912)
    const int size = sh_size;
    unsigned int *best = reinterpret_cast<unsigned int *>(cc);
    for (int p = tid; p < size; p += NT) best[p] = 0u;
    __syncthreads();
    for (int i = tid; i < n; i += NT)
        atomicMax(&best[knode[i]], ((ord[i] >> 24) << 24) | (0xFFFFFFu - (unsigned int)i));
    __syncthreads();
    uint32_t *out = P.lvlKp + (size_t)b * P.nodeTotal + g.nodeOff;
    for (int p = tid; p < size; p += NT) out[p] = ord[0xFFFFFFu - (best[p] & 0xFFFFFFu)];
    if (tid == 0) P.lvlCount[(size_t)b * P.nlevels + l] = size;
}

} // namespace orbb200
