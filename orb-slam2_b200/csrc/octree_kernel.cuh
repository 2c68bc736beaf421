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
