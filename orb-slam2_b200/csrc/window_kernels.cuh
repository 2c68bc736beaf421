// window_kernels.cuh -- F3 + F2 + the distance step of the windowed searches.
//   k_grid_build      Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:261-279, 444-456): counting sort of the
//                     keypoints by grid cell (posX * rows + posY), STABLE in keypoint index like the push_back loop.
//   k_window_count    Frame::GetFeaturesInArea (src/Frame.cc:386-442) per query window: number of hits.
//   k_window_fill     the same enumeration again, in the reference's order (ix outer, iy inner, insertion order inside a
//                     cell), writing the candidate index and its DescriptorDistance to the query descriptor.
// One warp per query; order inside a cell comes from ballot-ordered compaction, never from atomics.
#pragma once
#include "match_kernels.cuh"

namespace orbb200 {

struct GridParams {
    const orbb200_keypoint *kp; // target keypoints (the caller's mvKeysUn)
    int n;
    const int *nPtr;              // when not null: the keypoint count lives on the device (min(*nPtr, n) is used)
    float minX, minY, invW, invH; // mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv
    int cols, rows;               // FRAME_GRID_COLS, FRAME_GRID_ROWS
    int *cellStart;               // [cols*rows + 1]
    int *cellItems;               // [n]
};

// One block of GB_THREADS.  Shared: cols*rows + 1 ints, then sortLen u32 sort keys (sortLen = 0: no room, ordered fill by one warp).
// The stable counting sort is done as a SORT of (cell << 13 | keypoint index): ascending keys = cells ascending, insertion order
// inside a cell.  (The ordered fill by one warp -- 32 keypoints per step, each step a global load, a match.any and a dependent
// shared-memory update -- took 30 us for 2 000 keypoints; the bitonic sort of 2 048 keys in shared memory takes ~3.)
constexpr int GB_THREADS = 1024, GB_SORT_MAX = 8192;
__host__ __device__ inline size_t grid_build_smem(int nc, int n, int *sortLen)
{
    int S = 0;
    if (n <= GB_SORT_MAX && nc < (1 << 18)) { S = 32; while (S < n) S <<= 1; }
    *sortLen = S;
    return (size_t)(nc + 1) * 4 + (size_t)S * 4;
}

__global__ void __launch_bounds__(GB_THREADS) k_grid_build(const GridParams G, int sortLen)
{
    extern __shared__ int gsm[];
    __shared__ int scanTmp[34];
    const int nc = G.cols * G.rows, tid = threadIdx.x, NT = blockDim.x;
    unsigned int *skey = reinterpret_cast<unsigned int *>(gsm + nc + 1);
    const int n = G.nPtr ? min(max(*G.nPtr, 0), G.n) : G.n;
    const bool sorted = n <= sortLen;
    for (int c = tid; c <= nc; c += NT) gsm[c] = 0;
    __syncthreads();
    auto cell_of = [&](int i) {
        const orbb200_keypoint k = G.kp[i];
        const int px = (int)roundf(__fmul_rn(__fsub_rn(k.x, G.minX), G.invW));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(k.y, G.minY), G.invH));
        return (px < 0 || px >= G.cols || py < 0 || py >= G.rows) ? -1 : px * G.rows + py;
    };
    for (int i = tid; i < (sorted ? sortLen : n); i += NT) {
        const int c = i < n ? cell_of(i) : -1;
        if (sorted) skey[i] = c >= 0 ? ((unsigned int)c << 13) | (unsigned int)i : 0xFFFFFFFFu;
        if (c >= 0) atomicAdd(&gsm[c], 1);
    }
    __syncthreads();
    int run = 0;
    for (int base = 0; base < nc; base += NT) {
        const int c = base + tid;
        const int v = c < nc ? gsm[c] : 0;
        int tot;
        const int ex = block_excl_scan(v, scanTmp, &tot);
        if (c < nc) { G.cellStart[c] = run + ex; gsm[c] = run + ex; }
        run += tot;
    }
    if (tid == 0) G.cellStart[nc] = run;
    __syncthreads();
    if (sorted) {
        for (int k = 2; k <= sortLen; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = tid; i < sortLen; i += NT) {
                    const int ixj = i ^ j;
                    if (ixj > i) {
                        const unsigned int a = skey[i], c = skey[ixj];
                        const bool up = (i & k) == 0;
                        if (up ? a > c : a < c) { skey[i] = c; skey[ixj] = a; }
                    }
                }
                __syncthreads();
            }
        for (int j = tid; j < run; j += NT) G.cellItems[j] = (int)(skey[j] & 8191u);
        return;
    }
    // stable fill by ONE warp walking the keypoints in index order
    if (tid < 32) {
        const unsigned int lt = (1u << tid) - 1u;
        for (int base = 0; base < n; base += 32) {
            const int i = base + tid;
            const int c = i < n ? cell_of(i) : -1;
            const unsigned int act = __ballot_sync(0xffffffffu, c >= 0);
            if (c >= 0) {
                const unsigned int peers = __match_any_sync(act, c);
                const int pos = gsm[c] + __popc(peers & lt);
                G.cellItems[pos] = i;
                __syncwarp(act);
                if ((peers & lt) == 0) gsm[c] += __popc(peers);
            }
            __syncwarp();
        }
    }
}

struct WindowParams {
    GridParams G;
    const float *win;     // [nq][5]: x, y, r, minLevel, maxLevel (levels as floats holding small ints)
    const uint8_t *qdesc; // [nq][32]
    const uint8_t *tdesc; // [n][32]
    int nq;
    int *counts;          // [nq + 1] counts, scanned in place into offsets
    int *candIdx;
    unsigned short *dist;
    int cap;
};

template <bool FILL> __device__ __forceinline__ void window_walk(const WindowParams &W, int q, int lane)
{
    const GridParams &G = W.G;
    const float x = W.win[5 * q], y = W.win[5 * q + 1], r = W.win[5 * q + 2];
    const int minLevel = (int)W.win[5 * q + 3], maxLevel = (int)W.win[5 * q + 4];
    int n = 0;
    const int out0 = FILL ? W.counts[q] : 0;
    Desc Q;
    if (FILL) Q = load_desc(W.qdesc + (size_t)q * 32);
    const int c0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, G.minX), r), G.invW)));
    const int c1 = min(G.cols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, G.minX), r), G.invW)));
    const int r0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, G.minY), r), G.invH)));
    const int r1 = min(G.rows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, G.minY), r), G.invH)));
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    const unsigned int lt = (1u << lane) - 1u;
    if (c0 < G.cols && c1 >= 0 && r0 < G.rows && r1 >= 0) {
        for (int ix = c0; ix <= c1; ix++) {
            // cells (ix, r0..r1) are adjacent in the cell-major item array: one contiguous, already ordered run
            const int j0 = G.cellStart[ix * G.rows + r0], j1 = G.cellStart[ix * G.rows + r1 + 1];
            for (int base = j0; base < j1; base += 32) {
                const int j = base + lane;
                bool hit = false;
                int idx = -1;
                if (j < j1) {
                    idx = G.cellItems[j];
                    const orbb200_keypoint k = G.kp[idx];
                    hit = true;
                    if (bCheckLevels) {
                        if (k.octave < minLevel) hit = false;
                        if (maxLevel >= 0 && k.octave > maxLevel) hit = false;
                    }
                    const float distx = __fsub_rn(k.x, x), disty = __fsub_rn(k.y, y);
                    if (!(fabsf(distx) < r && fabsf(disty) < r)) hit = false;
                }
                const unsigned int m = __ballot_sync(0xffffffffu, hit);
                if (FILL && hit) {
                    const int pos = out0 + n + __popc(m & lt);
                    if (pos < W.cap) {
                        W.candIdx[pos] = idx;
                        W.dist[pos] = (unsigned short)hamming(Q, load_desc(W.tdesc + (size_t)idx * 32));
                    }
                }
                n += __popc(m);
            }
        }
    }
    if (!FILL && lane == 0) W.counts[q] = n;
}

__global__ void __launch_bounds__(256) k_window_count(const WindowParams W)
{
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= W.nq) return;
    window_walk<false>(W, q, lane);
}

__global__ void __launch_bounds__(256) k_window_fill(const WindowParams W)
{
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= W.nq) return;
    window_walk<true>(W, q, lane);
}

// in-place exclusive scan of counts[0..n) by one block; counts[n] = total.  mirror (optional): a second copy of the offsets,
// e.g. in mapped host memory, so that the caller needs no copy of its own
__global__ void __launch_bounds__(1024) k_scan_counts(int *counts, int n, int *mirror)
{
    __shared__ int scanTmp[34];
    int run = 0;
    for (int base = 0; base < n; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const int v = i < n ? counts[i] : 0;
        int tot;
        const int ex = block_excl_scan(v, scanTmp, &tot);
        if (i < n) {
            counts[i] = run + ex;
            if (mirror) mirror[i] = run + ex;
        }
        run += tot;
    }
    if (threadIdx.x == 0) {
        counts[n] = run;
        if (mirror) mirror[n] = run;
    }
}

} // namespace orbb200
