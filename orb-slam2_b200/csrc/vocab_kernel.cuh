// vocab_kernel.cuh -- the DBoW2 vocabulary-tree descent for a batch of ORB descriptors
// (TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup), Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1259;
// distance = FORB::distance, FORB.cpp:81-99).
//
// Tree layout in HBM (built by orbb200_vocabulary_create): nodes renumbered breadth-first so that the children of a node
// are CONTIGUOUS and in the reference's child order; per node a 32-byte descriptor (two uint4) and an int2
// {first child, child count}; a word (no children) keeps {word id, 0}.  ORBvoc.txt (k = 10, L = 6, 1.08 M nodes) is 35 MB of
// descriptors + 8.7 MB of node records: resident in the 126 MB L2 after the first pass over it.
//
// One group of G lanes (16, or 32 when a node has more than 16 children) walks one feature: lane c of the group loads
// child c's descriptor (2 x LDG.128, a group's lanes read consecutive 32-byte records) and its node record, 8 POPC, and
// one redux.sync.min over (distance << 8 | child index) picks the nearest child, the FIRST on ties like the reference's
// strict `<` scan; the winner's node record comes over by shuffle, so a level costs one dependent load round.
#pragma once
#include "common.cuh"

namespace orbb200 {

constexpr int VT_THREADS = 256;

// one 32-byte node descriptor in ONE request (LDG.E.256, sm_100): half the L1 wavefronts of two 16-byte loads
__device__ __forceinline__ void ldg_desc256(const uint4 *p, uint4 &a, uint4 &b)
{
    asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
        : "l"(p));
}

template <int G>
__global__ void __launch_bounds__(VT_THREADS)
k_vocab_transform(const uint4 *__restrict__ nodeDesc, const int2 *__restrict__ nodeInfo, const int *__restrict__ nodeOrig,
                  const double *__restrict__ nodeWeight, const uint4 *__restrict__ feats, int n, int nidLevel,
                  int *__restrict__ wordOut, double *__restrict__ weightOut, int *__restrict__ nidOut,
                  const int *__restrict__ counts, int stride)
{
    const unsigned lane = threadIdx.x & 31u, gl = lane & (G - 1), gbase = lane & ~(unsigned)(G - 1);
    const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << (G & 31)) - 1u) << gbase);
    const int f = blockIdx.x * (VT_THREADS / G) + (int)(threadIdx.x / G);
    if (f >= n) return; // group-uniform
    if (counts) { // strided batch ([image][stride] rows, counts[image] of them valid): skip the unused rows
        const int img = f / stride;
        if (f - img * stride >= __ldg(counts + img)) return;
    }

    const uint4 f0 = __ldg(feats + 2 * (size_t)f), f1 = __ldg(feats + 2 * (size_t)f + 1);
    const int2 root = __ldg(nodeInfo);
    int base = root.x, cnt = root.y, cur = 0, level = 0;
    int nidNode = nidLevel <= 0 ? 0 : -1; // if(nid_level <= 0 && nid != NULL) *nid = 0; (:1227)
    while (cnt > 0) {                     // do { ... } while(!m_nodes[final_id].isLeaf()) (:1232-1253)
        ++level;
        unsigned best = 0xffffffffu;
        int2 mine = make_int2(0, 0);
#pragma unroll 3
        for (int c = (int)gl; c < cnt; c += G) {
            const int id = base + c;
            uint4 a, b;
            ldg_desc256(nodeDesc + 2 * (size_t)id, a, b);
            const int2 inf = __ldg(nodeInfo + id);
            const int d = __popc(a.x ^ f0.x) + __popc(a.y ^ f0.y) + __popc(a.z ^ f0.z) + __popc(a.w ^ f0.w) +
                          __popc(b.x ^ f1.x) + __popc(b.y ^ f1.y) + __popc(b.z ^ f1.z) + __popc(b.w ^ f1.w);
            const unsigned key = ((unsigned)d << 8) | (unsigned)c;
            if (key < best) { best = key; mine = inf; }
        }
        const unsigned win = __reduce_min_sync(gmask, best);
        const int wc = (int)(win & 255u);
        const int src = (int)gbase + (wc & (G - 1)); // the lane that scanned child wc holds it as its own minimum
        cur = base + wc;
        base = __shfl_sync(gmask, mine.x, src);
        cnt = __shfl_sync(gmask, mine.y, src);
        if (level == nidLevel) nidNode = cur; // if(nid != NULL && current_level == nid_level) *nid = final_id; (:1250-1251)
    }
    if (gl == 0) {
        wordOut[f] = base;                     // word_id = m_nodes[final_id].word_id (:1256)
        weightOut[f] = __ldg(nodeWeight + cur); // weight = m_nodes[final_id].weight (:1257)
        // a word above the nid level: *nid is never written by the reference (indeterminate); filed under the word's own node
        nidOut[f] = __ldg(nodeOrig + (nidNode < 0 ? cur : nidNode));
    }
}

// Thread-per-feature variant: one thread scans all children of a node itself (the reference's loop, strict `<` keeps the
// first minimum), so a level costs no shuffles / redux and every issued POPC is useful; the children's 32-byte records
// are whole sectors, so per-thread loads waste no bandwidth.
constexpr int VW_THREADS = 128;

__global__ void __launch_bounds__(VW_THREADS)
k_vocab_walk(const uint4 *__restrict__ nodeDesc, const int2 *__restrict__ nodeInfo, const int *__restrict__ nodeOrig,
             const double *__restrict__ nodeWeight, const uint4 *__restrict__ feats, int n, int nidLevel,
             int *__restrict__ wordOut, double *__restrict__ weightOut, int *__restrict__ nidOut,
             const int *__restrict__ counts, int stride)
{
    const int f = blockIdx.x * VW_THREADS + (int)threadIdx.x;
    if (f >= n) return;
    if (counts) {
        const int img = f / stride;
        if (f - img * stride >= __ldg(counts + img)) return;
    }
    const uint4 f0 = __ldg(feats + 2 * (size_t)f), f1 = __ldg(feats + 2 * (size_t)f + 1);
    const int2 root = __ldg(nodeInfo);
    int base = root.x, cnt = root.y, cur = 0, level = 0;
    int nidNode = nidLevel <= 0 ? 0 : -1;
    while (cnt > 0) {
        ++level;
        int best = 1 << 30, bi = 0;
#pragma unroll 5
        for (int c = 0; c < cnt; c++) {
            uint4 a, b;
            ldg_desc256(nodeDesc + 2 * (size_t)(base + c), a, b);
            const int d = __popc(a.x ^ f0.x) + __popc(a.y ^ f0.y) + __popc(a.z ^ f0.z) + __popc(a.w ^ f0.w) +
                          __popc(b.x ^ f1.x) + __popc(b.y ^ f1.y) + __popc(b.z ^ f1.z) + __popc(b.w ^ f1.w);
            if (d < best) { best = d; bi = c; } // if(d < best_d) (:1243)
        }
        cur = base + bi;
        const int2 inf = __ldg(nodeInfo + cur);
        base = inf.x;
        cnt = inf.y;
        if (level == nidLevel) nidNode = cur;
    }
    wordOut[f] = base;
    weightOut[f] = __ldg(nodeWeight + cur);
    nidOut[f] = __ldg(nodeOrig + (nidNode < 0 ? cur : nidNode));
}

// ---- BowVector / FeatureVector fold (transform(features, v, fv, levelsup), TemplatedVocabulary.h:1127-1190) ----
// One block per image.  The image's features with weight > 0 ("not stopped") are sorted in shared memory by
// (word, feature index) -- the order std::map<WordId, ...> + addWeight's feature-order accumulation implies -- run heads
// become BowVector entries whose values are accumulated term by term in that order (BowVector.cpp:34-58), thread 0 then
// takes the norm in word order exactly like BowVector::normalize (:62-84), and a second sort by (node, feature index)
// yields the FeatureVector's node groups with ascending feature lists (FeatureVector.cpp:31-45).  All doubles are
// IEEE add / div / sqrt (no FMA: the library is built with -fmad=false), so the values are bit-identical to the host code.
constexpr int BF_THREADS = 512;

__device__ __forceinline__ void bitonic_sort_u64(unsigned long long *key, int npad)
{
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < (npad >> 1); t += BF_THREADS) {
                const int i = 2 * t - (t & (j - 1)), p = i + j;
                const unsigned long long a = key[i], b = key[p];
                const bool up = (i & k) == 0;
                if ((a > b) == up) { key[i] = b; key[p] = a; }
            }
            __syncthreads();
        }
    }
}

struct BowFoldParams {
    const int *word, *nid;   // per feature, all images back to back
    const double *weight;
    const int *imgOff;       // [nimg + 1] prefix sums of the per-image feature counts (packed batch, stride == 0)
    const int *counts;       // strided batch: image i's features are rows i * stride .. + counts[i]
    int stride;
    int tf, must, norm;      // weighting is TF / TF_IDF; scoring normalises; 1 = L1, 2 = L2
    int *bowId;              // image i's entries at its offset .. + nBow[i]
    double *bowVal;
    unsigned long long *fvKey; // [total]: (node << 32 | feature index) ascending, image i's at its offset .. + nKept[i]
    int *nBow, *nKept;       // [nimg]
};

__global__ void __launch_bounds__(BF_THREADS) k_bow_fold(BowFoldParams P, int npad)
{
    extern __shared__ __align__(16) unsigned char bf_smem[];
    unsigned long long *key = (unsigned long long *)bf_smem; // [npad]
    double *val = (double *)(key + npad);                     // [npad]
    __shared__ int scanTmp[34];
    __shared__ int sCount;
    __shared__ double sNorm;
    const int img = blockIdx.x, off = P.stride ? img * P.stride : P.imgOff[img];
    const int n = P.stride ? min(P.counts[img], P.stride) : P.imgOff[img + 1] - off;
    const int *word = P.word + off, *nid = P.nid + off;
    const double *w = P.weight + off;
    if (threadIdx.x == 0) sCount = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += BF_THREADS)
        if (w[i] > 0) key[atomicAdd(&sCount, 1)] = ((unsigned long long)(unsigned)word[i] << 32) | (unsigned)i; // if(w > 0) // not stopped
    __syncthreads();
    const int m = sCount;
    for (int i = m + threadIdx.x; i < npad; i += BF_THREADS) key[i] = ~0ull;
    __syncthreads();
    bitonic_sort_u64(key, npad);
    // run heads -> BowVector entries
    int nb = 0;
    for (int base = 0; base < m; base += BF_THREADS) {
        const int i = base + threadIdx.x;
        const bool head = i < m && (i == 0 || (unsigned)(key[i] >> 32) != (unsigned)(key[i - 1] >> 32));
        int total;
        const int pos = nb + block_excl_scan(head ? 1 : 0, scanTmp, &total);
        if (head) {
            const unsigned id = (unsigned)(key[i] >> 32);
            double v = w[(unsigned)key[i]];
            if (P.tf) // addWeight: vit->second += v in feature order; addIfNotExist keeps the first value
                for (int r = i + 1; r < m && (unsigned)(key[r] >> 32) == id; r++) v += w[(unsigned)key[r]];
            P.bowId[off + pos] = (int)id;
            val[pos] = v;
        }
        nb += total;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        if (P.must) { // BowVector::normalize
            if (P.norm == 1) {
                for (int j = 0; j < nb; j++) s += fabs(val[j]);
            } else {
                for (int j = 0; j < nb; j++) s += val[j] * val[j];
                s = sqrt(s);
            }
        }
        sNorm = s;
        P.nBow[img] = nb;
        P.nKept[img] = m;
    }
    __syncthreads();
    {
        const double nd = (double)nb, s = sNorm;
        for (int j = threadIdx.x; j < nb; j += BF_THREADS) {
            double v = val[j];
            if (P.tf && !P.must) v /= nd; // :1163-1169
            if (P.must && s > 0.0) v /= s;
            P.bowVal[off + j] = v;
        }
    }
    // FeatureVector: regroup the same features by node
    for (int i = threadIdx.x; i < m; i += BF_THREADS) {
        const unsigned f = (unsigned)key[i];
        key[i] = ((unsigned long long)(unsigned)nid[f] << 32) | f;
    }
    __syncthreads();
    bitonic_sort_u64(key, npad);
    for (int i = threadIdx.x; i < m; i += BF_THREADS) P.fvKey[off + i] = key[i];
}

} // namespace orbb200
