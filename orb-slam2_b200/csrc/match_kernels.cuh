// match_kernels.cuh -- sm_100a kernels of the ORBmatcher descriptor-distance searches.
//
// A descriptor is one 32-byte row (256 bits).  DescriptorDistance (src/ORBmatcher.cc:1733-1749) is 8 x
// (XOR, popcount) on 32-bit words; here 2 x 16-byte loads, 8 LOP3 + 8 POPC + adds.
// The reference's running best / second-best update (strict '<', :229-239) yields exactly the two smallest
// elements under the lexicographic key (distance, visit position) -- positions only grow, so '<' on the
// distance alone keeps the earlier one.  Every kernel below reproduces that order, so tie-breaking equals the
// sequential loop's.
#pragma once
#include "common.cuh"

namespace orbb200 {

struct Desc {
    uint4 lo, hi;
};

__device__ __forceinline__ Desc load_desc(const uint8_t *p)
{
    Desc d;
    d.lo = __ldg(reinterpret_cast<const uint4 *>(p));
    d.hi = __ldg(reinterpret_cast<const uint4 *>(p) + 1);
    return d;
}

__device__ __forceinline__ int hamming(const Desc &a, const Desc &b)
{
    return (__popc(a.lo.x ^ b.lo.x) + __popc(a.lo.y ^ b.lo.y)) + (__popc(a.lo.z ^ b.lo.z) + __popc(a.lo.w ^ b.lo.w)) +
           (__popc(a.hi.x ^ b.hi.x) + __popc(a.hi.y ^ b.hi.y)) + (__popc(a.hi.z ^ b.hi.z) + __popc(a.hi.w ^ b.hi.w));
}

struct Top2 {
    int d1, i1, d2, i2;
};

__device__ __forceinline__ void top2_init(Top2 &t) { t.d1 = 256; t.i1 = -1; t.d2 = 256; t.i2 = -1; }

// sequential update, candidates arrive in increasing position (src/ORBmatcher.cc:229-239)
__device__ __forceinline__ void top2_push(Top2 &t, int d, int i)
{
    if (d < t.d2) {
        if (d < t.d1) { t.d2 = t.d1; t.i2 = t.i1; t.d1 = d; t.i1 = i; }
        else { t.d2 = d; t.i2 = i; }
    }
}

// lexicographic (distance, position) order; position -1 (absent) sorts last
__device__ __forceinline__ bool lex_less(int da, int ia, int db, int ib)
{
    return da < db || (da == db && (unsigned int)ia < (unsigned int)ib);
}

// merge two top-2 sets (each internally ordered) into the two lexicographically smallest of their union
__device__ __forceinline__ Top2 top2_merge(const Top2 &a, const Top2 &b)
{
    Top2 r;
    if (lex_less(b.d1, b.i1, a.d1, a.i1)) {
        r.d1 = b.d1; r.i1 = b.i1;
        if (lex_less(a.d1, a.i1, b.d2, b.i2)) { r.d2 = a.d1; r.i2 = a.i1; } else { r.d2 = b.d2; r.i2 = b.i2; }
    } else {
        r.d1 = a.d1; r.i1 = a.i1;
        if (lex_less(b.d1, b.i1, a.d2, a.i2)) { r.d2 = b.d1; r.i2 = b.i1; } else { r.d2 = a.d2; r.i2 = a.i2; }
    }
    return r;
}

__device__ __forceinline__ Top2 top2_warp_merge(Top2 t)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Top2 u;
        u.d1 = __shfl_xor_sync(0xffffffffu, t.d1, o); u.i1 = __shfl_xor_sync(0xffffffffu, t.i1, o);
        u.d2 = __shfl_xor_sync(0xffffffffu, t.d2, o); u.i2 = __shfl_xor_sync(0xffffffffu, t.i2, o);
        t = top2_merge(t, u);
    }
    return t;
}

// M1 on arrays of pairs
__global__ void k_pair_distance(const uint8_t *__restrict__ a, const uint8_t *__restrict__ b, int n, int *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = hamming(load_desc(a + (size_t)i * 32), load_desc(b + (size_t)i * 32));
}

// ---------------------------------------------------------------------------------------------------
// X1  bf_top2: every query against a flat database (the loop of src/ORBmatcher.cc:214-240).
//   block = BF_THREADS threads x BF_QPT queries per thread (queries live in registers);
//   grid  = (query groups, database chunks); a block streams its chunk through shared memory in
//   BF_TILE-row tiles (cp.async double buffer) and every thread reads the rows as broadcast LDS.128.
//   Output: per-(chunk, query) partial top-2 with GLOBAL row indices; k_merge_top2 folds the chunks.
// ---------------------------------------------------------------------------------------------------
constexpr int BF_THREADS = 128, BF_QPT = 4, BF_TILE = 128;

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem)
{
    const unsigned int s = (unsigned int)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

__global__ void __launch_bounds__(BF_THREADS) k_bf_top2(const uint8_t *__restrict__ q, int nq, const uint8_t *__restrict__ db,
                                                        int ndb, int rowsPerChunk, int indexOffset,
                                                        orbb200_top2 *__restrict__ partial)
{
    __shared__ __align__(16) uint4 tile[2][BF_TILE * 2];
    const int q0 = (blockIdx.x * BF_THREADS + threadIdx.x) * BF_QPT;
    const int r0 = blockIdx.y * rowsPerChunk, r1 = min(r0 + rowsPerChunk, ndb);
    Desc Q[BF_QPT];
    Top2 T[BF_QPT];
#pragma unroll
    for (int k = 0; k < BF_QPT; k++) {
        const int qi = min(q0 + k, nq - 1);
        Q[k] = load_desc(q + (size_t)qi * 32);
        top2_init(T[k]);
    }
    const int ntiles = (r1 - r0 + BF_TILE - 1) / BF_TILE;
    auto issue = [&](int t, int buf) {
        const int base = r0 + t * BF_TILE;
        const int rows = min(BF_TILE, r1 - base);
        for (int i = threadIdx.x; i < rows * 2; i += BF_THREADS)
            cp_async16(&tile[buf][i], db + (size_t)base * 32 + (size_t)i * 16);
        cp_async_commit();
    };
    if (ntiles > 0) issue(0, 0);
    for (int t = 0; t < ntiles; t++) {
        if (t + 1 < ntiles) { issue(t + 1, (t + 1) & 1); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        __syncthreads();
        const int base = r0 + t * BF_TILE;
        const int rows = min(BF_TILE, r1 - base);
        const uint4 *rowp = tile[t & 1];
#pragma unroll 4
        for (int r = 0; r < rows; r++) {
            Desc D;
            D.lo = rowp[2 * r];
            D.hi = rowp[2 * r + 1];
#pragma unroll
            for (int k = 0; k < BF_QPT; k++) top2_push(T[k], hamming(Q[k], D), indexOffset + base + r);
        }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < BF_QPT; k++)
        if (q0 + k < nq) {
            orbb200_top2 o;
            o.best_dist = T[k].d1; o.best_idx = T[k].i1; o.second_dist = T[k].d2; o.second_idx = T[k].i2;
            partial[(size_t)blockIdx.y * nq + q0 + k] = o;
        }
}

// fold nshards partial results per query; shards are ordered by database position, so the lexicographic
// (distance, index) rule reproduces the first-seen tie-break of the sequential loop.
__global__ void k_merge_top2(const orbb200_top2 *__restrict__ parts, int nshards, int nq, orbb200_top2 *__restrict__ out)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= nq) return;
    Top2 t;
    top2_init(t);
    for (int s = lane; s < nshards; s += 32) {
        const orbb200_top2 p = parts[(size_t)s * nq + warp];
        Top2 u;
        u.d1 = p.best_dist; u.i1 = p.best_idx; u.d2 = p.second_dist; u.i2 = p.second_idx;
        t = top2_merge(t, u);
    }
    t = top2_warp_merge(t);
    if (lane == 0) {
        orbb200_top2 o;
        o.best_dist = t.d1; o.best_idx = t.i1; o.second_dist = t.d2; o.second_idx = t.i2;
        out[warp] = o;
    }
}

// ---------------------------------------------------------------------------------------------------
// M2-M9 core  match_candidates: one warp per query, lanes stride the query's CSR candidate list.
// Position = place in the list; the result carries the candidate's target index.
// ---------------------------------------------------------------------------------------------------
__global__ void k_match_candidates(const uint8_t *__restrict__ q, int nq, const uint8_t *__restrict__ t,
                                   const int *__restrict__ offs, const int *__restrict__ idx, orbb200_top2 *__restrict__ out)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= nq) return;
    const Desc Q = load_desc(q + (size_t)warp * 32);
    const int c0 = offs[warp], c1 = offs[warp + 1];
    Top2 r;
    top2_init(r);
    for (int c = c0 + lane; c < c1; c += 32) top2_push(r, hamming(Q, load_desc(t + (size_t)__ldg(idx + c) * 32)), c - c0);
    r = top2_warp_merge(r);
    if (lane == 0) {
        orbb200_top2 o;
        o.best_dist = r.d1; o.best_idx = r.i1 >= 0 ? idx[c0 + r.i1] : -1;
        o.second_dist = r.d2; o.second_idx = r.i2 >= 0 ? idx[c0 + r.i2] : -1;
        out[warp] = o;
    }
}

// all (query, candidate) distances of a CSR candidate set: dist[c] for every list entry c.  The host epilogues of
// the Search* methods that skip targets claimed earlier in the same call (src/ORBmatcher.cc:93-95, 222, 1485-1487)
// need every distance, not only the top-2.
__global__ void k_candidate_distances(const uint8_t *__restrict__ q, int nq, const uint8_t *__restrict__ t,
                                      const int *__restrict__ offs, const int *__restrict__ idx, unsigned short *__restrict__ dist)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= nq) return;
    const Desc Q = load_desc(q + (size_t)warp * 32);
    const int c0 = offs[warp], c1 = offs[warp + 1];
    for (int c = c0 + lane; c < c1; c += 32)
        dist[c] = (unsigned short)hamming(Q, load_desc(t + (size_t)__ldg(idx + c) * 32));
}

// POPC-pipe peak probe for the matching roofline: dependent-free popc chains, 8 per iteration.
__global__ void k_popc_peak(unsigned int *out, int iters)
{
    unsigned int a[8];
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = threadIdx.x * 2654435761u + k * 40503u + blockIdx.x;
    unsigned int acc = 0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const unsigned int p = __popc(a[k]);
            acc += p;
            a[k] = a[k] * 3u + p; // keeps the chain data dependent on the popc so it cannot be hoisted
        }
    }
    if (acc == 0xFFFFFFFFu) out[0] = acc;
}

} // namespace orbb200
