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

// full adder on 32 bit columns at once: sum = a ^ b ^ c, carry = majority(a, b, c) -- one LOP3 each
__device__ __forceinline__ void csa(unsigned int a, unsigned int b, unsigned int c, unsigned int &sum, unsigned int &carry)
{
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(sum) : "r"(a), "r"(b), "r"(c));
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(carry) : "r"(a), "r"(b), "r"(c));
}

// Hamming distance of two 256-bit rows.  The straight form is 8 XOR + 8 POPC; POPC issues at 16 lanes / clk / SM (a quarter of
// the logic rate), which is what bounded the brute-force search (92 % of the POPC peak).  Three carry-save adders fold seven of the
// eight XOR words into one word of weight 1 and three of weight 2: 5 POPC + 14 LOP3 per pair put the POPC pipe (10 clk per
// warp-pair) and the ALU pipe (~10 clk) in balance.  (A fourth adder -> 4 POPC makes the ALU pipe the longer one again.)
__device__ __forceinline__ int hamming(const Desc &a, const Desc &b)
{
#ifdef ORBB200_HAMMING_PLAIN
    return (__popc(a.lo.x ^ b.lo.x) + __popc(a.lo.y ^ b.lo.y)) + (__popc(a.lo.z ^ b.lo.z) + __popc(a.lo.w ^ b.lo.w)) +
           (__popc(a.hi.x ^ b.hi.x) + __popc(a.hi.y ^ b.hi.y)) + (__popc(a.hi.z ^ b.hi.z) + __popc(a.hi.w ^ b.hi.w));
#else
    unsigned int s0, c0, s1, c1, s2, c2;
    csa(a.lo.x ^ b.lo.x, a.lo.y ^ b.lo.y, a.lo.z ^ b.lo.z, s0, c0);
    csa(a.lo.w ^ b.lo.w, a.hi.x ^ b.hi.x, a.hi.y ^ b.hi.y, s1, c1);
    csa(s0, s1, a.hi.z ^ b.hi.z, s2, c2);
#ifdef ORBB200_HAMMING_4POPC
    unsigned int s3, c3;
    csa(c0, c1, c2, s3, c3);
    return __popc(s2) + __popc(a.hi.w ^ b.hi.w) + 2 * __popc(s3) + 4 * __popc(c3);
#else
    return (__popc(s2) + __popc(a.hi.w ^ b.hi.w)) + 2 * (__popc(c0) + __popc(c1) + __popc(c2));
#endif
#endif
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

// ---- config 5 on N GPUs: the exchange step fused into the search kernel -------------------------------
// Every rank owns a contiguous slice of database rows.  Every block folds its chunk's top-2 into per-query accumulators
// at L2 (two 64-bit atomicMin keys); the LAST block to finish a query group (512 queries) reads the group's shard result,
// stores it straight into every peer's gather buffer over
// NVLink (cudaIpc-mapped peer memory, 16-byte stores), publishes one release flag per (rank, group) on every peer, waits
// for the peers' flags of the same group and folds the world's shard results in rank order (= database order, so the
// lexicographic (distance, index) rule is the sequential first-seen rule).  One launch per step, no NCCL call; groups
// that finish early are exchanged while the other blocks are still searching.
constexpr int EX_MAX_WORLD = 16;
struct ExchangeParams {
    int world, rank, epoch, nqMax, maxGroups; // epoch: 1, 2, 3, ... identical on every rank
    int *groupCounter;                        // [maxGroups] local arrival counters (reset by the last block)
    unsigned long long *acc;                  // [nqMax][2] local per-query accumulators of the shard's top-2 (packed keys)
    int *status;                              // [1] local: set to 1 when a peer's flag did not arrive in time
    orbb200_top2 *out;                        // [nq] local result
    // per rank p: its exchange region = gather [2][world][nqMax] records, then flags [world][maxGroups]
    unsigned char *peer[EX_MAX_WORLD];
    long long timeoutNs;
};

__device__ __forceinline__ orbb200_top2 *ex_gather(unsigned char *base, const ExchangeParams &X, int parity, int src)
{
    return reinterpret_cast<orbb200_top2 *>(base) + ((size_t)parity * X.world + src) * X.nqMax;
}
__device__ __forceinline__ int *ex_flags(unsigned char *base, const ExchangeParams &X)
{
    return reinterpret_cast<int *>(base + sizeof(orbb200_top2) * 2 * (size_t)X.world * X.nqMax);
}
__device__ __forceinline__ void st_release_sys(int *p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int ld_acquire_sys(const int *p)
{
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ long long global_ns()
{
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ Top2 ex_load(const orbb200_top2 *p)
{
    const int4 v = __ldcg(reinterpret_cast<const int4 *>(p)); // written by other blocks / other GPUs: read at L2
    Top2 t;
    t.d1 = v.x; t.i1 = v.y; t.d2 = v.z; t.i2 = v.w;
    return t;
}

__device__ __forceinline__ unsigned long long ex_pack(int d, int i) { return ((unsigned long long)(unsigned int)d << 32) | (unsigned int)i; }
constexpr unsigned long long EX_ABSENT = (256ull << 32) | 0xFFFFFFFFull; // (256, -1): sorts after every real candidate

// Tail of k_bf_top2<true>.  T = the block's top-2 of its database chunk for the thread's BF_QPT queries.
__device__ __forceinline__ void bf_exchange_tail(const ExchangeParams &X, const Top2 (&Tin)[BF_QPT], int nq, int q0, int nchunks)
{
    __shared__ int sLast;
    const int g = blockIdx.x;
    // (0) fold the chunk result into the group's accumulators at L2: acc[q] = {smallest, second smallest} packed
    //     (distance, global index) key.  64-bit atomicMin on key 1 returns what it displaced; the loser of that duel and
    //     the chunk's own second both go to key 2 -- min is commutative, so the arrival order of the blocks is irrelevant
    //     and the final pair is the two lexicographically smallest keys of the whole shard.
#pragma unroll
    for (int k = 0; k < BF_QPT; k++) {
        if (q0 + k >= nq) continue;
        unsigned long long *a = X.acc + 2 * (size_t)(q0 + k);
        const unsigned long long k1 = ex_pack(Tin[k].d1, Tin[k].i1), k2 = ex_pack(Tin[k].d2, Tin[k].i2);
        if (k1 == EX_ABSENT) continue;
        const unsigned long long old = atomicMin(a, k1);
        const unsigned long long loser = old > k1 ? old : k1;
        if (loser != EX_ABSENT) atomicMin(a + 1, loser);
        if (k2 != EX_ABSENT) atomicMin(a + 1, k2);
    }
    __threadfence(); // the atomics are performed before the arrival is counted
    __syncthreads();
    if (threadIdx.x == 0) {
        const int prev = atomicAdd(&X.groupCounter[g], 1);
        sLast = prev == nchunks - 1;
        if (sLast) X.groupCounter[g] = 0; // every block of the group has arrived: ready for the next step
    }
    __syncthreads();
    if (!sLast) return;
    __threadfence();
    // (1) the last block reads the shard result and re-arms the accumulators for the next step
    Top2 T[BF_QPT];
#pragma unroll
    for (int k = 0; k < BF_QPT; k++) {
        top2_init(T[k]);
        if (q0 + k >= nq) continue;
        ulonglong2 *a = reinterpret_cast<ulonglong2 *>(X.acc + 2 * (size_t)(q0 + k));
        const ulonglong2 v = __ldcg(a);
        T[k].d1 = (int)(v.x >> 32); T[k].i1 = (int)(unsigned int)v.x;
        T[k].d2 = (int)(v.y >> 32); T[k].i2 = (int)(unsigned int)v.y;
        *a = make_ulonglong2(EX_ABSENT, EX_ABSENT);
    }
    // (2) 16-byte stores into every rank's gather buffer (own included), slot [epoch parity][this rank][query]
    const int parity = X.epoch & 1;
    for (int p = 0; p < X.world; p++) {
        orbb200_top2 *dst = ex_gather(X.peer[p], X, parity, X.rank);
#pragma unroll
        for (int k = 0; k < BF_QPT; k++)
            if (q0 + k < nq) *reinterpret_cast<int4 *>(dst + q0 + k) = make_int4(T[k].d1, T[k].i1, T[k].d2, T[k].i2);
    }
    __syncthreads();
    // (3) publish: one release store per peer, after a system-scope fence that covers the whole block's stores
    if (threadIdx.x == 0) __threadfence_system();
    __syncthreads();
    if (threadIdx.x < X.world) st_release_sys(ex_flags(X.peer[threadIdx.x], X) + X.rank * X.maxGroups + g, X.epoch);
    // (4) wait for the same group of every rank (bounded: a missing peer becomes an error, not a hang)
    if (threadIdx.x < X.world) {
        const int *f = ex_flags(X.peer[X.rank], X) + threadIdx.x * X.maxGroups + g;
        const long long t0 = global_ns();
        while (ld_acquire_sys(f) < X.epoch) {
            if (global_ns() - t0 > X.timeoutNs) { atomicExch(X.status, 1); break; }
            __nanosleep(32);
        }
    }
    __syncthreads();
    // (5) fold the world's shard results in rank order
#pragma unroll
    for (int k = 0; k < BF_QPT; k++) top2_init(T[k]);
    for (int s = 0; s < X.world; s++) {
        const orbb200_top2 *src = ex_gather(X.peer[X.rank], X, parity, s);
#pragma unroll
        for (int k = 0; k < BF_QPT; k++)
            if (q0 + k < nq) T[k] = top2_merge(T[k], ex_load(src + q0 + k));
    }
#pragma unroll
    for (int k = 0; k < BF_QPT; k++)
        if (q0 + k < nq) *reinterpret_cast<int4 *>(X.out + q0 + k) = make_int4(T[k].d1, T[k].i1, T[k].d2, T[k].i2);
}

template <bool EXCHANGE>
__global__ void __launch_bounds__(BF_THREADS) k_bf_top2(const uint8_t *__restrict__ q, int nq, const uint8_t *__restrict__ db,
                                                          int ndb, int rowsPerChunk, int indexOffset,
                                                          orbb200_top2 *__restrict__ partial, const ExchangeParams X)
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
    if constexpr (EXCHANGE) {
        bf_exchange_tail(X, T, nq, q0, (int)gridDim.y);
    } else {
#pragma unroll
        for (int k = 0; k < BF_QPT; k++)
            if (q0 + k < nq) {
                orbb200_top2 o;
                o.best_dist = T[k].d1; o.best_idx = T[k].i1; o.second_dist = T[k].d2; o.second_idx = T[k].i2;
                partial[(size_t)blockIdx.y * nq + q0 + k] = o;
            }
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

// ---------------------------------------------------------------------------------------------------
// M8  SearchForTriangulation, resolved on the device (src/ORBmatcher.cc:757-815, CheckDistEpipolarLine :147-164).
// In this code base vbMatched2 is never set (:725 / :773), so the queries are independent: per query the reference keeps the
// candidate with the smallest distance <= TH_LOW among those that pass the epipole gate (:791-797) and the epipolar chi-square
// test, a LATER candidate replacing an earlier one at equal distance (`dist > bestDist` continues, :784).  One warp per query:
// lanes stride the CSR list in order with `<=`, the warp keeps (min distance, max position).  Float expressions in the
// reference's order with explicit round-to-nearest operations (no FMA contraction); 3.84 * sigma2 is a double product (:163).
// ---------------------------------------------------------------------------------------------------
struct TriangulationParams {
    float F[9];         // F12 row-major
    float ex, ey;       // epipole of camera 1 in image 2
    float sf[ORBB200_MAX_LEVELS], sig2[ORBB200_MAX_LEVELS];
    int thLow;
};

__global__ void k_triangulation_search(const uint8_t *__restrict__ q, const float2 *__restrict__ qxy, const uint8_t *__restrict__ qstereo, int nq,
                                       const uint8_t *__restrict__ t, const orbb200_keypoint *__restrict__ tkp, const uint8_t *__restrict__ tstereo,
                                       const int *__restrict__ offs, const int *__restrict__ idx, const __grid_constant__ TriangulationParams P,
                                       int *__restrict__ bestIdx2, int *__restrict__ bestDist)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= nq) return;
    const Desc Q = load_desc(q + (size_t)warp * 32);
    const float2 p1 = qxy[warp];
    const bool st1 = qstereo[warp] != 0;
    // epipolar line of kp1 in image 2: l = x1' F12 = [a b c] (:150-152)
    const float a = __fadd_rn(__fadd_rn(__fmul_rn(p1.x, P.F[0]), __fmul_rn(p1.y, P.F[3])), P.F[6]);
    const float b = __fadd_rn(__fadd_rn(__fmul_rn(p1.x, P.F[1]), __fmul_rn(p1.y, P.F[4])), P.F[7]);
    const float c = __fadd_rn(__fadd_rn(__fmul_rn(p1.x, P.F[2]), __fmul_rn(p1.y, P.F[5])), P.F[8]);
    const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
    const int c0 = offs[warp], c1 = offs[warp + 1];
    int bd = P.thLow, bpos = -1;
    for (int cc = c0 + lane; cc < c1; cc += 32) {
        const int i2 = __ldg(idx + cc);
        const int d = hamming(Q, load_desc(t + (size_t)i2 * 32));
        if (d > P.thLow || d > bd) continue;
        const orbb200_keypoint k2 = tkp[i2];
        if (!st1 && tstereo[i2] == 0) {
            const float dx = __fsub_rn(P.ex, k2.x), dy = __fsub_rn(P.ey, k2.y);
            if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, P.sf[k2.octave])) continue;
        }
        const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, k2.x), __fmul_rn(b, k2.y)), c);
        if (den == 0.f) continue;
        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
        if (!((double)dsqr < 3.84 * (double)P.sig2[k2.octave])) continue;
        bd = d;
        bpos = cc;
    }
    // smallest distance, then the latest position
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const int od = __shfl_xor_sync(0xffffffffu, bd, o), op = __shfl_xor_sync(0xffffffffu, bpos, o);
        if (op >= 0 && (bpos < 0 || od < bd || (od == bd && op > bpos))) { bd = od; bpos = op; }
    }
    if (lane == 0) {
        bestIdx2[warp] = bpos >= 0 ? idx[bpos] : -1;
        bestDist[warp] = bpos >= 0 ? bd : -1;
    }
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
