// orbb200_match.cu -- C-ABI of the matcher searches (include/orbb200.h) over match_kernels.cuh / stereo_kernel.cuh.
// Mirrors the arithmetic of ORB_SLAM2::ORBmatcher (reference src/ORBmatcher.cc) on POD arrays; the
// MapPoint / KeyFrame bookkeeping of the Search* methods stays in the C++ facade (include/orbb200/ORBmatcher.h).
#include <atomic>
#include <mutex>
#include <algorithm>
#include <vector>

#include "extractor_host.cuh"
#include "distinct_kernel.cuh"
#include "match_kernels.cuh"
#include "stereo_kernel.cuh"
#include "window_kernels.cuh"

namespace orbb200 {

// grow-only per-thread device scratch, so the host-buffer entry points do not cudaMalloc on every call
struct Scratch {
    int device = -1;
    cudaStream_t stream = nullptr;
    void *buf[12] = {};
    size_t cap[12] = {};
    // the per-chunk partial top-2 buffer (slot 5) is shared by every brute-force search of this host thread, also by searches on
    // different caller streams: the event orders each search behind the previous user of the buffer
    cudaEvent_t partialDone = nullptr;
    bool partialUsed = false;
    // frame-sized searches: all inputs of a call travel in ONE copy from a pinned staging block, all outputs come back in one
    // (or are written by the kernel straight into the mapped pinned block), so a call costs one synchronisation
    unsigned char *hIn = nullptr, *dIn = nullptr, *hOut = nullptr, *dOut = nullptr, *hOutDev = nullptr;
    size_t hInCap = 0, dInCap = 0, hOutCap = 0, dOutCap = 0;
};
static thread_local Scratch g_scratch;

// bytes this translation unit's entry points copied host -> device (descriptor / keypoint / query uploads of the searches): lets a
// test prove that a device-resident frame is searched without re-uploading its descriptors (orbb200_h2d_bytes_total)
static std::atomic<unsigned long long> g_h2dBytes(0);
static inline void count_h2d(size_t bytes) { g_h2dBytes.fetch_add((unsigned long long)bytes); }

static int scratch_bind(int device)
{
    if (device < 0) ORB_CUDA_OK(cudaGetDevice(&device));
    ORB_CUDA_OK(cudaSetDevice(device));
    Scratch &s = g_scratch;
    if (s.device != device) {
        for (int i = 0; i < 12; i++) { s.buf[i] = nullptr; s.cap[i] = 0; } // buffers of another device are leaked on purpose
        s.stream = nullptr;
        s.partialDone = nullptr;
        s.partialUsed = false;
        s.hIn = s.dIn = s.hOut = s.dOut = s.hOutDev = nullptr;
        s.hInCap = s.dInCap = s.hOutCap = s.dOutCap = 0;
        s.device = device;
    }
    if (!s.stream) ORB_CUDA_OK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    return ORBB200_OK;
}

static int scratch_get(int slot, size_t bytes, void **out)
{
    Scratch &s = g_scratch;
    if (s.cap[slot] < bytes) {
        if (s.buf[slot]) { ORB_CUDA_OK(cudaStreamSynchronize(s.stream)); cudaFree(s.buf[slot]); s.buf[slot] = nullptr; s.cap[slot] = 0; }
        const size_t nb = std::max<size_t>(align_up(bytes, 256), 4096);
        ORB_CUDA_OK(cudaMalloc(&s.buf[slot], nb));
        s.cap[slot] = nb;
    }
    *out = s.buf[slot];
    return ORBB200_OK;
}

static inline size_t pad256(size_t b) { return (b + 255) & ~(size_t)255; }

// Packed transfers of the frame-sized searches.  Usage: sizes first (in(bytes) / out(bytes) return offsets), reserve(), then
// put() the inputs, upload(), launch on dIn + offset / dOut + offset, download(), synchronise, read hOut + offset.
struct Pack {
    size_t inBytes = 0, outBytes = 0;
    size_t in(size_t bytes) { const size_t o = inBytes; inBytes += pad256(bytes); return o; }
    size_t out(size_t bytes) { const size_t o = outBytes; outBytes += pad256(bytes); return o; }
    int reserve()
    {
        Scratch &s = g_scratch;
        if (s.hInCap < inBytes || s.dInCap < inBytes || s.hOutCap < outBytes || s.dOutCap < outBytes) ORB_CUDA_OK(cudaStreamSynchronize(s.stream));
        if (s.hInCap < inBytes) {
            if (s.hIn) cudaFreeHost(s.hIn);
            s.hIn = nullptr; s.hInCap = 0;
            const size_t nb = std::max<size_t>(pad256(inBytes + inBytes / 2), 1 << 16);
            ORB_CUDA_OK(cudaHostAlloc((void **)&s.hIn, nb, cudaHostAllocDefault));
            s.hInCap = nb;
        }
        if (s.dInCap < inBytes) {
            if (s.dIn) cudaFree(s.dIn);
            s.dIn = nullptr; s.dInCap = 0;
            const size_t nb = std::max<size_t>(pad256(inBytes + inBytes / 2), 1 << 16);
            ORB_CUDA_OK(cudaMalloc((void **)&s.dIn, nb));
            s.dInCap = nb;
        }
        if (s.hOutCap < outBytes) {
            if (s.hOut) cudaFreeHost(s.hOut);
            s.hOut = nullptr; s.hOutCap = 0; s.hOutDev = nullptr;
            const size_t nb = std::max<size_t>(pad256(outBytes + outBytes / 2), 1 << 16);
            ORB_CUDA_OK(cudaHostAlloc((void **)&s.hOut, nb, cudaHostAllocMapped));
            ORB_CUDA_OK(cudaHostGetDevicePointer((void **)&s.hOutDev, s.hOut, 0));
            s.hOutCap = nb;
        }
        if (s.dOutCap < outBytes) {
            if (s.dOut) cudaFree(s.dOut);
            s.dOut = nullptr; s.dOutCap = 0;
            const size_t nb = std::max<size_t>(pad256(outBytes + outBytes / 2), 1 << 16);
            ORB_CUDA_OK(cudaMalloc((void **)&s.dOut, nb));
            s.dOutCap = nb;
        }
        return ORBB200_OK;
    }
    template <class T> const T *put(size_t off, const void *src, size_t bytes) const
    {
        if (bytes) memcpy(g_scratch.hIn + off, src, bytes);
        return reinterpret_cast<const T *>(g_scratch.dIn + off);
    }
    template <class T> T *dout(size_t off) const { return reinterpret_cast<T *>(g_scratch.dOut + off); }
    template <class T> T *mapped(size_t off) const { return reinterpret_cast<T *>(g_scratch.hOutDev + off); } // device view of hOut
    template <class T> const T *hout(size_t off) const { return reinterpret_cast<const T *>(g_scratch.hOut + off); }
    int upload(size_t bytes) const // bytes: the prefix of the input block in use
    {
        ORB_CUDA_OK(cudaMemcpyAsync(g_scratch.dIn, g_scratch.hIn, bytes, cudaMemcpyHostToDevice, g_scratch.stream));
        return ORBB200_OK;
    }
    int download(size_t bytes) const
    {
        ORB_CUDA_OK(cudaMemcpyAsync(g_scratch.hOut, g_scratch.dOut, bytes, cudaMemcpyDeviceToHost, g_scratch.stream));
        return ORBB200_OK;
    }
};

// chunking of the database so that the grid is about 4 blocks per SM on 148 SMs
static void bf_plan(int nq, int ndb, int *groups, int *chunks, int *rowsPerChunk)
{
    *groups = div_up(nq, BF_THREADS * BF_QPT);
    int c = std::max(1, (148 * 4) / *groups);
    c = std::min(c, div_up(ndb, BF_TILE));
    int rows = div_up(ndb, c);
    rows = div_up(rows, BF_TILE) * BF_TILE;
    *rowsPerChunk = rows;
    *chunks = div_up(ndb, rows);
}

static int bf_launch(const uint8_t *q_d, int nq, const uint8_t *db_d, int ndb, orbb200_top2 *out_d, int indexOffset,
                     cudaStream_t st)
{
    int groups, chunks, rows;
    bf_plan(nq, ndb, &groups, &chunks, &rows);
    void *partial;
    Scratch &sc = g_scratch;
    if (!sc.partialDone) ORB_CUDA_OK(cudaEventCreateWithFlags(&sc.partialDone, cudaEventDisableTiming));
    const size_t need = sizeof(orbb200_top2) * (size_t)chunks * nq;
    if (sc.partialUsed && sc.cap[5] < need) ORB_CUDA_OK(cudaEventSynchronize(sc.partialDone)); // regrow: the last search may still read it
    int rc = scratch_get(5, need, &partial);
    if (rc) return rc;
    if (sc.partialUsed) ORB_CUDA_OK(cudaStreamWaitEvent(st, sc.partialDone, 0)); // previous search (any stream) is done with the buffer
    k_bf_top2<false><<<dim3(groups, chunks), BF_THREADS, 0, st>>>(q_d, nq, db_d, ndb, rows, indexOffset, (orbb200_top2 *)partial,
                                                                  ExchangeParams());
    k_merge_top2<<<div_up(nq * 32, 256), 256, 0, st>>>((const orbb200_top2 *)partial, chunks, nq, out_d);
    ORB_CUDA_OK(cudaGetLastError());
    ORB_CUDA_OK(cudaEventRecord(sc.partialDone, st));
    sc.partialUsed = true;
    return ORBB200_OK;
}

// ---- orbb200_match_bruteforce_top2_device_noalloc: the per-chunk partial results live in a static device array --------------------
// Every GPU that loads the library holds these 8 MiB until exit.  All calls on a device share the array.
constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)8 << 20;
constexpr size_t NOALLOC_MAX_PARTIALS = NOALLOC_SCRATCH_BYTES / sizeof(orbb200_top2); // 524 288 (chunk, query) records
static __device__ __align__(256) unsigned char g_noallocScratch[NOALLOC_SCRATCH_BYTES];
static std::mutex g_noallocMu;
static void *g_noallocPtr[64] = {};

// device address of g_noallocScratch on the CURRENT device: in host code the array's name is only a stand-in for the symbol, so
// the address is asked for once per device, with that device current, and cached
static int noalloc_scratch_ptr(orbb200_top2 **out)
{
    int device = -1;
    ORB_CUDA_OK(cudaGetDevice(&device));
    ORB_ARG_OK(device >= 0 && device < 64, "device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_noallocMu);
    if (!g_noallocPtr[device]) {
        void *p = nullptr;
        ORB_CUDA_OK(cudaGetSymbolAddress(&p, g_noallocScratch));
        ORB_ARG_OK(p != nullptr && ((size_t)p & 255) == 0, "cudaGetSymbolAddress returned an unusable address");
        g_noallocPtr[device] = p;
    }
    *out = (orbb200_top2 *)g_noallocPtr[device];
    return ORBB200_OK;
}

} // namespace orbb200

// config 5 on N GPUs: one exchange region per rank (gather slots + flags), mapped into every peer with CUDA IPC
struct orbb200_exchange {
    int rank = 0, world = 1, nqMax = 0, maxGroups = 0, device = 0, epoch = 0;
    bool connected = false;
    unsigned char *region = nullptr; // own region
    size_t regionBytes = 0;
    int *groupCounter = nullptr, *status = nullptr;
    unsigned long long *acc = nullptr;
    unsigned char *peer[orbb200::EX_MAX_WORLD] = {};
    cudaStream_t stream = nullptr;
};

#include "stereo_host.inc"

using namespace orbb200;

extern "C" {

int orbb200_descriptor_distance(const uint8_t *a, const uint8_t *b, int n, int *dist, int device)
{
    if (n <= 0) return ORBB200_OK;
    ORB_ARG_OK(a && b && dist, "null pointer");
    int rc = scratch_bind(device);
    if (rc) return rc;
    void *da, *db, *dd;
    if ((rc = scratch_get(0, (size_t)n * 32, &da)) || (rc = scratch_get(1, (size_t)n * 32, &db)) ||
        (rc = scratch_get(2, (size_t)n * 4, &dd)))
        return rc;
    cudaStream_t st = g_scratch.stream;
    ORB_CUDA_OK(cudaMemcpyAsync(da, a, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    ORB_CUDA_OK(cudaMemcpyAsync(db, b, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    k_pair_distance<<<div_up(n, 256), 256, 0, st>>>((const uint8_t *)da, (const uint8_t *)db, n, (int *)dd);
    ORB_CUDA_OK(cudaMemcpyAsync(dist, dd, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    return ORBB200_OK;
}

int orbb200_match_bruteforce_top2_device(const uint8_t *q_d, int nq, const uint8_t *db_d, int ndb, orbb200_top2 *out_d,
                                         int index_offset, void *stream)
{
    ORB_ARG_OK(q_d && db_d && out_d && nq > 0 && ndb > 0, "bad arguments");
    ORB_ARG_OK(((size_t)q_d & 15) == 0 && ((size_t)db_d & 15) == 0, "descriptor arrays must be 16-byte aligned");
    int rc = scratch_bind(-1);
    if (rc) return rc;
    return bf_launch(q_d, nq, db_d, ndb, out_d, index_offset, stream ? (cudaStream_t)stream : g_scratch.stream);
}

// No counterpart in the reference; added for callers that cannot allocate (CUDA graph capture,
// serving loops, harnesses that forbid cudaMalloc).  Same kernels and the same result as orbb200_match_bruteforce_top2_device; the
// per-chunk partial top-2 records go to the static device array above instead of a cudaMalloc'ed scratch buffer.  Nothing here
// allocates: no cudaMalloc / cudaFree, no stream or event creation (stream == NULL is the default stream of the current device).
// The array keeps the last call's contents; nothing in it has to start at zero because k_bf_top2 writes every (chunk, query) record
// that k_merge_top2 of the same call reads.  Calls that overlap on the GPU (different streams, different host threads) race on it.
int orbb200_match_bruteforce_top2_device_noalloc(const uint8_t *q_d, int nq, const uint8_t *db_d, int ndb, orbb200_top2 *out_d,
                                                 int index_offset, void *stream)
{
    ORB_ARG_OK(q_d && db_d && out_d && nq > 0 && ndb > 0, "bad arguments");
    ORB_ARG_OK(((size_t)q_d & 15) == 0 && ((size_t)db_d & 15) == 0, "descriptor arrays must be 16-byte aligned");
    int groups, chunks, rows;
    bf_plan(nq, ndb, &groups, &chunks, &rows);
    if ((size_t)chunks * (size_t)nq > NOALLOC_MAX_PARTIALS) {
        set_error("orbb200_match_bruteforce_top2_device_noalloc: %d queries x %d database chunks = %zu partial records exceed the static "
                  "scratch of %zu records (%zu bytes); use orbb200_match_bruteforce_top2_device or split the queries",
                  nq, chunks, (size_t)chunks * (size_t)nq, NOALLOC_MAX_PARTIALS, NOALLOC_SCRATCH_BYTES);
        return ORBB200_ERR_CAPACITY;
    }
    orbb200_top2 *partial = nullptr;
    int rc = noalloc_scratch_ptr(&partial);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    k_bf_top2<false><<<dim3(groups, chunks), BF_THREADS, 0, st>>>(q_d, nq, db_d, ndb, rows, index_offset, partial, ExchangeParams());
    k_merge_top2<<<div_up(nq * 32, 256), 256, 0, st>>>(partial, chunks, nq, out_d);
    ORB_CUDA_OK(cudaGetLastError());
    return ORBB200_OK;
}

int orbb200_match_bruteforce_top2(const uint8_t *q, int nq, const uint8_t *db, int ndb, orbb200_top2 *out, int device)
{
    if (nq <= 0) return ORBB200_OK;
    ORB_ARG_OK(q && out && (db || ndb == 0), "null pointer");
    if (ndb <= 0) {
        for (int i = 0; i < nq; i++) out[i] = orbb200_top2{256, -1, 256, -1};
        return ORBB200_OK;
    }
    int rc = scratch_bind(device);
    if (rc) return rc;
    void *dq, *ddb, *dout;
    if ((rc = scratch_get(0, (size_t)nq * 32, &dq)) || (rc = scratch_get(1, (size_t)ndb * 32, &ddb)) ||
        (rc = scratch_get(2, (size_t)nq * sizeof(orbb200_top2), &dout)))
        return rc;
    cudaStream_t st = g_scratch.stream;
    ORB_CUDA_OK(cudaMemcpyAsync(dq, q, (size_t)nq * 32, cudaMemcpyHostToDevice, st));
    ORB_CUDA_OK(cudaMemcpyAsync(ddb, db, (size_t)ndb * 32, cudaMemcpyHostToDevice, st));
    rc = bf_launch((const uint8_t *)dq, nq, (const uint8_t *)ddb, ndb, (orbb200_top2 *)dout, 0, st);
    if (rc) return rc;
    ORB_CUDA_OK(cudaMemcpyAsync(out, dout, (size_t)nq * sizeof(orbb200_top2), cudaMemcpyDeviceToHost, st));
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    return ORBB200_OK;
}

/* CSR candidate lists come from the caller's feature vectors: a bad index would read outside the target matrix on the device,
   so the lists are validated on the host (O(total), negligible beside the copy) */
static int csr_valid(const int *offs, int nq, const int *idx, int nt)
{
    if (offs[0] != 0) return 0;
    for (int i = 0; i < nq; i++)
        if (offs[i + 1] < offs[i]) return 0;
    const int total = offs[nq];
    for (int c = 0; c < total; c++)
        if ((unsigned)idx[c] >= (unsigned)nt) return 0;
    return 1;
}

int orbb200_match_candidates(const uint8_t *q, int nq, const uint8_t *t, int nt, const int *cand_offsets, const int *cand_idx,
                             orbb200_top2 *out, int device)
{
    if (nq <= 0) return ORBB200_OK;
    ORB_ARG_OK(q && cand_offsets && out, "null pointer");
    const int total = cand_offsets[nq];
    ORB_ARG_OK(total == 0 || (t && cand_idx && nt > 0), "candidates without targets");
    ORB_ARG_OK(total >= 0 && (total == 0 ? cand_offsets[0] == 0 : csr_valid(cand_offsets, nq, cand_idx, nt)), "candidate lists: offsets not ascending from 0 or index outside [0, nt)");
    int rc = scratch_bind(device);
    if (rc) return rc;
    Pack K;
    const size_t oq = K.in((size_t)nq * 32), ot = K.in((size_t)std::max(nt, 1) * 32), oo = K.in((size_t)(nq + 1) * 4), oi = K.in((size_t)std::max(total, 1) * 4);
    const size_t ro = K.out((size_t)nq * sizeof(orbb200_top2));
    if ((rc = K.reserve())) return rc;
    cudaStream_t st = g_scratch.stream;
    const uint8_t *dq = K.put<uint8_t>(oq, q, (size_t)nq * 32), *dt = K.put<uint8_t>(ot, t, (size_t)std::max(nt, 0) * 32);
    const int *doffs = K.put<int>(oo, cand_offsets, (size_t)(nq + 1) * 4), *didx = K.put<int>(oi, cand_idx, (size_t)total * 4);
    if ((rc = K.upload(K.inBytes))) return rc;
    count_h2d(K.inBytes);
    k_match_candidates<<<div_up(nq * 32, 256), 256, 0, st>>>(dq, nq, dt, doffs, didx, K.dout<orbb200_top2>(ro));
    if ((rc = K.download(K.outBytes))) return rc;
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    memcpy(out, K.hout<orbb200_top2>(ro), (size_t)nq * sizeof(orbb200_top2));
    return ORBB200_OK;
}

int orbb200_candidate_distances(const uint8_t *q, int nq, const uint8_t *t, int nt, const int *cand_offsets, const int *cand_idx,
                                uint16_t *dist, int device)
{
    if (nq <= 0) return ORBB200_OK;
    ORB_ARG_OK(q && cand_offsets, "null pointer");
    const int total = cand_offsets[nq];
    if (total <= 0) return ORBB200_OK;
    ORB_ARG_OK(t && cand_idx && dist && nt > 0, "candidates without targets");
    ORB_ARG_OK(csr_valid(cand_offsets, nq, cand_idx, nt), "candidate lists: offsets not ascending from 0 or index outside [0, nt)");
    int rc = scratch_bind(device);
    if (rc) return rc;
    Pack K;
    const size_t oq = K.in((size_t)nq * 32), ot = K.in((size_t)nt * 32), oo = K.in((size_t)(nq + 1) * 4), oi = K.in((size_t)total * 4);
    const size_t ro = K.out((size_t)total * 2);
    if ((rc = K.reserve())) return rc;
    cudaStream_t st = g_scratch.stream;
    const uint8_t *dq = K.put<uint8_t>(oq, q, (size_t)nq * 32), *dt = K.put<uint8_t>(ot, t, (size_t)nt * 32);
    const int *doffs = K.put<int>(oo, cand_offsets, (size_t)(nq + 1) * 4), *didx = K.put<int>(oi, cand_idx, (size_t)total * 4);
    if ((rc = K.upload(K.inBytes))) return rc;
    count_h2d(K.inBytes);
    k_candidate_distances<<<div_up(nq * 32, 256), 256, 0, st>>>(dq, nq, dt, doffs, didx, K.dout<unsigned short>(ro));
    if ((rc = K.download(K.outBytes))) return rc;
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    memcpy(dist, K.hout<uint16_t>(ro), (size_t)total * 2);
    return ORBB200_OK;
}

int orbb200_search_triangulation(const uint8_t *q, const float *qxy, const uint8_t *qstereo, int nq, const uint8_t *t,
                                 const orbb200_keypoint *tkp, const uint8_t *tstereo, int nt, const int *cand_offsets, const int *cand_idx,
                                 const float *F12, float ex, float ey, const float *scale_factors, const float *level_sigma2, int nlevels,
                                 int th_low, int *best_idx2, int *best_dist, int device)
{
    if (nq <= 0) return ORBB200_OK;
    ORB_ARG_OK(q && qxy && qstereo && cand_offsets && best_idx2 && F12 && scale_factors && level_sigma2, "null pointer");
    ORB_ARG_OK(nlevels > 0 && nlevels <= ORBB200_MAX_LEVELS, "bad level count");
    const int total = cand_offsets[nq];
    if (total <= 0) {
        ORB_ARG_OK(cand_offsets[0] == 0 && total == 0, "candidate offsets must start at 0");
        for (int i = 0; i < nq; i++) { best_idx2[i] = -1; if (best_dist) best_dist[i] = -1; }
        return ORBB200_OK;
    }
    ORB_ARG_OK(t && tkp && tstereo && cand_idx && nt > 0, "candidates without targets");
    ORB_ARG_OK(csr_valid(cand_offsets, nq, cand_idx, nt), "candidate lists: offsets not ascending from 0 or index outside [0, nt)");
    for (int i = 0; i < nt; i++) ORB_ARG_OK(tkp[i].octave >= 0 && tkp[i].octave < nlevels, "target keypoint octave outside the level tables");
    int rc = scratch_bind(device);
    if (rc) return rc;
    Pack K;
    const size_t oq = K.in((size_t)nq * 32), ot = K.in((size_t)nt * 32), ok = K.in((size_t)nt * sizeof(orbb200_keypoint)), ots = K.in((size_t)nt),
                 oxy = K.in((size_t)nq * 8), oqs = K.in((size_t)nq), oo = K.in((size_t)(nq + 1) * 4), oi = K.in((size_t)total * 4);
    const size_t rbi = K.out((size_t)nq * 4), rbd = K.out((size_t)nq * 4);
    if ((rc = K.reserve())) return rc;
    cudaStream_t st = g_scratch.stream;
    const uint8_t *dq = K.put<uint8_t>(oq, q, (size_t)nq * 32), *dt = K.put<uint8_t>(ot, t, (size_t)nt * 32);
    const orbb200_keypoint *dkp = K.put<orbb200_keypoint>(ok, tkp, (size_t)nt * sizeof(orbb200_keypoint));
    const uint8_t *dts = K.put<uint8_t>(ots, tstereo, (size_t)nt), *dqs = K.put<uint8_t>(oqs, qstereo, (size_t)nq);
    const float2 *dqxy = K.put<float2>(oxy, qxy, (size_t)nq * 8);
    const int *doffs = K.put<int>(oo, cand_offsets, (size_t)(nq + 1) * 4), *didx = K.put<int>(oi, cand_idx, (size_t)total * 4);
    if ((rc = K.upload(K.inBytes))) return rc;
    count_h2d(K.inBytes);
    TriangulationParams P;
    memset(&P, 0, sizeof(P));
    for (int i = 0; i < 9; i++) P.F[i] = F12[i];
    P.ex = ex; P.ey = ey; P.thLow = th_low;
    for (int l = 0; l < nlevels; l++) { P.sf[l] = scale_factors[l]; P.sig2[l] = level_sigma2[l]; }
    k_triangulation_search<<<div_up(nq * 32, 256), 256, 0, st>>>(dq, dqxy, dqs, nq, dt, dkp, dts, doffs, didx, P, K.dout<int>(rbi), K.dout<int>(rbd));
    if ((rc = K.download(K.outBytes))) return rc;
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    memcpy(best_idx2, K.hout<int>(rbi), (size_t)nq * 4);
    if (best_dist) memcpy(best_dist, K.hout<int>(rbd), (size_t)nq * 4);
    return ORBB200_OK;
}

// host queries / host results; the targets (keypoints + descriptors) either come from the host (uploaded here) or are already
// resident on the device (orbb200_window_search_resident: a frame that keeps its extraction outputs in HBM)
static int window_search_impl(const orbb200_keypoint *tkp, const uint8_t *tdesc, bool resident, int nt, float minX, float minY, float invW,
                              float invH, int cols, int rows, const float *windows, const uint8_t *qdesc, int nq, int *cand_offsets,
                              int *cand_idx, uint16_t *dist, int cand_cap, int device)
{
    ORB_ARG_OK(cand_offsets != nullptr, "cand_offsets is null");
    for (int i = 0; i <= std::max(nq, 0); i++) cand_offsets[i] = 0;
    if (nq <= 0 || nt <= 0) return ORBB200_OK;
    ORB_ARG_OK(tkp && tdesc && windows && qdesc && cols > 0 && rows > 0 && cols * rows <= 11000, "bad arguments");
    int rc = scratch_bind(device);
    if (rc) return rc;
    const int nc = cols * rows;
    void *dcs, *dci, *dcnt;
    if ((rc = scratch_get(8, (size_t)(nc + 1) * 4, &dcs)) || (rc = scratch_get(9, (size_t)nt * 4, &dci)) || (rc = scratch_get(3, (size_t)(nq + 1) * 4, &dcnt)))
        return rc;
    const int cap = std::max(cand_cap, 0);
    ORB_ARG_OK(cap == 0 || (cand_idx && dist), "null output");
    Pack K;
    const size_t ow = K.in((size_t)nq * 20), oqd = K.in((size_t)nq * 32);
    const size_t inQueries = K.inBytes;
    const size_t okp = K.in(resident ? 0 : (size_t)nt * sizeof(orbb200_keypoint)), otd = K.in(resident ? 0 : (size_t)nt * 32);
    // results are written by the kernels straight into the mapped pinned block: offsets, then candidates and distances up to cap
    const size_t ro = K.out((size_t)(nq + 1) * 4), rc_ = K.out((size_t)cap * 4), rd = K.out((size_t)cap * 2);
    if ((rc = K.reserve())) return rc;
    cudaStream_t st = g_scratch.stream;
    const float *dwin = K.put<float>(ow, windows, (size_t)nq * 20);
    const uint8_t *dqd = K.put<uint8_t>(oqd, qdesc, (size_t)nq * 32);
    const orbb200_keypoint *dkp = tkp;
    const uint8_t *dtd = tdesc;
    if (resident) {
        ORB_ARG_OK(((size_t)tdesc & 15) == 0, "resident descriptors must be 16-byte aligned");
        if ((rc = K.upload(inQueries))) return rc;
        count_h2d((size_t)nq * 52);
    } else {
        dkp = K.put<orbb200_keypoint>(okp, tkp, (size_t)nt * sizeof(orbb200_keypoint));
        dtd = K.put<uint8_t>(otd, tdesc, (size_t)nt * 32);
        if ((rc = K.upload(K.inBytes))) return rc;
        count_h2d((size_t)nt * (sizeof(orbb200_keypoint) + 32));
        count_h2d((size_t)nq * 52);
    }
    WindowParams W;
    memset(&W, 0, sizeof(W));
    W.G.kp = dkp; W.G.n = nt;
    W.G.minX = minX; W.G.minY = minY; W.G.invW = invW; W.G.invH = invH; W.G.cols = cols; W.G.rows = rows;
    W.G.cellStart = (int *)dcs; W.G.cellItems = (int *)dci;
    W.win = dwin; W.qdesc = dqd; W.tdesc = dtd; W.nq = nq;
    W.counts = (int *)dcnt;
    W.candIdx = K.mapped<int>(rc_); W.dist = K.mapped<unsigned short>(rd); W.cap = cap;
    {
        int sortLen;
        size_t gsmBytes = grid_build_smem(nc, nt, &sortLen);
        if (gsmBytes > 48 * 1024) { sortLen = 0; gsmBytes = (size_t)(nc + 1) * 4; } // no room for the sort keys: ordered fill by one warp
        k_grid_build<<<1, GB_THREADS, gsmBytes, st>>>(W.G, sortLen);
    }
    k_window_count<<<div_up(nq * 32, 256), 256, 0, st>>>(W);
    k_scan_counts<<<1, 1024, 0, st>>>((int *)dcnt, nq, K.mapped<int>(ro));
    k_window_fill<<<div_up(nq * 32, 256), 256, 0, st>>>(W); // positions at or beyond cap are dropped
    ORB_CUDA_OK(cudaGetLastError());
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    memcpy(cand_offsets, K.hout<int>(ro), (size_t)(nq + 1) * 4);
    const int total = cand_offsets[nq];
    if (total > cand_cap) {
        set_error("orbb200_window_search: %d candidates, capacity is %d", total, cand_cap);
        return ORBB200_ERR_CAPACITY; // cand_offsets is valid: the caller can size the buffers and call again
    }
    if (total == 0) return ORBB200_OK;
    memcpy(cand_idx, K.hout<int>(rc_), (size_t)total * 4);
    memcpy(dist, K.hout<uint16_t>(rd), (size_t)total * 2);
    return ORBB200_OK;
}

int orbb200_window_search(const orbb200_keypoint *tkp, const uint8_t *tdesc, int nt, float minX, float minY, float invW, float invH,
                          int cols, int rows, const float *windows, const uint8_t *qdesc, int nq, int *cand_offsets, int *cand_idx,
                          uint16_t *dist, int cand_cap, int device)
{
    return window_search_impl(tkp, tdesc, false, nt, minX, minY, invW, invH, cols, rows, windows, qdesc, nq, cand_offsets, cand_idx, dist, cand_cap,
                              device);
}

int orbb200_window_search_resident(const orbb200_keypoint *tkp_d, const uint8_t *tdesc_d, int nt, float minX, float minY, float invW,
                                   float invH, int cols, int rows, const float *windows, const uint8_t *qdesc, int nq, int *cand_offsets,
                                   int *cand_idx, uint16_t *dist, int cand_cap, int device)
{
    return window_search_impl(tkp_d, tdesc_d, true, nt, minX, minY, invW, invH, cols, rows, windows, qdesc, nq, cand_offsets, cand_idx, dist,
                              cand_cap, device);
}

/* ---- plain device-memory helpers for host objects that keep data resident (include/orbb200/DeviceFrame.h) ---- */
int orbb200_device_alloc(void **p, size_t bytes, int device)
{
    ORB_ARG_OK(p != nullptr && bytes > 0, "bad arguments");
    *p = nullptr;
    int rc = scratch_bind(device);
    if (rc) return rc;
    ORB_CUDA_OK(cudaMalloc(p, bytes));
    return ORBB200_OK;
}
int orbb200_device_free(void *p)
{
    if (p) ORB_CUDA_OK(cudaFree(p));
    return ORBB200_OK;
}
int orbb200_copy_to_device(void *dst_d, const void *src, size_t bytes)
{
    if (bytes == 0) return ORBB200_OK;
    ORB_ARG_OK(dst_d && src, "null pointer");
    ORB_CUDA_OK(cudaMemcpy(dst_d, src, bytes, cudaMemcpyHostToDevice));
    count_h2d(bytes);
    return ORBB200_OK;
}
int orbb200_copy_to_host(void *dst, const void *src_d, size_t bytes)
{
    if (bytes == 0) return ORBB200_OK;
    ORB_ARG_OK(dst && src_d, "null pointer");
    ORB_CUDA_OK(cudaMemcpy(dst, src_d, bytes, cudaMemcpyDeviceToHost));
    return ORBB200_OK;
}
int orbb200_copy_on_device(void *dst_d, const void *src_d, size_t bytes)
{
    if (bytes == 0) return ORBB200_OK;
    ORB_ARG_OK(dst_d && src_d, "null pointer");
    ORB_CUDA_OK(cudaMemcpy(dst_d, src_d, bytes, cudaMemcpyDeviceToDevice));
    return ORBB200_OK;
}
unsigned long long orbb200_h2d_bytes_total(void) { return g_h2dBytes.load(); }
int orbb200_device_sync(int device)
{
    int rc = scratch_bind(device);
    if (rc) return rc;
    ORB_CUDA_OK(cudaDeviceSynchronize());
    return ORBB200_OK;
}

int orbb200_window_search_device(const orbb200_keypoint *tkp_d, const uint8_t *tdesc_d, int nt, const int *nt_d, float minX, float minY,
                                 float invW, float invH, int cols, int rows, const float *windows_d, const uint8_t *qdesc_d, int nq,
                                 int *cand_offsets_d, int *cand_idx_d, uint16_t *dist_d, int cand_cap, int *workspace_d, void *stream)
{
    ORB_ARG_OK(cand_offsets_d != nullptr && stream != nullptr, "cand_offsets_d / stream is null");
    cudaStream_t st = (cudaStream_t)stream;
    if (nq <= 0 || nt <= 0) {
        ORB_CUDA_OK(cudaMemsetAsync(cand_offsets_d, 0, sizeof(int) * (size_t)(std::max(nq, 0) + 1), st));
        return ORBB200_OK;
    }
    ORB_ARG_OK(tkp_d && tdesc_d && windows_d && qdesc_d && workspace_d && cols > 0 && rows > 0 && cols * rows <= 11000, "bad arguments");
    ORB_ARG_OK(cand_cap >= 0 && (cand_cap == 0 || (cand_idx_d && dist_d)), "null candidate outputs");
    ORB_ARG_OK((((size_t)tdesc_d | (size_t)qdesc_d) & 15) == 0, "descriptor arrays must be 16-byte aligned");
    const int nc = cols * rows;
    WindowParams W;
    memset(&W, 0, sizeof(W));
    W.G.kp = tkp_d; W.G.n = nt; W.G.nPtr = nt_d;
    W.G.minX = minX; W.G.minY = minY; W.G.invW = invW; W.G.invH = invH; W.G.cols = cols; W.G.rows = rows;
    W.G.cellStart = workspace_d; W.G.cellItems = workspace_d + nc + 1;
    W.win = windows_d; W.qdesc = qdesc_d; W.tdesc = tdesc_d; W.nq = nq;
    W.counts = cand_offsets_d; W.candIdx = cand_idx_d; W.dist = dist_d; W.cap = cand_cap;
    {
        int sortLen;
        size_t gsmBytes = grid_build_smem(nc, nt, &sortLen);
        if (gsmBytes > 48 * 1024) { sortLen = 0; gsmBytes = (size_t)(nc + 1) * 4; } // no room for the sort keys: ordered fill by one warp
        k_grid_build<<<1, GB_THREADS, gsmBytes, st>>>(W.G, sortLen);
    }
    k_window_count<<<div_up(nq * 32, 256), 256, 0, st>>>(W);
    k_scan_counts<<<1, 1024, 0, st>>>(cand_offsets_d, nq, nullptr);
    if (cand_cap > 0) k_window_fill<<<div_up(nq * 32, 256), 256, 0, st>>>(W);
    ORB_CUDA_OK(cudaGetLastError());
    return ORBB200_OK;
}

int orbb200_merge_top2_device(const orbb200_top2 *parts_d, int nshards, int nq, orbb200_top2 *out_d, void *stream)
{
    ORB_ARG_OK(parts_d && out_d && nshards > 0 && nq > 0, "bad arguments");
    int rc = scratch_bind(-1);
    if (rc) return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : g_scratch.stream;
    k_merge_top2<<<div_up(nq * 32, 256), 256, 0, st>>>(parts_d, nshards, nq, out_d);
    ORB_CUDA_OK(cudaGetLastError());
    return ORBB200_OK;
}

int orbb200_merge_top2(const orbb200_top2 *parts, int nshards, int nq, orbb200_top2 *out, int device)
{
    if (nq <= 0) return ORBB200_OK;
    ORB_ARG_OK(parts && out && nshards > 0, "bad arguments");
    int rc = scratch_bind(device);
    if (rc) return rc;
    void *dp, *dout;
    const size_t nb = sizeof(orbb200_top2) * (size_t)nshards * nq;
    if ((rc = scratch_get(0, nb, &dp)) || (rc = scratch_get(2, sizeof(orbb200_top2) * (size_t)nq, &dout))) return rc;
    cudaStream_t st = g_scratch.stream;
    ORB_CUDA_OK(cudaMemcpyAsync(dp, parts, nb, cudaMemcpyHostToDevice, st));
    k_merge_top2<<<div_up(nq * 32, 256), 256, 0, st>>>((const orbb200_top2 *)dp, nshards, nq, (orbb200_top2 *)dout);
    ORB_CUDA_OK(cudaMemcpyAsync(out, dout, sizeof(orbb200_top2) * (size_t)nq, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    return ORBB200_OK;
}

int orbb200_distinctive_descriptors_device(const uint8_t *desc_d, const int *offsets_d, int ngroups, int *best_idx_d,
                                           int *best_median_d, void *stream)
{
    if (ngroups <= 0) return ORBB200_OK;
    ORB_ARG_OK(desc_d && offsets_d && best_idx_d && best_median_d, "null pointer");
    ORB_ARG_OK(((size_t)desc_d & 15) == 0, "descriptor array must be 16-byte aligned");
    int rc = scratch_bind(-1);
    if (rc) return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : g_scratch.stream;
    k_distinctive<<<div_up(ngroups, DD_WARPS), DD_WARPS * 32, 0, st>>>((const uint4 *)desc_d, offsets_d, ngroups, best_idx_d,
                                                                      best_median_d);
    ORB_CUDA_OK(cudaGetLastError());
    return ORBB200_OK;
}

int orbb200_distinctive_descriptors(const uint8_t *desc, const int *offsets, int ngroups, int *best_idx, int *best_median,
                                    int device)
{
    if (ngroups <= 0) return ORBB200_OK;
    ORB_ARG_OK(offsets && best_idx, "null pointer");
    ORB_ARG_OK(offsets[0] == 0, "offsets[0] must be 0");
    for (int g = 0; g < ngroups; g++) ORB_ARG_OK(offsets[g + 1] >= offsets[g], "offsets must not decrease");
    const int total = offsets[ngroups];
    ORB_ARG_OK(total == 0 || desc, "null descriptors");
    int rc = scratch_bind(device);
    if (rc) return rc;
    void *dd, *doffs, *dout;
    if ((rc = scratch_get(0, (size_t)std::max(total, 1) * 32, &dd)) || (rc = scratch_get(3, (size_t)(ngroups + 1) * 4, &doffs)) ||
        (rc = scratch_get(2, (size_t)ngroups * 8, &dout)))
        return rc;
    cudaStream_t st = g_scratch.stream;
    if (total > 0) ORB_CUDA_OK(cudaMemcpyAsync(dd, desc, (size_t)total * 32, cudaMemcpyHostToDevice, st));
    ORB_CUDA_OK(cudaMemcpyAsync(doffs, offsets, (size_t)(ngroups + 1) * 4, cudaMemcpyHostToDevice, st));
    k_distinctive<<<div_up(ngroups, DD_WARPS), DD_WARPS * 32, 0, st>>>((const uint4 *)dd, (const int *)doffs, ngroups, (int *)dout,
                                                                      (int *)dout + ngroups);
    ORB_CUDA_OK(cudaMemcpyAsync(best_idx, dout, (size_t)ngroups * 4, cudaMemcpyDeviceToHost, st));
    if (best_median)
        ORB_CUDA_OK(cudaMemcpyAsync(best_median, (int *)dout + ngroups, (size_t)ngroups * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    return ORBB200_OK;
}

int orbb200_exchange_create(orbb200_exchange **out, int rank, int world, int nq_max, int device, void *ipc_handle_out)
{
    ORB_ARG_OK(out && ipc_handle_out, "null pointer");
    ORB_ARG_OK(world >= 1 && world <= EX_MAX_WORLD && rank >= 0 && rank < world && nq_max > 0, "bad rank / world / nq_max");
    static_assert(sizeof(cudaIpcMemHandle_t) == ORBB200_IPC_HANDLE_BYTES, "IPC handle size");
    if (device < 0) ORB_CUDA_OK(cudaGetDevice(&device));
    ORB_CUDA_OK(cudaSetDevice(device));
    orbb200_exchange *x = new orbb200_exchange();
    x->rank = rank; x->world = world; x->device = device;
    x->nqMax = (int)align_up((size_t)nq_max, BF_THREADS * BF_QPT);
    x->maxGroups = x->nqMax / (BF_THREADS * BF_QPT);
    x->regionBytes = align_up(sizeof(orbb200_top2) * 2 * (size_t)world * x->nqMax + sizeof(int) * (size_t)world * x->maxGroups, 256);
    cudaError_t e = cudaMalloc(&x->region, x->regionBytes);
    if (e == cudaSuccess) e = cudaMemset(x->region, 0, x->regionBytes);
    if (e == cudaSuccess) e = cudaMalloc(&x->groupCounter, sizeof(int) * (size_t)(x->maxGroups + 1));
    if (e == cudaSuccess) e = cudaMemset(x->groupCounter, 0, sizeof(int) * (size_t)(x->maxGroups + 1));
    if (e == cudaSuccess) e = cudaMalloc(&x->acc, sizeof(unsigned long long) * 2 * (size_t)x->nqMax);
    if (e == cudaSuccess) {
        std::vector<unsigned long long> init(2 * (size_t)x->nqMax, EX_ABSENT);
        e = cudaMemcpy(x->acc, init.data(), sizeof(unsigned long long) * init.size(), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&x->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle((cudaIpcMemHandle_t *)ipc_handle_out, x->region);
    if (e == cudaSuccess) e = cudaDeviceSynchronize(); // the zeroed flags are in place before any peer can map the region
    if (e != cudaSuccess) {
        set_error("orbb200_exchange_create: %s", cudaGetErrorString(e));
        orbb200_exchange_destroy(x);
        return ORBB200_ERR_CUDA;
    }
    x->status = x->groupCounter + x->maxGroups;
    x->peer[rank] = x->region;
    x->connected = world == 1;
    *out = x;
    return ORBB200_OK;
}

int orbb200_exchange_connect(orbb200_exchange *x, const void *all_handles)
{
    ORB_ARG_OK(x && all_handles, "null pointer");
    ORB_CUDA_OK(cudaSetDevice(x->device));
    const cudaIpcMemHandle_t *h = (const cudaIpcMemHandle_t *)all_handles;
    for (int p = 0; p < x->world; p++) {
        if (p == x->rank || x->peer[p]) continue;
        void *ptr = nullptr;
        ORB_CUDA_OK(cudaIpcOpenMemHandle(&ptr, h[p], cudaIpcMemLazyEnablePeerAccess));
        x->peer[p] = (unsigned char *)ptr;
    }
    x->connected = true;
    return ORBB200_OK;
}

void orbb200_exchange_destroy(orbb200_exchange *x)
{
    if (!x) return;
    cudaSetDevice(x->device);
    if (x->stream) cudaStreamSynchronize(x->stream);
    for (int p = 0; p < x->world; p++)
        if (p != x->rank && x->peer[p]) cudaIpcCloseMemHandle(x->peer[p]);
    if (x->region) cudaFree(x->region);
    if (x->groupCounter) cudaFree(x->groupCounter);
    if (x->acc) cudaFree(x->acc);
    if (x->stream) cudaStreamDestroy(x->stream);
    delete x;
}

int orbb200_match_sharded_top2_device(orbb200_exchange *x, const uint8_t *q_d, int nq, const uint8_t *db_d, int ndb_local,
                                      int index_offset, orbb200_top2 *out_d, void *stream)
{
    ORB_ARG_OK(x && q_d && db_d && out_d && nq > 0 && ndb_local > 0, "bad arguments");
    ORB_ARG_OK(x->connected, "orbb200_exchange_connect has not been called");
    ORB_ARG_OK(nq <= x->nqMax, "nq exceeds the exchange's nq_max");
    ORB_ARG_OK(((size_t)q_d & 15) == 0 && ((size_t)db_d & 15) == 0 && ((size_t)out_d & 15) == 0, "arrays must be 16-byte aligned");
    ORB_CUDA_OK(cudaSetDevice(x->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : x->stream;
    int groups, chunks, rows;
    bf_plan(nq, ndb_local, &groups, &chunks, &rows);
    ExchangeParams X;
    memset(&X, 0, sizeof(X));
    X.world = x->world; X.rank = x->rank; X.epoch = ++x->epoch; X.nqMax = x->nqMax; X.maxGroups = x->maxGroups;
    X.groupCounter = x->groupCounter; X.acc = x->acc; X.status = x->status; X.out = out_d;
    for (int p = 0; p < x->world; p++) X.peer[p] = x->peer[p];
    X.timeoutNs = 10LL * 1000 * 1000 * 1000;
    k_bf_top2<true><<<dim3(groups, chunks), BF_THREADS, 0, st>>>(q_d, nq, db_d, ndb_local, rows, index_offset, nullptr, X);
    ORB_CUDA_OK(cudaGetLastError());
    return ORBB200_OK;
}

int orbb200_exchange_status(orbb200_exchange *x, void *stream)
{
    ORB_ARG_OK(x != nullptr, "null pointer");
    ORB_CUDA_OK(cudaSetDevice(x->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : x->stream;
    int s = 0;
    ORB_CUDA_OK(cudaMemcpyAsync(&s, x->status, sizeof(int), cudaMemcpyDeviceToHost, st));
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    if (s != 0) {
        set_error("orbb200_match_sharded_top2_device: a peer's shard result did not arrive within the timeout");
        return ORBB200_ERR_CUDA;
    }
    return ORBB200_OK;
}

int orbb200_popc_peak(int device, double *popc32_per_s)
{
    ORB_ARG_OK(popc32_per_s != nullptr, "null pointer");
    int rc = scratch_bind(device);
    if (rc) return rc;
    void *d;
    if ((rc = scratch_get(2, 256, &d))) return rc;
    cudaStream_t st = g_scratch.stream;
    cudaEvent_t e0, e1;
    ORB_CUDA_OK(cudaEventCreate(&e0));
    ORB_CUDA_OK(cudaEventCreate(&e1));
    const int blocks = 148 * 8, threads = 256, iters = 4096;
    double best = 0;
    for (int rep = 0; rep < 5; rep++) {
        ORB_CUDA_OK(cudaEventRecord(e0, st));
        k_popc_peak<<<blocks, threads, 0, st>>>((unsigned int *)d, iters);
        ORB_CUDA_OK(cudaEventRecord(e1, st));
        ORB_CUDA_OK(cudaEventSynchronize(e1));
        float ms = 0;
        ORB_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms > 0) best = std::max(best, (double)blocks * threads * iters * 8.0 / (ms * 1e-3));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *popc32_per_s = best;
    return ORBB200_OK;
}

int orbb200_stereo_match(orbb200_extractor *left, orbb200_extractor *right, float bf, float b, float *uRight, float *depth,
                         int *n_valid)
{
    return stereo_match_impl(left, right, bf, b, uRight, depth, n_valid);
}

int orbb200_stereo_match_batch_device(orbb200_extractor *e, int npairs, float bf, float b, float *uRight_d, float *depth_d,
                                      int *n_valid_d)
{
    return stereo_batch_device_impl(e, npairs, bf, b, uRight_d, depth_d, n_valid_d);
}

int orbb200_stereo_match_batch(orbb200_extractor *e, int npairs, float bf, float b, float *uRight, float *depth, int *n_valid)
{
    return stereo_batch_host_impl(e, npairs, bf, b, uRight, depth, n_valid);
}

} // extern "C"
