// extractor_host.cuh -- the handle behind orbb200_extractor (one CUDA stream + workspaces per instance;
// the reference runs two extractor instances concurrently on two host threads, src/Frame.cc:82-85).
#pragma once
#include <vector>

#include "extract_kernels.cuh"

struct orbb200_extractor {
    int device = 0;
    cudaStream_t stream = nullptr;
    // single frames / pairs: the blur of the pyramid runs on sideStream beside FAST + quadtree (fork after the last resize, join
    // before the descriptor stage); batches keep one stream (both sides are issue bound there)
    cudaStream_t sideStream = nullptr;
    cudaEvent_t evFork = nullptr, evJoin = nullptr;
    // E1 tables (src/ORBextractor.cc:429-534)
    int nfeatures = 0, nlevels = 0, iniTh = 0, minTh = 0;
    double scaleFactor = 0;
    float sf[ORBB200_MAX_LEVELS], isf[ORBB200_MAX_LEVELS], sig2[ORBB200_MAX_LEVELS], isig2[ORBB200_MAX_LEVELS];
    int nfeat[ORBB200_MAX_LEVELS];
    int umax[16];
    // shape-dependent state
    int w = 0, h = 0, batchCap = 0;
    orbb200::LevelGeom geom[ORBB200_MAX_LEVELS];
    size_t pyrStride = 0, candStride = 0, nodeWsStride = 0;
    int totalCells = 0, nodeTotal = 0, totalBlurTiles = 0, coefTotal = 0;
    size_t octreeSmem = 0;
    int octDepthCap = 7; // OCT_MAX_DEPTH
    int fastRows = 7; // staged ROI rows of the tallest cell row (k_fast_strips dynamic shared memory)
    int totalStrips = 0;
    // device buffers
    orbb200::LevelGeom *d_geom = nullptr;
    uint4 *d_fastStrips = nullptr;
    int *d_umax = nullptr;
    int *d_rzXofs = nullptr, *d_rzYofs = nullptr;
    short2 *d_rzXa = nullptr, *d_rzYb = nullptr;
    uint8_t *d_pyr = nullptr, *d_blur = nullptr;
    uint32_t *d_cand = nullptr, *d_ord = nullptr, *d_lvlKp = nullptr;
    int *d_knode = nullptr, *d_cellCount = nullptr, *d_cellScan = nullptr, *d_nodeWs = nullptr, *d_lvlCount = nullptr;
    // outputs owned by the handle (host API, stereo matcher)
    orbb200_keypoint *d_kps = nullptr;
    uint8_t *d_desc = nullptr;
    int *d_nOut = nullptr;
    int outCap = 0, outBatch = 0;
    // batched stereo matcher scratch / outputs (orbb200_stereo_match_batch*)
    float *d_stU = nullptr, *d_stD = nullptr;
    int *d_stSad = nullptr, *d_stN = nullptr, *d_stRows = nullptr;
    size_t stElems = 0, stRowElems = 0;
    int stPairs = 0, stOutCap = 0;
    // host-API pipeline (orbb200_extract_batch): copy / D2H streams, multi-buffered packed staging
    // and up to PIPE_STREAMS compute streams so that the latency-bound tail of one chunk (quadtree passes) overlaps
    // the throughput-bound head of the next (all per-image workspaces are disjoint between chunks)
    static constexpr int PIPE_STREAMS = 8;
    cudaStream_t copyStream = nullptr, d2hStream = nullptr, workStream[PIPE_STREAMS] = {};
    static constexpr int PIPE_STAGES = 6; // staging buffers: the H2D copies run up to 6 chunks ahead of the stage kernels
    cudaEvent_t evH2D[PIPE_STAGES] = {}, evDone[PIPE_STAGES] = {}, evStaged[PIPE_STAGES] = {};
    cudaEvent_t evWork[PIPE_STREAMS] = {}, evBegin = nullptr;
    // begin / wait form of the host-buffer batch calls: up to two calls in flight on a handle, so that the copies of call k + 1 run
    // under the tail of call k.  evChunkOut[ci]: the download of chunk ci of the previous call is complete (its device output block
    // may be overwritten by chunk ci of the next call)
    struct PendingCall { bool active = false; int batch = 0, cap = 0, chunk = 0, width = 0, height = 0, stereo = 0; int *n_out = nullptr; cudaEvent_t done = nullptr;
                         int result = 0; bool unreported = false; }; // result of a call completed by a later begin, kept for its wait
    PendingCall pend[2];
    int pendNext = 0;
    std::vector<cudaEvent_t> evChunkOut;
    size_t chunkOutRecorded = 0;
    uint8_t *d_stage[PIPE_STAGES] = {};
    size_t stageBytes = 0;
    int pipeChunk = 16, pipeStreams = 8, devChunk = 0;
    // begun (streamed) calls have no head / tail to keep short, so they take larger chunks, whose kernels fill the GPU better:
    // 160 KITTI pairs per step, two steps in flight: 4.02 ms at 16 images per chunk, 3.60 ms at 64.  orbb200_extractor_set_pipeline
    // fixes the chunk size for both forms.
    int pipeChunkStreamed = 64;
    // single-image latency path (orbb200_extract): pinned bounce buffers + the whole launch sequence (H2D, level-0 staging,
    // pyramid, FAST, quadtree, blur, describe, D2H) captured once per (shape, capacity) in a CUDA graph
    uint8_t *h_in = nullptr, *h_out = nullptr, *d_in1 = nullptr;
    size_t hInBytes = 0, hOutBytes = 0, dIn1Bytes = 0;
    uint8_t *h_pyr = nullptr; // pinned copy of one image's pyramid slab (orbb200_pyramid_map)
    size_t hPyrBytes = 0;
    cudaGraphExec_t graph1 = nullptr;
    int graphW = 0, graphH = 0, graphCap = 0, graphCh = 0, graphRFirst = 0;
    unsigned long long graphEpoch = 0, allocEpoch = 1; // allocEpoch moves whenever a device buffer the graph points at is reallocated
    int useGraph = 1;
    // last call
    int lastBatch = 0, lastCap = 0;
    const uint8_t *lastL0 = nullptr;
    size_t lastL0ImgStride = 0;
    int lastL0Pitch = 0;
    const orbb200_keypoint *lastKps = nullptr;
    const uint8_t *lastDesc = nullptr;
    const int *lastNOut = nullptr;
    int lastLaunches = 0;
    // profiling
    int profiling = 0;
    static constexpr int EV_RING = 64;
    // ring of per-call stage event sets, created when profiling is switched on: 0 start, 1 pyramid done, 2 FAST done,
    // 3 quadtree done, 4 describe start, 5 describe done, 6 / 7 blur start / done
    static constexpr int EV_N = 8;
    cudaEvent_t ev[EV_RING][EV_N] = {};
    int evCount = 0;                 // calls recorded since profiling was enabled
};

namespace orbb200 {
int extractor_ensure_shape(orbb200_extractor *e, int w, int h, int batch);
int extractor_ensure_outputs(orbb200_extractor *e, int batch, int cap);
int extractor_ensure_pipeline(orbb200_extractor *e, size_t stageBytes);
int extractor_run(orbb200_extractor *e, const uint8_t *l0, size_t l0ImgStride, int l0Pitch, int l0Valid, int batch,
                  orbb200_keypoint *kps_d, uint8_t *desc_d, int cap, int *nOut_d);
// batched stereo matcher pieces shared by orbb200_match.cu (owner) and the fused extract + stereo pipeline
int stereo_ensure_scratch(orbb200_extractor *e, int npairs, int cap);
int stereo_ensure_outputs(orbb200_extractor *e, int npairs, int cap);
int stereo_batch_range(orbb200_extractor *e, const orbb200_keypoint *kps, const uint8_t *desc, const int *nOut, const uint8_t *l0,
                       int l0Pitch, size_t l0ImgStride, int cap, int pair0, int npairs, int totalPairs, float bf, float b,
                       float *uRight_d, float *depth_d, int *n_valid_d, cudaStream_t st);
} // namespace orbb200
