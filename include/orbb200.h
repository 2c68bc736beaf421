/*
 * orbb200.h -- C-ABI of liborbb200.so: the B200 (sm_100a) drop-in for ORB-SLAM2's front-end hot path.
 *
 * The reference has no FFI/plugin layer for this path: callers instantiate two C++ classes directly
 * (ORB_SLAM2::ORBextractor, include/ORBextractor.h:53-135; ORB_SLAM2::ORBmatcher, include/ORBmatcher.h:38-100).
 * This header is the boundary a replacement library exports underneath source-compatible facades of those
 * classes (include/orbb200/ORBextractor.h, include/orbb200/ORBmatcher.h).  Each entry point cites the
 * reference interface it replaces as file:line relative to /root/reference.
 *
 * Conventions: plain pointers and sizes only.  Pointers are HOST memory unless the name ends in `_d`
 * (device memory on the handle's GPU).  Every function returns 0 on success, <0 on error; the message is
 * available from orbb200_last_error() (thread-local).  There is no CPU fallback: without a CUDA device every
 * compute entry point fails with ORBB200_ERR_CUDA.
 */
#ifndef ORBB200_H
#define ORBB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBB200_OK 0
#define ORBB200_ERR_ARG (-1)      /* bad argument (null, non-positive size, unsupported shape) */
#define ORBB200_ERR_CUDA (-2)     /* CUDA runtime error / no device */
#define ORBB200_ERR_CAPACITY (-3) /* caller's output capacity too small */
#define ORBB200_ERR_STATE (-4)    /* call order (e.g. pyramid requested before any extract) */

#define ORBB200_MAX_LEVELS 32

/* == cv::KeyPoint layout (7 x 4 bytes): pt.x, pt.y, size, angle, response, octave, class_id */
typedef struct {
    float x, y, size, angle, response;
    int octave, class_id;
} orbb200_keypoint;

typedef struct orbb200_extractor orbb200_extractor; /* opaque; one CUDA stream + workspace per handle */

/* ---- extractor ------------------------------------------------------------------------------------ */

/* replaces ORBextractor::ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST,
 * int minThFAST)  (include/ORBextractor.h:66-67, src/ORBextractor.cc:429-534).  device = -1: current. */
int orbb200_extractor_create(orbb200_extractor **out, int nfeatures, float scaleFactor, int nlevels, int iniThFAST,
                             int minThFAST, int device);
void orbb200_extractor_destroy(orbb200_extractor *e);

/* replaces GetLevels / GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares /
 * GetInverseScaleSigmaSquares (include/ORBextractor.h:78-98) and exposes mnFeaturesPerLevel (:123).
 * Any output pointer may be NULL.  Arrays have nlevels entries. */
int orbb200_level_tables(const orbb200_extractor *e, float *scale, float *inv_scale, float *sigma2, float *inv_sigma2,
                         int *nfeat);

/* the same tables without a handle or a GPU (host arithmetic only): mvScaleFactor, mvInvScaleFactor, mvLevelSigma2,
 * mvInvLevelSigma2, mnFeaturesPerLevel (nlevels entries each) and umax[16] (src/ORBextractor.cc:441-461, 468-493,
 * 510-533).  Any output pointer may be NULL. */
int orbb200_compute_level_tables(int nfeatures, float scaleFactor, int nlevels, float *scale, float *inv_scale,
                                 float *sigma2, float *inv_sigma2, int *nfeat, int *umax);

/* keypoints per image when every level stays within its quota: sum of nfeat[l] + 3 (DistributeOctTree can exceed the quota by
 * up to 3 nodes: src/ORBextractor.cc:879).  NOT a bound for wide images with a small quota -- see the shape-aware call below. */
int orbb200_max_keypoints(const orbb200_extractor *e);

/* upper bound on keypoints per image of this shape: per level max(nfeat[l] + 3, 4 * nIni_l), nIni_l = round(width / height) of
 * the level's bordered area -- the first DistributeOctTree pass splits all nIni roots whatever the quota is
 * (src/ORBextractor.cc:624-760) and the reference never truncates.  Use it to size kps / desc. */
int orbb200_max_keypoints_for_shape(const orbb200_extractor *e, int width, int height);

/* the `cap` of the last extraction on this handle (0 before the first): the row length of its device-resident outputs and the
 * number of floats orbb200_stereo_match may write per side */
int orbb200_last_capacity(const orbb200_extractor *e);

/* replaces ORBextractor::operator()(image, mask, keypoints, descriptors) (include/ORBextractor.h:74-76,
 * src/ORBextractor.cc:1261-1339) for one 8-bit single-channel image.  kps[cap], desc[cap*32] are host
 * buffers; *n_out receives the keypoint count.  An empty image (img NULL or w/h 0) returns OK with
 * *n_out = 0, like the reference's early return (:1264).  The mask argument of the reference is ignored
 * there and has no counterpart here. */
int orbb200_extract(orbb200_extractor *e, const uint8_t *img, int width, int height, int stride,
                    orbb200_keypoint *kps, uint8_t *desc, int cap, int *n_out);

/* the same for `batch` images of one shape (frame-sharded configs); image b's outputs start at
 * kps + b*cap and desc + b*cap*32; n_out[batch].  Host buffers; H2D/D2H copies are inside the call. */
int orbb200_extract_batch(orbb200_extractor *e, const uint8_t *const *imgs, int batch, int width, int height,
                          int stride, orbb200_keypoint *kps, uint8_t *desc, int cap, int *n_out);

/* tuning of orbb200_extract_batch's copy/compute pipeline (no reference counterpart: the reference has no
 * batching): the batch is cut into chunks of `chunk_images`; chunk c's upload, its kernels (rotating over
 * `compute_streams` <= 8 CUDA streams) and chunk c-1's download overlap.  Defaults 16 and 8.
 * device_chunk_images > 0 makes orbb200_extract_batch_device cut its batch the same way (chunks joined back
 * into the handle's stream before it returns, so the call stays stream-ordered for the caller); 0 = off, the
 * default: resident batches fill the GPU kernel by kernel and chunking buys < 4 %.  With stage profiling on
 * the device variant always runs unchunked on the handle's stream.
 * Results do not depend on any of these values. */
int orbb200_extractor_set_pipeline(orbb200_extractor *e, int chunk_images, int compute_streams, int device_chunk_images);

/* device-resident variant: imgs_d holds `batch` images, image b at imgs_d + b*image_stride, rows `stride`
 * bytes apart; outputs are device buffers laid out as above.  Asynchronous on the handle's stream (see
 * orbb200_extractor_stream); no host synchronisation inside. */
int orbb200_extract_batch_device(orbb200_extractor *e, const uint8_t *imgs_d, int batch, int width, int height,
                                 int stride, size_t image_stride, orbb200_keypoint *kps_d, uint8_t *desc_d, int cap,
                                 int *n_out_d);

/* replaces reads of the public member mvImagePyramid[level] (include/ORBextractor.h:100; read by
 * src/Frame.cc:584,674,686,691): copies the UN-blurred level of image `image_index` of the last call
 * to dst (rows dst_stride apart).  dst may be NULL to query the size only. */
int orbb200_pyramid_level(orbb200_extractor *e, int image_index, int level, uint8_t *dst, int dst_stride, int *w,
                          int *h);

/* mvImagePyramid in ONE copy: the un-blurred levels of an image live in one device slab (level l at offsets[l], rows pitches[l]
 * bytes apart, widths[l] x heights[l] pixels); orbb200_pyramid_download copies the slab of image `image_index` of the last call to
 * dst (slab_bytes bytes) with a single D2H transfer -- what the facade's lazy mvImagePyramid does on first access. */
int orbb200_pyramid_layout(orbb200_extractor *e, int *offsets, int *pitches, int *widths, int *heights, size_t *slab_bytes);
int orbb200_pyramid_download(orbb200_extractor *e, int image_index, uint8_t *dst, size_t dst_bytes);
/* same, into a pinned buffer owned by the handle (full-speed copy); *host_slab stays valid until the next orbb200_pyramid_map call
 * on this handle or its destruction */
int orbb200_pyramid_map(orbb200_extractor *e, int image_index, const uint8_t **host_slab);

/* pins the calling host thread to the CPUs local to `device` (-1: current) -- /sys/bus/pci/devices/<bus id>/local_cpulist -- so that
 * pinned buffers allocated afterwards live on the GPU's NUMA node and copies are submitted from a near core (multi-GPU hosts:
 * one call per rank before it allocates).  Returns the number of CPUs in the new mask, 0 when the topology cannot be read. */
int orbb200_bind_host_to_device(int device);

/* page-locked host memory for the buffers handed to the *_batch entry points (the dataset loader decodes straight into it):
 * copies from pinned memory run at full PCIe speed and really overlap the kernels */
int orbb200_host_alloc(void **p, size_t bytes);
int orbb200_host_free(void *p);

/* Tuning / test knob of the quadtree stage (DistributeOctTree, src/ORBextractor.cc:617-915): the kernel builds the tree from a
 * histogram of the keys down to a fixed depth (7 by default, less when shared memory is short) and falls back to an iterative
 * formulation without depth limit when a node has to be split below it.  max_depth caps that depth; 0 = iterative formulation only.
 * The result is the same for every value (tests/test_gpu_extract.py runs 0, 2 and the default against the oracle). */
int orbb200_extractor_set_octree_depth(orbb200_extractor *e, int max_depth);

/* single-image calls (orbb200_extract / orbb200_extract_color) replay a CUDA graph captured per (shape, capacity): pinned bounce
 * buffers, one graph launch, one synchronisation.  on = 0 sends them through the batch pipeline instead (A/B measurements). */
int orbb200_extractor_set_graph(orbb200_extractor *e, int on);

/* cudaStream_t of the handle (as void*), so callers can order their own work / events against it. */
void *orbb200_extractor_stream(orbb200_extractor *e);
int orbb200_extractor_sync(orbb200_extractor *e);

/* per-stage device time of the LAST extract call, measured with CUDA events on the handle's stream when
 * profiling is enabled (off by default; the events serialise nothing but add ~1 us per stage).
 * stages: 0 pyramid, 1 fast_cells, 2 octree_select, 3 blur7, 4 orient_describe; with n >= 6 also 5 = the whole
 * extraction, first launch to last.  Each stage is timed by its own event pair on the handle's stream; the stages run
 * back to back on that one stream, so the five stage times add up to ms[5] (up to event overhead). */
int orbb200_extractor_set_profiling(orbb200_extractor *e, int on);
int orbb200_extractor_stage_ms(orbb200_extractor *e, float *ms, int n);
/* sum of the per-stage times over the calls recorded since profiling was switched on (at most the last 64);
 * *ncalls = how many calls the sums cover.  Synchronises the handle's stream. */
int orbb200_extractor_stage_ms_sum(orbb200_extractor *e, float *ms, int n, int *ncalls);
/* number of kernel launches issued by the last extract call */
int orbb200_extractor_last_launches(const orbb200_extractor *e);

/* test/debug taps on the last call (image_index within the batch): pre-quadtree candidates of a level in the
 * reference's vToDistributeKeys order (src/ORBextractor.cc:997-1006), as (x, y, response) int triples with
 * coordinates relative to the 16-px border; and the blurred level (:1312-1314). */
int orbb200_debug_candidates(orbb200_extractor *e, int image_index, int level, int *xyr, int cap, int *n);
int orbb200_debug_blurred_level(orbb200_extractor *e, int image_index, int level, uint8_t *dst, int dst_stride);

/* ---- prologue of the path -------------------------------------------------------------------
 * Colour input: Tracking::GrabImageStereo / RGBD / Monocular convert with cv::cvtColor(im, gray, CV_RGB2GRAY | CV_BGR2GRAY |
 * CV_RGBA2GRAY | CV_BGRA2GRAY) (src/Tracking.cc:177-202, 218-231, 255-266) before building the Frame.  Here the conversion
 * is fused into the level-0 staging of the extraction: interleaved `channels` (3 or 4) bytes per pixel in, r_first != 0 when
 * channel 0 is red (mbRGB).  gray = (R*9798 + G*19235 + B*3735 + 16384) >> 15, bit-exact with OpenCV's 8-bit path; the
 * gray image (mImGray) is level 0 of the pyramid: orbb200_pyramid_level(e, image, 0, ...). */
int orbb200_extract_color(orbb200_extractor *e, const uint8_t *img, int width, int height, int stride, int channels,
                          int r_first, orbb200_keypoint *kps, uint8_t *desc, int cap, int *n_out);
int orbb200_extract_color_batch(orbb200_extractor *e, const uint8_t *const *imgs, int batch, int width, int height,
                                int stride, int channels, int r_first, orbb200_keypoint *kps, uint8_t *desc, int cap,
                                int *n_out);
/* the conversion alone (host buffers) */
int orbb200_cvt_gray(const uint8_t *img, int width, int height, int stride, int channels, int r_first, uint8_t *gray,
                     int gray_stride, int device);
/* replaces Frame::UndistortKeyPoints (src/Frame.cc:470-515) = cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK);
 * also serves the four corners of Frame::ComputeImageBounds (:517-550).  cam = {fx, fy, cx, cy} of mK and
 * dist_coef = {k1, k2, p1, p2, k3} of mDistCoef as the reference stores them (float).  k1 == 0 copies (:473-477).
 * Only x, y change; double arithmetic, 5 iterations, bit-exact with OpenCV.  out may alias kps. */
int orbb200_undistort_keypoints(const orbb200_keypoint *kps, int n, const float *cam, const float *dist_coef,
                                orbb200_keypoint *out, int device);
int orbb200_undistort_keypoints_device(const orbb200_keypoint *kps_d, int n, const float *cam, const float *dist_coef,
                                       orbb200_keypoint *out_d, void *stream);

/* ---- matcher ---------------------------------------------------------------------------------------- */

/* running best / second-best of ORBmatcher's searches (src/ORBmatcher.cc:214-240): best_idx is the first
 * candidate attaining best_dist; second_dist counts duplicates; 256 / -1 when absent. */
typedef struct {
    int best_dist, best_idx, second_dist, second_idx;
} orbb200_top2;

/* replaces static int ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&)
 * (include/ORBmatcher.h:44, src/ORBmatcher.cc:1733-1749) for arrays of pairs: dist[i] = H(a[i], b[i]). */
int orbb200_descriptor_distance(const uint8_t *a, const uint8_t *b, int n, int *dist, int device);

/* X1 (BASELINE config 5): the best/second-best loop of src/ORBmatcher.cc:214-244 over a flat database. */
int orbb200_match_bruteforce_top2(const uint8_t *q, int nq, const uint8_t *db, int ndb, orbb200_top2 *out,
                                  int device);
int orbb200_match_bruteforce_top2_device(const uint8_t *q_d, int nq, const uint8_t *db_d, int ndb,
                                         orbb200_top2 *out_d, int index_offset, void *stream);
/* NO counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Same search, same kernels, same result as
 * orbb200_match_bruteforce_top2_device on the CURRENT device, but it allocates nothing: the per-chunk partial results live in a
 * static 8 MiB device array that every GPU loading the library holds until exit.  Limit: (database chunks) x nq <= 524 288 partial
 * records of 16 bytes.  The launch plan uses at most floor(592 / ceil(nq / 512)) chunks, i.e. at most 303 104 records while the
 * database is chunked (BASELINE config 5, 2048 x 1 048 576: 148 chunks, 303 104 records) and nq records once it is not: the limit
 * is nq <= 524 288 queries per call, any database size.  Beyond it the call returns ORBB200_ERR_CAPACITY and launches nothing.
 * stream == NULL is the default stream (no stream is created).  All calls on a device share the array, which keeps the last call's
 * contents (nothing needs clearing: every record a call reads it has written first); calls that OVERLAP on the GPU race on it. */
int orbb200_match_bruteforce_top2_device_noalloc(const uint8_t *q_d, int nq, const uint8_t *db_d, int ndb,
                                                 orbb200_top2 *out_d, int index_offset, void *stream);

/* M2-M9 core: per-query candidate lists (CSR: cand_offsets[nq+1], cand_idx) visited in list order -- the
 * inner loops of SearchByProjection (src/ORBmatcher.cc:90-123, 1482-1508), SearchByBoW (:214-240),
 * SearchForTriangulation (:768-815), Fuse (:960-1003), SearchBySim3 (:1268-1283). */
int orbb200_match_candidates(const uint8_t *q, int nq, const uint8_t *t, int nt, const int *cand_offsets,
                             const int *cand_idx, orbb200_top2 *out, int device);

/* every distance of the same CSR candidate set: dist[c] = H(q[i], t[cand_idx[c]]) for cand_offsets[i] <= c <
 * cand_offsets[i+1].  For the searches whose best/second-best loop skips targets claimed earlier in the same call
 * (src/ORBmatcher.cc:93-95, 222, 475, 622, 1485-1487, 1624): the host replays the ordered loop on these distances. */
int orbb200_candidate_distances(const uint8_t *q, int nq, const uint8_t *t, int nt, const int *cand_offsets,
                                const int *cand_idx, uint16_t *dist, int device);

/* SearchForTriangulation resolved on the device (src/ORBmatcher.cc:757-815 with CheckDistEpipolarLine :147-164): for every query
 * (descriptor q[i], undistorted position qxy[i] = {x, y}, qstereo[i] = mvuRight >= 0) over its CSR candidate list into the targets
 * (t, tkp = mvKeysUn of keyframe 2, tstereo): the candidate with the smallest distance <= th_low that passes the epipole gate
 * (both monocular: squared distance to (ex, ey) >= 100 * scale_factors[octave]) and the epipolar test
 * ((a x2 + b y2 + c)^2 / (a^2 + b^2) < 3.84 * level_sigma2[octave], [a b c] = [x1 y1 1] F12), a later candidate winning ties.
 * best_idx2[i] = target index or -1, best_dist[i] (may be NULL) its distance.  The callers' gates that do not depend on the
 * distance (existing map points, bOnlyStereo) are applied when the lists are built. */
int orbb200_search_triangulation(const uint8_t *q, const float *qxy, const uint8_t *qstereo, int nq, const uint8_t *t,
                                 const orbb200_keypoint *tkp, const uint8_t *tstereo, int nt, const int *cand_offsets, const int *cand_idx,
                                 const float *F12, float ex, float ey, const float *scale_factors, const float *level_sigma2, int nlevels,
                                 int th_low, int *best_idx2, int *best_dist, int device);

/* F3 + F2 + distances in one call: Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:261-279, 444-456) on the
 * target keypoints tkp[nt] (the caller's mvKeysUn; grid = mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv,
 * FRAME_GRID_COLS x FRAME_GRID_ROWS), then Frame::GetFeaturesInArea (:386-442) for every query window
 * windows[nq][5] = {x, y, r, minLevel, maxLevel}, candidates in the reference's order (ix outer, iy inner, insertion
 * order inside a cell), and the DescriptorDistance of every candidate to qdesc[q].  Outputs: CSR cand_offsets[nq+1],
 * cand_idx[], dist[].  Returns ORBB200_ERR_CAPACITY with cand_offsets filled when cand_cap is too small.  This is the
 * device side of SearchByProjection / Fuse / SearchBySim3 / SearchForInitialization (src/ORBmatcher.cc:48-136, 305-418,
 * 430-563, 873-1164, 1166-1390, 1396-1682); the ordered accept / claim loop stays in the host facade. */
int orbb200_window_search(const orbb200_keypoint *tkp, const uint8_t *tdesc, int nt, float minX, float minY, float invW,
                          float invH, int cols, int rows, const float *windows, const uint8_t *qdesc, int nq,
                          int *cand_offsets, int *cand_idx, uint16_t *dist, int cand_cap, int device);

/* The same with everything resident in HBM and asynchronous on `stream` (the windowed searches chained
 * behind the extraction without a host round trip).  tkp_d / tdesc_d may be one image's slice of the block
 * orbb200_extract_batch_device wrote (tkp_d = kps_d + i*cap, tdesc_d = desc_d + i*cap*32, nt = cap, nt_d = n_out_d + i:
 * when nt_d is not NULL the keypoint count is read on the device, clipped to nt).  workspace_d: cols*rows + 1 + nt ints.
 * cand_offsets_d[nq + 1] receives the CSR offsets (cand_offsets_d[nq] = total number of candidates); candidates beyond
 * cand_cap are not written -- the caller compares the total with cand_cap after reading the offsets.  Same order and
 * same distances as orbb200_window_search. */
int orbb200_window_search_device(const orbb200_keypoint *tkp_d, const uint8_t *tdesc_d, int nt, const int *nt_d, float minX,
                                 float minY, float invW, float invH, int cols, int rows, const float *windows_d,
                                 const uint8_t *qdesc_d, int nq, int *cand_offsets_d, int *cand_idx_d, uint16_t *dist_d,
                                 int cand_cap, int *workspace_d, void *stream);

/* orbb200_window_search with the TARGET frame already resident on the device (tkp_d / tdesc_d = device pointers, e.g. the
 * outputs of the extraction or a copy a host object keeps -- include/orbb200/DeviceFrame.h); windows, query descriptors and results
 * are host buffers.  Only the queries travel: 52 bytes per window instead of 60 bytes per target keypoint per search. */
int orbb200_window_search_resident(const orbb200_keypoint *tkp_d, const uint8_t *tdesc_d, int nt, float minX, float minY, float invW,
                                   float invH, int cols, int rows, const float *windows, const uint8_t *qdesc, int nq, int *cand_offsets,
                                   int *cand_idx, uint16_t *dist, int cand_cap, int device);

/* device pointers of the outputs of the LAST extraction on this handle (kps [image][cap], desc [image][cap][32], counts [image]);
 * valid until the next extraction on the handle -- a frame that wants to keep them copies them (orbb200_copy_on_device) */
int orbb200_extractor_outputs_device(orbb200_extractor *e, const orbb200_keypoint **kps_d, const uint8_t **desc_d, const int **n_out_d, int *cap);

/* plain device-memory helpers for host-side objects that keep data resident; all synchronous */
int orbb200_device_alloc(void **p, size_t bytes, int device);
int orbb200_device_free(void *p);
int orbb200_copy_to_device(void *dst_d, const void *src, size_t bytes);
int orbb200_copy_to_host(void *dst, const void *src_d, size_t bytes);
int orbb200_copy_on_device(void *dst_d, const void *src_d, size_t bytes);
/* bytes the matcher entry points (and orbb200_copy_to_device) have copied host -> device so far, process-wide */
unsigned long long orbb200_h2d_bytes_total(void);
/* cudaDeviceSynchronize on `device` (-1: current): orders work issued on caller streams before the library's own streams */
int orbb200_device_sync(int device);

/* merge of per-shard top-2 results (config 5, database sharded by rows; shard s's indices already global):
 * parts[s*nq + i]; ties on best_dist go to the lowest shard (= lowest global index, first-seen rule). */
int orbb200_merge_top2(const orbb200_top2 *parts, int nshards, int nq, orbb200_top2 *out, int device);
int orbb200_merge_top2_device(const orbb200_top2 *parts_d, int nshards, int nq, orbb200_top2 *out_d, void *stream);

/* ---- config 5 on N GPUs of one box: database rows sharded over ranks, one process per GPU --------------------
 * The exchange step is fused into the search kernel: the last block to finish a group of 512 queries folds the rank's
 * partial results, stores the shard's top-2 straight into every peer's gather buffer over NVLink (CUDA-IPC mapped peer
 * memory), publishes a release flag, waits for the peers' flags and folds the shards in rank (= database) order, which
 * reproduces the first-seen tie-break of the sequential loop (src/ORBmatcher.cc:229-235).  No NCCL on the data path;
 * the 64-byte IPC handles travel through whatever the host uses for rendezvous (torch.distributed all_gather here).
 *   create:  allocates the rank's exchange region, returns its IPC handle in ipc_handle_out[ORBB200_IPC_HANDLE_BYTES]
 *   connect: all_handles = world handles in rank order (every rank's create has returned before any connect)
 *   match:   asynchronous on `stream` (NULL = the exchange's own); EVERY rank must call it the same number of times with
 *            the same nq; index_offset = global index of the rank's first row; out_d[nq] on every rank = the global top-2
 *   status:  synchronises `stream`; ORBB200_ERR_CUDA when a peer failed to deliver within 10 s (no hang) */
#define ORBB200_IPC_HANDLE_BYTES 64
typedef struct orbb200_exchange orbb200_exchange;
int orbb200_exchange_create(orbb200_exchange **out, int rank, int world, int nq_max, int device, void *ipc_handle_out);
int orbb200_exchange_connect(orbb200_exchange *x, const void *all_handles);
void orbb200_exchange_destroy(orbb200_exchange *x);
int orbb200_match_sharded_top2_device(orbb200_exchange *x, const uint8_t *q_d, int nq, const uint8_t *db_d, int ndb_local,
                                      int index_offset, orbb200_top2 *out_d, void *stream);
int orbb200_exchange_status(orbb200_exchange *x, void *stream);

/* replaces the arithmetic of MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:266-340),
 * batched over map points (callers: LocalMapping.cc:172,448,602,637; LoopClosing.cc:562; Tracking.cc:592,1232).
 * Group g = the descriptors of the non-bad keyframes observing map point g, in the iteration order of
 * mObservations (:288-294): rows [offsets[g], offsets[g+1]) of desc (32 bytes each), offsets[0] == 0.
 * best_idx[g] = row (relative to the group) with the least median Hamming distance to the group, median =
 * sorted row of Distances[i][*] (diagonal 0 included) at index (int)(0.5*(N-1)) (:327-331), first row wins ties
 * (strict <, :333); -1 for an empty group (the reference returns early, :281,296).  best_median may be NULL. */
int orbb200_distinctive_descriptors(const uint8_t *desc, const int *offsets, int ngroups, int *best_idx,
                                    int *best_median, int device);
/* the same with everything resident in HBM (desc_d 16-byte aligned), asynchronous on `stream` (NULL = internal) */
int orbb200_distinctive_descriptors_device(const uint8_t *desc_d, const int *offsets_d, int ngroups, int *best_idx_d,
                                           int *best_median_d, void *stream);

/* F1: replaces Frame::ComputeStereoMatches (src/Frame.cc:577-751) using the device-resident pyramids and
 * outputs of the LAST single-image extract on `left` and `right`.  uRight[nL], depth[nL] host outputs
 * (-1 where unmatched).  bf = mbf, b = mb (baseline in metres). */
int orbb200_stereo_match(orbb200_extractor *left, orbb200_extractor *right, float bf, float b, float *uRight,
                         float *depth, int *n_valid);

/* F1 for a batch of stereo pairs held by ONE handle: the last orbb200_extract_batch[_device] call on `e` carried
 * the pairs interleaved (image 2p = left, 2p+1 = right view of pair p).  Outputs are [npairs][cap] with cap the
 * capacity of that extract call; entries past a pair's left keypoint count are unspecified.  The _device variant
 * is asynchronous on the handle's stream. */
int orbb200_stereo_match_batch(orbb200_extractor *e, int npairs, float bf, float b, float *uRight, float *depth,
                               int *n_valid);
int orbb200_stereo_match_batch_device(orbb200_extractor *e, int npairs, float bf, float b, float *uRight_d,
                                      float *depth_d, int *n_valid_d);

/* replaces the stereo Frame constructor's feature work as ONE call (src/Frame.cc:60-117: ExtractORB left / right on
 * two threads, :82-85, then ComputeStereoMatches, :94) for `npairs` stereo pairs: imgs[2p] is the left and
 * imgs[2p+1] the right image of pair p (2*npairs host images of one shape).  Extraction outputs as
 * orbb200_extract_batch (kps, desc, n_out for 2*npairs images); uRight / depth are host [npairs][cap], n_valid
 * [npairs].  Same results as orbb200_extract_batch followed by orbb200_stereo_match_batch; here each chunk's pairs
 * are matched on the chunk's stream right behind its extraction, inside the copy/compute pipeline. */
int orbb200_extract_stereo_batch(orbb200_extractor *e, const uint8_t *const *imgs, int npairs, int width, int height,
                                 int stride, orbb200_keypoint *kps, uint8_t *desc, int cap, int *n_out, float bf, float b,
                                 float *uRight, float *depth, int *n_valid);

/* begin / wait form of the two host-buffer batch calls, for callers that stream a sequence through one handle (a dataset
 * replay, a serving loop): *_begin submits the whole pipelined call and returns at once with a ticket; orbb200_extract_batch_wait
 * blocks until that call's results are in its host buffers (and reports ORBB200_ERR_CAPACITY like the synchronous call).  Up to
 * TWO calls may be in flight on a handle, so the input copies of call k + 1 run under the last kernels and downloads of call k
 * (a synchronous call pays its first copy and its last chunk's kernels as pure latency, and keeps its chunks small to keep them
 * short; begun calls use 64-image chunks: on one B200 the step of 160 KITTI pairs takes 4.4 ms synchronously and 3.6 ms streamed,
 * against 3.5 ms of kernel time with everything resident).  Rules: the host buffers of a begun call (inputs
 * and outputs) must not be touched until its wait; consecutive begun calls need different output buffers; a third begin first
 * completes the oldest call; calls of a different shape / batch / capacity complete the calls in flight before they start; any
 * other entry point taking the handle (orbb200_extract, *_device, orbb200_extractor_sync, destroy) completes them too.  The
 * synchronous calls are begin + wait.  (No counterpart in the reference, whose Frame constructor is synchronous.) */
int orbb200_extract_batch_begin(orbb200_extractor *e, const uint8_t *const *imgs, int batch, int width, int height, int stride,
                                orbb200_keypoint *kps, uint8_t *desc, int cap, int *n_out, int *ticket);
int orbb200_extract_stereo_batch_begin(orbb200_extractor *e, const uint8_t *const *imgs, int npairs, int width, int height,
                                       int stride, orbb200_keypoint *kps, uint8_t *desc, int cap, int *n_out, float bf, float b,
                                       float *uRight, float *depth, int *n_valid, int *ticket);
int orbb200_extract_batch_wait(orbb200_extractor *e, int ticket);

/* ---- DBoW2 vocabulary transform (Frame::ComputeBoW, src/Frame.cc:459-466; KeyFrame::ComputeBoW, src/KeyFrame.cc:59-69) ----
 * ORBVocabulary = DBoW2::TemplatedVocabulary<cv::Mat, FORB> (include/ORBVocabulary.h:31-32); file:line below are under
 * Thirdparty/DBoW2/DBoW2/.  scoring / weighting use the reference's enum values (BowVector.h:36-53):
 * weighting 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY; scoring 0 L1_NORM ... 5 DOT_PRODUCT. */
typedef struct orbb200_vocabulary orbb200_vocabulary; /* opaque; the tree resident in HBM + one CUDA stream */

/* replaces TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1358-1443): first line "k L scoring weighting",
 * then one line per node, ids 1, 2, ... in line order: "parent isLeaf b0 .. b31 weight".  Same header checks as the
 * reference (:1385-1389 -> ORBB200_ERR_ARG).  Blank lines are skipped: the reference turns the empty string after the
 * final newline into a phantom child of the root with an uninitialised descriptor (while(!f.eof()) + getline). */
int orbb200_vocabulary_load_text(orbb200_vocabulary **out, const char *path, int device);
/* the same tree from arrays (what the loader builds): node 0 is the root; for n in 1..nnodes-1 parent[n] < n, children
 * are ordered by id (m_nodes[pid].children.push_back(nid), :1411), is_leaf[n] != 0 marks a word and words are numbered
 * in id order (:1430-1437); desc is [nnodes][32], weight [nnodes] (rows 0 are ignored).  A word with children or an
 * inner node without children is rejected (the reference would walk off an empty child list). */
int orbb200_vocabulary_create(orbb200_vocabulary **out, int k, int L, int scoring, int weighting, int nnodes,
                              const int *parent, const uint8_t *is_leaf, const uint8_t *desc, const double *weight,
                              int device);
void orbb200_vocabulary_destroy(orbb200_vocabulary *v);
/* getBranchingFactor / getDepthLevels / m_scoring / m_weighting / m_nodes.size() / size() (:83, :146-152) */
int orbb200_vocabulary_info(const orbb200_vocabulary *v, int *k, int *L, int *scoring, int *weighting, int *nnodes,
                            int *nwords);

/* replaces transform(feature, word_id, weight, nid, levelsup) (:1218-1259) for n descriptors (host [n][32]): the walk
 * from the root to a word, nearest child first-on-ties at every node.  node_id[i] is the node of feature i's path at
 * level L - levelsup (0 when that level is <= 0); where the word lies above that level the reference leaves *nid
 * unwritten, here it is the word's own node id. */
int orbb200_vocabulary_transform(orbb200_vocabulary *v, const uint8_t *desc, int n, int levelsup, int *word_id,
                                 double *weight, int *node_id);
/* same with device pointers (desc_d 16-byte aligned, e.g. the descriptor block orbb200_extract_batch_device wrote),
 * asynchronous on `stream` (a cudaStream_t; NULL = the vocabulary's own stream). */
int orbb200_vocabulary_transform_device(orbb200_vocabulary *v, const uint8_t *desc_d, int n, int levelsup,
                                        int *word_id_d, double *weight_d, int *node_id_d, void *stream);

/* replaces transform(features, BowVector, FeatureVector, levelsup) (:1127-1190) -- what Frame::ComputeBoW calls with
 * levelsup = 4 -- for nimg images in one call: desc holds the images' descriptor rows back to back, counts[i] rows for
 * image i.  The tree walks of all images run as one launch; the std::map contents come back flattened in key order:
 *   BowVector of image i    = (bow_id[j], bow_val[j])                       for j in [bow_off[i], bow_off[i+1])
 *   FeatureVector of image i = fv_node[j] -> fv_feat[fv_feat_off[j] .. fv_feat_off[j+1])  for j in [fv_off[i], fv_off[i+1])
 * with feature indices local to the image and ascending (addFeature's push_back order, FeatureVector.cpp:31-45).
 * Values follow BowVector.cpp:34-84 in double, term by term in the reference's order (addWeight in feature order,
 * normalize in word order).  Every output array needs room for sum(counts) entries (+1 for the *_off arrays). */
int orbb200_compute_bow_batch(orbb200_vocabulary *v, const uint8_t *desc, const int *counts, int nimg, int levelsup,
                              int *bow_off, int *bow_id, double *bow_val, int *fv_off, int *fv_node, int *fv_feat_off,
                              int *fv_feat);

/* the same fold with everything resident in HBM, asynchronous on `stream` (a cudaStream_t; NULL = the vocabulary's
 * own): desc_d is [nimg][cap][32] with n_d[i] valid rows per image -- exactly what orbb200_extract_batch_device wrote --
 * so extraction -> ComputeStereoMatches -> ComputeBoW chain on one stream without a host round trip.  Outputs are
 * [nimg][cap] blocks: image i's BowVector is (bow_id_d, bow_val_d)[i*cap .. + n_bow_d[i]) in word order; its
 * FeatureVector is fv_key_d[i*cap .. + n_kept_d[i]) = (node id << 32 | feature index) ascending, i.e. node groups in
 * map order with ascending feature lists.  cap <= 8192.  The handle's per-feature scratch is shared: keep one call in
 * flight per vocabulary. */
int orbb200_compute_bow_batch_device(orbb200_vocabulary *v, const uint8_t *desc_d, const int *n_d, int nimg, int cap,
                                     int levelsup, int *bow_id_d, double *bow_val_d, int *n_bow_d,
                                     unsigned long long *fv_key_d, int *n_kept_d, void *stream);

/* ---- EuRoC prologue: stereo rectification (Examples/Stereo/stereo_euroc.cc:60-97, 131-132) ----
 * The example rectifies both images of every frame with cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) before TrackStereo, M1 / M2
 * from cv::initUndistortRectifyMap(K, D, R, P.rowRange(0,3).colRange(0,3), size, CV_32F, M1, M2) once per run. */
typedef struct orbb200_remap orbb200_remap; /* opaque: one (M1, M2) pair converted to OpenCV's fixed-point form, resident in HBM */
/* map_x / map_y: host [dst_height][dst_width] float (CV_32FC1 maps as initUndistortRectifyMap writes them) */
int orbb200_remap_create(orbb200_remap **out, const float *map_x, const float *map_y, int dst_width, int dst_height, int device);
void orbb200_remap_destroy(orbb200_remap *r);
/* replaces cv::remap(src, dst, M1, M2, cv::INTER_LINEAR) (BORDER_CONSTANT, value 0) for `batch` 8-bit one-channel images of one shape,
 * bit-exact with OpenCV's fixed-point path (5 fraction bits, 15-bit weights).  Host images in, host images out: */
int orbb200_remap_batch(orbb200_remap *r, const uint8_t *const *imgs, int batch, int src_width, int src_height, int src_stride,
                        uint8_t *const *out, int out_stride);
/* the same on device memory, asynchronous on `stream` (a cudaStream_t; NULL = the handle's own): image b at imgs_d + b *
 * src_image_stride, result b at out_d + b * out_image_stride -- chains in front of orbb200_extract_batch_device without a host
 * round trip */
int orbb200_remap_batch_device(orbb200_remap *r, const uint8_t *imgs_d, int batch, int src_width, int src_height, int src_stride,
                               size_t src_image_stride, uint8_t *out_d, int out_stride, size_t out_image_stride, void *stream);
/* replaces cv::initUndistortRectifyMap(K, D, R, P(3x3), (width, height), CV_32F, map_x, map_y): K, R, P row-major 3x3 doubles, D =
 * (k1, k2, p1, p2, k3) with n_dist <= 5 entries given.  Host arithmetic (runs once per sequence, needs no device). */
int orbb200_init_undistort_rectify_map(const double *K, const double *D, int n_dist, const double *R, const double *P, int width,
                                       int height, float *map_x, float *map_y);

/* measurement utility: sustained 32-bit POPC throughput of the device (ops/s), the denominator of the matching
 * roofline; runs a dependent-free popc microkernel for a few milliseconds. */
int orbb200_popc_peak(int device, double *popc32_per_s);

const char *orbb200_last_error(void);
const char *orbb200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBB200_H */
