// stereo_kernel.cuh -- F1: Frame::ComputeStereoMatches (src/Frame.cc:577-751) on the device-resident outputs and
// pyramids of a left and a right extractor handle.
//   k_stereo_match   one warp per left keypoint: lanes stride the right keypoints in index order (the row-band
//                    lists of :594-604 hold right indices in increasing order, so the first-wins arg-min of
//                    :639-660 is the lexicographic (distance, iR) minimum); then 11 SAD windows of 11x11 on the
//                    un-blurred level (:672-703), parabola refinement (:709-716) and the disparity gate (:719-733).
//   k_stereo_median  one block: median of the SAD distances and the 1.5*1.4*median outlier cut (:737-750).
#pragma once
#include "extractor_host.cuh"
#include "match_kernels.cuh"

namespace orbb200 {

struct StereoSide {
    const orbb200_keypoint *kps;
    const uint8_t *desc;
    const int *n;
    const uint8_t *l0;
    int l0Pitch;
    const uint8_t *pyr;
    // batch strides from pair p to pair p+1 (0 for the single-pair call)
    size_t kpStride, nStride, l0Stride, pyrStride;
};

__device__ __forceinline__ StereoSide stereo_side(const StereoSide &s, int pair)
{
    StereoSide r = s;
    r.kps += pair * s.kpStride;
    r.desc += pair * s.kpStride * 32;
    r.n += pair * s.nStride;
    r.l0 += pair * s.l0Stride;
    r.pyr += pair * s.pyrStride;
    return r;
}

struct StereoParams {
    StereoSide L, R;
    const LevelGeom *geom;
    int nlevels, nRows;
    float bf, b;
    float sf[ORBB200_MAX_LEVELS], isf[ORBB200_MAX_LEVELS];
    float *uRight, *depth;
    int *sad; // SAD distance per left keypoint, -1 when no depth
    // row buckets of the right keypoints (:594-604): rowStart[pair][nRows+1], rowList[pair][rowListStride]
    int *rowStart;
    int2 *rowList; // entry = (iR | octave << 24, float bits of the keypoint's x): the matcher's gates need no second gather
    size_t rowListStride;
    size_t outStride; // per-pair stride of uRight / depth / sad
    int cap;          // keypoint capacity per image (counts are clamped to it)
};

__device__ __forceinline__ const uint8_t *stereo_level(const StereoSide &s, const LevelGeom &g, int l, int &pitch)
{
    if (l == 0) { pitch = s.l0Pitch; return s.l0; }
    pitch = g.pitch;
    return s.pyr + g.imgOff;
}

// Row buckets of the right keypoints (src/Frame.cc:594-604): keypoint iR is listed in every image row of
// [floor(y - 2*scale[octave]), ceil(y + 2*scale[octave])].  One block per pair; counts and fill cursors live in
// shared memory.  The order inside a row is not the reference's push_back order -- the matcher below takes the
// lexicographic (distance, iR) minimum, which is order independent and equals the reference's first-wins scan of
// a list sorted by iR.
__global__ void __launch_bounds__(512) k_stereo_rows(const __grid_constant__ StereoParams P)
{
    extern __shared__ int srow[]; // [nRows] counts, then cursors
    __shared__ int scanTmp[34];
    const int pair = blockIdx.x, tid = threadIdx.x, NT = blockDim.x;
    const StereoSide SR = stereo_side(P.R, pair);
    const int nR = min(*SR.n, P.cap), nRows = P.nRows;
    int *rowStart = P.rowStart + (size_t)pair * (nRows + 1);
    int2 *rowList = P.rowList + (size_t)pair * P.rowListStride;
    for (int r = tid; r < nRows; r += NT) srow[r] = 0;
    __syncthreads();
    for (int iR = tid; iR < nR; iR += NT) {
        const orbb200_keypoint kR = SR.kps[iR];
        const float r = __fmul_rn(2.0f, P.sf[kR.octave]);
        const int maxr = min((int)ceilf(__fadd_rn(kR.y, r)), nRows - 1), minr = max((int)floorf(__fsub_rn(kR.y, r)), 0);
        for (int yi = minr; yi <= maxr; yi++) atomicAdd(&srow[yi], 1);
    }
    __syncthreads();
    int run = 0;
    for (int base = 0; base < nRows; base += NT) {
        const int r = base + tid;
        const int c = r < nRows ? srow[r] : 0;
        int tot;
        const int ex = block_excl_scan(c, scanTmp, &tot);
        if (r < nRows) { rowStart[r] = run + ex; srow[r] = run + ex; }
        run += tot;
    }
    if (tid == 0) rowStart[nRows] = run;
    __syncthreads();
    for (int iR = tid; iR < nR; iR += NT) {
        const orbb200_keypoint kR = SR.kps[iR];
        const float r = __fmul_rn(2.0f, P.sf[kR.octave]);
        const int maxr = min((int)ceilf(__fadd_rn(kR.y, r)), nRows - 1), minr = max((int)floorf(__fsub_rn(kR.y, r)), 0);
        for (int yi = minr; yi <= maxr; yi++) {
            const int pos = atomicAdd(&srow[yi], 1);
            if ((size_t)pos < P.rowListStride) rowList[pos] = make_int2(iR | (kR.octave << 24), __float_as_int(kR.x));
        }
    }
}

__global__ void __launch_bounds__(256, 8) k_stereo_match(const __grid_constant__ StereoParams P)
{
    const int lane = threadIdx.x & 31;
    const int iL = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int pair = blockIdx.y;
    const StereoSide SL = stereo_side(P.L, pair), SR = stereo_side(P.R, pair);
    const int nL = min(*SL.n, P.cap);
    if (iL >= nL) return;
    const orbb200_keypoint kL = SL.kps[iL];
    float uRight = -1.f, depth = -1.f;
    int sadOut = -1;
    const int levelL = kL.octave;
    const float vL = kL.y, uL = kL.x;
    const int row = (int)vL;
    const float maxD = __fdiv_rn(P.bf, P.b);
    const float minU = __fsub_rn(uL, maxD), maxU = uL;
    Top2 t;
    top2_init(t);
    t.d1 = 100; // bestDist starts at TH_HIGH (:633)
    float bestX = 0.f;
    if (row >= 0 && row < P.nRows && !(maxU < 0)) {
        const Desc dL = load_desc(SL.desc + (size_t)iL * 32);
        const int *rowStart = P.rowStart + (size_t)pair * (P.nRows + 1);
        const int2 *rowList = P.rowList + (size_t)pair * P.rowListStride;
        const int c0 = rowStart[row], c1 = min(rowStart[row + 1], (int)P.rowListStride);
        for (int c = c0 + lane; c < c1; c += 32) {
            const int2 ent = rowList[c];
            const int iR = ent.x & 0xFFFFFF, octR = ent.x >> 24;
            if (octR < levelL - 1 || octR > levelL + 1) continue;
            const float uR = __int_as_float(ent.y);
            if (uR >= minU && uR <= maxU) {
                const int d = hamming(dL, load_desc(SR.desc + (size_t)iR * 32));
                if (d < t.d1 || (d == t.d1 && t.i1 >= 0 && iR < t.i1)) { t.d1 = d; t.i1 = iR; bestX = uR; }
            }
        }
    }
    // warp arg-min under (distance, iR); the winner's x travels with it
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const int od = __shfl_xor_sync(0xffffffffu, t.d1, o), oi = __shfl_xor_sync(0xffffffffu, t.i1, o);
        const float ox = __shfl_xor_sync(0xffffffffu, bestX, o);
        if (lex_less(od, oi, t.d1, t.i1)) { t.d1 = od; t.i1 = oi; bestX = ox; }
    }
    if (t.i1 >= 0 && t.d1 < 75) { // thOrbDist = (TH_HIGH+TH_LOW)/2 (:582,663)
        const float uR0 = bestX;
        const float scaleFactor = P.isf[levelL];
        const float scaleduL = roundf(__fmul_rn(uL, scaleFactor)), scaledvL = roundf(__fmul_rn(vL, scaleFactor));
        const float scaleduR0 = roundf(__fmul_rn(uR0, scaleFactor));
        const LevelGeom g = P.geom[levelL];
        const int W = g.w, H = g.h;
        const int cy = (int)scaledvL, cxl = (int)scaleduL, cxr = (int)scaleduR0;
        const float iniu = scaleduR0, endu = __fadd_rn(scaleduR0, 11.f);
        const bool inRef = !(iniu < 0 || endu >= (float)W);
        const bool inImg = !(cy - 5 < 0 || cy + 5 >= H || cxl - 5 < 0 || cxl + 5 >= W || cxr - 10 < 0 || cxr + 10 >= W);
        if (inRef && inImg) {
            int pl, pr;
            const uint8_t *IL = stereo_level(SL, g, levelL, pl), *IR = stereo_level(SR, g, levelL, pr);
            const int lc = IL[(size_t)cy * pl + cxl];
            // each lane owns up to 4 of the 121 patch pixels; its left values are inc-independent
            int lv[4], off[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int p = lane + 32 * k;
                const int dy = p / 11 - 5, dx = p % 11 - 5;
                lv[k] = p < 121 ? (int)IL[(size_t)(cy + dy) * pl + cxl + dx] - lc : 0;
                off[k] = p < 121 ? (cy + dy) * pr + cxr + dx : -1;
            }
            int best = 0x7fffffff, bestinc = 0, dists[11];
#pragma unroll
            for (int inc = -5; inc <= 5; inc++) {
                const int rc = IR[(size_t)cy * pr + cxr + inc];
                int s = 0;
#pragma unroll
                for (int k = 0; k < 4; k++)
                    if (off[k] >= 0) s += abs(lv[k] - ((int)IR[off[k] + inc] - rc));
                s = warp_sum(s);
                dists[inc + 5] = s;
                if (s < best) { best = s; bestinc = inc; }
            }
            if (bestinc != -5 && bestinc != 5) {
                float d1 = 0, d2 = 0, d3 = 0;
#pragma unroll
                for (int k = 1; k < 10; k++)
                    if (k == bestinc + 5) { d1 = (float)dists[k - 1]; d2 = (float)dists[k]; d3 = (float)dists[k + 1]; }
                const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
                if (!(deltaR < -1 || deltaR > 1)) {
                    float bestuR = __fmul_rn(P.sf[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));
                    float disparity = __fsub_rn(uL, bestuR);
                    if (disparity >= 0 && disparity < maxD) {
                        if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
                        depth = __fdiv_rn(P.bf, disparity);
                        uRight = bestuR;
                        sadOut = best;
                    }
                }
            }
        }
    }
    if (lane == 0) {
        const size_t o = (size_t)pair * P.outStride + iL;
        P.uRight[o] = uRight; P.depth[o] = depth; P.sad[o] = sadOut;
    }
}

__global__ void __launch_bounds__(1024) k_stereo_median(const __grid_constant__ StereoParams P, int *nValid)
{
    __shared__ int s_median, s_cnt, s_removed;
    const int pair = blockIdx.x;
    const int nL = min(P.L.n[pair * P.L.nStride], P.cap);
    int *sad = P.sad + (size_t)pair * P.outStride;
    float *uRight = P.uRight + (size_t)pair * P.outStride, *depth = P.depth + (size_t)pair * P.outStride;
    nValid += pair;
    if (threadIdx.x == 0) { s_median = -1; s_cnt = 0; s_removed = 0; }
    __syncthreads();
    int c = 0;
    for (int i = threadIdx.x; i < nL; i += blockDim.x) c += sad[i] >= 0;
    if (c) atomicAdd(&s_cnt, c);
    __syncthreads();
    const int n = s_cnt, k = n / 2;
    if (n == 0) { if (threadIdx.x == 0) *nValid = 0; return; }
    // k-th smallest SAD (0-based) by a two-level 256-bin histogram select (values < 2^16: 121 * 510)
    __shared__ int hist[256];
    __shared__ int s_bin, s_rank;
    for (int pass = 0; pass < 2; pass++) {
        for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        for (int i = threadIdx.x; i < nL; i += blockDim.x) {
            const int v = sad[i];
            if (v < 0) continue;
            if (pass == 0) atomicAdd(&hist[min(v >> 8, 255)], 1);
            else if (min(v >> 8, 255) == s_bin) atomicAdd(&hist[v & 255], 1);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int rank = pass == 0 ? k : s_rank, bsel = 255;
            for (int i = 0, acc = 0; i < 256; i++) {
                if (acc + hist[i] > rank) { bsel = i; rank -= acc; break; }
                acc += hist[i];
            }
            if (pass == 0) { s_bin = bsel; s_rank = rank; }
            else s_median = (s_bin << 8) | bsel;
        }
        __syncthreads();
    }
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)s_median);
    int removed = 0;
    for (int i = threadIdx.x; i < nL; i += blockDim.x) {
        const int v = sad[i];
        if (v >= 0 && !((float)v < thDist)) { uRight[i] = -1.f; depth[i] = -1.f; removed++; }
    }
    if (removed) atomicAdd(&s_removed, removed);
    __syncthreads();
    if (threadIdx.x == 0) *nValid = n - s_removed;
}

int stereo_match_impl(orbb200_extractor *left, orbb200_extractor *right, float bf, float b, float *uRight, float *depth,
                      int *n_valid);
int stereo_batch_device_impl(orbb200_extractor *e, int npairs, float bf, float b, float *uRight_d, float *depth_d,
                             int *n_valid_d);
int stereo_batch_host_impl(orbb200_extractor *e, int npairs, float bf, float b, float *uRight, float *depth, int *n_valid);

} // namespace orbb200
