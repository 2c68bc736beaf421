// orbb200_remap.cu -- the EuRoC prologue: cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) of every stereo frame
// (Examples/Stereo/stereo_euroc.cc:131-132) and the once-per-run cv::initUndistortRectifyMap (:96-97).
//
// cv::remap with CV_32FC1 maps runs in fixed point (OpenCV imgwarp / remap): sx = cvRound(map1 * 32), sy = cvRound(map2 * 32),
// integer part sx >> 5 saturated to short, fraction index (sy & 31) * 32 + (sx & 31) into a table of four 15-bit weights
// cvRound(float((1 - fy)(1 - fx)) * 32768) ..., dst = (sum w_i p_i + 16384) >> 15, pixels outside the source = 0 (BORDER_CONSTANT).
//   k_remap_prepare  once per map: the two float maps -> one packed record per destination pixel (ix | iy << 16, fraction index)
//   k_remap_linear   per batch: a thread per 4 destination pixels of a row: 4 records, 16 byte gathers (L2 resident source rows:
//                    rectification moves a pixel by a few rows at most), weights from a shared-memory copy of the 8 KB table,
//                    one 32-bit store.  HBM bound: 8 B of record + 1 B written + ~1 B read per pixel.
#include <math.h>

#include <mutex>
#include <new>

#include "common.cuh"

namespace orbb200 {

struct RemapRec {
    int xy;  // ix (low 16, signed) | iy << 16
    int fi;  // fraction index 0..1023
};

__global__ void k_remap_prepare(const float *__restrict__ mapx, const float *__restrict__ mapy, int n, RemapRec *__restrict__ rec)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // cvRound = round half to even; saturate_cast<int> of the product first (values beyond int range are outside any image anyway)
    const float fx = __fmul_rn(mapx[i], 32.f), fy = __fmul_rn(mapy[i], 32.f);
    const int sx = __float2int_rn(fminf(fmaxf(fx, -2147483520.f), 2147483520.f)), sy = __float2int_rn(fminf(fmaxf(fy, -2147483520.f), 2147483520.f));
    const int ix = min(max(sx >> 5, -32768), 32767), iy = min(max(sy >> 5, -32768), 32767);
    RemapRec r;
    r.xy = (ix & 0xFFFF) | (iy << 16);
    r.fi = (sy & 31) * 32 + (sx & 31);
    rec[i] = r;
}

constexpr int RM_THREADS = 256;

__global__ void __launch_bounds__(RM_THREADS) k_remap_linear(const uint8_t *__restrict__ src, int sw, int sh, int sstride, size_t sImgStride,
                                                            const RemapRec *__restrict__ rec, const short4 *__restrict__ tab, int dw, int dh,
                                                            uint8_t *__restrict__ dst, int dstride, size_t dImgStride)
{
    __shared__ short4 sTab[1024];
    for (int i = threadIdx.x; i < 1024; i += RM_THREADS) sTab[i] = tab[i];
    __syncthreads();
    const int quads = (dw + 3) >> 2;
    const int q = blockIdx.x * RM_THREADS + threadIdx.x;
    if (q >= quads * dh) return;
    const int y = q / quads, x0 = (q - y * quads) << 2;
    const uint8_t *S = src + (size_t)blockIdx.y * sImgStride;
    uint8_t *D = dst + (size_t)blockIdx.y * dImgStride + (size_t)y * dstride + x0;
    unsigned int out = 0;
    // the four records of the thread: two 16-byte loads when the row is aligned and whole, else one by one
    RemapRec rr[4];
    const RemapRec *rp = rec + (size_t)y * dw + x0;
    if (x0 + 3 < dw && (((size_t)rp) & 15) == 0) {
        const int4 a = __ldg(reinterpret_cast<const int4 *>(rp)), c = __ldg(reinterpret_cast<const int4 *>(rp) + 1);
        rr[0].xy = a.x; rr[0].fi = a.y; rr[1].xy = a.z; rr[1].fi = a.w; rr[2].xy = c.x; rr[2].fi = c.y; rr[3].xy = c.z; rr[3].fi = c.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++) rr[k] = x0 + k < dw ? rp[k] : RemapRec{0, 0};
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (x0 + k >= dw) break;
        const RemapRec r = rr[k];
        const int ix = (int)(short)(r.xy & 0xFFFF), iy = r.xy >> 16;
        const short4 w = sTab[r.fi];
        const bool x0in = ix >= 0 && ix < sw, x1in = ix + 1 >= 0 && ix + 1 < sw, y0in = iy >= 0 && iy < sh, y1in = iy + 1 >= 0 && iy + 1 < sh;
        const uint8_t *p = S + (size_t)iy * sstride + ix;
        const int p00 = (x0in && y0in) ? p[0] : 0, p01 = (x1in && y0in) ? p[1] : 0;
        const int p10 = (x0in && y1in) ? p[sstride] : 0, p11 = (x1in && y1in) ? p[sstride + 1] : 0;
        int v = (p00 * w.x + p01 * w.y + p10 * w.z + p11 * w.w + 16384) >> 15;
        v = min(max(v, 0), 255);
        out |= (unsigned int)v << (8 * k);
    }
    if (x0 + 3 < dw && ((size_t)D & 3) == 0) *reinterpret_cast<unsigned int *>(D) = out;
    else
        for (int k = 0; k < 4 && x0 + k < dw; k++) D[k] = (uint8_t)(out >> (8 * k));
}

} // namespace orbb200

struct orbb200_remap {
    int device = 0, dw = 0, dh = 0;
    orbb200::RemapRec *d_rec = nullptr;
    short4 *d_tab = nullptr;
    cudaStream_t stream = nullptr;
    uint8_t *d_src = nullptr, *d_dst = nullptr; // staging of the host-buffer call, grown on demand
    size_t srcCap = 0, dstCap = 0;
};

using namespace orbb200;

extern "C" {

int orbb200_remap_create(orbb200_remap **out, const float *map_x, const float *map_y, int dst_width, int dst_height, int device)
{
    ORB_ARG_OK(out != nullptr, "out is null");
    *out = nullptr;
    ORB_ARG_OK(map_x && map_y && dst_width > 0 && dst_height > 0, "bad arguments");
    if (device < 0) ORB_CUDA_OK(cudaGetDevice(&device));
    ORB_CUDA_OK(cudaSetDevice(device));
    orbb200_remap *r = new (std::nothrow) orbb200_remap();
    ORB_ARG_OK(r != nullptr, "out of host memory");
    r->device = device; r->dw = dst_width; r->dh = dst_height;
    const size_t n = (size_t)dst_width * dst_height;
    // the weight table, exactly as initInterTab2D builds it for INTER_LINEAR (float products, cvRound of w * 32768)
    short4 tab[1024];
    float t1[32][2];
    for (int i = 0; i < 32; i++) { const float x = (float)i * (1.0f / 32); t1[i][0] = 1.0f - x; t1[i][1] = x; }
    for (int i = 0; i < 32; i++)
        for (int j = 0; j < 32; j++) {
            const float f[4] = {t1[i][0] * t1[j][0], t1[i][0] * t1[j][1], t1[i][1] * t1[j][0], t1[i][1] * t1[j][1]};
            short w[4];
            for (int k = 0; k < 4; k++) { const long v = lrintf(f[k] * 32768.f); w[k] = (short)(v > 32767 ? 32767 : v < -32768 ? -32768 : v); }
            tab[i * 32 + j] = make_short4(w[0], w[1], w[2], w[3]);
        }
    float *dmx = nullptr, *dmy = nullptr;
    cudaError_t ce = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaMalloc((void **)&r->d_rec, n * sizeof(RemapRec));
    if (ce == cudaSuccess) ce = cudaMalloc((void **)&r->d_tab, sizeof(tab));
    if (ce == cudaSuccess) ce = cudaMalloc((void **)&dmx, n * sizeof(float));
    if (ce == cudaSuccess) ce = cudaMalloc((void **)&dmy, n * sizeof(float));
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(r->d_tab, tab, sizeof(tab), cudaMemcpyHostToDevice, r->stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(dmx, map_x, n * sizeof(float), cudaMemcpyHostToDevice, r->stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(dmy, map_y, n * sizeof(float), cudaMemcpyHostToDevice, r->stream);
    if (ce == cudaSuccess) {
        k_remap_prepare<<<(unsigned int)((n + 255) / 256), 256, 0, r->stream>>>(dmx, dmy, (int)n, r->d_rec);
        ce = cudaStreamSynchronize(r->stream);
    }
    if (dmx) cudaFree(dmx);
    if (dmy) cudaFree(dmy);
    if (ce != cudaSuccess) {
        set_error("orbb200_remap_create: %s", cudaGetErrorString(ce));
        orbb200_remap_destroy(r);
        return ORBB200_ERR_CUDA;
    }
    *out = r;
    return ORBB200_OK;
}

void orbb200_remap_destroy(orbb200_remap *r)
{
    if (!r) return;
    cudaSetDevice(r->device);
    if (r->stream) { cudaStreamSynchronize(r->stream); cudaStreamDestroy(r->stream); }
    if (r->d_rec) cudaFree(r->d_rec);
    if (r->d_tab) cudaFree(r->d_tab);
    if (r->d_src) cudaFree(r->d_src);
    if (r->d_dst) cudaFree(r->d_dst);
    delete r;
}

int orbb200_remap_batch_device(orbb200_remap *r, const uint8_t *imgs_d, int batch, int src_width, int src_height, int src_stride,
                               size_t src_image_stride, uint8_t *out_d, int out_stride, size_t out_image_stride, void *stream)
{
    ORB_ARG_OK(r != nullptr, "null handle");
    if (batch <= 0) return ORBB200_OK;
    ORB_ARG_OK(imgs_d && out_d && src_width > 0 && src_height > 0 && src_stride >= src_width && out_stride >= r->dw && batch <= 65535,
               "bad arguments");
    ORB_CUDA_OK(cudaSetDevice(r->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : r->stream;
    const int quads = ((r->dw + 3) >> 2) * r->dh;
    k_remap_linear<<<dim3(div_up(quads, RM_THREADS), batch), RM_THREADS, 0, st>>>(imgs_d, src_width, src_height, src_stride, src_image_stride, r->d_rec,
                                                                                 r->d_tab, r->dw, r->dh, out_d, out_stride, out_image_stride);
    ORB_CUDA_OK(cudaGetLastError());
    return ORBB200_OK;
}

int orbb200_remap_batch(orbb200_remap *r, const uint8_t *const *imgs, int batch, int src_width, int src_height, int src_stride,
                        uint8_t *const *out, int out_stride)
{
    ORB_ARG_OK(r != nullptr, "null handle");
    if (batch <= 0) return ORBB200_OK;
    ORB_ARG_OK(imgs && out && src_width > 0 && src_height > 0 && src_stride >= src_width && out_stride >= r->dw, "bad arguments");
    for (int b = 0; b < batch; b++) ORB_ARG_OK(imgs[b] && out[b], "null image");
    ORB_CUDA_OK(cudaSetDevice(r->device));
    const size_t sImg = (size_t)src_width * src_height, dImg = (size_t)r->dw * r->dh;
    if (r->srcCap < sImg * batch) {
        ORB_CUDA_OK(cudaStreamSynchronize(r->stream));
        if (r->d_src) cudaFree(r->d_src);
        r->d_src = nullptr; r->srcCap = 0;
        ORB_CUDA_OK(cudaMalloc((void **)&r->d_src, sImg * batch + 256));
        r->srcCap = sImg * batch;
    }
    if (r->dstCap < dImg * batch) {
        ORB_CUDA_OK(cudaStreamSynchronize(r->stream));
        if (r->d_dst) cudaFree(r->d_dst);
        r->d_dst = nullptr; r->dstCap = 0;
        ORB_CUDA_OK(cudaMalloc((void **)&r->d_dst, dImg * batch + 256));
        r->dstCap = dImg * batch;
    }
    for (int b = 0; b < batch; b++)
        ORB_CUDA_OK(cudaMemcpy2DAsync(r->d_src + (size_t)b * sImg, src_width, imgs[b], src_stride, src_width, src_height, cudaMemcpyHostToDevice,
                                      r->stream));
    int rc = orbb200_remap_batch_device(r, r->d_src, batch, src_width, src_height, src_width, sImg, r->d_dst, r->dw, dImg, r->stream);
    if (rc) return rc;
    for (int b = 0; b < batch; b++)
        ORB_CUDA_OK(cudaMemcpy2DAsync(out[b], out_stride, r->d_dst + (size_t)b * dImg, r->dw, r->dw, r->dh, cudaMemcpyDeviceToHost, r->stream));
    ORB_CUDA_OK(cudaStreamSynchronize(r->stream));
    return ORBB200_OK;
}

/* host arithmetic only (doubles, one narrowing per entry): usable without a device */
int orbb200_init_undistort_rectify_map(const double *K, const double *D, int n_dist, const double *R, const double *P, int width, int height,
                                       float *map_x, float *map_y)
{
    ORB_ARG_OK(K && R && P && map_x && map_y && width > 0 && height > 0 && n_dist >= 0 && n_dist <= 5 && (D || n_dist == 0), "bad arguments");
    double d[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < n_dist; i++) d[i] = D[i];
    double M[9], ir[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) M[3 * i + j] = P[3 * i] * R[j] + P[3 * i + 1] * R[3 + j] + P[3 * i + 2] * R[6 + j];
    const double det = M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
    ORB_ARG_OK(det != 0.0, "P * R is singular");
    const double id = 1.0 / det;
    ir[0] = (M[4] * M[8] - M[5] * M[7]) * id; ir[1] = (M[2] * M[7] - M[1] * M[8]) * id; ir[2] = (M[1] * M[5] - M[2] * M[4]) * id;
    ir[3] = (M[5] * M[6] - M[3] * M[8]) * id; ir[4] = (M[0] * M[8] - M[2] * M[6]) * id; ir[5] = (M[2] * M[3] - M[0] * M[5]) * id;
    ir[6] = (M[3] * M[7] - M[4] * M[6]) * id; ir[7] = (M[1] * M[6] - M[0] * M[7]) * id; ir[8] = (M[0] * M[4] - M[1] * M[3]) * id;
    const double fx = K[0], fy = K[4], u0 = K[2], v0 = K[5];
    const double k1 = d[0], k2 = d[1], p1 = d[2], p2 = d[3], k3 = d[4];
    for (int i = 0; i < height; i++) {
        double _x = i * ir[1] + ir[2], _y = i * ir[4] + ir[5], _w = i * ir[7] + ir[8];
        for (int j = 0; j < width; j++, _x += ir[0], _y += ir[3], _w += ir[6]) {
            const double ww = 1. / _w, x = _x * ww, y = _y * ww;
            const double x2 = x * x, y2 = y * y, r2 = x2 + y2, _2xy = 2 * x * y;
            const double kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
            const double xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2), yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy;
            map_x[(size_t)i * width + j] = (float)(fx * xd + u0);
            map_y[(size_t)i * width + j] = (float)(fy * yd + v0);
        }
    }
    return ORBB200_OK;
}

} // extern "C"
