// tma_probe2.cu -- which tensor-map shapes does cp.async.bulk.tensor accept here?  argv: rank boxW boxH
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
struct alignas(64) Map { unsigned char m[128]; };
template <int RANK> __global__ void k(const __grid_constant__ Map T, int x, int y, int img, int bytes, unsigned char *out)
{
    extern __shared__ __align__(128) unsigned char buf[];
    __shared__ __align__(8) unsigned long long bar;
    const unsigned b = (unsigned)__cvta_generic_to_shared(&bar), d = (unsigned)__cvta_generic_to_shared(buf);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        if (RANK == 3)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(d), "l"(&T), "r"(x), "r"(y), "r"(img), "r"(b) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(d), "l"(&T), "r"(x), "r"(y), "r"(b) : "memory");
    }
    unsigned done = 0;
    while (!done) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(b), "r"(0) : "memory");
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = buf[i];
}
typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                        const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char **argv)
{
    const int rank = atoi(argv[1]), bw = atoi(argv[2]), bh = atoi(argv[3]);
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaFree(0);
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    const int pitch = 1248, H = 376, nimg = 3;
    const size_t istride = ((size_t)pitch * H + 255) / 256 * 256;
    std::vector<unsigned char> h(istride * nimg);
    for (size_t i = 0; i < h.size(); i++) h[i] = (unsigned char)((i * 2654435761u) >> 13);
    unsigned char *d, *dout;
    cudaMalloc(&d, h.size());
    cudaMalloc(&dout, 65536);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    Map M;
    const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)H, (cuuint64_t)nimg}, strides[2] = {(cuuint64_t)pitch, (cuuint64_t)istride};
    const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
    CUresult r = ((Enc)p)((CUtensorMap *)M.m, CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("rank %d box %dx%d: encode %d; ", rank, bw, bh, (int)r);
    if (r) { printf("\n"); return 1; }
    const int x = 100, y = 50, img = rank == 3 ? 1 : 0, bytes = bw * bh;
    if (rank == 3) k<3><<<1, 128, bytes + 128>>>(M, x, y, img, bytes, dout); else k<2><<<1, 128, bytes + 128>>>(M, x, y, img, bytes, dout);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); return 1; }
    std::vector<unsigned char> o(bytes);
    cudaMemcpy(o.data(), dout, bytes, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < bh; rr++)
        for (int c = 0; c < bw; c++) bad += o[rr * bw + c] != h[(size_t)img * istride + (size_t)(y + rr) * pitch + x + c];
    printf("%s (%d mismatches)\n", bad ? "WRONG" : "ok", bad);
    return bad != 0;
}
