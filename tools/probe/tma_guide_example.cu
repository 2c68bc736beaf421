// tma_probe3.cu -- the CUDA programming guide's own TMA example (libcu++ barrier + cde::cp_async_bulk_tensor_2d_global_to_shared)
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdio>
#include <vector>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
constexpr int W = 64, H = 32;
__global__ void k(const __grid_constant__ CUtensorMap tensor_map, int x, int y, unsigned char *out)
{
    __shared__ alignas(128) unsigned char smem_buffer[H][W];
#pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier bar;
    if (threadIdx.x == 0) {
        init(&bar, blockDim.x);
        cde::fence_proxy_async_shared_cta();
    }
    __syncthreads();
    barrier::arrival_token token;
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_2d_global_to_shared(&smem_buffer, &tensor_map, x, y, bar);
        token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(smem_buffer));
    } else {
        token = bar.arrive();
    }
    bar.wait(std::move(token));
    for (int i = threadIdx.x; i < W * H; i += blockDim.x) out[i] = smem_buffer[i / W][i % W];
}
typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                        const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main()
{
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaFree(0);
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    const int pitch = 1248, HH = 376;
    std::vector<unsigned char> h((size_t)pitch * HH);
    for (size_t i = 0; i < h.size(); i++) h[i] = (unsigned char)((i * 2654435761u) >> 13);
    unsigned char *d, *dout;
    cudaMalloc(&d, h.size());
    cudaMalloc(&dout, W * H);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    CUtensorMap M{};
    const cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)HH}, strides[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {W, H}, es[2] = {1, 1};
    CUresult r = ((Enc)p)(&M, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                          CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d; ", (int)r);
    k<<<1, 128>>>(M, 100, 50, dout);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); return 1; }
    std::vector<unsigned char> o(W * H);
    cudaMemcpy(o.data(), dout, o.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < H; rr++) for (int c = 0; c < W; c++) bad += o[rr * W + c] != h[(size_t)(50 + rr) * pitch + 100 + c];
    printf("%s (%d mismatches)\n", bad ? "WRONG" : "ok", bad);
    return bad != 0;
}
