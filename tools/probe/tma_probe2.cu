// Development probe: the CUDA programming guide's 2-D int32 TMA example, reduced.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include <cuda/barrier>
namespace cde = cuda::device::experimental;
using barrier_t = cuda::barrier<cuda::thread_scope_block>;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
template <typename T, int BW, int BH>
__global__ void k2(const __grid_constant__ CUtensorMap map, T* out, int x, int y) {
    __shared__ alignas(128) T tile[BH][BW];
    #pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier_t bar;
    if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
    __syncthreads();
    barrier_t::arrival_token token;
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_2d_global_to_shared(&tile, &map, x, y, bar);
        token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(tile));
    } else token = bar.arrive();
    bar.wait(std::move(token));
    for (int i = threadIdx.x; i < BW * BH; i += blockDim.x) out[i] = (&tile[0][0])[i];
}
template <typename T, int BW, int BH>
int run(CUtensorMapDataType dt, const char* name, int W, int H, int x, int y) {
    T* d; CK(cudaMalloc(&d, sizeof(T) * W * H));
    std::vector<T> h((size_t)W * H);
    for (size_t i = 0; i < h.size(); i++) h[i] = (T)(i * 2654435761u >> 13);
    CK(cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
    CUtensorMap m;
    cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H}, strides[1] = {(cuuint64_t)W * sizeof(T)};
    cuuint32_t box[2] = {BW, BH}, es[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(&m, dt, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                        CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    T* out; CK(cudaMalloc(&out, sizeof(T) * BW * BH));
    k2<T, BW, BH><<<1, 128>>>(m, out, x, y);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s %dx%d box %dx%d at (%d,%d): encode %d, run: %s\n", name, W, H, BW, BH, x, y, (int)r, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<T> ho(BW * BH);
    CK(cudaMemcpy(ho.data(), out, ho.size() * sizeof(T), cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int j = 0; j < BH; j++) for (int i = 0; i < BW; i++) { T want = (y + j < H && x + i < W) ? h[(size_t)(y + j) * W + x + i] : (T)0; bad += ho[j * BW + i] != want; }
    printf("  mismatches %d\n", bad);
    return 0;
}
int main(int argc, char** argv) {
    int which = argc > 1 ? atoi(argv[1]) : 0;
    if (which == 0) return run<int, 32, 32>(CU_TENSOR_MAP_DATA_TYPE_INT32, "int32", 1024, 1024, 64, 96);
    if (which == 1) return run<unsigned char, 64, 64>(CU_TENSOR_MAP_DATA_TYPE_UINT8, "u8", 1024, 1024, 64, 96);
    if (which == 2) return run<unsigned char, 80, 64>(CU_TENSOR_MAP_DATA_TYPE_UINT8, "u8", 816, 518, 220, 463);
    if (which == 3) return run<unsigned char, 80, 64>(CU_TENSOR_MAP_DATA_TYPE_UINT8, "u8", 816, 518, 220, 100);
    return 0;
}
