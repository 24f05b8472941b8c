#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdio>
#include <vector>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
__global__ void k3(const __grid_constant__ CUtensorMap map, int x, int y, int z, uint8_t* out) {
  __shared__ alignas(128) uint8_t smem[48 * 43];
#pragma nv_diag_suppress static_var_with_dynamic_init
  __shared__ barrier bar;
  if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
  __syncthreads();
  barrier::arrival_token token;
  if (threadIdx.x == 0) {
    cde::cp_async_bulk_tensor_3d_global_to_shared(&smem, &map, x, y, z, bar);
    token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(smem));
  } else token = bar.arrive();
  bar.wait(std::move(token));
  for (int i = threadIdx.x; i < 48 * 43; i += 32) out[i] = smem[i];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
  const int pitch = 1280, rows = 414, slots = 3; const size_t stride = (size_t)pitch * rows;
  std::vector<uint8_t> h(stride * slots); for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
  uint8_t *d, *dout; cudaMalloc(&d, h.size()); cudaMalloc(&dout, 48 * 43); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  void* p = nullptr; cudaDriverEntryPointQueryResult q; cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  EncodeTiledFn fn = (EncodeTiledFn)p; CUtensorMap m;
  cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)slots}; cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)stride};
  cuuint32_t box[3] = {48, 43, 1}, es[3] = {1, 1, 1};
  CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode %d\n", (int)r);
  const int xs[6] = {64, 1008, 1264, 64, 1264, 0}, ys[6] = {32, 101, 100, 400, 400, 0};
  for (int trial = 0; trial < 6; ++trial) {
    const int x = xs[trial], y = ys[trial], z = 2;
    k3<<<1, 32>>>(m, x, y, z, dout);
    cudaError_t e = cudaDeviceSynchronize(); if (e != cudaSuccess) { printf("trial %d (x=%d,y=%d) CUDA error %s\n", trial, x, y, cudaGetErrorString(e)); return 1; }
    std::vector<uint8_t> o(48 * 43); cudaMemcpy(o.data(), dout, o.size(), cudaMemcpyDeviceToHost);
    int bad = 0; for (int rr = 0; rr < 43; ++rr) for (int c = 0; c < 48; ++c) { uint8_t want = (x + c < pitch && y + rr < rows) ? h[(size_t)z * stride + (size_t)(y + rr) * pitch + x + c] : 0; bad += o[rr * 48 + c] != want; }
    printf("trial %d (x=%d,y=%d): %d mismatches\n", trial, x, y, bad);
  }
  return 0;
}
