// Stand-alone probe: which ways of handing a CUtensorMap to cp.async.bulk.tensor work on this box?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tma_probe tools/tma_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include "../orb_slam2_b200/csrc/tma_utils.cuh"
using namespace orbx;
struct Maps { CUtensorMap m[8]; };

__device__ void body(const CUtensorMap* map, int x, int y, int z, uint8_t* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 48 * 43 + 112);   // 2064+112 = 2176
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  __syncwarp();
  if (threadIdx.x == 0) { mbar_expect_tx(bar, 48 * 43); tma_load_3d(smem, map, x, y, z, bar); }
  mbar_wait(bar, 0);
  for (int i = threadIdx.x; i < 48 * 43; i += 32) out[i] = smem[i];
}
__global__ void k_single(const __grid_constant__ CUtensorMap map, int x, int y, int z, uint8_t* out) { body(&map, x, y, z, out); }
__global__ void k_array(int pad0, const __grid_constant__ Maps maps, int level, int x, int y, int z, uint8_t* out) { body(&maps.m[level], x, y, z, out); }
__global__ void k_global(const CUtensorMap* maps, int level, int x, int y, int z, uint8_t* out) { body(maps + level, x, y, z, out); }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int pitch = 1280, rows = 414, slots = 3;
  const size_t stride = (size_t)pitch * rows;   // multiple of 256? 1280*414 = 529920 = 2070*256 ok
  std::vector<uint8_t> h(stride * slots);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
  uint8_t *d, *dout; cudaMalloc(&d, h.size()); cudaMalloc(&dout, 48 * 43);
  cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  EncodeTiledFn fn = (EncodeTiledFn)p;
  Maps maps;
  for (int l = 0; l < 8; ++l) {
    cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)slots};
    cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)stride};
    cuuint32_t box[3] = {48, 43, 1}, es[3] = {1, 1, 1};
    CUresult r = fn(&maps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  }
  CUtensorMap* dmaps; cudaMalloc(&dmaps, sizeof(Maps)); cudaMemcpy(dmaps, &maps, sizeof(Maps), cudaMemcpyHostToDevice);
  const int x = 1237, y = 377, z = 2;   // unaligned start, box runs past the right edge (pitch 1280) and bottom? 377+43=420 > 414
  auto check = [&](const char* name) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(e)); return false; }
    std::vector<uint8_t> o(48 * 43); cudaMemcpy(o.data(), dout, o.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r = 0; r < 43; ++r) for (int c = 0; c < 48; ++c) {
      uint8_t want = (x + c < pitch && y + r < rows) ? h[(size_t)z * stride + (size_t)(y + r) * pitch + x + c] : 0;
      bad += o[r * 48 + c] != want;
    }
    printf("%s: %d mismatches\n", name, bad); return bad == 0;
  };
  const size_t smem = 2176 + 16;
  k_single<<<1, 32, smem>>>(maps.m[3], x, y, z, dout); bool a = check("grid_constant single map");
  if (a) { k_array<<<1, 32, smem>>>(0, maps, 5, x, y, z, dout); a = check("grid_constant struct array, dynamic index"); }
  cudaMemset(dout, 0, 48 * 43);
  k_global<<<1, 32, smem>>>(dmaps, 6, x, y, z, dout); check("global memory map array");
  return 0;
}
