// K1 — image pyramid: fused bilinear resize + reflect-101 border for a whole batch of images.
//
// Replaces ORBextractor::ComputePyramid (src/ORBextractor.cc:1107-1132): level 0 is the input
// with a 19-px BORDER_REFLECT_101 frame; level k is cv::resize(INTER_LINEAR) of level k-1's ROI
// (chained, :1120) followed by the same border (:1122).  OpenCV's 8-bit path is integer fixed
// point: 11-bit tap weights (generated on the host in float exactly as OpenCV does, see
// orbx_api.cu), int32 horizontal pass, ((b*(H>>4))>>16 ... +2)>>2 vertical pass.
//
// Design: one thread produces 4 (level 0: 16) consecutive bytes of one bordered row and stores
// them as one word.  A border pixel equals the interior pixel at the reflected coordinate, so it is
// computed with the same formula instead of being copied afterwards — no intra-level dependency,
// one launch per level for the whole batch.  HBM-bound in principle (reads L2-hot source rows).
#include "orbx_internal.h"

namespace orbx {

__global__ void __launch_bounds__(256) k_border_level0(DeviceLayout L, const uint8_t* __restrict__ src, int src_pitch,
                                                        size_t src_img_stride, int first_slot, int n) {
  const LevelGeom g = L.geom[0];
  const int vec_per_row = g.pitch >> 4;
  const long long per_img = (long long)vec_per_row * g.rows;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= per_img * n) return;
  const int img = (int)(tid / per_img);
  const int rem = (int)(tid - (long long)img * per_img);
  const int by = rem / vec_per_row, bx0 = (rem - by * vec_per_row) << 4;
  const uint8_t* s = src + (size_t)img * src_img_stride + (size_t)reflect101(by - kEdge, g.h) * src_pitch;
  uint32_t wds[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    uint32_t v = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int bx = bx0 + q * 4 + b;
      uint32_t px = 0;
      if (bx < g.w + 2 * kEdge) px = __ldg(s + reflect101(bx - kEdge, g.w));
      v |= px << (8 * b);
    }
    wds[q] = v;
  }
  uint8_t* dst = L.pyr + g.plane_off + (size_t)(first_slot + img) * g.plane_stride + (size_t)by * g.pitch + bx0;
  *reinterpret_cast<uint4*>(dst) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
}

// One thread = 4 consecutive output bytes (one 32-bit store): a warp writes 128 contiguous bytes and
// reads 32 contiguous tap entries (two 128-bit loads per thread), so every request is coalesced.
// The tap tables are indexed by BORDERED coordinates (reflection and the zero pad columns are
// baked in on the host), so the kernel has no branches.
__global__ void __launch_bounds__(256) k_resize_level(DeviceLayout L, int level, int first_slot, int n) {
  const LevelGeom g = L.geom[level];
  const LevelGeom gs = L.geom[level - 1];
  const int wpr = g.pitch >> 2;                       // words per bordered row
  const long long per_img = (long long)wpr * g.rows;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= per_img * n) return;
  const int img = (int)(tid / per_img);
  const int rem = (int)(tid - (long long)img * per_img);
  const int by = rem / wpr, bx0 = (rem - by * wpr) << 2;
  const ResizeTap ty = L.taps[g.ytab_off + by];
  // The second tap of an edge pixel has weight 0 (OpenCV clamps the index instead): reading one
  // column / row past the ROI lands in the plane's own border, so no index clamp is needed.
  const uint8_t* p0 = L.pyr + gs.plane_off + (size_t)(first_slot + img) * gs.plane_stride + (size_t)(kEdge + ty.ofs) * gs.pitch + kEdge;
  const uint8_t* p1 = p0 + gs.pitch;
  const uint4* xt = reinterpret_cast<const uint4*>(L.taps + g.xtab_off + bx0);
  const uint4 q0 = __ldg(xt), q1 = __ldg(xt + 1);
  const int ofs[4] = {(int)q0.x, (int)q0.z, (int)q1.x, (int)q1.z};
  const uint32_t wts[4] = {q0.y, q0.w, q1.y, q1.w};
  const int b0 = ty.a0, b1 = ty.a1;
  uint32_t v = 0;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    const int a0 = (int)(wts[b] & 0xffff), a1 = (int)(wts[b] >> 16);   // weights are in [0, 2048]
    const uint8_t* s0 = p0 + ofs[b];
    const uint8_t* s1 = p1 + ofs[b];
    const int h0 = (int)s0[0] * a0 + (int)s0[1] * a1;
    const int h1 = (int)s1[0] * a0 + (int)s1[1] * a1;
    // <= 255 by construction (weights sum to <= 2049), so no saturation is needed
    const int r = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
    v |= (uint32_t)r << (8 * b);
  }
  uint8_t* dst = L.pyr + g.plane_off + (size_t)(first_slot + img) * g.plane_stride + (size_t)by * g.pitch + bx0;
  *reinterpret_cast<uint32_t*>(dst) = v;
}

int launch_pyramid(const DeviceLayout& L, const LevelGeom* hg, const uint8_t* src, int src_pitch,
                   size_t src_img_stride, int first_slot, int n, cudaStream_t st) {
  int launches = 0;
  {
    long long total = (long long)(hg[0].pitch >> 4) * hg[0].rows * n;
    k_border_level0<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(L, src, src_pitch, src_img_stride, first_slot, n);
    ++launches;
  }
  for (int l = 1; l < L.nlevels; ++l) {
    long long total = (long long)(hg[l].pitch >> 2) * hg[l].rows * n;
    k_resize_level<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(L, l, first_slot, n);
    ++launches;
  }
  return launches;
}

}  // namespace orbx
