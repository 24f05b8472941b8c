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
#include <algorithm>

#include "orbx_internal.h"
#include "tma_utils.cuh"

namespace orbx {

// Level 0: one thread = 16 consecutive bytes of one bordered row.  blockIdx.z = image; blockIdx.y = 0 walks
// the interior vectors of all rows (16 source bytes at an arbitrary address: 5 aligned words + funnel shift),
// blockIdx.y = 1 the few vectors per row that touch the reflected border or the pad, byte by byte - kept in
// separate warps so that the common case never executes the slow branch.
__global__ void __launch_bounds__(256) k_border_level0(DeviceLayout L, const uint8_t* __restrict__ src, int src_pitch,
                                                        size_t src_img_stride, int first_slot, int v_lo, int n_int) {
  const LevelGeom g = L.geom[0];
  const int vec_per_row = g.pitch >> 4, n_edge = vec_per_row - n_int;
  const int img = blockIdx.z, edge = blockIdx.y;
  const int per_row = edge ? n_edge : n_int;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= per_row * g.rows) return;
  const int by = idx / per_row, k = idx - by * per_row;
  const int v = edge ? (k < v_lo ? k : k + n_int) : v_lo + k;
  const int bx0 = v << 4;
  const int sy = reflect101(by - kEdge, g.h), x0 = bx0 - kEdge;
  const uint8_t* s = src + (size_t)img * src_img_stride + (size_t)sy * src_pitch;
  uint32_t wds[4];
  if (!edge && (sy + 1 < g.h || x0 + 20 <= g.w)) {
    // the guard keeps the fifth word inside the image (the next row, or this row's own tail)
    const uintptr_t a = reinterpret_cast<uintptr_t>(s + x0);
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
    const int sh = (int)(a & 3) * 8;
    const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2), w3 = __ldg(wp + 3), w4 = __ldg(wp + 4);
    wds[0] = __funnelshift_r(w0, w1, sh); wds[1] = __funnelshift_r(w1, w2, sh);
    wds[2] = __funnelshift_r(w2, w3, sh); wds[3] = __funnelshift_r(w3, w4, sh);
  } else {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t wv = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int bx = bx0 + q * 4 + b;
        uint32_t px = 0;
        if (bx < g.w + 2 * kEdge) px = __ldg(s + reflect101(bx - kEdge, g.w));
        wv |= px << (8 * b);
      }
      wds[q] = wv;
    }
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

// Tiled variant (used when the scale factor keeps a tile's source footprint inside the static
// shared-memory budget, i.e. for the usual 1.2): one CTA = 32 x 128 bordered output pixels.
//   1. the source footprint (rows/cols spanned by the tile's taps, reflection included because the
//      tap tables are in bordered coordinates) is fetched by ONE TMA box load (192 x 48 bytes from
//      the 16-byte boundary below the footprint; the hardware zero-fills outside the plane)
//   2. horizontal pass once per SOURCE row (1.2x fewer than two per output row), (h >> 4) as uint16
//   3. vertical pass, 4 pixels per thread, (b*H) >> 16 as one multiply-high with the +2 rounding
//      riding in the accumulator; one aligned 32-bit store per thread, a warp writes 128 B
constexpr int kRtTH = kResizeTileH, kRtTW = kResizeTileW, kRtSH = kResizeBoxH, kRtSWP = kResizeBoxW;

__global__ void __launch_bounds__(256) k_resize_tiled(DeviceLayout L, const __grid_constant__ TmaMaps maps, const __grid_constant__ LevelGeom g, int level, int first_slot) {
  __shared__ __align__(128) uint8_t s_src[kRtSH * kRtSWP];
  __shared__ __align__(16) uint32_t s_h[kRtSH * kRtTW];
  __shared__ int4 s_ty[kRtTH];          // {row offset into s_h, a0 << 16, a1 << 16, -}
  __shared__ __align__(8) uint64_t s_bar;
  const int t = threadIdx.x;                         // g: this level's geometry as a kernel parameter (no global load ahead of the TMA request)
  const int txi = blockIdx.x, tyi = blockIdx.y, img = blockIdx.z;   // grid = (tiles_x, tiles_y, images)
  const int bx0 = txi * kRtTW, by0 = tyi * kRtTH;
  const int tw = min(kRtTW, g.pitch - bx0), th = min(kRtTH, g.rows - by0);
  // source footprint of this tile from the host-made tile table: first source column / row, row count
  const int sx_lo = L.taps[g.xtile_off + txi].ofs;
  const ResizeTap ytile = L.taps[g.ytile_off + tyi];
  const int sy_lo = ytile.ofs, SH = ytile.a0;
  const int phase = (kEdge + sx_lo) & 15;            // smem column of source column sx_lo
  if (t == 0) {
    mbar_init(&s_bar, 1);
    fence_barrier_init();
    mbar_expect_tx(&s_bar, kRtSH * kRtSWP);
    tma_load_3d(s_src, &maps.m[level - 1], (kEdge + sx_lo) & ~15, kEdge + sy_lo, first_slot + img, &s_bar);
  }
  const int c = t & (kRtTW - 1);
  ResizeTap tx; tx.ofs = 0; tx.a0 = 0; tx.a1 = 0;
  if (c < tw) tx = L.taps[g.xtab_off + bx0 + c];
  // mulhi(b << 16, H) == (b * H) >> 16: the vertical weights are kept pre-shifted
  if (t < th) {
    const ResizeTap tyr = L.taps[g.ytab_off + by0 + t];
    s_ty[t] = make_int4((tyr.ofs - sy_lo) * kRtTW, tyr.a0 << 16, tyr.a1 << 16, 0);
  }
  __syncthreads();                                   // barrier initialised, s_ty visible
  mbar_wait(&s_bar, 0);
  if (c < tw) {
    const int a0 = tx.a0, a1 = tx.a1;
    const uint8_t* p = s_src + (t >> 7) * kRtSWP + (tx.ofs - sx_lo + phase);
    uint32_t* o = s_h + (t >> 7) * kRtTW + c;
#pragma unroll
    for (int k = 0; k < kRtSH / 2; ++k) {            // source rows (t >> 7) + 2k; SH is CTA-uniform
      if ((k & 3) == 0 && 2 * k >= SH) break;
      o[2 * k * kRtTW] = (uint32_t)(((int)p[2 * k * kRtSWP] * a0 + (int)p[2 * k * kRtSWP + 1] * a1) >> 4);
    }
  }
  __syncthreads();
  const int c4 = (t & 31) * 4;
  if (c4 < tw) {
    uint8_t* dst = L.pyr + g.plane_off + (size_t)(first_slot + img) * g.plane_stride + (size_t)(by0 + (t >> 5)) * g.pitch + bx0 + c4;
    const int step = 8 * g.pitch;                      // a level plane is far below 2 GB: 32-bit row offsets
#pragma unroll
    for (int k = 0; k < kRtTH / 8; ++k) {
      const int r = (t >> 5) + 8 * k;
      if (r < th) {
        const int4 ty = s_ty[r];
        const uint4 u0 = *reinterpret_cast<const uint4*>(s_h + ty.x + c4), u1 = *reinterpret_cast<const uint4*>(s_h + ty.x + kRtTW + c4);
        const unsigned b0 = (unsigned)ty.y, b1 = (unsigned)ty.z;
        const unsigned r0 = (__umulhi(b1, u1.x) + (__umulhi(b0, u0.x) + 2u)) >> 2;
        const unsigned r1 = (__umulhi(b1, u1.y) + (__umulhi(b0, u0.y) + 2u)) >> 2;
        const unsigned r2 = (__umulhi(b1, u1.z) + (__umulhi(b0, u0.z) + 2u)) >> 2;
        const unsigned r3 = (__umulhi(b1, u1.w) + (__umulhi(b0, u0.w) + 2u)) >> 2;
        *reinterpret_cast<uint32_t*>(dst + k * step) = (r1 * 256u + r0) + ((r3 * 256u + r2) << 16);
      }
    }
  }
}

// Does every 32x128 tile of this level fit the static source-footprint budget?  (host, from the tap tables)
bool resize_tiled_fits(const LevelGeom& g, const ResizeTap* taps) {
  for (int bx0 = 0; bx0 < g.pitch; bx0 += kRtTW) {
    int lo = 1 << 30, hi = -1;
    for (int c = bx0; c < std::min(bx0 + kRtTW, g.pitch); ++c) { lo = std::min(lo, taps[g.xtab_off + c].ofs); hi = std::max(hi, taps[g.xtab_off + c].ofs + 1); }
    if (hi - lo + 1 + 15 > kRtSWP) return false;   // + up to 15 columns below the footprint for the 16-byte aligned TMA start
  }
  for (int by0 = 0; by0 < g.rows; by0 += kRtTH) {
    int lo = 1 << 30, hi = -1;
    for (int r = by0; r < std::min(by0 + kRtTH, g.rows); ++r) { lo = std::min(lo, taps[g.ytab_off + r].ofs); hi = std::max(hi, taps[g.ytab_off + r].ofs + 1); }
    if (hi - lo + 1 > kRtSH) return false;
  }
  return true;
}

int launch_pyramid(const DeviceLayout& L, const LevelGeom* hg, const uint8_t* src, int src_pitch,
                   size_t src_img_stride, int first_slot, int n, unsigned tiled_mask, const TmaMaps* resize_maps, cudaStream_t st) {
  int launches = 0;
  {
    // interior vectors: all 16 source bytes inside the row, x0 = 16 v - 19 >= 0 and x0 + 16 <= w
    const int vec_per_row = hg[0].pitch >> 4;
    const int v_lo = (kEdge + 15) / 16, v_hi = std::min((hg[0].w + kEdge - 16) / 16, vec_per_row - 1);
    const int n_int = std::max(v_hi - v_lo + 1, 0);
    const int per_row = std::max(n_int, vec_per_row - n_int);
    const unsigned blocks = (unsigned)(((long long)per_row * hg[0].rows + 255) / 256);
    k_border_level0<<<dim3(blocks, 2, (unsigned)n), 256, 0, st>>>(L, src, src_pitch, src_img_stride, first_slot, n_int > 0 ? v_lo : 0, n_int);
    ++launches;
  }
  for (int l = 1; l < L.nlevels; ++l) {
    if (tiled_mask & (1u << l)) {
      const int tiles_x = (hg[l].pitch + kRtTW - 1) / kRtTW, tiles_y = (hg[l].rows + kRtTH - 1) / kRtTH;
      k_resize_tiled<<<dim3((unsigned)tiles_x, (unsigned)tiles_y, (unsigned)n), 256, 0, st>>>(L, *resize_maps, hg[l], l, first_slot);
    } else {
      long long total = (long long)(hg[l].pitch >> 2) * hg[l].rows * n;
      k_resize_level<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(L, l, first_slot, n);
    }
    ++launches;
  }
  return launches;
}

}  // namespace orbx
