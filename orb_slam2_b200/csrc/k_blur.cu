// K4a — 7x7 sigma-2 Gaussian blur of every pyramid level of a batch, one launch.
//
// Replaces the per-level `clone + cv::GaussianBlur(7x7, 2, 2, BORDER_REFLECT_101)` of ORBextractor::operator()
// (src/ORBextractor.cc:1085-1086).  cv2's 8-bit path is a fixed separable integer stencil: taps {18,34,48,56,48,34,18}
// (8.8 fixed point, sum 256), horizontal pass exact in 16 bits, vertical pass (sum + 32768) >> 16.  The reference
// blurs an isolated clone of the level, i.e. with reflect-101 of the level itself at its edges; the pyramid's
// 19-px border already holds exactly those pixels, so the kernel reads the bordered plane and needs no edge case.
// The result goes to a second arena with the pyramid's own plane layout (same offsets, pitch and +19 origin), so
// one tensor-map set with another base address serves K4b's tap tiles.
//
// One CTA = a 128 x 32 tile of one (level, slot).  The 160 x 38 source box (3-px halo, from the 16-byte boundary
// below it) arrives by ONE TMA load.  Horizontal pass: a task is 4 columns x 2 rows from 3 aligned words per row
// and 10 dp4a per row (the 7 taps slide over the packed bytes by shifting the coefficient vectors); the two rows'
// 16-bit sums are packed into one word per column and leave as one 128-bit store: vertically adjacent H values
// share a word.  Vertical pass: a task is 4 columns x 8 rows walking down a sliding window of row-pair words,
// 4 dp2a per pixel (u16 x u8, the rounding constant rides in the accumulator), results packed four to a word:
// a warp stores 128 contiguous bytes of one output row.  Bound: HBM (reads 1.74 MB + writes 1.44 MB per KITTI
// eye) with ~10 warp instructions per 32 pixels beside it.
#include "orbx_internal.h"
#include "tma_utils.cuh"

namespace orbx {

constexpr int kBtW = 128, kBtH = 32;                   // output tile
constexpr int kBsW = kBtW + 32, kBsH = kBtH + 6;       // source box: columns X0-16 .. X0+143, rows Y0-3 .. Y0+34
constexpr int kBlurThreads = 128;
constexpr int kBlurOrigin = 16;                        // tiles start at bordered column / row 16 (16-byte aligned, just below the ROI origin 19)
static_assert(kBsW == kBlurBoxW && kBsH == kBlurBoxH, "tensor-map box of the blur kernel");

__device__ __forceinline__ unsigned bpk4(unsigned a, unsigned b, unsigned c, unsigned d) {
  unsigned r;   // kept in an ordinary register (see k_describe.cu: as literals the packed taps are re-materialised per use)
  asm volatile("mov.b32 %0, %1;" : "=r"(r) : "r"(a | (b << 8) | (c << 16) | (d << 24)));
  return r;
}

__global__ void __launch_bounds__(kBlurThreads) k_blur_levels(const __grid_constant__ TmaMaps maps, const __grid_constant__ BlurGeom G, uint8_t* __restrict__ blur,
                                                               int first_slot) {
  __shared__ __align__(128) uint8_t s_src[kBsH * kBsW];                  // 6080 B
  __shared__ __align__(16) uint32_t s_h[(kBsH / 2) * kBtW];             // [row pair][column]: H(2r) | H(2r+1) << 16, 9728 B
  __shared__ __align__(8) uint64_t s_bar;
  // grid = (tile columns of level 0, tile rows of all levels stacked, slots): the level and the tile row of a CTA come from one
  // table entry (no division, no search); levels narrower than level 0 leave their surplus CTAs at once
  const int t = threadIdx.x, slot = first_slot + blockIdx.z;
  const int rowinfo = G.row_info[blockIdx.y], level = rowinfo & 15, ty = rowinfo >> 4, tx = blockIdx.x;
  if (tx >= G.tiles_x[level]) return;
  const int X0 = kBlurOrigin + tx * kBtW, Y0 = kBlurOrigin + ty * kBtH;
  if (t == 0) {
    mbar_init(&s_bar, 1);
    fence_barrier_init();
    mbar_expect_tx(&s_bar, kBsH * kBsW);
    tma_load_3d(s_src, &maps.m[level], X0 - 16, Y0 - 3, slot, &s_bar);
  }
  __syncthreads();
  mbar_wait(&s_bar, 0);

  // horizontal pass.  Output column c = 4g + q of the tile reads box bytes 13 + c .. 19 + c: with w0, w1, w2 the aligned
  // words 3 + g, 4 + g, 5 + g of the row, q = 0 takes bytes 1..7 of (w0, w1), q = 1 bytes 2..8, q = 2 bytes 3..9, q = 3 bytes 0..6 of (w1, w2).
  {
    const unsigned A0 = bpk4(18, 34, 48, 56), A1 = bpk4(48, 34, 18, 0);
    const unsigned B0 = bpk4(0, 18, 34, 48), B1 = bpk4(56, 48, 34, 18);
    const unsigned C0 = bpk4(0, 0, 18, 34), C1 = bpk4(48, 56, 48, 34), C2 = bpk4(18, 0, 0, 0);
    const unsigned D0 = bpk4(0, 0, 0, 18), D1 = bpk4(34, 48, 56, 48), D2 = bpk4(34, 18, 0, 0);
    const int g = t & 31;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(s_src) + 3 + g;
#pragma unroll
    for (int r = t >> 5; r < kBsH / 2; r += kBlurThreads / 32) {          // row pairs r = warp, warp + 4, ...: 19 of them
      const uint32_t* p = src + (2 * r) * (kBsW / 4);
      const uint32_t a0 = p[0], a1 = p[1], a2 = p[2];
      const uint32_t b0 = p[kBsW / 4], b1 = p[kBsW / 4 + 1], b2 = p[kBsW / 4 + 2];
      uint4 o;                                                            // the sums fit 16 bits: one PRMT packs a row pair
      o.x = __byte_perm(__dp4a(a1, B1, __dp4a(a0, B0, 0u)), __dp4a(b1, B1, __dp4a(b0, B0, 0u)), 0x5410);
      o.y = __byte_perm(__dp4a(a2, C2, __dp4a(a1, C1, __dp4a(a0, C0, 0u))), __dp4a(b2, C2, __dp4a(b1, C1, __dp4a(b0, C0, 0u))), 0x5410);
      o.z = __byte_perm(__dp4a(a2, D2, __dp4a(a1, D1, __dp4a(a0, D0, 0u))), __dp4a(b2, D2, __dp4a(b1, D1, __dp4a(b0, D0, 0u))), 0x5410);
      o.w = __byte_perm(__dp4a(a2, A1, __dp4a(a1, A0, 0u)), __dp4a(b2, A1, __dp4a(b1, A0, 0u)), 0x5410);
      *reinterpret_cast<uint4*>(s_h + r * kBtW + 4 * g) = o;
    }
  }
  __syncthreads();

  // vertical pass: thread = 4 columns (4g ..) x 8 output rows (8s ..).  Output row o reads H rows o .. o + 6; with P[q] the
  // row-pair word (H(2q), H(2q+1)): even o = 2q -> P[q..q+3] x (18,34 | 48,56 | 48,34 | 18,0), odd o = 2q + 1 -> x (0,18 | 34,48 | 56,48 | 34,18)
  {
    const unsigned E0 = bpk4(18, 34, 0, 0), E1 = bpk4(48, 56, 0, 0), E2 = bpk4(48, 34, 0, 0), E3 = bpk4(18, 0, 0, 0);
    const unsigned F0 = bpk4(0, 18, 0, 0), F1 = bpk4(34, 48, 0, 0), F2 = bpk4(56, 48, 0, 0), F3 = bpk4(34, 18, 0, 0);
    const int g = t & 31, s = t >> 5;
    const int col = X0 + 4 * g, pitch = G.pitch[level], rows = G.rows[level];
    const uint4* hp = reinterpret_cast<const uint4*>(s_h + (4 * s) * kBtW + 4 * g);
    // a level plane is far below 2 GB: 32-bit offsets inside it
    uint8_t* dst = blur + G.plane_off[level] + (size_t)slot * G.plane_stride[level] + (unsigned)((Y0 + 8 * s) * pitch + col);
    uint4 W0 = hp[0], W1 = hp[kBtW / 4], W2 = hp[2 * (kBtW / 4)];
    // rows this thread may write: all 8 in an interior tile (CTA-uniform test first), fewer at the bottom edge; none right of the plane
    const int nrow = (X0 + kBtW <= pitch && Y0 + kBtH <= rows) ? 8 : (col < pitch ? min(8, rows - (Y0 + 8 * s)) : 0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint4 W3 = hp[(k + 3) * (kBtW / 4)];
      unsigned e[4], f[4];
#define ORBX_V(j, m) \
      e[j] = __dp2a_lo(W3.m, E3, __dp2a_lo(W2.m, E2, __dp2a_lo(W1.m, E1, __dp2a_lo(W0.m, E0, 32768u)))); \
      f[j] = __dp2a_lo(W3.m, F3, __dp2a_lo(W2.m, F2, __dp2a_lo(W1.m, F1, __dp2a_lo(W0.m, F0, 32768u))));
      ORBX_V(0, x) ORBX_V(1, y) ORBX_V(2, z) ORBX_V(3, w)
#undef ORBX_V
      // byte 2 of each sum is the pixel; four of them make one word of the output row
      const unsigned re = __byte_perm(__byte_perm(e[0], e[1], 0x0062), __byte_perm(e[2], e[3], 0x0062), 0x5410);
      const unsigned rf = __byte_perm(__byte_perm(f[0], f[1], 0x0062), __byte_perm(f[2], f[3], 0x0062), 0x5410);
      if (2 * k < nrow) *reinterpret_cast<uint32_t*>(dst) = re;
      if (2 * k + 1 < nrow) *reinterpret_cast<uint32_t*>(dst + pitch) = rf;
      dst += 2 * pitch;
      W0 = W1; W1 = W2; W2 = W3;
    }
  }
}

int launch_blur(const DeviceLayout& L, const LevelGeom* hg, const TmaMaps& maps, int first_slot, int n, cudaStream_t st) {
  BlurGeom G{};
  int rows_total = 0;
  for (int l = 0; l < L.nlevels; ++l) {
    // the ROI is columns 19 .. 19 + w - 1 of the bordered plane; tiles start at column / row 16
    G.tiles_x[l] = (hg[l].w + 3 + kBtW - 1) / kBtW;
    const int tiles_y = (hg[l].h + 3 + kBtH - 1) / kBtH;
    for (int ty = 0; ty < tiles_y && rows_total < kBlurMaxRows; ++ty) G.row_info[rows_total++] = l | (ty << 4);
    G.pitch[l] = hg[l].pitch; G.rows[l] = hg[l].rows; G.plane_off[l] = hg[l].plane_off; G.plane_stride[l] = hg[l].plane_stride;
  }
  k_blur_levels<<<dim3((unsigned)G.tiles_x[0], (unsigned)rows_total, (unsigned)n), kBlurThreads, 0, st>>>(maps, G, L.blur, first_slot);
  return 1;
}

}  // namespace orbx
