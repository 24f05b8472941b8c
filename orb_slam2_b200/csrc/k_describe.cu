// K4 — orientation + 7x7 Gaussian blur + 256-bit steered BRIEF, fused, one warp per keypoint.
//
// Replaces, for the keypoints the quad-tree kept:
//   IC_Angle / computeOrientation        src/ORBextractor.cc:77-104, 472-479  (raw level pixels)
//   clone + cv::GaussianBlur(7x7, s=2)   src/ORBextractor.cc:1085-1086       (whole level in the reference)
//   computeOrbDescriptor                 src/ORBextractor.cc:108-147         (blurred pixels)
//   the final pt *= scale, octave, size  src/ORBextractor.cc:837-847, 1095-1101
//
// Blur on demand is exact: cv2's 8-bit path is a fixed separable stencil ({18,34,48,56,48,34,18},
// 8.8 horizontal pass, (sum+32768)>>16 vertical pass), the descriptor taps lie within 18 px of the
// keypoint, so a 43x43 raw tile gives every blurred pixel a tap can touch.  The reference blurs an
// isolated clone of the level (reflect-101 of the level itself); the pyramid's own 19-px border
// holds exactly those reflected pixels, so the tile is read straight from the bordered plane.
//
// All float math is written with explicit round-to-nearest intrinsics (never contracted to FMA),
// sinf/cosf are the restated glibc routines (double polynomial), see DESIGN.md hazards H2/H3.
#include "orbx_internal.h"
#include "tma_utils.cuh"

namespace orbx {

constexpr int kDescWarps = 2;             // 2-warp CTAs: 12 per SM; measured 0.7 % faster than 6 x 4 warps (finer tail)
// keypoints per warp (launch argument): the next tile is in flight while this one is processed.  16 for large batches
// (fewer warp prologues: 1.426 -> 1.386 ms per 256 KITTI pairs; 4 / 8 / 12 / 16 / 32 were measured), 4 when the whole
// launch is small, so that a single image still spreads over the SMs
constexpr int kKpPerWarpBatch = 16, kKpPerWarpSmall = 4;
constexpr int kTileR = kDescTileH, kTileP = kDescTileW;   // raw tile rows / pitch (TMA box); the keypoint sits at (21, phase16+21)
constexpr int kHtC = 40, kHtP = 50;       // horizontal pass, TRANSPOSED: [column][row] uint16, 50-entry pitch = 25 words: odd, so 32 columns never share a bank, and (unlike 23) the four row pieces of the skewed columns collide only two-way
constexpr int kHtSkew = 4;                // columns 32+ sit two words further: the 10 four-column groups of a store sweep then hit 10 distinct banks
constexpr int kBlC = 40, kBlP = 44;       // blurred patch, TRANSPOSED: [column][row] bytes, rows 0..37 (row 37 is scratch), centre (18,18)
// per-warp shared memory: raw tile | Ht | blurred | mbarrier.  The raw tile is dead once the moments and
// the horizontal pass have read it, so the NEXT keypoint's tile is requested right then and lands while
// the vertical pass and the descriptor of this one run.  
constexpr int kOffHt = 2816, kOffBl = 6832, kOffBar = 8592, kWarpSmem = 8704;
static_assert(kTileR * kTileP <= kOffHt && kOffHt + (kHtC * kHtP + kHtSkew) * 2 <= kOffBl && kOffBl + kBlC * kBlP <= kOffBar, "smem carve-up");
static_assert(kWarpSmem % 128 == 0 && kOffBar % 8 == 0, "TMA destinations must be 128-byte aligned");

// 7x7 sigma-2 taps of cv2's 8-bit path, {18,34,48,56,48,34,18}/256, packed for dp4a / dp2a
__device__ __forceinline__ unsigned pk4(unsigned a, unsigned b, unsigned c, unsigned d) {
  // the asm keeps the packed taps in an ordinary register: as compile-time constants they are
  // re-materialised through the uniform datapath (one UMOV per dot product) inside the hot loops
  unsigned r;
  asm volatile("mov.b32 %0, %1;" : "=r"(r) : "r"(a | (b << 8) | (c << 16) | (d << 24)));
  return r;
}

// umax of the reference for HALF_PATCH_SIZE = 15 (src/ORBextractor.cc:454-469)
// The disc is symmetric (|u| <= umax[|v|]  <=>  |v| <= umax[|u|]), so a lane that owns column u just needs
// its own row limit umax[|u|]; the table rides in one 64-bit literal, one nibble per entry.
// umax[0..15] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3}; the host checks its table against this at create
constexpr unsigned long long kUmaxNibbles = 0x3689ABCDDEEEFFFFull;

struct SinCosTab { double sign[4], hpi_inv, hpi, c0, c1, c2, c3, c4, s1, s2, s3; };
__constant__ SinCosTab c_sincos[2] = {
    {{1.0, -1.0, -1.0, 1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, 0x1p0,
     -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10, 0x1.99343027bf8c3p-16,
     -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13},
    {{-1.0, 1.0, 1.0, -1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, -0x1p0,
     0x1.ffffffd0c621cp-2, -0x1.55553e1068f19p-5, 0x1.6c087e89a359dp-10, -0x1.99343027bf8c3p-16,
     -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13}};

__device__ __forceinline__ float sincos_poly(double x, double x2, const SinCosTab& p, int n) {
  if ((n & 1) == 0) {
    const double x3 = __dmul_rn(x, x2), s1 = __dadd_rn(p.s2, __dmul_rn(x2, p.s3)), x7 = __dmul_rn(x3, x2),
                 s = __dadd_rn(x, __dmul_rn(x3, p.s1));
    return __double2float_rn(__dadd_rn(s, __dmul_rn(x7, s1)));
  }
  const double x4 = __dmul_rn(x2, x2), c2 = __dadd_rn(p.c3, __dmul_rn(x2, p.c4)), c1 = __dadd_rn(p.c0, __dmul_rn(x2, p.c1)),
               x6 = __dmul_rn(x4, x2), c = __dadd_rn(c1, __dmul_rn(x4, p.c2));
  return __double2float_rn(__dadd_rn(c, __dmul_rn(x6, c2)));
}
// glibc 2.39 sinf (cosine=0) / cosf (cosine=1) for 0 <= y < 120
__device__ __forceinline__ float libm_sincos(float y, int cosine) {
  double x = (double)y;
  const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
  if (top < ((__float_as_uint(0x1.921FB6p-1f) >> 20) & 0x7ff)) {
    if (top < ((__float_as_uint(0x1p-12f) >> 20) & 0x7ff)) return cosine ? 1.0f : y;
    return sincos_poly(x, __dmul_rn(x, x), c_sincos[0], cosine);
  }
  const double r = __dmul_rn(x, c_sincos[0].hpi_inv);
  const int n = (__double2int_rz(r) + 0x800000) >> 24;
  x = __dsub_rn(x, __dmul_rn((double)n, c_sincos[0].hpi));
  const double s = c_sincos[0].sign[n & 3];
  return sincos_poly(__dmul_rn(x, s), __dmul_rn(x, x), c_sincos[(n >> 1) & 1], n ^ cosine);
}

// cv::fastAtan2, scalar path (OpenCV 4.x), degrees
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  const float p1 = 57.2836266f, p3 = -18.6674461f, p5 = 8.91400051f, p7 = -2.53972459f;
  const float eps = (float)2.2204460492503131e-16;
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

template <int N> struct IntC { static constexpr int value = N; };
struct KpRef { int level, px, py, resp; };

// keypoint j of the level-major concatenation of the per-level lists (src/ORBextractor.cc:1073-1103)
__device__ __forceinline__ KpRef locate_kp(const DeviceLayout& L, int slot, int j, int incl, int myc, int lane) {
  const uint32_t lvl_mask = __ballot_sync(0xffffffffu, lane < L.nlevels && j < incl);
  KpRef k;
  k.level = __ffs(lvl_mask) - 1;
  const int lvl_first = __shfl_sync(0xffffffffu, incl - myc, k.level);
  const uint32_t key = L.lvlkp[(size_t)slot * L.lvlkp_per_slot + L.geom[k.level].lvl_kp_off + (j - lvl_first)];
  k.px = unpack_x(key); k.py = unpack_y(key); k.resp = unpack_r(key);
  return k;
}

__global__ void __launch_bounds__(kDescWarps * 32, 12) k_describe(DeviceLayout L, const __grid_constant__ TmaMaps maps, int first_slot,
                                                              int n_img, int ctas_per_img, int kp_per_warp, const int* __restrict__ umax,
                                                              const char4* __restrict__ pattern) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* wbase = smem + (size_t)warp * kWarpSmem;
  uint16_t* ht = reinterpret_cast<uint16_t*>(wbase + kOffHt);
  uint8_t* bl = wbase + kOffBl;
  uint64_t* bar = reinterpret_cast<uint64_t*>(wbase + kOffBar);

  const int img = blockIdx.x / ctas_per_img;
  if (img >= n_img) return;
  const int slot = first_slot + img;
  const int j0 = ((blockIdx.x - img * ctas_per_img) * kDescWarps + warp) * kp_per_warp;   // first output keypoint of this warp

  const int myc = lane < L.nlevels ? L.lvlcnt[slot * kMaxLevels + lane] : 0;
  int incl = myc;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const int total = min(__shfl_sync(0xffffffffu, incl, 31), L.kp_cap);
  if (j0 == 0 && lane == 0) L.nkp[slot] = total;
  if (j0 >= total) return;
  const int nk = min(kp_per_warp, total - j0);

  if (lane == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  __syncwarp();
  // TMA: the 64x43 box that holds the 43x43 tile centred on the keypoint, straight from the bordered plane (+19 offset)
  KpRef nxt = locate_kp(L, slot, j0, incl, myc, lane);
  if (lane == 0) {
    mbar_expect_tx(bar, kTileR * kTileP);
    tma_load_3d(wbase, &maps.m[nxt.level], (nxt.px + kEdge - 21) & ~15, nxt.py + kEdge - 21, slot, bar);
  }
  const uint8_t* raw = wbase;

  // this lane's 8 tests {x0, y0, x1, y1} as floats, resident in registers for all keypoints of the warp
  float4 pat[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    const char4 q = __ldg(pattern + lane * 8 + t);
    pat[t] = make_float4((float)q.x, (float)q.y, (float)q.z, (float)q.w);
  }

  for (int t = 0; t < nk; ++t) {
  const KpRef cur = nxt;
  const int j = j0 + t;
  const int level = cur.level, px = cur.px, py = cur.py, resp = cur.resp;
  const LevelGeom& g = L.geom[level];
  // the TMA box starts at the 16-byte boundary below the tile: tile column 0 is smem column phase16
  const int phase16 = (px + kEdge - 21) & 15, g0 = phase16 >> 2, phase = phase16 & 3;
  mbar_wait(bar, t & 1);

  // intensity-centroid moments over the radius-15 disc, int32 (src/ORBextractor.cc:77-104).
  // lane = column u (31 of them), rows unrolled: every load is independent, the disc half-width
  // of each row is a compile-time constant (kUmax, checked against the host table at create).
  int m10, m01;
  {
    const int u = lane - kHalfPatch, au = u < 0 ? -u : u;
    const int vlim = lane < 31 ? (int)((kUmaxNibbles >> (4 * (au & 15))) & 15ull) : -1;
    const uint8_t* colp = raw + 21 * kTileP + phase16 + 21 + (lane < 31 ? u : 0);
    int s_all = lane < 31 ? (int)colp[0] : 0, s_v = 0;
#pragma unroll
    for (int v = 1; v <= kHalfPatch; ++v) {
      const int a = colp[v * kTileP], b = colp[-v * kTileP];   // always inside the tile: load first, then decide
      // if (v <= vlim) { s_all += a + b; s_v += v * (a - b); } with the predication spelled out: left to itself
      // the compiler turns these 15 rows into branches or predicates depending on unrelated code (7 % of the kernel)
      asm("{\n\t.reg .pred p;\n\tsetp.ge.s32 p, %2, %3;\n\t@p add.s32 %0, %0, %4;\n\t@p mad.lo.s32 %1, %5, %3, %1;\n\t}"
          : "+r"(s_all), "+r"(s_v) : "r"(vlim), "r"(v), "r"(a + b), "r"(a - b));
    }
    m10 = u * s_all;
    m01 = s_v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m10 += __shfl_xor_sync(0xffffffffu, m10, o);
    m01 += __shfl_xor_sync(0xffffffffu, m01, o);
  }
  const float angle = fast_atan2_deg((float)m01, (float)m10);

  // separable fixed-point blur of the region the descriptor can reach.
  // Horizontal pass: one task = 4 adjacent smem columns of one row from 3 aligned words, 10 dp4a
  // (the 7 taps slide over the packed bytes by shifting the coefficient vectors, not the data);
  // results go to a transposed uint16 array so the vertical pass can pair rows in one word.
  {
    const unsigned A0 = pk4(18, 34, 48, 56), A1 = pk4(48, 34, 18, 0);
    const unsigned B0 = pk4(0, 18, 34, 48), B1 = pk4(56, 48, 34, 18);
    const unsigned C0 = pk4(0, 0, 18, 34), C1 = pk4(48, 56, 48, 34), C2 = pk4(18, 0, 0, 0);
    const unsigned D0 = pk4(0, 0, 0, 18), D1 = pk4(34, 48, 56, 48), D2 = pk4(34, 18, 0, 0);
    // lanes 0..29: 3 rows x 10 four-column groups per sweep, 15 sweeps (43 rows); all offsets are immediates
    const int lr = lane / 10, gq = lane - lr * 10;
    const uint32_t* rp = reinterpret_cast<const uint32_t*>(raw) + lr * (kTileP / 4) + g0 + gq;
    uint16_t* o = ht + (4 * gq) * kHtP + lr + (gq >= 8 ? kHtSkew : 0);
#pragma unroll
    for (int it = 0; it < 15; ++it) {
      if (lane < 30 && 3 * it + lr < kTileR) {
        const uint32_t w0 = rp[it * 3 * (kTileP / 4)], w1 = rp[it * 3 * (kTileP / 4) + 1], w2 = rp[it * 3 * (kTileP / 4) + 2];
        o[3 * it] = (uint16_t)__dp4a(w1, A1, __dp4a(w0, A0, 0u));
        o[3 * it + kHtP] = (uint16_t)__dp4a(w1, B1, __dp4a(w0, B0, 0u));
        o[3 * it + 2 * kHtP] = (uint16_t)__dp4a(w2, C2, __dp4a(w1, C1, __dp4a(w0, C0, 0u)));
        o[3 * it + 3 * kHtP] = (uint16_t)__dp4a(w2, D2, __dp4a(w1, D1, __dp4a(w0, D0, 0u)));
      }
    }
  }
  __syncwarp();
  if (t + 1 < nk) {                                   // warp-uniform: the raw tile is dead, fetch the next keypoint's
    nxt = locate_kp(L, slot, j + 1, incl, myc, lane);
    if (lane == 0) {
      fence_proxy_async();
      mbar_expect_tx(bar, kTileR * kTileP);
      tma_load_3d(wbase, &maps.m[nxt.level], (nxt.px + kEdge - 21) & ~15, nxt.py + kEdge - 21, slot, bar);
    }
  }
  // Vertical pass: one task = one column and a run of row pairs, walking down two output rows per
  // step with a 4-word sliding window and 8 dp2a (u16 x u8); the rounding constant rides in the
  // accumulator and the two result bytes leave as one 16-bit store into the transposed patch.
  {
    const unsigned E0 = pk4(18, 34, 0, 0), E1 = pk4(48, 56, 0, 0), E2 = pk4(48, 34, 0, 0), E3 = pk4(18, 0, 0, 0);
    const unsigned F0 = pk4(0, 18, 0, 0), F1 = pk4(34, 48, 0, 0), F2 = pk4(56, 48, 0, 0), F3 = pk4(34, 18, 0, 0);
    auto vrun = [&](const int col, const int rbase, auto nsteps) {
      const uint32_t* hc = reinterpret_cast<const uint32_t*>(ht + col * kHtP + rbase + (col >= 32 ? kHtSkew : 0));
      uint16_t* o = reinterpret_cast<uint16_t*>(bl + col * kBlP + rbase);
      uint32_t W0 = hc[0], W1 = hc[1], W2 = hc[2];
#pragma unroll
      for (int sidx = 0; sidx < decltype(nsteps)::value; ++sidx) {
        const uint32_t W3 = hc[sidx + 3];
        const unsigned e = __dp2a_lo(W3, E3, __dp2a_lo(W2, E2, __dp2a_lo(W1, E1, __dp2a_lo(W0, E0, 32768u))));
        const unsigned f = __dp2a_lo(W3, F3, __dp2a_lo(W2, F2, __dp2a_lo(W1, F1, __dp2a_lo(W0, F0, 32768u))));
        o[sidx] = (uint16_t)__byte_perm(e, f, 0x0062);   // byte 2 of e (row), byte 2 of f (row + 1)
        W0 = W1; W1 = W2; W2 = W3;
      }
    };
    // Sweeps 1 and 2: lane = absolute column 0..31 (the pitches 23 and 11 words are odd, so 32 consecutive
    // columns at one row offset never share a bank; columns below `phase` are computed for nothing, which is
    // cheaper than the conflicts a phase-relative cut runs into at the skewed columns), rows 0-19 and 18-37.
    // Sweep 3: the 8 skewed columns 32..39 in four 5-step pieces (rows 0-9, 10-19, 18-27, 28-37).
    vrun(lane, 0, IntC<10>{});
    vrun(lane, 18, IntC<10>{});
    {
      const int piece = lane & 3;
      vrun(32 + (lane >> 2), piece == 0 ? 0 : piece == 1 ? 10 : piece == 2 ? 18 : 28, IntC<5>{});
    }
  }
  __syncwarp();

  // steered BRIEF: lane i produces descriptor byte i (tests 8i .. 8i+7)
  const float rad = __fmul_rn(angle, (float)(3.1415926535897932384626433832795 / 180.f));
  const float ca = libm_sincos(rad, 1), sb = libm_sincos(rad, 0);
  // cvRound (round half to even) by the 1.5*2^23 trick: adding the constant rounds the sum to an integer
  // exactly like lrintf, and the integer then sits in the low mantissa bits
  const float kMagic = 12582912.0f;
  const uint32_t kBias = 0x4B400000u * (uint32_t)(kBlP + 1);
  const uint8_t* ctr = bl + (phase + 18) * kBlP + 18;
  uint32_t byte = 0;
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    const float4 q = pat[t];
    const uint32_t r0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(q.x, sb), __fmul_rn(q.y, ca)), kMagic));
    const uint32_t c0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(q.x, ca), __fmul_rn(q.y, sb)), kMagic));
    const uint32_t r1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(q.z, sb), __fmul_rn(q.w, ca)), kMagic));
    const uint32_t c1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(q.z, ca), __fmul_rn(q.w, sb)), kMagic));
    const int v0 = ctr[(int)(c0 * kBlP + r0 - kBias)], v1 = ctr[(int)(c1 * kBlP + r1 - kBias)];
    byte |= (uint32_t)(v0 < v1) << t;
  }
  const size_t oi = (size_t)slot * L.kp_cap + j;
  L.desc[oi * 32 + lane] = (uint8_t)byte;
  if (lane == 0) {
    orbx_keypoint kp;
    kp.x = level ? __fmul_rn((float)px, g.scale) : (float)px;
    kp.y = level ? __fmul_rn((float)py, g.scale) : (float)py;
    kp.size = (float)g.patch_size;
    kp.angle = angle;
    kp.response = (float)resp;
    kp.octave = level;
    kp.class_id = -1;
    L.kps[oi] = kp;
    L.kpk[oi] = make_uint2(pack_xyr(px, py, resp), __float_as_uint(angle));
  }
  __syncwarp();                                       // every lane is done with this tile, Ht and the blurred patch
  }
}

// ------------------------------------------------------------------------------------------------------------------
// K4b — the same stage on top of the blurred level planes of K4a (k_blur.cu): no blur work per keypoint.
//
// The fused kernel above blurs a 43 x 40 / 38 x 40 region around every keypoint: 3.2 k pixels x ~2000 keypoints = 4.5x the
// pixels of the level planes themselves, and those two passes were three quarters of its shared-memory traffic.  Here a
// keypoint costs two TMA boxes - 48 x 31 raw bytes for the orientation moments (IC_Angle reads the UNBLURRED level,
// src/ORBextractor.cc:77-104) and 64 x 37 blurred bytes for the 512 taps (:108-147) - the moments, fastAtan2, sinf / cosf
// and the 16 rotated taps per lane.  Two stages per warp: the boxes of keypoint t + 1 are requested before keypoint t is
// touched.  Arithmetic, rounding and output are those of k_describe, line for line.
constexpr int kMomBytes = kMomBoxW * kMomBoxH, kTapBytes = kTapBoxW * kTapBoxH;       // 1488, 2368
constexpr int kOffTap = 1536, kStageBytes = 3968, kOffBar2 = 2 * kStageBytes, kWarpSmem2 = 8064;
static_assert(kMomBytes <= kOffTap && kOffTap + kTapBytes <= kStageBytes && kOffBar2 + 16 <= kWarpSmem2, "smem carve-up");
static_assert(kOffTap % 128 == 0 && kStageBytes % 128 == 0 && kWarpSmem2 % 128 == 0, "TMA destinations must be 128-byte aligned");

__global__ void __launch_bounds__(kDescWarps * 32, 12) k_describe_planes(DeviceLayout L, const __grid_constant__ TmaMaps mom_maps,
                                                                         const __grid_constant__ TmaMaps tap_maps, int first_slot, int n_img,
                                                                         int ctas_per_img, int kp_per_warp, const char4* __restrict__ pattern) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t s_mask[8][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* wbase = smem + (size_t)warp * kWarpSmem2;
  uint64_t* bar = reinterpret_cast<uint64_t*>(wbase + kOffBar2);     // bar[0], bar[1]: one per stage

  const int img = blockIdx.x / ctas_per_img;
  if (img >= n_img) return;
  const int slot = first_slot + img;
  const int j0 = ((blockIdx.x - img * ctas_per_img) * kDescWarps + warp) * kp_per_warp;   // first output keypoint of this warp

  const int myc = lane < L.nlevels ? L.lvlcnt[slot * kMaxLevels + lane] : 0;
  int incl = myc;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const int total = min(__shfl_sync(0xffffffffu, incl, 31), L.kp_cap);
  if (j0 == 0 && lane == 0) L.nkp[slot] = total;
  if (j0 >= total) return;
  const int nk = min(kp_per_warp, total - j0);

  if (lane == 0) { mbar_init(bar, 1); mbar_init(bar + 1, 1); fence_barrier_init(); }
  // byte masks of the orientation disc, per row (= lane) and word: byte b of word k is column u = -15 + 4k + b, inside the
  // disc iff |u| <= umax[|v|]; lane 31 (no row) masks everything.  Each warp fills the table with the same values.
  {
    const int v = lane - kHalfPatch, av = v < 0 ? -v : v;
    const int half = lane < 31 ? (int)((kUmaxNibbles >> (4 * (av & 15))) & 15ull) : -1;
#pragma unroll
    for (int k2 = 0; k2 < 8; ++k2) {
      uint32_t mk = 0u;
#pragma unroll
      for (int b = 0; b < 4; ++b) { const int u = -15 + 4 * k2 + b; if ((u < 0 ? -u : u) <= half) mk |= 0xffu << (8 * b); }
      s_mask[k2][lane] = mk;
    }
  }
  __syncwarp();
  auto request = [&](const KpRef& k, int stage) {            // lane 0: both boxes of one keypoint on the stage's barrier
    uint8_t* dst = wbase + stage * kStageBytes;
    mbar_expect_tx(bar + stage, kMomBytes + kTapBytes);
    tma_load_3d(dst, &mom_maps.m[k.level], (k.px + kEdge - 15) & ~15, k.py + kEdge - 15, slot, bar + stage);
    tma_load_3d(dst + kOffTap, &tap_maps.m[k.level], (k.px + kEdge - 18) & ~15, k.py + kEdge - 18, slot, bar + stage);
  };
  KpRef nxt = locate_kp(L, slot, j0, incl, myc, lane);
  if (lane == 0) request(nxt, 0);

  float4 pat[8];                                              // this lane's 8 tests {x0, y0, x1, y1}
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    const char4 q = __ldg(pattern + lane * 8 + t);
    pat[t] = make_float4((float)q.x, (float)q.y, (float)q.z, (float)q.w);
  }

  for (int t = 0; t < nk; ++t) {
    const KpRef cur = nxt;
    const int j = j0 + t, stage = t & 1;
    if (t + 1 < nk) {                                         // warp-uniform: the other stage was released at the end of the last iteration
      nxt = locate_kp(L, slot, j + 1, incl, myc, lane);
      if (lane == 0) { fence_proxy_async(); request(nxt, stage ^ 1); }
    }
    const int level = cur.level, px = cur.px, py = cur.py;
    const LevelGeom& g = L.geom[level];
    const uint8_t* raw = wbase + stage * kStageBytes;
    const uint8_t* bl = raw + kOffTap;
    mbar_wait(bar + stage, (t >> 1) & 1);

    // intensity-centroid moments over the radius-15 disc (src/ORBextractor.cc:77-104).  lane = ROW v of the 31 x 31 patch: its
    // 31 bytes come as 9 aligned words, byte-aligned by funnel shifts, masked to the row's half-width umax[|v|] (s_mask, made
    // once per CTA); two dp4a per word then give the row's sum of pixels and its sum of u * pixel (u = -15..15 as signed-byte
    // coefficients).  m10 = sum of the latter, m01 = sum of v * row sum: ~45 instructions per keypoint instead of ~105 with
    // one column per lane and one load per pixel.
    int m10, m01;
    {
      const int phM = (px + kEdge - 15) & 15, sh = (phM & 3) * 8;
      const uint32_t* rowp = reinterpret_cast<const uint32_t*>(raw + (lane < 31 ? lane : 0) * kMomBoxW + (phM & ~3));
      uint32_t W[9];
#pragma unroll
      for (int k2 = 0; k2 < 9; ++k2) W[k2] = rowp[k2];
      int rsum = 0, usum = 0;
#pragma unroll
      for (int k2 = 0; k2 < 8; ++k2) {
        const uint32_t dmask = __funnelshift_r(W[k2], W[k2 + 1], sh) & s_mask[k2][lane];
        // u of the four bytes: -15 + 4 k2 + {0, 1, 2, 3}, as signed bytes (the 32nd byte, u = 16, is always masked out)
        const int u0 = -15 + 4 * k2;
        const uint32_t ucoef = (uint32_t)(u0 & 0xff) | ((uint32_t)((u0 + 1) & 0xff) << 8) | ((uint32_t)((u0 + 2) & 0xff) << 16) | ((uint32_t)((u0 + 3) & 0xff) << 24);
        rsum = (int)__dp4a(dmask, 0x01010101u, (uint32_t)rsum);
        asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(usum) : "r"(dmask), "r"(ucoef));
      }
      m10 = usum;
      m01 = (lane - kHalfPatch) * rsum;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m10 += __shfl_xor_sync(0xffffffffu, m10, o);
      m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // steered BRIEF on the blurred box: lane i produces descriptor byte i; tap (row r, column c) = ctr[r * 64 + c]
    const float rad = __fmul_rn(angle, (float)(3.1415926535897932384626433832795 / 180.f));
    const float ca = libm_sincos(rad, 1), sb = libm_sincos(rad, 0);
    const float kMagic = 12582912.0f;                         // cvRound by the 1.5 * 2^23 trick, see k_describe
    const uint32_t kBias = 0x4B400000u * (uint32_t)(kTapBoxW + 1);
    const uint8_t* ctr = bl + 18 * kTapBoxW + ((px + kEdge - 18) & 15) + 18;
    uint32_t byte = 0;
#pragma unroll
    for (int q8 = 0; q8 < 8; ++q8) {
      const float4 q = pat[q8];
      const uint32_t r0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(q.x, sb), __fmul_rn(q.y, ca)), kMagic));
      const uint32_t c0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(q.x, ca), __fmul_rn(q.y, sb)), kMagic));
      const uint32_t r1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(q.z, sb), __fmul_rn(q.w, ca)), kMagic));
      const uint32_t c1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(q.z, ca), __fmul_rn(q.w, sb)), kMagic));
      const int v0 = ctr[(int)(r0 * kTapBoxW + c0 - kBias)], v1 = ctr[(int)(r1 * kTapBoxW + c1 - kBias)];
      byte |= (uint32_t)(v0 < v1) << q8;
    }
    const size_t oi = (size_t)slot * L.kp_cap + j;
    L.desc[oi * 32 + lane] = (uint8_t)byte;
    if (lane == 0) {
      orbx_keypoint kp;
      kp.x = level ? __fmul_rn((float)px, g.scale) : (float)px;
      kp.y = level ? __fmul_rn((float)py, g.scale) : (float)py;
      kp.size = (float)g.patch_size;
      kp.angle = angle;
      kp.response = (float)cur.resp;
      kp.octave = level;
      kp.class_id = -1;
      L.kps[oi] = kp;
      L.kpk[oi] = make_uint2(pack_xyr(px, py, cur.resp), __float_as_uint(angle));
    }
    __syncwarp();                                             // every lane is done with this stage
  }
}

int launch_describe_planes(const DeviceLayout& L, const TmaMaps& mom_maps, const TmaMaps& tap_maps, int first_slot, int n, const char4* d_pattern,
                           cudaStream_t st) {
  const int kp_per_warp = (long long)n * L.kp_cap >= 148LL * 24 * kKpPerWarpBatch ? kKpPerWarpBatch : kKpPerWarpSmall;
  const int per_cta = kDescWarps * kp_per_warp;
  const int ctas_per_img = (L.kp_cap + per_cta - 1) / per_cta;
  const size_t smem = (size_t)kDescWarps * kWarpSmem2;
  k_describe_planes<<<n * ctas_per_img, kDescWarps * 32, smem, st>>>(L, mom_maps, tap_maps, first_slot, n, ctas_per_img, kp_per_warp, d_pattern);
  return 1;
}

int launch_describe(const DeviceLayout& L, const TmaMaps& maps, int first_slot, int n, const int* d_umax, const char4* d_pattern, cudaStream_t st) {
  const int kp_per_warp = (long long)n * L.kp_cap >= 148LL * 24 * kKpPerWarpBatch ? kKpPerWarpBatch : kKpPerWarpSmall;
  const int per_cta = kDescWarps * kp_per_warp;
  const int ctas_per_img = (L.kp_cap + per_cta - 1) / per_cta;
  const size_t smem = (size_t)kDescWarps * kWarpSmem;
  k_describe<<<n * ctas_per_img, kDescWarps * 32, smem, st>>>(L, maps, first_slot, n, ctas_per_img, kp_per_warp, d_umax, d_pattern);
  return 1;
}

}  // namespace orbx
