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

namespace orbx {

constexpr int kDescWarps = 4;
constexpr int kTileR = 43, kTileP = 48;   // raw tile rows / pitch, centre (21,21)
constexpr int kHpP = 40;                  // horizontal-pass pitch (uint16), 43 rows x 37 cols
constexpr int kBlP = 40, kBlR = 37;       // blurred tile pitch / rows, centre (18,18)
constexpr int kWarpSmem = kTileR * kTileP + kTileR * kHpP * 2 + kBlR * kBlP;   // 2064 + 3440 + 1480
static_assert(kWarpSmem % 8 == 0, "per-warp smem must keep 8-byte alignment");

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

__global__ void __launch_bounds__(kDescWarps * 32) k_describe(DeviceLayout L, int first_slot, int n_img, int ctas_per_img,
                                                              const int* __restrict__ umax, const char4* __restrict__ pattern) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* raw = smem + (size_t)warp * kWarpSmem;
  uint16_t* hp = reinterpret_cast<uint16_t*>(raw + kTileR * kTileP);
  uint8_t* bl = raw + kTileR * kTileP + kTileR * kHpP * 2;

  const int img = blockIdx.x / ctas_per_img;
  if (img >= n_img) return;
  const int slot = first_slot + img;
  const int j = (blockIdx.x - img * ctas_per_img) * kDescWarps + warp;   // output keypoint index

  // level-major concatenation of the per-level lists (src/ORBextractor.cc:1073-1103)
  const int myc = lane < L.nlevels ? L.lvlcnt[slot * kMaxLevels + lane] : 0;
  int incl = myc;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  if (j == 0 && lane == 0) L.nkp[slot] = min(total, L.kp_cap);
  if (j >= total || j >= L.kp_cap) return;
  const uint32_t lvl_mask = __ballot_sync(0xffffffffu, lane < L.nlevels && j < incl);
  const int level = __ffs(lvl_mask) - 1;
  const int lvl_first = __shfl_sync(0xffffffffu, incl - myc, level);
  const LevelGeom& g = L.geom[level];
  const uint32_t key = L.lvlkp[(size_t)slot * L.lvlkp_per_slot + g.lvl_kp_off + (j - lvl_first)];
  const int px = unpack_x(key), py = unpack_y(key), resp = unpack_r(key);

  // stage the 43x43 raw tile whose centre is the keypoint (bordered-plane coords: +19)
  const uint8_t* plane = L.pyr + g.plane_off + (size_t)slot * g.plane_stride;
  const uint8_t* src = plane + (size_t)(py + kEdge - 21) * g.pitch + (px + kEdge - 21);
  for (int r = 0; r < kTileR; ++r) {
    const uint8_t* s = src + (size_t)r * g.pitch;
    raw[r * kTileP + lane] = __ldg(s + lane);
    if (lane < kTileR - 32) raw[r * kTileP + 32 + lane] = __ldg(s + 32 + lane);
  }
  __syncwarp();

  // intensity-centroid moments over the radius-15 disc, int32 (src/ORBextractor.cc:77-104)
  int m10 = 0, m01 = 0;
  if (lane < 2 * kHalfPatch + 1) {
    const int v = lane - kHalfPatch;
    const int d = umax[v < 0 ? -v : v];
    const uint8_t* row = raw + (21 + v) * kTileP + 21;
    int si = 0;
    for (int u = -d; u <= d; ++u) {
      const int I = row[u];
      si += I;
      m10 += u * I;
    }
    m01 = v * si;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m10 += __shfl_xor_sync(0xffffffffu, m10, o);
    m01 += __shfl_xor_sync(0xffffffffu, m01, o);
  }
  const float angle = fast_atan2_deg((float)m01, (float)m10);

  // separable fixed-point blur of the 37x37 region the descriptor can reach
  for (int idx = lane; idx < kTileR * 37; idx += 32) {
    const int r = idx / 37, c = idx - r * 37;
    const uint8_t* p = raw + r * kTileP + c;
    hp[r * kHpP + c] = (uint16_t)(18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3]);
  }
  __syncwarp();
  for (int idx = lane; idx < kBlR * 37; idx += 32) {
    const int r = idx / 37, c = idx - r * 37;
    const uint16_t* p = hp + r * kHpP + c;
    const uint32_t acc = 18u * (p[0] + p[6 * kHpP]) + 34u * (p[kHpP] + p[5 * kHpP]) + 48u * (p[2 * kHpP] + p[4 * kHpP]) +
                         56u * p[3 * kHpP] + 32768u;
    bl[r * kBlP + c] = (uint8_t)(acc >> 16);
  }
  __syncwarp();

  // steered BRIEF: lane i produces descriptor byte i (tests 8i .. 8i+7)
  const float rad = __fmul_rn(angle, (float)(3.1415926535897932384626433832795 / 180.f));
  const float ca = libm_sincos(rad, 1), sb = libm_sincos(rad, 0);
  const uint8_t* ctr = bl + 18 * kBlP + 18;
  uint32_t byte = 0;
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    const char4 q = __ldg(pattern + lane * 8 + t);
    const float x0 = (float)q.x, y0 = (float)q.y, x1 = (float)q.z, y1 = (float)q.w;
    const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sb), __fmul_rn(y0, ca)));
    const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sb)));
    const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sb), __fmul_rn(y1, ca)));
    const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sb)));
    const int v0 = ctr[r0 * kBlP + c0], v1 = ctr[r1 * kBlP + c1];
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
  }
}

int launch_describe(const DeviceLayout& L, int first_slot, int n, const int* d_umax, const char4* d_pattern, cudaStream_t st) {
  const int ctas_per_img = (L.kp_cap + kDescWarps - 1) / kDescWarps;
  const size_t smem = (size_t)kDescWarps * kWarpSmem;
  k_describe<<<n * ctas_per_img, kDescWarps * 32, smem, st>>>(L, first_slot, n, ctas_per_img, d_umax, d_pattern);
  return 1;
}

}  // namespace orbx
