// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the shipped library;
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may build, load or call it, and only as the checker or the timed CPU baseline.
//
// CPU restatements of the OpenCV primitives the reference's front end calls.  OpenCV is an
// external, un-vendored dependency of the reference (CMakeLists.txt:30-36 asks for >=2.4.3);
// the semantics restated here are those of the cv2 4.13.0 wheel, the only OpenCV on the box,
// and tests/test_oracle_cv2.py pins every function below against it bit for bit.
//
//   resize_linear_u8   <- cv::resize(INTER_LINEAR)        called at src/ORBextractor.cc:1120
//   border_reflect101  <- cv::copyMakeBorder(REFLECT_101) called at src/ORBextractor.cc:1122,1127
//   fast9_16           <- cv::FAST(img,kps,t,true)        called at src/ORBextractor.cc:809,814
//   gauss7_sigma2      <- cv::GaussianBlur(7x7, 2, 2)     called at src/ORBextractor.cc:1086
//   fast_atan2_deg     <- cv::fastAtan2                   called at src/ORBextractor.cc:103
//   cv_round           <- cvRound                         called at src/ORBextractor.cc:81,115,119,442,460,1112
//   libm_sinf/cosf     <- glibc sinf/cosf                 called at src/ORBextractor.cc:113
#pragma once
#include <emmintrin.h>

#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace orbo {

// cvRound: round half to even (cvtsd2si under the default rounding mode).
static inline int cv_round(double v) { return (int)std::lrint(v); }
static inline int cv_round(float v) { return (int)std::lrintf(v); }

// ---------------------------------------------------------------------------------------
// glibc 2.39 sinf / cosf, restated (double-precision polynomial after a one-multiply range
// reduction; the published "optimized-routines" sincosf algorithm).  The reference calls the
// host libm (src/ORBextractor.cc:112-113), which is not correctly rounded, so a device port has
// to follow this exact evaluation.  tools/sweep_sincos.c proves equality with the host libm
// for every float in [0, 2*pi*1.01] (1 087 050 388 inputs, 0 mismatches, with and without FMA).
// Valid for |x| < 120; the front end only feeds angles in [0, 2*pi].
// ---------------------------------------------------------------------------------------
struct SinCosTab {
  double sign[4], hpi_inv, hpi, c0, c1, c2, c3, c4, s1, s2, s3;
};
static const SinCosTab kSinCosTab[2] = {
    {{1.0, -1.0, -1.0, 1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, 0x1p0,
     -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10, 0x1.99343027bf8c3p-16,
     -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13},
    {{-1.0, 1.0, 1.0, -1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, -0x1p0,
     0x1.ffffffd0c621cp-2, -0x1.55553e1068f19p-5, 0x1.6c087e89a359dp-10, -0x1.99343027bf8c3p-16,
     -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13}};

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline uint32_t top12(float f) { return (f2u(f) >> 20) & 0x7ff; }

static inline float sincos_poly(double x, double x2, const SinCosTab& p, int n) {
  if ((n & 1) == 0) {
    double x3 = x * x2, s1 = p.s2 + x2 * p.s3, x7 = x3 * x2, s = x + x3 * p.s1;
    return (float)(s + x7 * s1);
  }
  double x4 = x2 * x2, c2 = p.c3 + x2 * p.c4, c1 = p.c0 + x2 * p.c1, x6 = x4 * x2,
         c = c1 + x4 * p.c2;
  return (float)(c + x6 * c2);
}
static inline float sincos_eval(float y, int cosine) {
  double x = y;
  const SinCosTab* p = &kSinCosTab[0];
  if (top12(y) < top12(0x1.921FB6p-1f)) {
    if (top12(y) < top12(0x1p-12f)) return cosine ? 1.0f : y;
    return sincos_poly(x, x * x, *p, cosine);
  }
  // |y| < 120 assumed (callers guarantee it)
  double r = x * p->hpi_inv;
  int n = ((int32_t)r + 0x800000) >> 24;
  x = x - n * p->hpi;
  double s = p->sign[n & 3];
  if (n & 2) p = &kSinCosTab[1];
  return sincos_poly(x * s, x * x, *p, n ^ cosine);
}
static inline float libm_sinf(float y) { return sincos_eval(y, 0); }
static inline float libm_cosf(float y) { return sincos_eval(y, 1); }

// cv::fastAtan2 (scalar path): degree-valued 7th-order odd polynomial, every operation rounded
// separately (this file is compiled with -ffp-contract=off).
static inline float fast_atan2_deg(float y, float x) {
  const float p1 = 57.2836266f, p3 = -18.6674461f, p5 = 8.91400051f, p7 = -2.53972459f;
  const float eps = (float)2.2204460492503131e-16;
  float ax = std::fabs(x), ay = std::fabs(y), a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + eps);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + eps);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

static inline int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = (i < 0) ? -i : 2 * n - 2 - i;
  return i;
}

// dst is (w+2b) x (h+2b) with row pitch dpitch; src is w x h with pitch spitch (isolated).
// src may be the interior of dst itself (the reference borders every level in place,
// src/ORBextractor.cc:1122): interior rows are then left alone and only read.
inline void border_reflect101(const uint8_t* src, int w, int h, int spitch, uint8_t* dst, int dpitch,
                              int b) {
  const bool in_place = src == dst + (size_t)b * dpitch + b;
  for (int y = 0; y < h; ++y) {                       // interior rows first: copy + left/right margins
    const uint8_t* s = src + (size_t)y * spitch;
    uint8_t* d = dst + (size_t)(y + b) * dpitch;
    if (!in_place) std::memcpy(d + b, s, w);
    for (int x = 1; x <= b; ++x) { d[b - x] = s[reflect101(-x, w)]; d[b + w - 1 + x] = s[reflect101(w - 1 + x, w)]; }
  }
  for (int y = 1; y <= b; ++y) {                      // top/bottom margins are copies of finished rows
    std::memcpy(dst + (size_t)(b - y) * dpitch, dst + (size_t)(b + reflect101(-y, h)) * dpitch, w + 2 * b);
    std::memcpy(dst + (size_t)(b + h - 1 + y) * dpitch, dst + (size_t)(b + reflect101(h - 1 + y, h)) * dpitch, w + 2 * b);
  }
}

// cv::resize INTER_LINEAR, 8-bit, one channel: 11-bit fixed-point coefficients generated in
// float, int32 horizontal pass, two-step shifted vertical pass.
inline void resize_linear_u8(const uint8_t* src, int sw, int sh, int spitch, uint8_t* dst, int dw,
                             int dh, int dpitch) {
  std::vector<int> xo(dw), yo(dh);
  std::vector<short> xa(2 * dw), ya(2 * dh);
  auto coeffs = [](int dn, int sn, std::vector<int>& ofs, std::vector<short>& al) {
    double inv = (double)dn / sn, sc = 1.0 / inv;
    for (int d = 0; d < dn; ++d) {
      float f = (float)((d + 0.5) * sc - 0.5);
      int s = (int)std::floor(f);
      f -= s;
      if (s < 0) { s = 0; f = 0.f; }
      if (s >= sn - 1) { s = sn - 1; f = 0.f; }
      ofs[d] = s;
      al[2 * d] = (short)cv_round((1.f - f) * 2048.f);
      al[2 * d + 1] = (short)cv_round(f * 2048.f);
    }
  };
  coeffs(dw, sw, xo, xa);
  coeffs(dh, sh, yo, ya);
  // the horizontal pass of a source row is computed once and reused by the (up to two)
  // destination rows that tap it, as OpenCV's resize does with its row ring buffer
  std::vector<int> hb[2] = {std::vector<int>(dw), std::vector<int>(dw)};
  int have[2] = {-1, -1};
  auto hrow = [&](int sy) -> const int* {
    for (int k = 0; k < 2; ++k) if (have[k] == sy) return hb[k].data();
    const int k = (have[0] < have[1]) ? 0 : 1;       // evict the older (smaller) row
    const uint8_t* p = src + (size_t)sy * spitch;
    int* o = hb[k].data();
    for (int x = 0; x < dw; ++x) {
      const int sx0 = xo[x], sx1 = std::min(sx0 + 1, sw - 1);
      o[x] = p[sx0] * xa[2 * x] + p[sx1] * xa[2 * x + 1];
    }
    have[k] = sy;
    return o;
  };
  for (int y = 0; y < dh; ++y) {
    const int sy0 = yo[y], sy1 = std::min(sy0 + 1, sh - 1);
    const int* r0 = hrow(sy0);
    const int* r1 = hrow(sy1);
    r0 = hrow(sy0);                                   // (r0 may have been evicted by the second call when both are new)
    const int b0 = ya[2 * y], b1 = ya[2 * y + 1];
    uint8_t* d = dst + (size_t)y * dpitch;
    for (int x = 0; x < dw; ++x) {
      int v = (((b0 * (r0[x] >> 4)) >> 16) + ((b1 * (r1[x] >> 4)) >> 16) + 2) >> 2;
      d[x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
    }
  }
}

// cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on an isolated 8-bit image, cv2 4.x
// fixed-point path: taps {18,34,48,56,48,34,18}/256, 8.8 horizontal pass, rounded 16-bit shift.
inline void gauss7_sigma2(const uint8_t* src, int w, int h, int spitch, uint8_t* dst, int dpitch) {
  static const int k[7] = {18, 34, 48, 56, 48, 34, 18};
  std::vector<uint16_t> hp((size_t)w * h);
  std::vector<uint8_t> pad(w + 6);
  for (int y = 0; y < h; ++y) {
    const uint8_t* s = src + (size_t)y * spitch;
    for (int x = -3; x < w + 3; ++x) pad[x + 3] = (x >= 0 && x < w) ? s[x] : s[reflect101(x, w)];
    uint16_t* o = &hp[(size_t)y * w];
    const uint8_t* p = pad.data();
    for (int x = 0; x < w; ++x)
      o[x] = (uint16_t)(18 * (p[x] + p[x + 6]) + 34 * (p[x + 1] + p[x + 5]) + 48 * (p[x + 2] + p[x + 4]) + 56 * p[x + 3]);
  }
  for (int y = 0; y < h; ++y) {
    uint8_t* d = dst + (size_t)y * dpitch;
    const uint16_t* r[7];
    for (int j = 0; j < 7; ++j) r[j] = &hp[(size_t)reflect101(y + j - 3, h) * w];
    for (int x = 0; x < w; ++x) {
      uint32_t acc = 32768;
      for (int j = 0; j < 7; ++j) acc += (uint32_t)k[j] * r[j][x];
      d[x] = (uint8_t)(acc >> 16);
    }
  }
}

struct FastPt { int x, y, score; };

// FAST-9/16 arc score S(p): the largest t' such that p is still a corner at threshold t'-1,
// i.e. max over the 16 arcs of 9 contiguous ring pixels of max(min d_k, min -d_k).
static inline int fast_arc_score(const uint8_t* p, int pitch) {
  const int ox[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
  const int oy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  int d[25];
  for (int k = 0; k < 16; ++k) d[k] = (int)p[0] - (int)p[oy[k] * pitch + ox[k]];
  for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
  int best = -256;
  for (int k = 0; k < 16; ++k) {
    int mn = d[k], mx = d[k];
    for (int j = 1; j < 9; ++j) {
      mn = std::min(mn, d[k + j]);
      mx = std::max(mx, d[k + j]);
    }
    best = std::max(best, std::max(mn, -mx));
  }
  return best;
}

// cv::FAST(img, kps, t, nonmaxSuppression=true), TYPE_9_16.  Pixels closer than 3 to the window
// edge are never tested; NMS is strict against the 8 neighbours inside the same window, where a
// neighbour that is not a corner at t counts as 0.  Output is row-major.
// Sixteen pixels at a time with SSE2, the way OpenCV's own SIMD path works (compass pre-test on
// ring pixels 0/4/8/12, then a running count of consecutive brighter / darker ring pixels over
// 25 positions); the score of the survivors comes from fast_arc_score.  This keeps the CPU
// baseline honest: a scalar FAST would flatter the GPU.
inline void fast9_16(const uint8_t* img, int w, int h, int pitch, int t, std::vector<FastPt>& out) {
  out.clear();
  if (w < 7 || h < 7) return;
  static thread_local std::vector<int> sc;
  sc.assign((size_t)w * h, 0);
  const int ox[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
  const int oy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  int ofs[25];
  for (int k = 0; k < 25; ++k) ofs[k] = oy[k & 15] * pitch + ox[k & 15];
  const __m128i delta = _mm_set1_epi8((char)0x80), tt = _mm_set1_epi8((char)t), k8 = _mm_set1_epi8(8);
  bool any = false;
  for (int y = 3; y < h - 3; ++y) {
    const uint8_t* row = img + (size_t)y * pitch;
    int x = 3;
    for (; x + 16 <= w - 3; x += 16) {
      const uint8_t* p = row + x;
      const __m128i m = _mm_loadu_si128((const __m128i*)p);
      const __m128i v0 = _mm_xor_si128(_mm_adds_epu8(m, tt), delta), v1 = _mm_xor_si128(_mm_subs_epu8(m, tt), delta);
      const __m128i x0 = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(p + ofs[0])), delta);
      const __m128i x1 = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(p + ofs[4])), delta);
      const __m128i x2 = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(p + ofs[8])), delta);
      const __m128i x3 = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(p + ofs[12])), delta);
      const __m128i b0 = _mm_cmpgt_epi8(x0, v0), b1 = _mm_cmpgt_epi8(x1, v0), b2 = _mm_cmpgt_epi8(x2, v0), b3 = _mm_cmpgt_epi8(x3, v0);
      const __m128i d0 = _mm_cmpgt_epi8(v1, x0), d1 = _mm_cmpgt_epi8(v1, x1), d2 = _mm_cmpgt_epi8(v1, x2), d3 = _mm_cmpgt_epi8(v1, x3);
      __m128i pre = _mm_or_si128(_mm_or_si128(_mm_and_si128(b0, b1), _mm_and_si128(b1, b2)), _mm_or_si128(_mm_and_si128(b2, b3), _mm_and_si128(b3, b0)));
      pre = _mm_or_si128(pre, _mm_or_si128(_mm_or_si128(_mm_and_si128(d0, d1), _mm_and_si128(d1, d2)), _mm_or_si128(_mm_and_si128(d2, d3), _mm_and_si128(d3, d0))));
      if (!_mm_movemask_epi8(pre)) continue;
      __m128i c0 = _mm_setzero_si128(), c1 = c0, mx0 = c0, mx1 = c0;
      for (int k = 0; k < 25; ++k) {
        const __m128i xx = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(p + ofs[k])), delta);
        const __m128i g0 = _mm_cmpgt_epi8(xx, v0), g1 = _mm_cmpgt_epi8(v1, xx);
        c0 = _mm_and_si128(_mm_sub_epi8(c0, g0), g0);
        c1 = _mm_and_si128(_mm_sub_epi8(c1, g1), g1);
        mx0 = _mm_max_epu8(mx0, c0);
        mx1 = _mm_max_epu8(mx1, c1);
      }
      int mask = _mm_movemask_epi8(_mm_cmpgt_epi8(_mm_max_epu8(mx0, mx1), k8));
      for (int b = 0; mask; ++b, mask >>= 1)
        if (mask & 1) { sc[(size_t)y * w + x + b] = fast_arc_score(p + b, pitch) - 1; any = true; }
    }
    const int pa[8] = {ofs[0], ofs[1], ofs[2], ofs[3], ofs[4], ofs[5], ofs[6], ofs[7]};
    for (; x < w - 3; ++x) {        // scalar tail: opposite-pair pre-test, then the score
      const uint8_t* p = row + x;
      const int v = p[0], lo = v - t, hi = v + t;
      int m = 3;
      for (int k = 0; k < 8 && m; ++k) {
        const int a = p[pa[k]], b = p[-pa[k]];
        m &= ((a < lo) | (b < lo)) | (((a > hi) | (b > hi)) << 1);
      }
      if (!m) continue;
      const int s = fast_arc_score(p, pitch);
      if (s > t) { sc[(size_t)y * w + x] = s - 1; any = true; }
    }
  }
  if (!any) return;
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) {
      const int s = sc[(size_t)y * w + x];
      if (s == 0) continue;  // not a corner (for t >= 1 a corner has score S-1 >= t >= 1)
      const int* c = &sc[(size_t)y * w + x];
      if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] &&
          s > c[w] && s > c[w + 1])
        out.push_back({x, y, s});
    }
}

// cv::undistortPoints(src, dst, K, distCoef, noArray(), K) as Frame::UndistortKeyPoints and
// Frame::ComputeImageBounds call it (src/Frame.cc:404-434, 436-464): OpenCV 4.13 normalises with K,
// runs exactly 5 fixed-point iterations of the radial/tangential model in double (the 6-argument
// overload passes TermCriteria(MAX_ITER, 5, 0.01), so there is no early exit on the error), and
// re-projects with P = K.  d = {k1, k2, p1, p2, k3}.  Pinned against cv2.undistortPoints in
// tests/test_oracle_cv2.py.
struct Camera { float fx, fy, cx, cy, d[5]; };
inline void undistort_point(const Camera& c, float u, float v, float* ox, float* oy) {
  const double fx = c.fx, fy = c.fy, cx = c.cx, cy = c.cy, ifx = 1. / fx, ify = 1. / fy;
  double k[12] = {c.d[0], c.d[1], c.d[2], c.d[3], c.d[4], 0, 0, 0, 0, 0, 0, 0};
  double x = ((double)u - cx) * ifx, y = ((double)v - cy) * ify;
  const double x0 = x, y0 = y;
  for (int j = 0; j < 5; ++j) {
    const double r2 = x * x + y * y;
    const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
    if (icdist < 0) { x = ((double)u - cx) * ifx; y = ((double)v - cy) * ify; break; }
    const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
    const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
    x = (x0 - deltaX) * icdist;
    y = (y0 - deltaY) * icdist;
  }
  const double xx = fx * x + 0. * y + cx, yy = 0. * x + fy * y + cy, ww = 1. / (0. * x + 0. * y + 1.);
  *ox = (float)(xx * ww); *oy = (float)(yy * ww);
}
// glibc logf (sysdeps/ieee754/flt-32/e_logf.c, table of 16 and a cubic in double) for positive normal x.
// MapPoint::PredictScale (src/MapPoint.cc:385-417) calls the float overload of log, so the predicted pyramid level
// depends on these bits.  tools/sweep_logf.c checks the restatement against the host libm over all 2 130 706 432
// positive normal floats, with and without FMA contraction: 0 mismatches.
inline float libm_logf(float x) {
  static const double T[16][2] = {
      {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
      {0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2},  {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
      {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3},
      {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
      {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
      {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4},
      {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3},
      {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}};
  uint32_t ix; std::memcpy(&ix, &x, 4);
  if (ix == 0x3f800000u) return 0.f;
  const uint32_t tmp = ix - 0x3f330000u;
  const int i = (int)((tmp >> 19) % 16), k = (int32_t)tmp >> 23;
  const uint32_t iz = ix - (tmp & 0xff800000u);
  float zf; std::memcpy(&zf, &iz, 4);
  const double z = (double)zf, r = z * T[i][0] - 1, y0 = T[i][1] + (double)k * 0x1.62e42fefa39efp-1, r2 = r * r;
  double y = 0x1.5575b0be00b6ap-2 * r + -0x1.ffffef20a4123p-2;
  y = -0x1.00ea348b88334p-2 * r2 + y;
  y = y * r2 + (y0 + r);
  return (float)y;
}

}  // namespace orbo
