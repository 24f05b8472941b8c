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
inline void border_reflect101(const uint8_t* src, int w, int h, int spitch, uint8_t* dst, int dpitch,
                              int b) {
  for (int y = -b; y < h + b; ++y) {
    const uint8_t* s = src + (size_t)reflect101(y, h) * spitch;
    uint8_t* d = dst + (size_t)(y + b) * dpitch;
    for (int x = -b; x < w + b; ++x) d[x + b] = s[reflect101(x, w)];
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
  std::vector<int> r0(dw), r1(dw);
  for (int y = 0; y < dh; ++y) {
    int sy0 = yo[y], sy1 = std::min(sy0 + 1, sh - 1);
    const uint8_t* p0 = src + (size_t)sy0 * spitch;
    const uint8_t* p1 = src + (size_t)sy1 * spitch;
    for (int x = 0; x < dw; ++x) {
      int sx0 = xo[x], sx1 = std::min(sx0 + 1, sw - 1);
      r0[x] = p0[sx0] * xa[2 * x] + p0[sx1] * xa[2 * x + 1];
      r1[x] = p1[sx0] * xa[2 * x] + p1[sx1] * xa[2 * x + 1];
    }
    int b0 = ya[2 * y], b1 = ya[2 * y + 1];
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
  for (int y = 0; y < h; ++y) {
    const uint8_t* s = src + (size_t)y * spitch;
    uint16_t* o = &hp[(size_t)y * w];
    for (int x = 0; x < w; ++x) {
      int acc = 0;
      if (x >= 3 && x < w - 3) {
        for (int i = 0; i < 7; ++i) acc += k[i] * s[x + i - 3];
      } else {
        for (int i = 0; i < 7; ++i) acc += k[i] * s[reflect101(x + i - 3, w)];
      }
      o[x] = (uint16_t)acc;
    }
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
inline void fast9_16(const uint8_t* img, int w, int h, int pitch, int t, std::vector<FastPt>& out) {
  out.clear();
  if (w < 7 || h < 7) return;
  std::vector<int> sc((size_t)w * h, 0);
  // ring offsets of the 8 opposite pairs (k, k+8): any 9-arc holds one pixel of every pair, so a
  // corner needs every pair to have a pixel on the same side of the [v-t, v+t] band.
  const int pa[8] = {3 * pitch, 3 * pitch + 1, 2 * pitch + 2, pitch + 3, 3, -pitch + 3, -2 * pitch + 2, -3 * pitch + 1};
  bool any = false;
  for (int y = 3; y < h - 3; ++y) {
    const uint8_t* row = img + (size_t)y * pitch;
    for (int x = 3; x < w - 3; ++x) {
      const uint8_t* p = row + x;
      const int v = p[0], lo = v - t, hi = v + t;
      int m = 3;
      for (int k = 0; k < 8 && m; ++k) {
        const int a = p[pa[k]], b = p[-pa[k]];
        m &= ((a < lo) | (b < lo)) | (((a > hi) | (b > hi)) << 1);
      }
      if (!m) continue;
      int s = fast_arc_score(p, pitch);
      if (s > t) { sc[(size_t)y * w + x] = s - 1; any = true; }
    }
  }
  if (!any) return;
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) {
      int s = sc[(size_t)y * w + x];
      if (s == 0) continue;  // not a corner (for t >= 1 a corner has score S-1 >= t >= 1)
      const int* c = &sc[(size_t)y * w + x];
      if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] &&
          s > c[w] && s > c[w + 1])
        out.push_back({x, y, s});
    }
}

}  // namespace orbo
