// ORACLE — TEST INFRASTRUCTURE ONLY (see cv_prims.hpp).  Plain-C entry points so that pytest
// (ctypes) and bench.py's CPU-baseline leg can drive the restatement in orb_oracle.hpp.
#include <chrono>
#include <thread>

#include "orb_oracle.hpp"

using namespace orbo;

extern "C" {

void* orbo_create(int nf, float sf, int nl, int ini, int mn) { return new Extractor(nf, sf, nl, ini, mn); }
void orbo_destroy(void* e) { delete (Extractor*)e; }

void orbo_tables(void* e, float* scale, float* inv, float* s2, float* is2, int* quota, int* umax) {
  Extractor* E = (Extractor*)e;
  for (int i = 0; i < E->nlevels; ++i) {
    scale[i] = E->scale[i]; inv[i] = E->invScale[i]; s2[i] = E->sigma2[i]; is2[i] = E->invSigma2[i];
    quota[i] = E->quota[i];
  }
  for (int i = 0; i <= kHalfPatch; ++i) umax[i] = E->umax[i];
}

// returns the keypoint count, <0 on degenerate geometry, -100 if cap is too small
int orbo_extract(void* e, const uint8_t* img, int w, int h, int pitch, KeyPt* kps, uint8_t* desc, int cap) {
  Extractor* E = (Extractor*)e;
  std::vector<KeyPt> k; std::vector<uint8_t> d;
  int n = E->extract(img, w, h, pitch, k, d);
  if (n < 0) return n;
  if (n > cap) return -100;
  if (n) { std::memcpy(kps, k.data(), sizeof(KeyPt) * n); std::memcpy(desc, d.data(), 32 * (size_t)n); }
  return n;
}

void orbo_level_size(void* e, int l, int* w, int* h) { Extractor* E = (Extractor*)e; *w = E->pyr[l].w; *h = E->pyr[l].h; }
// copies the bordered plane, (w+38) x (h+38), tightly packed
void orbo_level_copy(void* e, int l, uint8_t* dst) {
  Extractor* E = (Extractor*)e;
  std::memcpy(dst, E->pyr[l].buf.data(), E->pyr[l].buf.size());
}
int orbo_level_candidates(void* e, int l, KeyPt* out, int cap) {
  Extractor* E = (Extractor*)e;
  int n = (int)E->candidates[l].size();
  for (int i = 0; i < n && i < cap; ++i) out[i] = E->candidates[l][i];
  return n;
}
int orbo_level_keys(void* e, int l, KeyPt* out, int cap) {
  Extractor* E = (Extractor*)e;
  int n = (int)E->levelKeys[l].size();
  for (int i = 0; i < n && i < cap; ++i) out[i] = E->levelKeys[l][i];
  return n;
}

int orbo_stereo(void* eL, void* eR, const KeyPt* kL, const uint8_t* dL, int nL, const KeyPt* kR,
                const uint8_t* dR, int nR, float bf, float fx, float* uRight, float* depth, int* sad) {
  std::vector<KeyPt> a(kL, kL + nL), b(kR, kR + nR);
  return stereo_match(*(Extractor*)eL, *(Extractor*)eR, a, dL, b, dR, bf, fx, uRight, depth, sad);
}

int orbo_search_init(const KeyPt* k1, const uint8_t* d1, int n1, const KeyPt* k2, const uint8_t* d2,
                     int n2, const float* bounds, float* prev, float ratio, int window, int checkOri,
                     int* m12) {
  return search_for_init(k1, d1, n1, k2, d2, n2, bounds, prev, ratio, window, checkOri != 0, m12);
}

int orbo_knn_ratio(const uint8_t* dA, const float* angA, int nA, const uint8_t* dB, const float* angB,
                   int nB, int th, float ratio, int checkOri, int strict, int* m12) {
  return knn_ratio_bucket(dA, angA, nA, dB, angB, nB, th, ratio, checkOri != 0, strict != 0, m12);
}

int orbo_search_by_bow(const uint8_t* dA, const float* angA, const uint8_t* validA, const int* nodeA, int nA, const uint8_t* dB,
                       const float* angB, const uint8_t* validB, const int* nodeB, int nB, int th, float ratio, int checkOri,
                       int strict, int* m12) {
  return search_by_bow(dA, angA, validA, nodeA, nA, dB, angB, validB, nodeB, nB, th, ratio, checkOri != 0, strict != 0, m12);
}

int orbo_search_for_triangulation(const KeyPt* k1, const uint8_t* d1, const uint8_t* hasMp1, const float* uR1, const int* node1, int n1,
                                  const KeyPt* k2, const uint8_t* d2, const uint8_t* hasMp2, const float* uR2, const int* node2, int n2,
                                  const float* F12, const float* pose2, const float* Cw, const float* cam2, const float* scale2,
                                  const float* sigma2_2, int onlyStereo, int checkOri, int* m12) {
  return search_for_triangulation(k1, d1, hasMp1, uR1, node1, n1, k2, d2, hasMp2, uR2, node2, n2, F12, pose2, Cw, cam2, scale2, sigma2_2,
                                  onlyStereo != 0, checkOri != 0, m12);
}

int orbo_voc_transform(const uint8_t* vdesc, const int* child_first, const int* child_count, const int* children, const int* word_id,
                       const double* weight, int L, const uint8_t* desc, int n, int levelsup, int weighting, int norm, int* word_of,
                       int* node_of, int* bow_word, double* bow_val) {
  VocabView V{vdesc, child_first, child_count, children, word_id, weight, L};
  return voc_transform(V, desc, n, levelsup, weighting, norm, word_of, node_of, bow_word, bow_val);
}

// kf = {fx, fy, cx, cy, bf, logScale, minX, maxX, minY, maxY}; xf = R1(9) t1(3) R2(9) t2(3) Ow(3); mode bits: 1 two_step,
// 2 dist_from_centre, 4 angle_check, 8 stereo_check, 16 invz_double
void orbo_project_search(const KeyPt* k, const uint8_t* d, const float* uRight, int n, const float* kf, const float* scale,
                         const float* invSigma2, int nlevels, const float* xf, const uint8_t* valid, const float* xw,
                         const float* normal, const float* dmin, const float* dmax, const uint8_t* mpDesc, int m, float th, int mode,
                         int thDist, int* best) {
  KFSearchView K{k, d, uRight, n, {kf[6], kf[7], kf[8], kf[9]}, scale, invSigma2, nlevels, kf[5], kf[0], kf[1], kf[2], kf[3], kf[4]};
  FrameGrid grid(k, n, K.bounds);
  ProjMode M{(mode & 1) != 0, (mode & 2) != 0, (mode & 4) != 0, (mode & 8) != 0, (mode & 16) != 0, thDist};
  for (int i = 0; i < m; ++i)
    best[i] = valid[i] ? project_search_one(K, grid, xf, xf + 9, xf + 12, xf + 21, xf + 24, xw + 3 * i, normal + 3 * i, dmin[i], dmax[i],
                                            mpDesc + 32 * (size_t)i, th, M)
                       : -1;
}
float orbo_logf(float x) { return libm_logf(x); }
// kf as in orbo_project_search; gridBounds = the Frame's {minX, maxX, minY, maxY}; xf = Rcw (9), tcw (3), Ow (3)
int orbo_project_match_ordered(const KeyPt* k, const uint8_t* d, int n, const float* kf, const float* gridBounds, const float* scale, int nlevels,
                               int reloc, const float* xf, const uint8_t* valid, const float* xw, const float* normal, const float* dmin,
                               const float* dmax, const uint8_t* mpDesc, const float* srcAngle, int m, float th, int thDist, int checkOri,
                               uint8_t* occupied, int* assigned) {
  KFSearchView K{k, d, nullptr, n, {kf[6], kf[7], kf[8], kf[9]}, scale, nullptr, nlevels, kf[5], kf[0], kf[1], kf[2], kf[3], kf[4]};
  return project_match_ordered(K, gridBounds, reloc != 0, xf, xf + 9, xf + 12, valid, xw, normal, dmin, dmax, mpDesc, srcAngle, m, th, thDist,
                               checkOri != 0, occupied, assigned);
}

int orbo_distinctive_descriptor(const uint8_t* d, int n) { return distinctive_descriptor(d, n); }

int orbo_search_projection(const KeyPt* k, const uint8_t* d, const float* uRight, int n, const float* bounds, const float* scale,
                            const uint8_t* inView, const float* projX, const float* projY, const float* projXR, const int* level,
                            const float* viewCos, const uint8_t* mpDesc, const int* mpObs, int m, float th, float nnratio,
                            int* frameObs, int* assigned) {
  return search_by_projection_points(k, d, uRight, n, bounds, scale, inView, projX, projY, projXR, level, viewCos, mpDesc,
                                     mpObs, m, th, nnratio, frameObs, assigned);
}

int orbo_search_projection_frame(const KeyPt* k, const uint8_t* d, const float* uRight, int n, const float* bounds, const float* scale,
                                  const float* Rcw, const float* tcw, float fx, float fy, float cx, float cy, float bf,
                                  const uint8_t* lastValid, const float* lastXw, const int* lastOctave, const float* lastAngle,
                                  const uint8_t* mpDesc, const int* mpObs, int nLast, float th, int mode, int checkOri, int* frameObs,
                                  int* assigned) {
  return search_by_projection_frame(k, d, uRight, n, bounds, scale, Rcw, tcw, fx, fy, cx, cy, bf, lastValid, lastXw, lastOctave,
                                    lastAngle, mpDesc, mpObs, nLast, th, mode, checkOri != 0, frameObs, assigned);
}

void orbo_transform_points(const float* Rcw, const float* tcw, const float* xw, float* out, int n) {
  for (int i = 0; i < n; ++i)
    for (int r = 0; r < 3; ++r)
      out[3 * i + r] = ((Rcw[3 * r] * xw[3 * i] + Rcw[3 * r + 1] * xw[3 * i + 1]) + Rcw[3 * r + 2] * xw[3 * i + 2]) + tcw[r];
}

int orbo_hamming256(const uint8_t* a, const uint8_t* b) { return hamming256(a, b); }

// cam = {fx, fy, cx, cy, k1, k2, p1, p2, k3}
void orbo_undistort_points(const float* cam, const float* xy, int n, float* out) {
  Camera c; std::memcpy(&c, cam, sizeof(c));
  for (int i = 0; i < n; ++i) undistort_point(c, xy[2 * i], xy[2 * i + 1], out + 2 * i, out + 2 * i + 1);
}
// Frame::UndistortKeyPoints + ComputeImageBounds + AssignFeaturesToGrid: kun, bounds {minX,maxX,minY,maxY,invW,invH},
// cell_start[64*48+1] / items[n] (cell = ix*48 + iy, insertion order inside a cell)
void orbo_undistort_grid(const float* cam, const KeyPt* k, int n, int cols, int rows, KeyPt* kun, float* bounds,
                         int* cell_start, int* items) {
  Camera c; std::memcpy(&c, cam, sizeof(c));
  undistort_keypoints(c, k, n, kun);
  image_bounds(c, cols, rows, bounds);
  FrameGrid g(kun, n, bounds);
  bounds[4] = g.invW; bounds[5] = g.invH;
  int pos = 0;
  for (int ix = 0; ix < kGridCols; ++ix)
    for (int iy = 0; iy < kGridRows; ++iy) {
      cell_start[ix * kGridRows + iy] = pos;
      for (int idx : g.cell[ix][iy]) items[pos++] = idx;
    }
  cell_start[kGridCols * kGridRows] = pos;
}
int orbo_features_in_area(const KeyPt* kun, int n, const float* bounds, float x, float y, float r, int minLevel,
                          int maxLevel, int* out, int cap) {
  FrameGrid g(kun, n, bounds);
  std::vector<int> v;
  g.in_area(kun, x, y, r, minLevel, maxLevel, v);
  for (int i = 0; i < (int)v.size() && i < cap; ++i) out[i] = v[i];
  return (int)v.size();
}

// --- primitives, exported so tests can pin them against cv2 -------------------------------
void orbo_resize(const uint8_t* s, int sw, int sh, int sp, uint8_t* d, int dw, int dh, int dp) {
  resize_linear_u8(s, sw, sh, sp, d, dw, dh, dp);
}
void orbo_border(const uint8_t* s, int w, int h, int sp, uint8_t* d, int dp, int b) { border_reflect101(s, w, h, sp, d, dp, b); }
void orbo_blur(const uint8_t* s, int w, int h, int sp, uint8_t* d, int dp) { gauss7_sigma2(s, w, h, sp, d, dp); }
int orbo_fast(const uint8_t* s, int w, int h, int sp, int t, int* xys, int cap) {
  std::vector<FastPt> v;
  fast9_16(s, w, h, sp, t, v);
  for (int i = 0; i < (int)v.size() && i < cap; ++i) { xys[3 * i] = v[i].x; xys[3 * i + 1] = v[i].y; xys[3 * i + 2] = v[i].score; }
  return (int)v.size();
}
float orbo_atan2(float y, float x) { return fast_atan2_deg(y, x); }
void orbo_atan2_array(const float* y, const float* x, float* out, int n) { for (int i = 0; i < n; ++i) out[i] = fast_atan2_deg(y[i], x[i]); }
void orbo_sincos_array(const float* a, float* s, float* c, int n) { for (int i = 0; i < n; ++i) { s[i] = libm_sinf(a[i]); c[i] = libm_cosf(a[i]); } }
void orbo_libm_sincos_array(const float* a, float* s, float* c, int n) { for (int i = 0; i < n; ++i) { s[i] = sinf(a[i]); c[i] = cosf(a[i]); } }
int orbo_quadtree(const KeyPt* cand, int n, int minX, int maxX, int minY, int maxY, int N, KeyPt* out, int cap) {
  std::vector<KeyPt> c(cand, cand + n), o;
  if (!distribute_quadtree(c, minX, maxX, minY, maxY, N, o)) return -1;
  for (int i = 0; i < (int)o.size() && i < cap; ++i) out[i] = o[i];
  return (int)o.size();
}

// --- CPU baseline: the reference's structure (one extractor per eye, per-frame serial
// pipeline), one stereo frame per std::thread as Examples/Stereo/stereo_kitti.cc times it.
// imgsL/imgsR: nframes tightly packed w*h images.  Each thread owns its extractors and walks
// frames t, t+T, ...  `reps` passes over the set.  Returns wall seconds; totals via out params.
double orbo_bench_stereo(const uint8_t* imgsL, const uint8_t* imgsR, int nframes, int w, int h,
                         int nf, float sf, int nl, int ini, int mn, float bf, float fx, int threads,
                         int reps, long long* totalKeys, long long* totalMatches) {
  std::vector<long long> keys(threads, 0), matches(threads, 0);
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      Extractor L(nf, sf, nl, ini, mn), R(nf, sf, nl, ini, mn);
      std::vector<KeyPt> kl, kr; std::vector<uint8_t> dl, dr; std::vector<float> ur, dp;
      for (int r = 0; r < reps; ++r)
        for (int f = t; f < nframes; f += threads) {
          int a = L.extract(imgsL + (size_t)f * w * h, w, h, w, kl, dl);
          if (imgsR) {
            int b = R.extract(imgsR + (size_t)f * w * h, w, h, w, kr, dr);
            if (a > 0 && b > 0) {
              ur.resize(a); dp.resize(a);
              matches[t] += stereo_match(L, R, kl, dl.data(), kr, dr.data(), bf, fx, ur.data(), dp.data());
            }
            keys[t] += (b > 0 ? b : 0);
          }
          keys[t] += (a > 0 ? a : 0);
        }
    });
  for (auto& th : pool) th.join();
  double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  long long k = 0, m = 0;
  for (int t = 0; t < threads; ++t) { k += keys[t]; m += matches[t]; }
  if (totalKeys) *totalKeys = k;
  if (totalMatches) *totalMatches = m;
  return s;
}
// Full results for a batch of frames on `threads` std::threads (soak tests, bench.py's self-check): frame f goes to
// rows f (left) and nframes + f (right, when imgsR is given) of kps [.][cap] / desc [.][cap][32] / counts [.];
// uRight / depth [nframes][cap] and nMatches [nframes] for stereo.  Untouched entries keep the caller's fill.
// Returns 0, or the first negative extract() status / -100 when cap is too small.
int orbo_batch_stereo(const uint8_t* imgsL, const uint8_t* imgsR, int nframes, int w, int h, int nf, float sf, int nl,
                      int ini, int mn, float bf, float fx, int threads, int cap, KeyPt* kps, uint8_t* desc, int* counts,
                      float* uRight, float* depth, int* nMatches) {
  std::vector<int> status(threads, 0);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      Extractor L(nf, sf, nl, ini, mn), R(nf, sf, nl, ini, mn);
      std::vector<KeyPt> kl, kr; std::vector<uint8_t> dl, dr;
      for (int f = t; f < nframes && status[t] == 0; f += threads) {
        const int a = L.extract(imgsL + (size_t)f * w * h, w, h, w, kl, dl);
        if (a < 0 || a > cap) { status[t] = a < 0 ? a : -100; break; }
        counts[f] = a;
        if (a) { std::memcpy(kps + (size_t)f * cap, kl.data(), sizeof(KeyPt) * a); std::memcpy(desc + (size_t)f * cap * 32, dl.data(), 32 * (size_t)a); }
        if (!imgsR) continue;
        const int b = R.extract(imgsR + (size_t)f * w * h, w, h, w, kr, dr);
        if (b < 0 || b > cap) { status[t] = b < 0 ? b : -100; break; }
        const size_t fr = (size_t)nframes + f;
        counts[fr] = b;
        if (b) { std::memcpy(kps + fr * cap, kr.data(), sizeof(KeyPt) * b); std::memcpy(desc + fr * cap * 32, dr.data(), 32 * (size_t)b); }
        nMatches[f] = (a > 0 && b > 0) ? stereo_match(L, R, kl, dl.data(), kr, dr.data(), bf, fx, uRight + (size_t)f * cap, depth + (size_t)f * cap) : 0;
      }
    });
  for (auto& th : pool) th.join();
  for (int t = 0; t < threads; ++t) if (status[t]) return status[t];
  return 0;
}

int orbo_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
