// ORACLE — TEST INFRASTRUCTURE ONLY (see cv_prims.hpp for the rule on who may use oracle/).
//
// CPU restatement of the reference's per-frame feature front end:
//   Extractor            <- ORB_SLAM2::ORBextractor            src/ORBextractor.cc:410-470, 1043-1132
//   Extractor::pyramid   <- ORBextractor::ComputePyramid       src/ORBextractor.cc:1107-1132
//   Extractor::detect    <- ComputeKeyPointsOctTree            src/ORBextractor.cc:765-853
//   distribute_quadtree  <- DistributeOctTree / DivideNode     src/ORBextractor.cc:539-763, 481-537
//   patch_angle          <- IC_Angle                           src/ORBextractor.cc:77-104
//   steered_brief        <- computeOrbDescriptor               src/ORBextractor.cc:108-147
//   stereo_match         <- Frame::ComputeStereoMatches        src/Frame.cc:466-640
//   search_for_init      <- ORBmatcher::SearchForInitialization src/ORBmatcher.cc:405-520
//                           (+ Frame::GetFeaturesInArea / PosInGrid, src/Frame.cc:327-392)
//   knn_ratio_bucket     <- ORBmatcher::SearchByBoW(KF,KF), one bucket, src/ORBmatcher.cc:522-655
//   three_maxima         <- ORBmatcher::ComputeThreeMaxima     src/ORBmatcher.cc:1601-1642
//   hamming256           <- ORBmatcher::DescriptorDistance     src/ORBmatcher.cc:1647-1663
//
// Deliberate normalisations (DESIGN.md "hazards"):
//   H1 quad-tree ties in the careful phase are broken by node creation order (the reference
//      sorts on heap addresses, src/ORBextractor.cc:684) = the reference under a bump allocator.
//   H2 sinf/cosf are the restated glibc 2.39 routines (cv_prims.hpp), proven equal to the host's.
//   H3 compiled with -ffp-contract=off.
//   H4 Frame::mb is read before it is set (src/Frame.cc:90 vs :114); defined as mb = bf/fx.
//   H7 degenerate geometry (a level narrower than 62 px, portrait levels with round(W/H)==0, an
//      empty stereo match list) returns an error / leaves outputs untouched instead of UB.
#pragma once
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <list>
#include <utility>
#include <vector>

#include "cv_prims.hpp"

namespace orbo {

constexpr int kPatch = 31, kHalfPatch = 15, kEdge = 19;

struct KeyPt {  // same 28-byte layout as cv::KeyPoint
  float x, y, size, angle, response;
  int octave, class_id;
};

struct Plane {  // one bordered pyramid level; roi() is what mvImagePyramid[l] points at
  int w = 0, h = 0, pitch = 0;
  std::vector<uint8_t> buf;
  uint8_t* roi() { return buf.data() + (size_t)kEdge * pitch + kEdge; }
  const uint8_t* roi() const { return buf.data() + (size_t)kEdge * pitch + kEdge; }
};

static const int8_t kBrief[1024] = {
#include "brief_pattern.inc"
};

static inline int hamming256(const uint8_t* a, const uint8_t* b) {
  int d = 0;
  for (int i = 0; i < 8; ++i) {
    uint32_t u, v;
    std::memcpy(&u, a + 4 * i, 4);
    std::memcpy(&v, b + 4 * i, 4);
    d += __builtin_popcount(u ^ v);
  }
  return d;
}

// ----------------------------------------------------------------------------------------
struct QNode {
  int x0, y0, x1, y1;  // UL = (x0,y0), BR = (x1,y1)
  std::vector<KeyPt> keys;
  bool leaf = false;
  int seq = 0;  // creation counter: the H1 stand-in for the heap address
  std::list<QNode>::iterator self;
};

inline void split4(const QNode& n, QNode c[4]) {
  const int hx = (int)std::ceil((float)(n.x1 - n.x0) / 2);
  const int hy = (int)std::ceil((float)(n.y1 - n.y0) / 2);
  const int mx = n.x0 + hx, my = n.y0 + hy;
  c[0].x0 = n.x0; c[0].y0 = n.y0; c[0].x1 = mx;   c[0].y1 = my;
  c[1].x0 = mx;   c[1].y0 = n.y0; c[1].x1 = n.x1; c[1].y1 = my;
  c[2].x0 = n.x0; c[2].y0 = my;   c[2].x1 = mx;   c[2].y1 = n.y1;
  c[3].x0 = mx;   c[3].y0 = my;   c[3].x1 = n.x1; c[3].y1 = n.y1;
  for (const KeyPt& k : n.keys) {
    int q = (k.x < (float)mx) ? ((k.y < (float)my) ? 0 : 2) : ((k.y < (float)my) ? 1 : 3);
    c[q].keys.push_back(k);
  }
  for (int q = 0; q < 4; ++q) c[q].leaf = c[q].keys.size() == 1;
}

// Returns false on degenerate geometry (H7).
inline bool distribute_quadtree(const std::vector<KeyPt>& cand, int minX, int maxX, int minY,
                                int maxY, int N, std::vector<KeyPt>& out) {
  out.clear();
  const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));
  if (nIni < 1) return false;
  const float hX = (float)(maxX - minX) / nIni;
  std::list<QNode> L;
  std::vector<QNode*> roots(nIni);
  int seq = 0;
  for (int i = 0; i < nIni; ++i) {
    QNode r;
    r.x0 = (int)(hX * (float)i);
    r.x1 = (int)(hX * (float)(i + 1));
    r.y0 = 0;
    r.y1 = maxY - minY;
    r.seq = seq++;
    L.push_back(r);
    roots[i] = &L.back();
  }
  for (const KeyPt& k : cand) {
    int r = (int)(k.x / hX);
    if (r < 0 || r >= nIni) return false;
    roots[r]->keys.push_back(k);
  }
  for (auto it = L.begin(); it != L.end();) {
    if (it->keys.empty()) it = L.erase(it);
    else { it->leaf = it->keys.size() == 1; ++it; }
  }
  typedef std::pair<int, int> SizeSeq;           // (key count, creation seq)
  std::vector<std::pair<SizeSeq, QNode*>> open;  // children that can still be split
  auto push_children = [&](QNode c[4]) {
    for (int q = 0; q < 4; ++q) {
      if (c[q].keys.empty()) continue;
      c[q].seq = seq++;
      L.push_front(c[q]);
      L.front().self = L.begin();
      if (c[q].keys.size() > 1) open.push_back({{(int)c[q].keys.size(), c[q].seq}, &L.front()});
    }
  };
  bool done = false;
  while (!done) {
    const int before = (int)L.size();
    open.clear();
    int expandable = 0;
    for (auto it = L.begin(); it != L.end();) {
      if (it->leaf) { ++it; continue; }
      QNode c[4];
      split4(*it, c);
      size_t o0 = open.size();
      push_children(c);
      expandable += (int)(open.size() - o0);
      it = L.erase(it);
    }
    if ((int)L.size() >= N || (int)L.size() == before) {
      done = true;
    } else if ((int)L.size() + 3 * expandable > N) {
      while (!done) {
        const int before2 = (int)L.size();
        auto todo = open;
        open.clear();
        std::sort(todo.begin(), todo.end(),
                  [](const std::pair<SizeSeq, QNode*>& a, const std::pair<SizeSeq, QNode*>& b) {
                    return a.first < b.first;
                  });
        for (int j = (int)todo.size() - 1; j >= 0; --j) {
          QNode c[4];
          split4(*todo[j].second, c);
          push_children(c);
          L.erase(todo[j].second->self);
          if ((int)L.size() >= N) break;
        }
        if ((int)L.size() >= N || (int)L.size() == before2) done = true;
      }
    }
  }
  out.reserve(L.size());
  for (const QNode& n : L) {
    const KeyPt* best = &n.keys[0];
    for (size_t k = 1; k < n.keys.size(); ++k)
      if (n.keys[k].response > best->response) best = &n.keys[k];
    out.push_back(*best);
  }
  return true;
}

// ----------------------------------------------------------------------------------------
static inline float patch_angle(const uint8_t* roi, int pitch, float px, float py, const int* umax) {
  const uint8_t* c = roi + (size_t)cv_round(py) * pitch + cv_round(px);
  int m01 = 0, m10 = 0;
  for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
  for (int v = 1; v <= kHalfPatch; ++v) {
    int vs = 0, d = umax[v];
    for (int u = -d; u <= d; ++u) {
      int a = c[u + v * pitch], b = c[u - v * pitch];
      vs += a - b;
      m10 += u * (a + b);
    }
    m01 += v * vs;
  }
  return fast_atan2_deg((float)m01, (float)m10);
}

static inline void steered_brief(const uint8_t* img, int pitch, const KeyPt& kp, uint8_t* desc) {
  const float deg2rad = (float)(3.1415926535897932384626433832795 / 180.f);
  float ang = kp.angle * deg2rad;
  float a = libm_cosf(ang), b = libm_sinf(ang);
  const uint8_t* c = img + (size_t)cv_round(kp.y) * pitch + cv_round(kp.x);
  for (int i = 0; i < 32; ++i) {
    int byte = 0;
    for (int j = 0; j < 8; ++j) {
      const int8_t* t = kBrief + (i * 8 + j) * 4;
      float x0 = t[0], y0 = t[1], x1 = t[2], y1 = t[3];
      int v0 = c[cv_round(x0 * b + y0 * a) * pitch + cv_round(x0 * a - y0 * b)];
      int v1 = c[cv_round(x1 * b + y1 * a) * pitch + cv_round(x1 * a - y1 * b)];
      byte |= (v0 < v1) << j;
    }
    desc[i] = (uint8_t)byte;
  }
}

// ----------------------------------------------------------------------------------------
class Extractor {
 public:
  int nfeatures, nlevels, iniTh, minTh;
  double scaleFactor;
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> quota;
  int umax[kHalfPatch + 1];
  std::vector<Plane> pyr;
  std::vector<std::vector<KeyPt>> candidates;  // per level, before the quad-tree (window coords)
  std::vector<std::vector<KeyPt>> levelKeys;   // per level, after the quad-tree (level coords)

  Extractor(int nf, float sf, int nl, int ini, int mn)
      : nfeatures(nf), nlevels(nl), iniTh(ini), minTh(mn), scaleFactor(sf) {
    scale.assign(nl, 1.f); sigma2.assign(nl, 1.f); invScale.resize(nl); invSigma2.resize(nl);
    for (int i = 1; i < nl; ++i) {
      scale[i] = (float)(scale[i - 1] * scaleFactor);
      sigma2[i] = scale[i] * scale[i];
    }
    for (int i = 0; i < nl; ++i) { invScale[i] = 1.0f / scale[i]; invSigma2[i] = 1.0f / sigma2[i]; }
    pyr.resize(nl);
    quota.resize(nl);
    float factor = (float)(1.0f / scaleFactor);
    float want = nf * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
      quota[l] = cv_round(want);
      sum += quota[l];
      want *= factor;
    }
    quota[nl - 1] = std::max(nf - sum, 0);
    // disc half-widths for the orientation moments
    int vmax = (int)std::floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
    const double r2 = kHalfPatch * kHalfPatch;
    for (int v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(r2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
      while (umax[v0] == umax[v0 + 1]) ++v0;
      umax[v] = v0;
      ++v0;
    }
  }

  // 0 ok, <0 degenerate geometry
  int pyramid(const uint8_t* img, int w, int h, int pitch) {
    for (int l = 0; l < nlevels; ++l) {
      Plane& p = pyr[l];
      p.w = cv_round((float)w * invScale[l]);
      p.h = cv_round((float)h * invScale[l]);
      if (p.w < 2 * (kEdge - 3) + 30 || p.h < 2 * (kEdge - 3) + 30) return -1;
      if (l > 0 && pyr[l - 1].w == 2 * p.w && pyr[l - 1].h == 2 * p.h) return -3;   // OpenCV would use INTER_AREA here
      p.pitch = p.w + 2 * kEdge;
      p.buf.assign((size_t)p.pitch * (p.h + 2 * kEdge), 0);
      if (l == 0) {
        border_reflect101(img, w, h, pitch, p.buf.data(), p.pitch, kEdge);
      } else {
        const Plane& q = pyr[l - 1];
        resize_linear_u8(q.roi(), q.w, q.h, q.pitch, p.roi(), p.w, p.h, p.pitch);
        border_reflect101(p.roi(), p.w, p.h, p.pitch, p.buf.data(), p.pitch, kEdge);   // in place, as the reference
      }
    }
    return 0;
  }

  int detect() {
    candidates.assign(nlevels, {});
    levelKeys.assign(nlevels, {});
    const float W = 30;
    std::vector<FastPt> cell;
    for (int l = 0; l < nlevels; ++l) {
      const Plane& p = pyr[l];
      const int minBX = kEdge - 3, minBY = minBX, maxBX = p.w - kEdge + 3, maxBY = p.h - kEdge + 3;
      const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
      const int nCols = (int)(width / W), nRows = (int)(height / W);
      if (nCols < 1 || nRows < 1) return -1;
      const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
      std::vector<KeyPt>& cand = candidates[l];
      for (int i = 0; i < nRows; ++i) {
        const float iniY = (float)(minBY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBY - 3) continue;
        if (maxY > maxBY) maxY = (float)maxBY;
        for (int j = 0; j < nCols; ++j) {
          const float iniX = (float)(minBX + j * wCell);
          float maxX = iniX + wCell + 6;
          if (iniX >= maxBX - 6) continue;
          if (maxX > maxBX) maxX = (float)maxBX;
          const int x0 = (int)iniX, y0 = (int)iniY, ww = (int)maxX - x0, hh = (int)maxY - y0;
          const uint8_t* win = p.roi() + (size_t)y0 * p.pitch + x0;
          fast9_16(win, ww, hh, p.pitch, iniTh, cell);
          if (cell.empty()) fast9_16(win, ww, hh, p.pitch, minTh, cell);
          for (const FastPt& f : cell) {
            KeyPt k;
            k.x = (float)f.x + j * wCell;
            k.y = (float)f.y + i * hCell;
            k.size = 7.f; k.angle = -1.f; k.response = (float)f.score; k.octave = 0; k.class_id = -1;
            cand.push_back(k);
          }
        }
      }
      if (!distribute_quadtree(cand, minBX, maxBX, minBY, maxBY, quota[l], levelKeys[l])) return -2;
      const int patch = (int)(kPatch * scale[l]);
      for (KeyPt& k : levelKeys[l]) {
        k.x += minBX; k.y += minBY; k.octave = l; k.size = (float)patch;
      }
    }
    for (int l = 0; l < nlevels; ++l)
      for (KeyPt& k : levelKeys[l]) k.angle = patch_angle(pyr[l].roi(), pyr[l].pitch, k.x, k.y, umax);
    return 0;
  }

  // Full operator(): returns keypoint count (>=0) or a negative error.
  int extract(const uint8_t* img, int w, int h, int pitch, std::vector<KeyPt>& kps,
              std::vector<uint8_t>& desc) {
    kps.clear(); desc.clear();
    if (!img || w <= 0 || h <= 0) return 0;  // empty image: silent return (src/ORBextractor.cc:1046)
    int rc = pyramid(img, w, h, pitch);
    if (rc < 0) return rc;
    rc = detect();
    if (rc < 0) return rc * 10;
    size_t total = 0;
    for (auto& v : levelKeys) total += v.size();
    kps.reserve(total); desc.resize(total * 32);
    size_t off = 0;
    std::vector<uint8_t> blurred;
    for (int l = 0; l < nlevels; ++l) {
      std::vector<KeyPt>& ks = levelKeys[l];
      if (ks.empty()) continue;
      const Plane& p = pyr[l];
      blurred.resize((size_t)p.w * p.h);
      gauss7_sigma2(p.roi(), p.w, p.h, p.pitch, blurred.data(), p.w);
      for (size_t i = 0; i < ks.size(); ++i) steered_brief(blurred.data(), p.w, ks[i], &desc[(off + i) * 32]);
      off += ks.size();
      for (const KeyPt& k0 : ks) {
        KeyPt k = k0;
        if (l != 0) { k.x *= scale[l]; k.y *= scale[l]; }
        kps.push_back(k);
      }
    }
    return (int)total;
  }
};

// ----------------------------------------------------------------------------------------
// Frame::ComputeStereoMatches.  EL/ER hold the left/right pyramids of the preceding extract().
// uRight/depth get -1 where there is no match.  Returns the number of surviving matches.
inline int stereo_match(const Extractor& EL, const Extractor& ER, const std::vector<KeyPt>& kL,
                        const uint8_t* dL, const std::vector<KeyPt>& kR, const uint8_t* dR, float bf,
                        float fx, float* uRight, float* depth, int* sadOut = nullptr) {
  const int N = (int)kL.size();
  for (int i = 0; i < N; ++i) { uRight[i] = -1.f; depth[i] = -1.f; if (sadOut) sadOut[i] = -1; }
  const int thOrb = (100 + 50) / 2;
  const int nRows = EL.pyr[0].h;
  std::vector<std::vector<int>> rows(nRows);
  for (int iR = 0; iR < (int)kR.size(); ++iR) {
    const float y = kR[iR].y, r = 2.0f * EL.scale[kR[iR].octave];
    const int hi = (int)std::ceil(y + r), lo = (int)std::floor(y - r);
    for (int yi = lo; yi <= hi; ++yi)
      if (yi >= 0 && yi < nRows) rows[yi].push_back(iR);
  }
  const float mb = bf / fx;  // H4
  const float minD = 0, maxD = bf / mb;
  std::vector<std::pair<int, int>> sadIdx;
  for (int iL = 0; iL < N; ++iL) {
    const KeyPt& kl = kL[iL];
    const int lvl = kl.octave;
    const float vL = kl.y, uL = kl.x;
    const int row = (int)vL;
    if (row < 0 || row >= nRows) continue;
    const std::vector<int>& cands = rows[row];
    if (cands.empty()) continue;
    const float minU = uL - maxD, maxU = uL - minD;
    if (maxU < 0) continue;
    int best = 100, bestR = 0;
    for (int iR : cands) {
      const KeyPt& kr = kR[iR];
      if (kr.octave < lvl - 1 || kr.octave > lvl + 1) continue;
      if (kr.x >= minU && kr.x <= maxU) {
        int d = hamming256(dL + 32 * (size_t)iL, dR + 32 * (size_t)iR);
        if (d < best) { best = d; bestR = iR; }
      }
    }
    if (best >= thOrb) continue;
    const float uR0 = kR[bestR].x, inv = EL.invScale[lvl];
    const float su = std::round(kl.x * inv), sv = std::round(kl.y * inv), sr = std::round(uR0 * inv);
    const int w = 5, Ls = 5;
    const Plane& PL = EL.pyr[lvl];
    const Plane& PR = ER.pyr[lvl];
    const float iniu = sr + Ls - w, endu = sr + Ls + w + 1;
    if (iniu < 0 || endu >= PR.w) continue;
    const int cy = (int)sv, cxl = (int)su, cxr = (int)sr;
    const uint8_t* pl = PL.roi();
    const uint8_t* pr = PR.roi();
    const float cl = (float)pl[(size_t)cy * PL.pitch + cxl];
    int bestSad = INT_MAX, bestInc = 0;
    float dist[11];
    for (int inc = -Ls; inc <= Ls; ++inc) {
      const float cr = (float)pr[(size_t)cy * PR.pitch + cxr + inc];
      double acc = 0;
      for (int dy = -w; dy <= w; ++dy)
        for (int dx = -w; dx <= w; ++dx) {
          float a = (float)pl[(size_t)(cy + dy) * PL.pitch + cxl + dx] - cl;
          float b = (float)pr[(size_t)(cy + dy) * PR.pitch + cxr + inc + dx] - cr;
          acc += std::fabs(a - b);
        }
      float d = (float)acc;
      if (d < (float)bestSad) { bestSad = (int)d; bestInc = inc; }
      dist[Ls + inc] = d;
    }
    if (bestInc == -Ls || bestInc == Ls) continue;
    const float d1 = dist[Ls + bestInc - 1], d2 = dist[Ls + bestInc], d3 = dist[Ls + bestInc + 1];
    const float delta = (d1 - d3) / (2.0f * (d1 + d3 - 2.0f * d2));
    if (delta < -1 || delta > 1) continue;
    float bestu = EL.scale[lvl] * ((float)sr + (float)bestInc + delta);
    float disp = uL - bestu;
    if (disp >= minD && disp < maxD) {
      if (disp <= 0) { disp = 0.01; bestu = uL - 0.01; }
      depth[iL] = bf / disp;
      uRight[iL] = bestu;
      if (sadOut) sadOut[iL] = bestSad;
      sadIdx.push_back({bestSad, iL});
    }
  }
  if (sadIdx.empty()) return 0;  // H7: the reference indexes an empty vector here
  std::sort(sadIdx.begin(), sadIdx.end());
  const float median = (float)sadIdx[sadIdx.size() / 2].first;
  const float th = 1.5f * 1.4f * median;
  int kept = (int)sadIdx.size();
  for (int i = (int)sadIdx.size() - 1; i >= 0; --i) {
    if ((float)sadIdx[i].first < th) break;
    uRight[sadIdx[i].second] = -1;
    depth[sadIdx[i].second] = -1;
    --kept;
  }
  return kept;
}

// ----------------------------------------------------------------------------------------
inline void three_maxima(const std::vector<int>* histo, int L, int& i1, int& i2, int& i3) {
  int m1 = 0, m2 = 0, m3 = 0;
  for (int i = 0; i < L; ++i) {
    const int s = (int)histo[i].size();
    if (s > m1) { m3 = m2; m2 = m1; m1 = s; i3 = i2; i2 = i1; i1 = i; }
    else if (s > m2) { m3 = m2; m2 = s; i3 = i2; i2 = i; }
    else if (s > m3) { m3 = s; i3 = i; }
  }
  if (m2 < 0.1f * (float)m1) { i2 = -1; i3 = -1; }
  else if (m3 < 0.1f * (float)m1) { i3 = -1; }
}

constexpr int kGridCols = 64, kGridRows = 48, kHisto = 30;

// The 64x48 bucket grid of a frame (Frame::AssignFeaturesToGrid, src/Frame.cc:230-245) for an
// undistorted camera (k1 == 0): bounds are [0,cols) x [0,rows).
// Frame::UndistortKeyPoints (src/Frame.cc:404-434): a copy when k1 == 0
inline void undistort_keypoints(const Camera& c, const KeyPt* k, int n, KeyPt* out) {
  for (int i = 0; i < n; ++i) {
    out[i] = k[i];
    if (c.d[0] != 0.f) undistort_point(c, k[i].x, k[i].y, &out[i].x, &out[i].y);
  }
}
// Frame::ComputeImageBounds (src/Frame.cc:436-464) -> {mnMinX, mnMaxX, mnMinY, mnMaxY}
inline void image_bounds(const Camera& c, int cols, int rows, float b[4]) {
  if (c.d[0] != 0.f) {
    float px[4], py[4];
    const float cu[4] = {0.f, (float)cols, 0.f, (float)cols}, cv[4] = {0.f, 0.f, (float)rows, (float)rows};
    for (int i = 0; i < 4; ++i) undistort_point(c, cu[i], cv[i], &px[i], &py[i]);
    b[0] = std::min(px[0], px[2]); b[1] = std::max(px[1], px[3]);
    b[2] = std::min(py[0], py[1]); b[3] = std::max(py[2], py[3]);
  } else {
    b[0] = 0.f; b[1] = (float)cols; b[2] = 0.f; b[3] = (float)rows;
  }
}

// Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea (src/Frame.cc:230-245, 327-392)
struct FrameGrid {
  float minX = 0.f, minY = 0.f, invW, invH;
  std::vector<int> cell[kGridCols][kGridRows];
  FrameGrid(const KeyPt* k, int n, int cols, int rows) {
    invW = (float)kGridCols / (float)cols;
    invH = (float)kGridRows / (float)rows;
    assign(k, n);
  }
  FrameGrid(const KeyPt* k, int n, const float b[4]) {          // undistorted bounds, src/Frame.cc:101-102
    minX = b[0]; minY = b[2];
    invW = (float)kGridCols / (b[1] - b[0]);
    invH = (float)kGridRows / (b[3] - b[2]);
    assign(k, n);
  }
  void assign(const KeyPt* k, int n) {
    for (int i = 0; i < n; ++i) {
      int gx = (int)std::round((k[i].x - minX) * invW), gy = (int)std::round((k[i].y - minY) * invH);
      if (gx < 0 || gx >= kGridCols || gy < 0 || gy >= kGridRows) continue;
      cell[gx][gy].push_back(i);
    }
  }
  void in_area(const KeyPt* k, float x, float y, float r, int minLevel, int maxLevel,
               std::vector<int>& out) const {
    out.clear();
    const int cx0 = std::max(0, (int)std::floor((x - minX - r) * invW));
    if (cx0 >= kGridCols) return;
    const int cx1 = std::min(kGridCols - 1, (int)std::ceil((x - minX + r) * invW));
    if (cx1 < 0) return;
    const int cy0 = std::max(0, (int)std::floor((y - minY - r) * invH));
    if (cy0 >= kGridRows) return;
    const int cy1 = std::min(kGridRows - 1, (int)std::ceil((y - minY + r) * invH));
    if (cy1 < 0) return;
    const bool chk = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = cx0; ix <= cx1; ++ix)
      for (int iy = cy0; iy <= cy1; ++iy)
        for (int idx : cell[ix][iy]) {
          if (chk) {
            if (k[idx].octave < minLevel) continue;
            if (maxLevel >= 0 && k[idx].octave > maxLevel) continue;
          }
          if (std::fabs(k[idx].x - x) < r && std::fabs(k[idx].y - y) < r) out.push_back(idx);
        }
  }
};

static inline int rot_bin(float a1, float a2) {
  const float factor = 1.0f / kHisto;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * factor);
  if (bin == kHisto) bin = 0;
  return bin;
}

// ORBmatcher::SearchForInitialization.  prev is vbPrevMatched (in/out, xy pairs).
inline int search_for_init(const KeyPt* k1, const uint8_t* d1, int n1, const KeyPt* k2,
                           const uint8_t* d2, int n2, const float* bounds, float* prev, float ratio,
                           int window, bool checkOri, int* m12) {   // bounds = frame 2's {mnMinX, mnMaxX, mnMinY, mnMaxY}
  int nm = 0;
  for (int i = 0; i < n1; ++i) m12[i] = -1;
  std::vector<int> hist[kHisto];
  std::vector<int> dist2(n2, INT_MAX), m21(n2, -1);
  FrameGrid grid(k2, n2, bounds);
  std::vector<int> cand;
  for (int i1 = 0; i1 < n1; ++i1) {
    if (k1[i1].octave > 0) continue;
    grid.in_area(k2, prev[2 * i1], prev[2 * i1 + 1], (float)window, 0, 0, cand);
    if (cand.empty()) continue;
    int b1 = INT_MAX, b2 = INT_MAX, bi = -1;
    for (int i2 : cand) {
      int d = hamming256(d1 + 32 * (size_t)i1, d2 + 32 * (size_t)i2);
      if (dist2[i2] <= d) continue;
      if (d < b1) { b2 = b1; b1 = d; bi = i2; }
      else if (d < b2) b2 = d;
    }
    if (b1 <= 50 && (float)b1 < (float)b2 * ratio) {
      if (m21[bi] >= 0) { m12[m21[bi]] = -1; --nm; }
      m12[i1] = bi; m21[bi] = i1; dist2[bi] = b1; ++nm;
      if (checkOri) hist[rot_bin(k1[i1].angle, k2[bi].angle)].push_back(i1);
    }
  }
  if (checkOri) {
    int a = -1, b = -1, c = -1;
    three_maxima(hist, kHisto, a, b, c);
    for (int i = 0; i < kHisto; ++i) {
      if (i == a || i == b || i == c) continue;
      for (int idx : hist[i])
        if (m12[idx] >= 0) { m12[idx] = -1; --nm; }
    }
  }
  for (int i1 = 0; i1 < n1; ++i1)
    if (m12[i1] >= 0) { prev[2 * i1] = k2[m12[i1]].x; prev[2 * i1 + 1] = k2[m12[i1]].y; }
  return nm;
}

// ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th)  src/ORBmatcher.cc:45-129,
// the local-map search Tracking::SearchLocalPoints runs on every frame (src/Tracking.cc:1192).
// The MapPoint fields it reads are what Frame::isInFrustum left behind (src/Frame.cc:316-322):
// inView (mbTrackInView && !isBad), projX/projY/projXR, level (mnTrackScaleLevel), viewCos, the
// point's descriptor and Observations().  frameObs[i] is the state of F.mvpMapPoints[i] on entry:
// -1 = none, else that point's Observations(); assigned[i] receives the index of the map point
// matched to keypoint i by this call (-1 = none).  Returns nmatches.
inline int search_by_projection_points(const KeyPt* k, const uint8_t* d, const float* uRight, int n,
                                       const float* bounds, const float* scale, const uint8_t* inView, const float* projX,
                                       const float* projY, const float* projXR, const int* level,
                                       const float* viewCos, const uint8_t* mpDesc, const int* mpObs, int m,
                                       float th, float nnratio, int* frameObs, int* assigned) {
  for (int i = 0; i < n; ++i) assigned[i] = -1;
  FrameGrid grid(k, n, bounds);
  const bool bFactor = th != 1.0;
  int nm = 0;
  std::vector<int> cand;
  for (int iMP = 0; iMP < m; ++iMP) {
    if (!inView[iMP]) continue;
    const int lvl = level[iMP];
    float r = (viewCos[iMP] > 0.998) ? 2.5f : 4.0f;     // RadiusByViewingCos, :131-137
    if (bFactor) r *= th;
    grid.in_area(k, projX[iMP], projY[iMP], r * scale[lvl], lvl - 1, lvl, cand);
    if (cand.empty()) continue;
    int best = 256, bestLvl = -1, best2 = 256, bestLvl2 = -1, bestIdx = -1;
    for (int idx : cand) {
      if (frameObs[idx] > 0) continue;                    // F.mvpMapPoints[idx] && Observations() > 0
      if (uRight[idx] > 0) {
        const float er = std::fabs(projXR[iMP] - uRight[idx]);
        if (er > r * scale[lvl]) continue;
      }
      const int dist = hamming256(mpDesc + 32 * (size_t)iMP, d + 32 * (size_t)idx);
      if (dist < best) { best2 = best; best = dist; bestLvl2 = bestLvl; bestLvl = k[idx].octave; bestIdx = idx; }
      else if (dist < best2) { bestLvl2 = k[idx].octave; best2 = dist; }
    }
    if (best <= 100) {
      if (bestLvl == bestLvl2 && (float)best > nnratio * (float)best2) continue;
      assigned[bestIdx] = iMP;
      frameObs[bestIdx] = mpObs[iMP];
      ++nm;
    }
  }
  return nm;
}

// ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono)  src/ORBmatcher.cc:1328-1470,
// the motion-model search of every tracked frame (src/Tracking.cc:886,892).  Flat form: per keypoint
// of the last frame, lastValid = (mvpMapPoints[i] && !mvbOutlier[i]), the map point's world position,
// descriptor and Observations(), the keypoint's octave and angle.  Rcw/tcw are CurrentFrame.mTcw
// (row-major 3x3 and 3x1).  mode: 0 = neither, 1 = bForward, 2 = bBackward (the caller evaluates
// :1347-1348 from tlc and mb).  Camera-frame coordinates use OpenCV's small-matrix gemm arithmetic,
// which tests/test_oracle_cv2.py pins against cv2.gemm: float products and sums, left to right.
inline int search_by_projection_frame(const KeyPt* k, const uint8_t* d, const float* uRight, int n,
                                      const float* bounds, const float* scale, const float* Rcw, const float* tcw, float fx,
                                      float fy, float cx, float cy, float bf, const uint8_t* lastValid,
                                      const float* lastXw, const int* lastOctave, const float* lastAngle,
                                      const uint8_t* mpDesc, const int* mpObs, int nLast, float th, int mode,
                                      bool checkOri, int* frameObs, int* assigned) {
  for (int i = 0; i < n; ++i) assigned[i] = -1;
  FrameGrid grid(k, n, bounds);
  std::vector<int> hist[kHisto];
  std::vector<int> cand;
  int nm = 0;
  for (int i = 0; i < nLast; ++i) {
    if (!lastValid[i]) continue;
    const float* X = lastXw + 3 * (size_t)i;
    float c3[3];
    for (int r = 0; r < 3; ++r) c3[r] = ((Rcw[3 * r] * X[0] + Rcw[3 * r + 1] * X[1]) + Rcw[3 * r + 2] * X[2]) + tcw[r];
    const float xc = c3[0], yc = c3[1];
    const float invzc = (float)(1.0 / c3[2]);
    if (invzc < 0) continue;
    const float u = fx * xc * invzc + cx, v = fy * yc * invzc + cy;
    if (u < bounds[0] || u > bounds[1]) continue;      // CurrentFrame.mnMinX / mnMaxX, :1373
    if (v < bounds[2] || v > bounds[3]) continue;
    const int oct = lastOctave[i];
    const float radius = th * scale[oct];
    if (mode == 1) grid.in_area(k, u, v, radius, oct, -1, cand);
    else if (mode == 2) grid.in_area(k, u, v, radius, 0, oct, cand);
    else grid.in_area(k, u, v, radius, oct - 1, oct + 1, cand);
    if (cand.empty()) continue;
    int best = 256, bestIdx = -1;
    for (int i2 : cand) {
      if (frameObs[i2] > 0) continue;
      if (uRight[i2] > 0) {
        const float ur = u - bf * invzc;
        const float er = std::fabs(ur - uRight[i2]);
        if (er > radius) continue;
      }
      const int dist = hamming256(mpDesc + 32 * (size_t)i, d + 32 * (size_t)i2);
      if (dist < best) { best = dist; bestIdx = i2; }
    }
    if (best <= 100) {
      assigned[bestIdx] = i;
      frameObs[bestIdx] = mpObs[i];
      ++nm;
      if (checkOri) hist[rot_bin(lastAngle[i], k[bestIdx].angle)].push_back(bestIdx);
    }
  }
  if (checkOri) {
    int a = -1, b = -1, c = -1;
    three_maxima(hist, kHisto, a, b, c);
    for (int i = 0; i < kHisto; ++i) {
      if (i == a || i == b || i == c) continue;
      for (int idx : hist[i]) { assigned[idx] = -1; --nm; }
    }
  }
  return nm;
}

// ORBmatcher::SearchByBoW(KF,KF) with every feature valid and a single vocabulary bucket holding
// all features of each side in index order (BASELINE config 5).  strict: best < th (KF,KF form,
// src/ORBmatcher.cc:598) vs best <= th (KF,Frame form, :228).
inline int knn_ratio_bucket(const uint8_t* dA, const float* angA, int nA, const uint8_t* dB,
                            const float* angB, int nB, int th, float ratio, bool checkOri,
                            bool strict, int* m12) {
  for (int i = 0; i < nA; ++i) m12[i] = -1;
  std::vector<char> used(nB, 0);
  std::vector<int> hist[kHisto];
  int nm = 0;
  for (int i = 0; i < nA; ++i) {
    int b1 = 256, b2 = 256, bi = -1;
    for (int j = 0; j < nB; ++j) {
      if (used[j]) continue;
      int d = hamming256(dA + 32 * (size_t)i, dB + 32 * (size_t)j);
      if (d < b1) { b2 = b1; b1 = d; bi = j; }
      else if (d < b2) b2 = d;
    }
    bool pass = strict ? (b1 < th) : (b1 <= th);
    if (pass && (float)b1 < ratio * (float)b2) {
      m12[i] = bi; used[bi] = 1;
      if (checkOri) hist[rot_bin(angA[i], angB[bi])].push_back(i);
      ++nm;
    }
  }
  if (checkOri) {
    int a = -1, b = -1, c = -1;
    three_maxima(hist, kHisto, a, b, c);
    for (int i = 0; i < kHisto; ++i) {
      if (i == a || i == b || i == c) continue;
      for (int idx : hist[i]) { m12[idx] = -1; --nm; }
    }
  }
  return nm;
}

// ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...) (src/ORBmatcher.cc:522-655, strict = true: bestDist1 < TH_LOW)
// and SearchByBoW(KeyFrame*, Frame&, ...) (src/ORBmatcher.cc:159-288, strict = false: bestDist1 <= TH_LOW) on flat
// arrays.  The DBoW2::FeatureVector of each side is given as the vocabulary node id of every feature
// (nodeA/nodeB; std::map iteration = ascending node id, the index list of a node is in ascending feature index,
// Thirdparty/DBoW2/DBoW2/FeatureVector.cpp:28-45).  validA = pMP1 && !pMP1->isBad(); validB = pMP2 && !isBad() for
// the keyframe form, all ones for the frame form.  m12[i] = matched B index or -1.
inline int search_by_bow(const uint8_t* dA, const float* angA, const uint8_t* validA, const int* nodeA, int nA,
                         const uint8_t* dB, const float* angB, const uint8_t* validB, const int* nodeB, int nB, int th,
                         float ratio, bool checkOri, bool strict, int* m12) {
  std::vector<std::pair<int, std::vector<int>>> fa, fb;                 // FeatureVector: (node, indices) ascending
  auto build = [](const int* node, int n, std::vector<std::pair<int, std::vector<int>>>& f) {
    std::vector<std::pair<int, int>> o(n);
    for (int i = 0; i < n; ++i) o[i] = {node[i], i};
    std::sort(o.begin(), o.end());
    for (auto& e : o) {
      if (f.empty() || f.back().first != e.first) f.push_back({e.first, {}});
      f.back().second.push_back(e.second);
    }
  };
  build(nodeA, nA, fa); build(nodeB, nB, fb);
  for (int i = 0; i < nA; ++i) m12[i] = -1;
  std::vector<char> used(nB, 0);
  std::vector<int> hist[kHisto];
  int nm = 0;
  size_t ia = 0, ib = 0;
  while (ia < fa.size() && ib < fb.size()) {
    if (fa[ia].first == fb[ib].first) {
      for (int i : fa[ia].second) {
        if (!validA[i]) continue;
        int b1 = 256, b2 = 256, bi = -1;
        for (int j : fb[ib].second) {
          if (used[j] || !validB[j]) continue;
          const int d = hamming256(dA + 32 * (size_t)i, dB + 32 * (size_t)j);
          if (d < b1) { b2 = b1; b1 = d; bi = j; }
          else if (d < b2) b2 = d;
        }
        const bool pass = strict ? (b1 < th) : (b1 <= th);
        if (pass && (float)b1 < ratio * (float)b2) {
          m12[i] = bi; used[bi] = 1;
          if (checkOri) hist[rot_bin(angA[i], angB[bi])].push_back(i);
          ++nm;
        }
      }
      ++ia; ++ib;
    } else if (fa[ia].first < fb[ib].first) {
      while (ia < fa.size() && fa[ia].first < fb[ib].first) ++ia;     // lower_bound
    } else {
      while (ib < fb.size() && fb[ib].first < fa[ia].first) ++ib;
    }
  }
  if (checkOri) {
    int a = -1, b = -1, c = -1;
    three_maxima(hist, kHisto, a, b, c);
    for (int i = 0; i < kHisto; ++i) {
      if (i == a || i == b || i == c) continue;
      for (int idx : hist[i]) { m12[idx] = -1; --nm; }
    }
  }
  return nm;
}

// ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-823) + CheckDistEpipolarLine (:140-157) on flat arrays.
// hasMp = GetMapPoint(idx) != NULL, uRight = mvuRight, node = FeatureVector node per feature (as in search_by_bow).
// pose2 = R2w row-major (9) then t2w (3), Cw = camera centre of KF1; cam2 = fx, fy, cx, cy of KF2.  The epipole uses
// OpenCV's small-matrix gemm arithmetic for R2w*Cw+t2w (float products and sums left to right).  Note that the
// reference never sets vbMatched2 (:686, :723): several features of KF1 may take the same feature of KF2, and the
// queries are independent.  m12[i] = matched KF2 index or -1; the caller lists the pairs in index order (:812-818).
inline int search_for_triangulation(const KeyPt* k1, const uint8_t* d1, const uint8_t* hasMp1, const float* uR1, const int* node1,
                                    int n1, const KeyPt* k2, const uint8_t* d2, const uint8_t* hasMp2, const float* uR2,
                                    const int* node2, int n2, const float* F12, const float* pose2, const float* Cw,
                                    const float* cam2, const float* scale2, const float* sigma2_2, bool onlyStereo,
                                    bool checkOri, int* m12) {
  float C2[3];
  for (int r = 0; r < 3; ++r) C2[r] = ((pose2[3 * r] * Cw[0] + pose2[3 * r + 1] * Cw[1]) + pose2[3 * r + 2] * Cw[2]) + pose2[9 + r];
  const float invz = 1.0f / C2[2];
  const float ex = cam2[0] * C2[0] * invz + cam2[2], ey = cam2[1] * C2[1] * invz + cam2[3];
  std::vector<std::pair<int, std::vector<int>>> fa, fb;
  auto build = [](const int* node, int n, std::vector<std::pair<int, std::vector<int>>>& f) {
    std::vector<std::pair<int, int>> o(n);
    for (int i = 0; i < n; ++i) o[i] = {node[i], i};
    std::sort(o.begin(), o.end());
    for (auto& e : o) {
      if (f.empty() || f.back().first != e.first) f.push_back({e.first, {}});
      f.back().second.push_back(e.second);
    }
  };
  build(node1, n1, fa); build(node2, n2, fb);
  for (int i = 0; i < n1; ++i) m12[i] = -1;
  std::vector<int> hist[kHisto];
  int nm = 0;
  size_t ia = 0, ib = 0;
  while (ia < fa.size() && ib < fb.size()) {
    if (fa[ia].first == fb[ib].first) {
      for (int idx1 : fa[ia].second) {
        if (hasMp1[idx1]) continue;
        const bool st1 = uR1[idx1] >= 0;
        if (onlyStereo && !st1) continue;
        const KeyPt& kp1 = k1[idx1];
        int bestDist = 50, bestIdx2 = -1;
        for (int idx2 : fb[ib].second) {
          if (hasMp2[idx2]) continue;
          const bool st2 = uR2[idx2] >= 0;
          if (onlyStereo && !st2) continue;
          const int dist = hamming256(d1 + 32 * (size_t)idx1, d2 + 32 * (size_t)idx2);
          if (dist > 50 || dist > bestDist) continue;
          const KeyPt& kp2 = k2[idx2];
          if (!st1 && !st2) {
            const float distex = ex - kp2.x, distey = ey - kp2.y;
            if (distex * distex + distey * distey < 100 * scale2[kp2.octave]) continue;
          }
          const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
          const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
          const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
          const float num = a * kp2.x + b * kp2.y + c, den = a * a + b * b;
          if (den == 0) continue;
          const float dsqr = num * num / den;
          if ((double)dsqr < 3.84 * (double)sigma2_2[kp2.octave]) { bestIdx2 = idx2; bestDist = dist; }
        }
        if (bestIdx2 >= 0) {
          m12[idx1] = bestIdx2; ++nm;
          if (checkOri) hist[rot_bin(kp1.angle, k2[bestIdx2].angle)].push_back(idx1);
        }
      }
      ++ia; ++ib;
    } else if (fa[ia].first < fb[ib].first) {
      while (ia < fa.size() && fa[ia].first < fb[ib].first) ++ia;
    } else {
      while (ib < fb.size() && fb[ib].first < fa[ia].first) ++ib;
    }
  }
  if (checkOri) {
    int a = -1, b = -1, c = -1;
    three_maxima(hist, kHisto, a, b, c);
    for (int i = 0; i < kHisto; ++i) {
      if (i == a || i == b || i == c) continue;
      for (int idx : hist[i]) { m12[idx] = -1; --nm; }
    }
  }
  return nm;
}

// DBoW2 TemplatedVocabulary::transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1196 for a feature set,
// :1216-1258 for one feature, FORB::distance FORB.cpp:81-101, BowVector::addWeight / addIfNotExist / normalize
// BowVector.cpp:32-92, FeatureVector::addFeature FeatureVector.cpp:28-45), as Frame::ComputeBoW calls it
// (src/Frame.cc:395-402, levelsup = 4).  The vocabulary tree is given flat: for node n its children are
// children[child_first[n] .. child_first[n] + child_count[n]) in m_nodes[n].children order, a leaf has
// child_count 0 and carries word_id / weight.  weighting: 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY; norm: 0 none, 1 L1, 2 L2
// (what ScoringObject::mustNormalize returns; ORBvoc.txt is TF_IDF + L1).
struct VocabView {
  const uint8_t* desc; const int* child_first; const int* child_count; const int* children;
  const int* word_id; const double* weight; int L;
};
inline void voc_transform_one(const VocabView& V, const uint8_t* f, int levelsup, int* word, double* w, int* nid) {
  const int nid_level = V.L - levelsup;
  if (nid_level <= 0) *nid = 0;
  int final_id = 0, level = 0;
  do {
    ++level;
    const int* ch = V.children + V.child_first[final_id];
    const int nc = V.child_count[final_id];
    final_id = ch[0];
    int best = hamming256(f, V.desc + 32 * (size_t)final_id);
    for (int c = 1; c < nc; ++c) {
      const int d = hamming256(f, V.desc + 32 * (size_t)ch[c]);
      if (d < best) { best = d; final_id = ch[c]; }
    }
    if (level == nid_level) *nid = final_id;
  } while (V.child_count[final_id] > 0);
  *word = V.word_id[final_id]; *w = V.weight[final_id];
}
// Returns the BowVector size; bow_word / bow_val in ascending word id; word_of / node_of per feature (-1 if stopped).
inline int voc_transform(const VocabView& V, const uint8_t* desc, int n, int levelsup, int weighting, int norm,
                         int* word_of, int* node_of, int* bow_word, double* bow_val) {
  std::vector<std::pair<int, double>> v;                               // BowVector (std::map) kept sorted
  auto find = [&](int id) { return std::lower_bound(v.begin(), v.end(), std::make_pair(id, -1e300)); };
  for (int i = 0; i < n; ++i) {
    int id, nid = 0; double w;
    voc_transform_one(V, desc + 32 * (size_t)i, levelsup, &id, &w, &nid);
    word_of[i] = -1; node_of[i] = -1;
    if (w > 0) {
      auto it = find(id);
      const bool exists = it != v.end() && it->first == id;
      if (weighting <= 1) { if (exists) it->second += w; else v.insert(it, {id, w}); }
      else if (!exists) v.insert(it, {id, w});
      word_of[i] = id; node_of[i] = nid;
    }
  }
  if (weighting <= 1 && !v.empty() && norm == 0) {
    const double nd = (double)v.size();
    for (auto& e : v) e.second /= nd;
  }
  if (norm != 0) {
    double nrm = 0.0;
    if (norm == 1) for (auto& e : v) nrm += std::fabs(e.second);
    else { for (auto& e : v) nrm += e.second * e.second; nrm = std::sqrt(nrm); }
    if (nrm > 0.0) for (auto& e : v) e.second /= nrm;
  }
  for (size_t k = 0; k < v.size(); ++k) { bow_word[k] = v[k].first; bow_val[k] = v[k].second; }
  return (int)v.size();
}

// The projection searches of the back end on flat arrays: ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:825-975)
// and the two passes of ORBmatcher::SearchBySim3 (:1102-1326).  A map point is carried into the keyframe's camera by one or
// two rigid / similarity steps (gemm arithmetic: float products, sums left to right), projected, range- and (Fuse) angle-
// checked, its pyramid level predicted with MapPoint::PredictScale (src/MapPoint.cc:385-400: ceil(logf(dmax/dist)/logScale)),
// and the best descriptor among the keyframe's features in the window (KeyFrame::GetFeaturesInArea order, levels
// [pred-1, pred]) is taken.  Returns the keypoint index or -1.
struct KFSearchView {
  const KeyPt* k; const uint8_t* d; const float* uRight; int n;
  float bounds[4];                                   // mnMinX, mnMaxX, mnMinY, mnMaxY
  const float* scale; const float* invSigma2; int nlevels; float logScale;
  float fx, fy, cx, cy, bf;
};
struct ProjMode {
  bool two_step;          // Sim3: p = (R2,t2) o (R1,t1);  Fuse: p = (R1,t1)
  bool dist_from_centre;  // Fuse: dist3D = |p3Dw - Ow|, else |p in the target camera|
  bool angle_check;       // Fuse: PO.dot(Pn) < 0.5*dist3D rejects
  bool stereo_check;      // Fuse: chi2 gate on the reprojection error (7.8 stereo / 5.99 mono)
  bool invz_double;       // 1.0/z (double, then float) vs 1/z (float)
  int th_dist;            // TH_LOW (Fuse) / TH_HIGH (SearchBySim3)
};
inline int project_search_one(const KFSearchView& K, const FrameGrid& grid, const float* R1, const float* t1, const float* R2,
                              const float* t2, const float* Ow, const float* xw, const float* normal, float dmin, float dmax,   // dmin / dmax = mfMinDistance / mfMaxDistance
                              const uint8_t* desc, float th, const ProjMode& M) {
  float p[3], q[3];
  for (int r = 0; r < 3; ++r) p[r] = ((R1[3 * r] * xw[0] + R1[3 * r + 1] * xw[1]) + R1[3 * r + 2] * xw[2]) + t1[r];
  if (M.two_step) { for (int r = 0; r < 3; ++r) q[r] = ((R2[3 * r] * p[0] + R2[3 * r + 1] * p[1]) + R2[3 * r + 2] * p[2]) + t2[r]; }
  else { q[0] = p[0]; q[1] = p[1]; q[2] = p[2]; }
  if (q[2] < 0.0f) return -1;
  const float invz = M.invz_double ? (float)(1.0 / (double)q[2]) : 1 / q[2];
  const float x = q[0] * invz, y = q[1] * invz;
  const float u = K.fx * x + K.cx, v = K.fy * y + K.cy;
  if (!(u >= K.bounds[0] && u < K.bounds[1] && v >= K.bounds[2] && v < K.bounds[3])) return -1;   // KeyFrame::IsInImage
  const float ur = u - K.bf * invz;
  float PO[3] = {q[0], q[1], q[2]};
  if (M.dist_from_centre) for (int r = 0; r < 3; ++r) PO[r] = xw[r] - Ow[r];
  const float dist3D = (float)std::sqrt((double)PO[0] * PO[0] + (double)PO[1] * PO[1] + (double)PO[2] * PO[2]);   // cv::norm: double sums
  if (dist3D < 0.8f * dmin || dist3D > 1.2f * dmax) return -1;      // Get{Min,Max}DistanceInvariance, src/MapPoint.cc:373-383
  if (M.angle_check) {
    const double dot = (double)PO[0] * normal[0] + (double)PO[1] * normal[1] + (double)PO[2] * normal[2];           // Mat::dot: double sums
    if (dot < 0.5 * dist3D) return -1;
  }
  const float ratio = dmax / dist3D;
  int level = (int)std::ceil(libm_logf(ratio) / K.logScale);
  if (level < 0) level = 0; else if (level >= K.nlevels) level = K.nlevels - 1;
  const float radius = th * K.scale[level];
  std::vector<int> cand;
  grid.in_area(K.k, u, v, radius, -1, -1, cand);
  int bestDist = INT_MAX, bestIdx = -1;
  for (int idx : cand) {
    const KeyPt& kp = K.k[idx];
    if (kp.octave < level - 1 || kp.octave > level) continue;
    if (M.stereo_check) {
      const float ex = u - kp.x, ey = v - kp.y;
      if (K.uRight[idx] >= 0) {
        const float er = ur - K.uRight[idx];
        const float e2 = ex * ex + ey * ey + er * er;
        if (e2 * K.invSigma2[kp.octave] > 7.8) continue;
      } else {
        const float e2 = ex * ex + ey * ey;
        if (e2 * K.invSigma2[kp.octave] > 5.99) continue;
      }
    }
    const int dist = hamming256(desc, K.d + 32 * (size_t)idx);
    if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
  }
  return bestDist <= M.th_dist ? bestIdx : -1;
}

// The two remaining SearchByProjection members, which claim keypoints as they go (hazard H6):
//   SearchByProjection(pKF, Scw, vpPoints, vpMatched, th)                         src/ORBmatcher.cc:290-403   (reloc = false)
//   SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)             src/ORBmatcher.cc:1472-1599 (reloc = true)
// Map points are visited in list order; occupied[i] != 0 marks a keypoint that already holds a map point (vpMatched[i] /
// CurrentFrame.mvpMapPoints[i]); an accepted match occupies its keypoint.  assigned[i] = position of the map point given to
// keypoint i by this call, or -1.  The two differ in how the projection is written, in the image test, the viewing-angle
// test, the level window ([pred-1, pred] vs [pred-1, pred+1]) and the rotation histogram (reloc only).
inline int project_match_ordered(const KFSearchView& K, const float* gridBounds, bool reloc, const float* Rcw, const float* tcw,
                                 const float* Ow, const uint8_t* valid, const float* xw, const float* normal, const float* dminA,
                                 const float* dmaxA, const uint8_t* desc, const float* srcAngle, int m, float th, int thDist,
                                 bool checkOri, uint8_t* occupied, int* assigned) {
  FrameGrid grid(K.k, K.n, gridBounds);
  // the query window uses the target's own bounds (KeyFrame: ints, Frame: floats); cells were assigned with the Frame's
  grid.minX = K.bounds[0]; grid.minY = K.bounds[2];
  for (int i = 0; i < K.n; ++i) assigned[i] = -1;
  std::vector<int> hist[kHisto];
  int nm = 0;
  std::vector<int> cand;
  for (int iMP = 0; iMP < m; ++iMP) {
    if (!valid[iMP]) continue;
    const float* X = xw + 3 * iMP;
    float p[3];
    for (int r = 0; r < 3; ++r) p[r] = ((Rcw[3 * r] * X[0] + Rcw[3 * r + 1] * X[1]) + Rcw[3 * r + 2] * X[2]) + tcw[r];
    float u, v;
    if (!reloc) {
      if (p[2] < 0.0f) continue;
      const float invz = 1 / p[2];
      const float x = p[0] * invz, y = p[1] * invz;
      u = K.fx * x + K.cx; v = K.fy * y + K.cy;
      if (!(u >= K.bounds[0] && u < K.bounds[1] && v >= K.bounds[2] && v < K.bounds[3])) continue;
    } else {
      const float invzc = (float)(1.0 / (double)p[2]);
      u = K.fx * p[0] * invzc + K.cx; v = K.fy * p[1] * invzc + K.cy;
      if (u < K.bounds[0] || u > K.bounds[1]) continue;
      if (v < K.bounds[2] || v > K.bounds[3]) continue;
    }
    const float PO[3] = {X[0] - Ow[0], X[1] - Ow[1], X[2] - Ow[2]};
    const float dist3D = (float)std::sqrt((double)PO[0] * PO[0] + (double)PO[1] * PO[1] + (double)PO[2] * PO[2]);
    if (dist3D < 0.8f * dminA[iMP] || dist3D > 1.2f * dmaxA[iMP]) continue;
    if (!reloc) {
      const float* nn = normal + 3 * iMP;
      const double dot = (double)PO[0] * nn[0] + (double)PO[1] * nn[1] + (double)PO[2] * nn[2];
      if (dot < 0.5 * dist3D) continue;
    }
    int level = (int)std::ceil(libm_logf(dmaxA[iMP] / dist3D) / K.logScale);
    if (level < 0) level = 0; else if (level >= K.nlevels) level = K.nlevels - 1;
    const float radius = th * K.scale[level];
    grid.in_area(K.k, u, v, radius, -1, -1, cand);
    int bestDist = 256, bestIdx = -1;
    for (int idx : cand) {
      if (occupied[idx]) continue;
      const int oct = K.k[idx].octave;
      if (oct < level - 1 || oct > (reloc ? level + 1 : level)) continue;
      const int dist = hamming256(desc + 32 * (size_t)iMP, K.d + 32 * (size_t)idx);
      if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
    if (bestDist <= thDist) {
      occupied[bestIdx] = 1; assigned[bestIdx] = iMP; ++nm;
      if (reloc && checkOri) hist[rot_bin(srcAngle[iMP], K.k[bestIdx].angle)].push_back(bestIdx);
    }
  }
  if (reloc && checkOri) {
    int a = -1, b = -1, c = -1;
    three_maxima(hist, kHisto, a, b, c);
    for (int i = 0; i < kHisto; ++i) {
      if (i == a || i == b || i == c) continue;
      for (int idx : hist[i]) { assigned[idx] = -1; --nm; }
    }
  }
  return nm;
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-305): among the N observed descriptors, the one
// whose median distance to all of them (itself included, sorted row element [0.5*(N-1)]) is least; first on ties.
inline int distinctive_descriptor(const uint8_t* d, int N) {
  if (N <= 0) return -1;
  int bestMedian = INT_MAX, bestIdx = 0;
  std::vector<int> row(N);
  for (int i = 0; i < N; ++i) {
    for (int j = 0; j < N; ++j) row[j] = i == j ? 0 : hamming256(d + 32 * (size_t)i, d + 32 * (size_t)j);
    std::sort(row.begin(), row.end());
    const int median = row[(int)(0.5 * (N - 1))];
    if (median < bestMedian) { bestMedian = median; bestIdx = i; }
  }
  return bestIdx;
}

}  // namespace orbo
