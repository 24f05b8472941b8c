// dropin/ORBmatcher.cc — takes the place of the reference's src/ORBmatcher.cc: every public member of
// ORB_SLAM2::ORBmatcher (include/ORBmatcher.h:41-86), implemented on the orbx C ABI (include/orbx.h).
//
// It is compiled inside the reference tree, against the reference's own Frame.h / KeyFrame.h / MapPoint.h; each member
// flattens what the kernel needs out of those objects, makes ONE call, and writes the result back the way the reference's
// loop does (including Fuse's map updates, which stay on the host).  oracle/_ref/orbref_slam_gpu is the reference's
// Frame / KeyFrame / MapPoint / Map sources linked with this file and dropin/ORBextractor.h instead of src/ORBmatcher.cc
// and src/ORBextractor.cc; tests/test_zz_dropin.py checks that it writes the same bytes as the all-reference binary.
//
// Two things the reference keeps protected are needed: MapPoint::mfMinDistance / mfMaxDistance (only their 0.8x / 1.2x
// products are public, and dividing those back is not exact in float).  A maintainer adds the two getters shown in
// INTEGRATION.md; so that this file also compiles against an unpatched tree it reads them through pointers to members
// named via a derived class (MapPointFields below), which is plain C++ and touches no declaration of the reference.
#include <algorithm>
#include <cmath>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <opencv2/core/core.hpp>
#include "MapPoint.h"
#include "Frame.h"
#include "KeyFrame.h"
#include "ORBmatcher.h"
#include "orbx.h"

using namespace std;

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = 100;
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;

namespace {

// &Derived::member for a protected member of the base has type `T MapPoint::*` and passes the access check.  Position, normal
// and descriptor are read in place under the reference's own mutexes (what GetWorldPos / GetNormal / GetDescriptor do, minus
// the three cv::Mat clones - heap allocations - per map point, which were most of the cost of flattening a few hundred points).
struct MapPointFields : MapPoint {
  static float MinDistance(MapPoint* p) { return p->*(&MapPointFields::mfMinDistance); }
  static float MaxDistance(MapPoint* p) { return p->*(&MapPointFields::mfMaxDistance); }
  static void PosNormal(MapPoint* p, float* xw, float* nn) {
    std::unique_lock<std::mutex> lock(p->*(&MapPointFields::mMutexPos));
    const cv::Mat& X = p->*(&MapPointFields::mWorldPos);
    const cv::Mat& N = p->*(&MapPointFields::mNormalVector);
    for (int a = 0; a < 3; ++a) { xw[a] = X.at<float>(a); if (nn) nn[a] = N.at<float>(a); }
  }
  static void Descriptor(MapPoint* p, unsigned char* out) {
    std::unique_lock<std::mutex> lock(p->*(&MapPointFields::mMutexFeatures));
    const cv::Mat& d = p->*(&MapPointFields::mDescriptor);
    std::copy(d.ptr(0), d.ptr(0) + 32, out);
  }
};

// Frame's static image bounds and inverse cell sizes (src/Frame.cc:207-213) as the C ABI's grid_bounds[6]
struct FrameBounds {
  float v[6];
  FrameBounds() {
    v[0] = Frame::mnMinX; v[1] = Frame::mnMaxX; v[2] = Frame::mnMinY; v[3] = Frame::mnMaxY;
    v[4] = Frame::mfGridElementWidthInv; v[5] = Frame::mfGridElementHeightInv;
  }
};

void check(orbx_status s) {
  if (s != ORBX_OK) throw std::runtime_error(std::string("orbx: ") + orbx_last_error_string());
}

// mvKeysUn / mDescriptors / mvuRight of a Frame or KeyFrame as the flat arrays the C ABI takes.  cv::KeyPoint and
// orbx_keypoint have the same layout and the descriptor matrix is N contiguous rows of 32 bytes, so when no padding is asked
// for (cap == N, the usual case) the three arrays are the object's own storage: nothing is copied.
template <class T> struct View {
  const T* p = nullptr; std::vector<T> own;
  const T* data() const { return p; }
  const T& operator[](size_t i) const { return p[i]; }
};
struct Flat {
  int n = 0, cap = 1;
  View<orbx_keypoint> kps; View<unsigned char> desc; View<float> uright;
  template <class F> explicit Flat(const F& f, int cap_ = 0) {
    n = f.N; cap = std::max(std::max(n, cap_), 1);
    static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");
    if (n > 0 && cap == n && f.mDescriptors.isContinuous() && f.mDescriptors.rows >= n && f.mDescriptors.cols == 32 && f.mDescriptors.type() == CV_8UC1 &&
        (int)f.mvKeysUn.size() >= n && (int)f.mvuRight.size() >= n) {
      kps.p = reinterpret_cast<const orbx_keypoint*>(f.mvKeysUn.data()); desc.p = f.mDescriptors.ptr(0); uright.p = f.mvuRight.data();
      return;
    }
    kps.own.resize(cap); desc.own.assign((size_t)cap * 32, 0); uright.own.assign(cap, -1.f);
    for (int i = 0; i < n; ++i) {
      kps.own[i] = reinterpret_cast<const orbx_keypoint&>(f.mvKeysUn[i]);
      std::copy(f.mDescriptors.ptr(i), f.mDescriptors.ptr(i) + 32, &desc.own[(size_t)i * 32]);
      uright.own[i] = f.mvuRight[i];
    }
    kps.p = kps.own.data(); desc.p = desc.own.data(); uright.p = uright.own.data();
  }
};

// The projection searches walk mGrid.  The library rebuilds it from mvKeysUn with Frame::PosInGrid's rule and the static bounds
// every Frame shares (src/Frame.cc:230-245, 382-392) inside the kernel, so no grid is flattened here: NULL is passed for it.
struct Grid {
  float bounds[6];
  const int* cell_start = nullptr; const int* cell_items = nullptr;
  explicit Grid(const Flat&) {
    bounds[0] = Frame::mnMinX; bounds[1] = Frame::mnMaxX; bounds[2] = Frame::mnMinY; bounds[3] = Frame::mnMaxY;
    bounds[4] = Frame::mfGridElementWidthInv; bounds[5] = Frame::mfGridElementHeightInv;
  }
};

// one vocabulary node per feature out of a DBoW2::FeatureVector
std::vector<int> nodes_of(const DBoW2::FeatureVector& fv, int cap) {
  std::vector<int> node(cap, 0x7fffffff);
  for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it)
    for (size_t k = 0; k < it->second.size(); ++k) node[it->second[k]] = (int)it->first;
  return node;
}

// the fields of a list of MapPoints the projection kernels read
struct MPs {
  int m = 0, cap = 1;
  std::vector<unsigned char> valid, desc; std::vector<float> xw, normal, dmin, dmax, angle; std::vector<int> obs;
  explicit MPs(size_t count) {
    m = (int)count; cap = std::max(m, 1);
    valid.assign(cap, 0); desc.assign((size_t)cap * 32, 0); xw.assign((size_t)cap * 3, 0.f); normal.assign((size_t)cap * 3, 0.f);
    dmin.assign(cap, 0.f); dmax.assign(cap, 0.f); angle.assign(cap, 0.f); obs.assign(cap, 0);
  }
  void set(int k, MapPoint* mp) {
    valid[k] = 1;
    MapPointFields::PosNormal(mp, &xw[3 * k], &normal[3 * k]);
    MapPointFields::Descriptor(mp, &desc[(size_t)k * 32]);
    dmin[k] = MapPointFields::MinDistance(mp); dmax[k] = MapPointFields::MaxDistance(mp);
    obs[k] = mp->Observations();
  }
};

void pose12(const cv::Mat& R, const cv::Mat& t, float* out) {
  for (int y = 0; y < 3; ++y) for (int x = 0; x < 3; ++x) out[3 * y + x] = R.at<float>(y, x);
  for (int y = 0; y < 3; ++y) out[9 + y] = t.at<float>(y);
}

}  // namespace

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbx_hamming256(a.ptr(0), b.ptr(0)); }

// Tracking::SearchLocalPoints -> the local-map search (src/ORBmatcher.cc:45-129)
int ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th) {
  const Flat f(F);
  const int m = (int)vpMapPoints.size(), mcap = std::max(m, 1);
  std::vector<unsigned char> inview(mcap, 0), desc((size_t)mcap * 32, 0);
  std::vector<float> px(mcap, 0.f), py(mcap, 0.f), pxr(mcap, 0.f), vcos(mcap, 0.f);
  std::vector<int> level(mcap, 0), obs(mcap, 0);
  for (int k = 0; k < m; ++k) {
    MapPoint* mp = vpMapPoints[k];
    if (!mp->mbTrackInView || mp->isBad()) continue;
    inview[k] = 1; px[k] = mp->mTrackProjX; py[k] = mp->mTrackProjY; pxr[k] = mp->mTrackProjXR;
    level[k] = mp->mnTrackScaleLevel; vcos[k] = mp->mTrackViewCos; obs[k] = mp->Observations();
    MapPointFields::Descriptor(mp, &desc[(size_t)k * 32]);
  }
  std::vector<int> frame_obs(f.cap, -1), assigned(f.cap, -1);
  for (int i = 0; i < f.n; ++i) if (F.mvpMapPoints[i]) frame_obs[i] = F.mvpMapPoints[i]->Observations();
  int nm = 0;
  const FrameBounds gb;
  check(orbx_search_projection_batch(f.kps.data(), f.desc.data(), f.uright.data(), &f.n, 1, f.cap, gb.v, F.mvScaleFactors.data(), F.mnScaleLevels, inview.data(), px.data(), py.data(),
                                     pxr.data(), level.data(), vcos.data(), desc.data(), obs.data(), &m, mcap, th, mfNNratio, frame_obs.data(),
                                     assigned.data(), &nm, 0, nullptr));
  for (int i = 0; i < f.n; ++i) if (assigned[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[assigned[i]];
  return nm;
}

// Tracking::TrackWithMotionModel (src/ORBmatcher.cc:1328-1470)
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
  const Flat f(CurrentFrame);
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3), tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  // forward / backward motion decides the level window (:1339-1348)
  const cv::Mat twc = -Rcw.t() * tcw;
  const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3), tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat tlc = Rlw * twc + tlw;
  const bool bForward = tlc.at<float>(2) > CurrentFrame.mb && !bMono, bBackward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  const int mode = bForward ? 1 : (bBackward ? 2 : 0);
  const int nl = LastFrame.N, lcap = std::max(nl, 1);
  std::vector<unsigned char> valid(lcap, 0), desc((size_t)lcap * 32, 0);
  std::vector<float> xw((size_t)lcap * 3, 0.f), angle(lcap, 0.f);
  std::vector<int> octave(lcap, 0), obs(lcap, 0);
  for (int i = 0; i < nl; ++i) {
    MapPoint* mp = LastFrame.mvpMapPoints[i];
    if (!mp || LastFrame.mvbOutlier[i]) continue;
    valid[i] = 1;
    MapPointFields::PosNormal(mp, &xw[3 * i], nullptr);
    MapPointFields::Descriptor(mp, &desc[(size_t)i * 32]);
    octave[i] = LastFrame.mvKeys[i].octave; angle[i] = LastFrame.mvKeysUn[i].angle; obs[i] = mp->Observations();
  }
  float pose[12];
  pose12(Rcw, tcw, pose);
  std::vector<int> frame_obs(f.cap, -1), assigned(f.cap, -1);
  for (int i = 0; i < f.n; ++i) if (CurrentFrame.mvpMapPoints[i]) frame_obs[i] = CurrentFrame.mvpMapPoints[i]->Observations();
  int nm = 0;
  const FrameBounds gb;
  check(orbx_search_projection_frame_batch(f.kps.data(), f.desc.data(), f.uright.data(), &f.n, 1, f.cap, gb.v, CurrentFrame.mvScaleFactors.data(), CurrentFrame.mnScaleLevels, pose, &mode,
                                           CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx, CurrentFrame.cy, CurrentFrame.mbf, valid.data(), xw.data(),
                                           octave.data(), angle.data(), desc.data(), obs.data(), &nl, lcap, th, mbCheckOrientation ? 1 : 0,
                                           frame_obs.data(), assigned.data(), &nm, 0, nullptr));
  for (int i = 0; i < f.n; ++i) if (assigned[i] >= 0) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[assigned[i]];
  return nm;
}

// Tracking::Relocalization (src/ORBmatcher.cc:1472-1599)
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, const float th, const int ORBdist) {
  const Flat f(CurrentFrame);
  const Grid g(f);
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3), tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat Ow = -Rcw.t() * tcw;
  const vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
  MPs mps(vpMPs.size());
  for (size_t i = 0; i < vpMPs.size(); ++i) {
    MapPoint* mp = vpMPs[i];
    if (!mp || mp->isBad() || sAlreadyFound.count(mp)) continue;
    mps.set((int)i, mp);
    mps.angle[i] = pKF->mvKeysUn[i].angle;
  }
  std::vector<unsigned char> occupied(f.cap, 0);
  for (int i = 0; i < f.n; ++i) occupied[i] = CurrentFrame.mvpMapPoints[i] ? 1 : 0;
  float pose[12], ow[3] = {Ow.at<float>(0), Ow.at<float>(1), Ow.at<float>(2)};
  pose12(Rcw, tcw, pose);
  std::vector<int> assigned(f.cap, -1);
  int nm = 0;
  check(orbx_search_projection_ordered_batch(f.kps.data(), f.desc.data(), &f.n, 1, f.cap, g.bounds, g.cell_start, g.cell_items, g.bounds, pose, ow,
                                             CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx, CurrentFrame.cy, CurrentFrame.mvScaleFactors.data(),
                                             CurrentFrame.mnScaleLevels, CurrentFrame.mfLogScaleFactor, occupied.data(), mps.valid.data(), mps.xw.data(),
                                             mps.normal.data(), mps.dmin.data(), mps.dmax.data(), mps.desc.data(), mps.angle.data(), &mps.m, mps.cap, th,
                                             ORBdist, 1, mbCheckOrientation ? 1 : 0, assigned.data(), &nm, 0, nullptr));
  for (int i = 0; i < f.n; ++i) if (assigned[i] >= 0) CurrentFrame.mvpMapPoints[i] = vpMPs[assigned[i]];
  return nm;
}

// LoopClosing::ComputeSim3 (src/ORBmatcher.cc:290-403)
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched, int th) {
  const Flat f(*pKF);
  const Grid g(f);
  cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
  const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
  cv::Mat Rcw = sRcw / scw, tcw = Scw.rowRange(0, 3).col(3) / scw, Ow = -Rcw.t() * tcw;
  set<MapPoint*> found(vpMatched.begin(), vpMatched.end());
  found.erase(static_cast<MapPoint*>(NULL));
  MPs mps(vpPoints.size());
  for (size_t k = 0; k < vpPoints.size(); ++k) if (!vpPoints[k]->isBad() && !found.count(vpPoints[k])) mps.set((int)k, vpPoints[k]);
  std::vector<unsigned char> occupied(f.cap, 0);
  for (int i = 0; i < f.n; ++i) occupied[i] = vpMatched[i] ? 1 : 0;
  const float kfb[4] = {(float)pKF->mnMinX, (float)pKF->mnMaxX, (float)pKF->mnMinY, (float)pKF->mnMaxY};
  float pose[12], ow[3] = {Ow.at<float>(0), Ow.at<float>(1), Ow.at<float>(2)};
  pose12(Rcw, tcw, pose);
  std::vector<int> assigned(f.cap, -1);
  int nm = 0;
  check(orbx_search_projection_ordered_batch(f.kps.data(), f.desc.data(), &f.n, 1, f.cap, g.bounds, g.cell_start, g.cell_items, kfb, pose, ow, pKF->fx,
                                             pKF->fy, pKF->cx, pKF->cy, pKF->mvScaleFactors.data(), pKF->mnScaleLevels, pKF->mfLogScaleFactor, occupied.data(),
                                             mps.valid.data(), mps.xw.data(), mps.normal.data(), mps.dmin.data(), mps.dmax.data(), mps.desc.data(),
                                             mps.angle.data(), &mps.m, mps.cap, (float)th, TH_LOW, 0, 0, assigned.data(), &nm, 0, nullptr));
  for (int i = 0; i < f.n; ++i) if (assigned[i] >= 0) vpMatched[i] = vpPoints[assigned[i]];
  return nm;
}

// both SearchByBoW forms share one call: the feature vectors become one node id per feature
static int bow_search(const Flat& a, const std::vector<unsigned char>& validA, const DBoW2::FeatureVector& fvA, const std::vector<float>& angA, const Flat& b,
                      const std::vector<unsigned char>& validB, const DBoW2::FeatureVector& fvB, const std::vector<float>& angB, float ratio, bool ori,
                      bool strict, std::vector<int>& m12) {
  const int cap = a.cap;                                                    // both sides were flattened with the common capacity
  const std::vector<int> nA = nodes_of(fvA, cap), nB = nodes_of(fvB, cap);
  m12.assign(cap, -1);
  int nm = 0;
  check(orbx_search_by_bow_batch(a.desc.data(), angA.data(), validA.data(), nA.data(), &a.n, b.desc.data(), angB.data(), validB.data(), nB.data(), &b.n, 1, cap, 50, ratio,
                                 ori ? 1 : 0, strict ? 1 : 0, m12.data(), &nm, 0, nullptr));
  return nm;
}

// Tracking::TrackReferenceKeyFrame / Relocalization (src/ORBmatcher.cc:159-288)
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches) {
  const vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
  vpMapPointMatches = vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
  const int cap = std::max(pKF->N, F.N);
  const Flat a(*pKF, cap), b(F, cap);
  std::vector<unsigned char> vA(a.cap, 0), vB(b.cap, 1);
  std::vector<float> angA(a.cap, 0.f), angB(b.cap, 0.f);
  for (int i = 0; i < a.n; ++i) { vA[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad(); angA[i] = pKF->mvKeysUn[i].angle; }
  for (int j = 0; j < b.n; ++j) angB[j] = F.mvKeys[j].angle;              // :242 reads mvKeys of the frame
  std::vector<int> m12;
  const int nm = bow_search(a, vA, pKF->mFeatVec, angA, b, vB, F.mFeatVec, angB, mfNNratio, mbCheckOrientation, false, m12);
  for (int i = 0; i < a.n; ++i) if (m12[i] >= 0) vpMapPointMatches[m12[i]] = vpMapPointsKF[i];
  return nm;
}

// LoopClosing::ComputeSim3 (src/ORBmatcher.cc:522-655)
int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12) {
  const vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
  vpMatches12 = vector<MapPoint*>(vpMapPoints1.size(), static_cast<MapPoint*>(NULL));
  const int cap = std::max(pKF1->N, pKF2->N);
  const Flat a(*pKF1, cap), b(*pKF2, cap);
  std::vector<unsigned char> vA(a.cap, 0), vB(b.cap, 0);
  std::vector<float> angA(a.cap, 0.f), angB(b.cap, 0.f);
  for (int i = 0; i < a.n; ++i) { vA[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad(); angA[i] = pKF1->mvKeysUn[i].angle; }
  for (int j = 0; j < b.n; ++j) { vB[j] = vpMapPoints2[j] && !vpMapPoints2[j]->isBad(); angB[j] = pKF2->mvKeysUn[j].angle; }
  std::vector<int> m12;
  const int nm = bow_search(a, vA, pKF1->mFeatVec, angA, b, vB, pKF2->mFeatVec, angB, mfNNratio, mbCheckOrientation, true, m12);
  for (int i = 0; i < a.n; ++i) if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];
  return nm;
}

// Tracking::MonocularInitialization (src/ORBmatcher.cc:405-520)
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched, vector<int>& vnMatches12, int windowSize) {
  const int cap = std::max(std::max(F1.N, F2.N), 1);
  const Flat a(F1, cap), b(F2, cap);
  std::vector<float> prev((size_t)cap * 2, 0.f);
  for (int i = 0; i < F1.N; ++i) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
  vnMatches12.assign(cap, -1);
  int nm = 0;
  const FrameBounds gb;
  check(orbx_search_init_batch(a.kps.data(), a.desc.data(), &a.n, b.kps.data(), b.desc.data(), &b.n, 1, cap, gb.v, prev.data(), mfNNratio, windowSize, mbCheckOrientation ? 1 : 0, vnMatches12.data(), &nm, 0,
                               nullptr));
  vnMatches12.resize(F1.N);
  for (int i = 0; i < F1.N; ++i) { vbPrevMatched[i].x = prev[2 * i]; vbPrevMatched[i].y = prev[2 * i + 1]; }
  return nm;
}

// LocalMapping::CreateNewMapPoints (src/ORBmatcher.cc:657-823)
int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, vector<pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo) {
  const int cap = std::max(std::max(pKF1->N, pKF2->N), 1);
  const Flat a(*pKF1, cap), b(*pKF2, cap);
  std::vector<unsigned char> has1(cap, 0), has2(cap, 0);
  for (int i = 0; i < a.n; ++i) has1[i] = pKF1->GetMapPoint(i) ? 1 : 0;
  for (int j = 0; j < b.n; ++j) has2[j] = pKF2->GetMapPoint(j) ? 1 : 0;
  const std::vector<int> n1 = nodes_of(pKF1->mFeatVec, cap), n2 = nodes_of(pKF2->mFeatVec, cap);
  float f12[9], pose2[12], cw[3];
  for (int y = 0; y < 3; ++y) for (int x = 0; x < 3; ++x) f12[3 * y + x] = F12.at<float>(y, x);
  pose12(pKF2->GetRotation(), pKF2->GetTranslation(), pose2);
  const cv::Mat Cw = pKF1->GetCameraCenter();
  for (int y = 0; y < 3; ++y) cw[y] = Cw.at<float>(y);
  const float cam2[4] = {pKF2->fx, pKF2->fy, pKF2->cx, pKF2->cy};
  std::vector<int> m12(cap, -1);
  int nm = 0;
  check(orbx_search_triangulation_batch(a.kps.data(), a.desc.data(), has1.data(), a.uright.data(), n1.data(), &a.n, b.kps.data(), b.desc.data(), has2.data(),
                                        b.uright.data(), n2.data(), &b.n, 1, cap, f12, pose2, cw, cam2, pKF2->mvScaleFactors.data(), pKF2->mvLevelSigma2.data(),
                                        pKF2->mnScaleLevels, bOnlyStereo ? 1 : 0, mbCheckOrientation ? 1 : 0, m12.data(), &nm, 0, nullptr));
  vMatchedPairs.clear();
  vMatchedPairs.reserve(nm);
  for (int i = 0; i < a.n; ++i) if (m12[i] >= 0) vMatchedPairs.push_back(make_pair((size_t)i, (size_t)m12[i]));
  return nm;
}

// LoopClosing::ComputeSim3 (src/ORBmatcher.cc:1102-1326)
int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12, const cv::Mat& t12, const float th) {
  const int cap = std::max(std::max(pKF1->N, pKF2->N), 1);
  const Flat a(*pKF1, cap), b(*pKF2, cap);
  const Grid g1(a), g2(b);
  // the Sim3 in both directions, as :1119-1122
  cv::Mat sR12 = s12 * R12, sR21 = (1.0 / s12) * R12.t(), t21 = -sR21 * t12;
  float xf[24];
  pose12(sR12, t12, xf); pose12(sR21, t21, xf + 12);
  float pose1[12], pose2[12];
  pose12(pKF1->GetRotation(), pKF1->GetTranslation(), pose1); pose12(pKF2->GetRotation(), pKF2->GetTranslation(), pose2);
  const vector<MapPoint*> vp1 = pKF1->GetMapPointMatches(), vp2 = pKF2->GetMapPointMatches();
  const int N1 = (int)vp1.size(), N2 = (int)vp2.size();
  std::vector<bool> done1(N1, false), done2(N2, false);
  for (int i = 0; i < N1; ++i) {
    MapPoint* mp = vpMatches12[i];
    if (!mp) continue;
    done1[i] = true;
    const int idx2 = mp->GetIndexInKeyFrame(pKF2);
    if (idx2 >= 0 && idx2 < N2) done2[idx2] = true;
  }
  // the map points that take part, listed with their feature index
  std::vector<int> idx1, idx2;
  for (int i = 0; i < N1; ++i) if (vp1[i] && !done1[i] && !vp1[i]->isBad()) idx1.push_back(i);
  for (int j = 0; j < N2; ++j) if (vp2[j] && !done2[j] && !vp2[j]->isBad()) idx2.push_back(j);
  const int mcap = std::max((int)std::max(idx1.size(), idx2.size()), 1);
  MPs m1(mcap), m2(mcap);
  m1.m = (int)idx1.size(); m2.m = (int)idx2.size();
  std::vector<int> i1(mcap, -1), i2(mcap, -1);
  for (size_t k = 0; k < idx1.size(); ++k) { i1[k] = idx1[k]; m1.set((int)k, vp1[idx1[k]]); }
  for (size_t k = 0; k < idx2.size(); ++k) { i2[k] = idx2[k]; m2.set((int)k, vp2[idx2[k]]); }
  const float kb1[4] = {(float)pKF1->mnMinX, (float)pKF1->mnMaxX, (float)pKF1->mnMinY, (float)pKF1->mnMaxY};
  const float kb2[4] = {(float)pKF2->mnMinX, (float)pKF2->mnMaxX, (float)pKF2->mnMinY, (float)pKF2->mnMaxY};
  std::vector<int> m12(cap, -1);
  int nfound = 0;
  check(orbx_search_by_sim3_batch(a.kps.data(), a.desc.data(), &a.n, g1.cell_start, g1.cell_items, kb1, pose1, b.kps.data(), b.desc.data(), &b.n,
                                  g2.cell_start, g2.cell_items, kb2, pose2, 1, cap, g1.bounds, pKF1->fx, pKF1->fy, pKF1->cx, pKF1->cy,
                                  pKF2->mvScaleFactors.data(), pKF2->mnScaleLevels, pKF2->mfLogScaleFactor, xf, i1.data(), m1.xw.data(), m1.dmin.data(),
                                  m1.dmax.data(), m1.desc.data(), &m1.m, i2.data(), m2.xw.data(), m2.dmin.data(), m2.dmax.data(), m2.desc.data(), &m2.m, mcap, th,
                                  m12.data(), &nfound, 0, nullptr));
  for (int i = 0; i < N1; ++i) if (m12[i] >= 0) vpMatches12[i] = vp2[m12[i]];
  return nfound;
}

// LocalMapping::SearchInNeighbors (src/ORBmatcher.cc:825-975): the search on the device, the map update here in candidate order
int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th) {
  const Flat f(*pKF);
  const Grid g(f);
  MPs mps(vpMapPoints.size());
  for (size_t k = 0; k < vpMapPoints.size(); ++k) {
    MapPoint* mp = vpMapPoints[k];
    if (mp && !mp->isBad() && !mp->IsInKeyFrame(pKF)) mps.set((int)k, mp);
  }
  float pose[12];
  pose12(pKF->GetRotation(), pKF->GetTranslation(), pose);
  const cv::Mat Ow = pKF->GetCameraCenter();
  const float ow[3] = {Ow.at<float>(0), Ow.at<float>(1), Ow.at<float>(2)};
  const float kfb[4] = {(float)pKF->mnMinX, (float)pKF->mnMaxX, (float)pKF->mnMinY, (float)pKF->mnMaxY};
  std::vector<int> best(mps.cap, -1);
  check(orbx_fuse_search_batch(f.kps.data(), f.desc.data(), f.uright.data(), &f.n, 1, f.cap, g.bounds, g.cell_start, g.cell_items, kfb, pose, ow, pKF->fx,
                               pKF->fy, pKF->cx, pKF->cy, pKF->mbf, pKF->mvScaleFactors.data(), pKF->mvInvLevelSigma2.data(), pKF->mnScaleLevels,
                               pKF->mfLogScaleFactor, mps.valid.data(), mps.xw.data(), mps.normal.data(), mps.dmin.data(), mps.dmax.data(), mps.desc.data(), &mps.m,
                               mps.cap, th, 0, best.data(), 0, nullptr));
  int nFused = 0;
  for (int k = 0; k < mps.m; ++k) {
    MapPoint* pMP = vpMapPoints[k];
    if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF) || best[k] < 0) continue;      // the loop head of :846-851 again, then :944-966
    MapPoint* pMPinKF = pKF->GetMapPoint(best[k]);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) {
        if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
        else pMPinKF->Replace(pMP);
      }
    } else {
      pMP->AddObservation(pKF, best[k]);
      pKF->AddMapPoint(pMP, best[k]);
    }
    nFused++;
  }
  return nFused;
}

// LoopClosing::SearchAndFuse (src/ORBmatcher.cc:977-1100)
int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, float th, vector<MapPoint*>& vpReplacePoint) {
  const Flat f(*pKF);
  const Grid g(f);
  cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
  const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
  cv::Mat Rcw = sRcw / scw, tcw = Scw.rowRange(0, 3).col(3) / scw, Ow = -Rcw.t() * tcw;
  const set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();
  MPs mps(vpPoints.size());
  for (size_t k = 0; k < vpPoints.size(); ++k) if (!vpPoints[k]->isBad() && !spAlreadyFound.count(vpPoints[k])) mps.set((int)k, vpPoints[k]);
  float pose[12];
  pose12(Rcw, tcw, pose);
  const float ow[3] = {Ow.at<float>(0), Ow.at<float>(1), Ow.at<float>(2)};
  const float kfb[4] = {(float)pKF->mnMinX, (float)pKF->mnMaxX, (float)pKF->mnMinY, (float)pKF->mnMaxY};
  std::vector<int> best(mps.cap, -1);
  check(orbx_fuse_search_batch(f.kps.data(), f.desc.data(), f.uright.data(), &f.n, 1, f.cap, g.bounds, g.cell_start, g.cell_items, kfb, pose, ow, pKF->fx,
                               pKF->fy, pKF->cx, pKF->cy, pKF->mbf, pKF->mvScaleFactors.data(), pKF->mvInvLevelSigma2.data(), pKF->mnScaleLevels,
                               pKF->mfLogScaleFactor, mps.valid.data(), mps.xw.data(), mps.normal.data(), mps.dmin.data(), mps.dmax.data(), mps.desc.data(), &mps.m,
                               mps.cap, th, 1, best.data(), 0, nullptr));
  int nFused = 0;
  for (int k = 0; k < mps.m; ++k) {
    if (best[k] < 0) continue;
    MapPoint* pMP = vpPoints[k];
    MapPoint* pMPinKF = pKF->GetMapPoint(best[k]);                    // :1078-1093
    if (pMPinKF) {
      if (!pMPinKF->isBad()) vpReplacePoint[k] = pMPinKF;
    } else {
      pMP->AddObservation(pKF, best[k]);
      pKF->AddMapPoint(pMP, best[k]);
    }
    nFused++;
  }
  return nFused;
}

}  // namespace ORB_SLAM2
