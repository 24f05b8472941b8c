// orbx_cxx.hpp — source-compatible C++ stand-ins for the two reference classes on the hot path,
// implemented on the C ABI in orbx.h.  Header only; compile it against the OpenCV headers the
// rest of ORB-SLAM2 already uses (it needs cv::Mat, cv::KeyPoint, cv::InputArray/OutputArray).
//
//   ORB_SLAM2::ORBextractor   replaces include/ORBextractor.h + src/ORBextractor.cc
//   ORB_SLAM2::ORBmatcher     the hot-path subset of include/ORBmatcher.h (DescriptorDistance,
//                             SearchForInitialization on flat arrays); the other searches keep
//                             using the reference implementation and call DescriptorDistance
//   ORB_SLAM2::ComputeStereoMatchesGPU   body for Frame::ComputeStereoMatches (src/Frame.cc:466)
//   ORB_SLAM2::FrameGridGPU              bodies for UndistortKeyPoints / ComputeImageBounds / AssignFeaturesToGrid /
//                                        GetFeaturesInArea (src/Frame.cc:230-245, 327-464)
//
// Same constructor arguments, same operator() signature, same getters, same public
// mvImagePyramid member (filled on demand by DownloadPyramid(), because the stereo matcher reads
// the pyramid on the device).  Errors that the reference turns into assert()/cv::Exception
// surface as std::runtime_error carrying orbx_last_error_string().
#ifndef ORBX_CXX_HPP_
#define ORBX_CXX_HPP_

#include <opencv2/core/core.hpp>

#include <algorithm>
#include <stdexcept>
#include <string>
#include <vector>

#include "orbx.h"

namespace ORB_SLAM2 {

inline void orbx_check(orbx_status s) {
  if (s != ORBX_OK) throw std::runtime_error(std::string("orbx: ") + orbx_last_error_string());
}

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

  ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0)
      : device_(device), handle_(nullptr), stream_(nullptr), max_w_(0), max_h_(0) {
    params_.nfeatures = nfeatures; params_.scale_factor = scaleFactor; params_.nlevels = nlevels;
    params_.ini_th_fast = iniThFAST; params_.min_th_fast = minThFAST;
    mvImagePyramid.resize(nlevels);
  }
  ~ORBextractor() { orbx_extractor_destroy(handle_); orbx_stream_destroy(stream_); }
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // Mask is ignored, as in the reference (include/ORBextractor.h:58).
  void operator()(cv::InputArray _image, cv::InputArray /*mask*/, std::vector<cv::KeyPoint>& keypoints,
                  cv::OutputArray _descriptors) {
    if (_image.empty()) return;                                   // src/ORBextractor.cc:1046
    cv::Mat image = _image.getMat();
    if (image.type() != CV_8UC1) throw std::runtime_error("orbx: image must be CV_8UC1");   // assert at :1050
    Ensure(image.cols, image.rows);
    static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");
    const int cap = orbx_max_keypoints(handle_);
    keypoints.resize(cap);
    std::vector<unsigned char> desc((size_t)cap * 32);
    int n = 0;
    orbx_check(orbx_extract(handle_, image.data, image.cols, image.rows, (int)image.step,
                            reinterpret_cast<orbx_keypoint*>(keypoints.data()), desc.data(), cap, &n, stream_));
    keypoints.resize(n);
    if (n == 0) { _descriptors.release(); return; }
    _descriptors.create(n, 32, CV_8U);
    cv::Mat out = _descriptors.getMat();
    for (int i = 0; i < n; ++i) std::copy(desc.begin() + (size_t)i * 32, desc.begin() + (size_t)(i + 1) * 32, out.ptr(i));
#ifdef ORBX_DROPIN_HOST_PYRAMID
    DownloadPyramid();   // for callers that still read mvImagePyramid on the host (unmodified Frame::ComputeStereoMatches)
#endif
  }

  int GetLevels() { return params_.nlevels; }
  float GetScaleFactor() { return params_.scale_factor; }
  std::vector<float> GetScaleFactors() { return Table(0); }
  std::vector<float> GetInverseScaleFactors() { return Table(1); }
  std::vector<float> GetScaleSigmaSquares() { return Table(2); }
  std::vector<float> GetInverseScaleSigmaSquares() { return Table(3); }

  // The reference's public pyramid (include/ORBextractor.h:85).  Host copies are made only when
  // somebody asks: level l is the inner ROI of a (w+38) x (h+38) bordered plane, as upstream.
  std::vector<cv::Mat> mvImagePyramid;
  void DownloadPyramid() {
    for (int l = 0; l < params_.nlevels; ++l) {
      int w = 0, h = 0;
      orbx_check(orbx_pyramid_level(handle_, 0, l, nullptr, &w, &h, nullptr));
      cv::Mat whole(h + 38, w + 38, CV_8UC1);
      orbx_check(orbx_download_level(handle_, 0, l, whole.data));
      mvImagePyramid[l] = whole(cv::Rect(19, 19, w, h));
    }
  }

  orbx_extractor* handle() { return handle_; }
  void* stream() { return stream_; }
  // Sizes the device arenas; called by operator() with every image's size.  The reference accepts any image size on
  // any call (it reallocates mvImagePyramid, src/ORBextractor.cc:1115), so a later image that is larger than the
  // arenas were sized for gets a new handle.  Each stand-in owns a non-blocking stream: the reference's two extractor
  // threads (src/Frame.cc:78-81) then overlap on the GPU instead of queueing on the legacy default stream.
  void Ensure(int w, int h, int max_batch = 1) {
    if (!stream_) orbx_check(orbx_stream_create(device_, &stream_));
    if (handle_ && w <= max_w_ && h <= max_h_) return;
    if (handle_) { orbx_extractor_destroy(handle_); handle_ = nullptr; }
    max_w_ = std::max(max_w_, w); max_h_ = std::max(max_h_, h);
    orbx_check(orbx_extractor_create(&params_, max_w_, max_h_, max_batch, device_, &handle_));
  }

 private:
  std::vector<float> Table(int which) {
    float t[4][8]; int n = 0;
    orbx_check(orbx_compute_tables(&params_, &n, t[0], t[1], t[2], t[3], nullptr));   // no handle needed: Frame reads these first
    return std::vector<float>(t[which], t[which] + n);
  }
  orbx_params params_;
  int device_;
  orbx_extractor* handle_;
  void* stream_;
  int max_w_, max_h_;
};

// Body of Frame::ComputeStereoMatches (src/Frame.cc:466-640): call right after both ExtractORB
// threads joined (src/Frame.cc:78-81; operator() returns with its stream drained, so both slots are complete).
// N = mvKeys.size(); fx = mK.at<float>(0,0) - the static Frame::fx is still unset when the FIRST stereo frame gets
// here (src/Frame.cc:90 runs before :105).
inline void ComputeStereoMatchesGPU(ORBextractor& left, ORBextractor& right, int N, float mbf, float fx,
                                    std::vector<float>& mvuRight, std::vector<float>& mvDepth) {
  const int cap = orbx_max_keypoints(left.handle());
  std::vector<float> ur(cap, -1.f), dp(cap, -1.f);
  int nmatches = 0;
  orbx_check(orbx_stereo_match_batch(left.handle(), 0, right.handle(), 0, 1, mbf, fx, ur.data(), dp.data(), &nmatches, cap, 0, left.stream()));
  orbx_check(orbx_stream_synchronize(left.stream()));
  mvuRight.assign(ur.begin(), ur.begin() + N);
  mvDepth.assign(dp.begin(), dp.begin() + N);
}

// Bodies for Frame::UndistortKeyPoints + ComputeImageBounds + AssignFeaturesToGrid (src/Frame.cc:404-464, 230-245)
// and Frame::GetFeaturesInArea (src/Frame.cc:327-380).  mK is the 3x3 CV_32F camera matrix, mDistCoef the 4x1 or
// 5x1 CV_32F coefficient vector of the Frame constructors.
struct FrameGridGPU {
  std::vector<cv::KeyPoint> mvKeysUn;
  float mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
  std::vector<int> cell_start, cell_items;   // mGrid[ix][iy] = cell_items[cell_start[ix*48+iy] .. cell_start[ix*48+iy+1])

  void Build(const std::vector<cv::KeyPoint>& mvKeys, const cv::Mat& mK, const cv::Mat& mDistCoef, int cols, int rows) {
    orbx_camera cam = {mK.at<float>(0, 0), mK.at<float>(1, 1), mK.at<float>(0, 2), mK.at<float>(1, 2), 0, 0, 0, 0, 0};
    float* d = &cam.k1;
    for (int i = 0; i < (int)mDistCoef.total() && i < 5; ++i) d[i] = mDistCoef.at<float>(i);
    const int n = (int)mvKeys.size(), cap = std::max(n, 1);
    mvKeysUn.resize(cap);
    cell_start.assign(ORBX_GRID_COLS * ORBX_GRID_ROWS + 1, 0); cell_items.assign(cap, 0);
    float b[6];
    std::vector<cv::KeyPoint> in(mvKeys); in.resize(cap);
    orbx_check(orbx_undistort_grid_batch(reinterpret_cast<const orbx_keypoint*>(in.data()), &n, 1, cap, &cam, cols, rows,
                                         reinterpret_cast<orbx_keypoint*>(mvKeysUn.data()), b, cell_start.data(), cell_items.data(), 0, nullptr));
    mvKeysUn.resize(n);
    mnMinX = b[0]; mnMaxX = b[1]; mnMinY = b[2]; mnMaxY = b[3]; mfGridElementWidthInv = b[4]; mfGridElementHeightInv = b[5];
  }
  std::vector<size_t> GetFeaturesInArea(float x, float y, float r, int minLevel = -1, int maxLevel = -1) const {
    const int n = (int)mvKeysUn.size(), cap = std::max(n, 1), q_frame = 0;
    const float b[6] = {mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv};
    std::vector<cv::KeyPoint> in(mvKeysUn); in.resize(cap);
    std::vector<int> idx(cap);
    int cnt = 0;
    orbx_check(orbx_features_in_area_batch(reinterpret_cast<const orbx_keypoint*>(in.data()), &n, 1, cap, b, cell_start.data(), cell_items.data(),
                                           &q_frame, &x, &y, &r, &minLevel, &maxLevel, 1, cap, idx.data(), &cnt, 0, nullptr));
    return std::vector<size_t>(idx.begin(), idx.begin() + cnt);
  }
};

#ifndef ORBX_CXX_EXTRACTOR_ONLY   // dropin/ORBextractor.h: inside the reference tree include/ORBmatcher.h declares the class
class ORBmatcher {
 public:
  static const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;   // src/ORBmatcher.cc:37-39
  ORBmatcher(float nnratio = 0.6f, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbx_hamming256(a.ptr(), b.ptr()); }

  // SearchForInitialization on the arrays Frame holds (mvKeysUn, mDescriptors of F1 and F2) and Frame's static
  // bounds {mnMinX, mnMaxX, mnMinY, mnMaxY} ({0, cols, 0, rows} without distortion); vbPrevMatched and vnMatches12
  // as in include/ORBmatcher.h:69.
  int SearchForInitialization(const std::vector<cv::KeyPoint>& keys1, const cv::Mat& desc1,
                              const std::vector<cv::KeyPoint>& keys2, const cv::Mat& desc2, const float bounds[4],
                              std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize = 10) {
    const float gb[6] = {bounds[0], bounds[1], bounds[2], bounds[3], 64.0f / (bounds[1] - bounds[0]), 48.0f / (bounds[3] - bounds[2])};   // src/Frame.cc:212-213
    const int n1 = (int)keys1.size(), n2 = (int)keys2.size(), cap = std::max(std::max(n1, n2), 1);
    std::vector<orbx_keypoint> k1(cap), k2(cap);
    std::vector<unsigned char> d1((size_t)cap * 32), d2((size_t)cap * 32);
    std::vector<float> prev((size_t)cap * 2);
    for (int i = 0; i < n1; ++i) { k1[i] = reinterpret_cast<const orbx_keypoint&>(keys1[i]); std::copy(desc1.ptr(i), desc1.ptr(i) + 32, &d1[(size_t)i * 32]);
                                   prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
    for (int i = 0; i < n2; ++i) { k2[i] = reinterpret_cast<const orbx_keypoint&>(keys2[i]); std::copy(desc2.ptr(i), desc2.ptr(i) + 32, &d2[(size_t)i * 32]); }
    vnMatches12.assign(cap, -1);
    int nm = 0;
    orbx_check(orbx_search_init_batch(k1.data(), d1.data(), &n1, k2.data(), d2.data(), &n2, 1, cap, gb, prev.data(), mfNNratio,
                                      windowSize, mbCheckOrientation ? 1 : 0, vnMatches12.data(), &nm, 0, nullptr));
    vnMatches12.resize(n1);
    for (int i = 0; i < n1; ++i) { vbPrevMatched[i].x = prev[2 * i]; vbPrevMatched[i].y = prev[2 * i + 1]; }
    return nm;
  }

 protected:
  float mfNNratio;
  bool mbCheckOrientation;
};

#endif  // ORBX_CXX_EXTRACTOR_ONLY

}  // namespace ORB_SLAM2
#endif  // ORBX_CXX_HPP_
