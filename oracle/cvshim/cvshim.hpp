// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// A minimal stand-in for the slice of the OpenCV C++ API that the reference's
// src/ORBextractor.cc touches, so that file can be compiled *where it lies* under
// /root/reference (never copied) into oracle/_ref/ and run against the restatement.
// OpenCV's C++ headers/libraries are not on this box; the arithmetic behind resize / FAST /
// GaussianBlur / copyMakeBorder / fastAtan2 is supplied by cv_prims.hpp, which the tests pin
// bit-for-bit against the cv2 4.13.0 Python wheel.
#pragma once
#include <algorithm>
#include <cassert>
#include <climits>
#include <iomanip>
#include <cmath>
#include <cstddef>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <stdexcept>
#include <vector>

#include "../cv_prims.hpp"

typedef unsigned char uchar;
#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_64F 6

static inline int cvRound(double v) { return orbo::cv_round(v); }
static inline int cvRound(float v) { return orbo::cv_round(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return (int)std::floor(v); }
static inline int cvCeil(double v) { return (int)std::ceil(v); }

namespace cv {

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T a, T b) : x(a), y(b) {}
  template <typename U>
  Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
static inline Point2f& operator*=(Point2f& p, float s) { p.x = p.x * s; p.y = p.y * s; return p; }

struct Size {
  int width, height;
  Size() : width(0), height(0) {}
  Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
  int x, y, width, height;
  Rect(int a, int b, int w, int h) : x(a), y(b), width(w), height(h) {}
};

struct KeyPoint {
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
      : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};

// stand-in for the MatExpr that Mat::zeros / ones / eye return: a shape, a type and a fill rule
struct MatInit { int rows, cols, type, kind; double scale; };   // kind 0 zeros, 1 ones, 2 eye; scale multiplies ones
typedef MatInit MatZeros;

static inline int cvshim_esz(int type) { return type == CV_32F ? 4 : (type == 6 /*CV_64F*/ ? 8 : 1); }

class Mat {
 public:
  int rows = 0, cols = 0;
  size_t step = 0;          // bytes per row
  uchar* data = nullptr;
  int tp = CV_8U, cn = 1;   // depth and channels (channels only matter to reshape / undistortPoints)
  std::shared_ptr<std::vector<uchar>> store;

  Mat() {}
  Mat(Size s, int type) { alloc(s.height, s.width, type); }
  Mat(int r, int c, int type) { alloc(r, c, type); }
  Mat(int r, int c, int type, void* d, size_t st) : rows(r), cols(c), step(st), data((uchar*)d), tp(type) {}
  int esz() const { return cvshim_esz(tp) * cn; }
  void alloc(int r, int c, int type = CV_8U) {
    rows = r; cols = c; tp = type; cn = 1; step = (size_t)c * esz();
    store = std::make_shared<std::vector<uchar>>((size_t)r * step);
    data = store->data();
  }
  void create(int r, int c, int type) { if (r != rows || c != cols || type != tp || !data) alloc(r, c, type); }
  void release() { rows = cols = 0; step = 0; data = nullptr; store.reset(); }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return tp; }
  int channels() const { return cn; }
  size_t step1() const { return step / cvshim_esz(tp); }
  Size size() const { return Size(cols, rows); }
  bool isContinuous() const { return step == (size_t)cols * esz(); }
  Mat operator()(const Rect& r) const {
    if (r.x < 0 || r.y < 0 || r.x + r.width > cols || r.y + r.height > rows) throw std::out_of_range("Mat roi");
    Mat m(*this); m.rows = r.height; m.cols = r.width; m.data = data + (size_t)r.y * step + (size_t)r.x * esz(); return m;
  }
  Mat rowRange(int a, int b) const {
    if (a < 0 || a > b || b > rows) throw std::out_of_range("rowRange");
    Mat m(*this); m.rows = b - a; m.data = data + (size_t)a * step; return m;
  }
  Mat colRange(int a, int b) const {
    if (a < 0 || a > b || b > cols) throw std::out_of_range("colRange");
    Mat m(*this); m.cols = b - a; m.data = data + (size_t)a * esz(); return m;
  }
  Mat row(int y) const { return rowRange(y, y + 1); }
  Mat col(int x) const { return colRange(x, x + 1); }
  Mat clone() const {
    Mat m; if (!data) return m;
    m.alloc(rows, cols, tp); m.cn = cn; m.step = (size_t)cols * esz(); m.store = std::make_shared<std::vector<uchar>>((size_t)rows * m.step); m.data = m.store->data();
    for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * esz());
    return m;
  }
  // copyTo into a destination of the right shape writes in place (views included), otherwise allocates
  void copyTo(Mat& dst) const {
    if (dst.rows != rows || dst.cols != cols || dst.tp != tp || dst.cn != cn || !dst.data) dst = clone();
    else for (int y = 0; y < rows; ++y) std::memmove(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols * esz());
  }
  void copyTo(const Mat& view) const {   // temporaries such as T.rowRange(0,3).colRange(0,3)
    if (view.rows != rows || view.cols != cols || view.tp != tp || !view.data) throw std::logic_error("copyTo: view of another shape");
    for (int y = 0; y < rows; ++y) std::memmove(view.data + (size_t)y * view.step, data + (size_t)y * step, (size_t)cols * esz());
  }
  template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
  template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }
  // one index: element i of a row or column vector (cv::Mat::at(int i0))
  template <typename T> T& at(int i) { return rows == 1 ? *(T*)(data + i * sizeof(T)) : *(T*)(data + (size_t)i * step); }
  template <typename T> const T& at(int i) const { return rows == 1 ? *(const T*)(data + i * sizeof(T)) : *(const T*)(data + (size_t)i * step); }
  size_t total() const { return (size_t)rows * cols; }
  uchar* ptr(int y = 0) { return data + (size_t)y * step; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
  static MatInit zeros(int r, int c, int type) { return MatInit{r, c, type, 0, 1.0}; }
  static MatInit ones(int r, int c, int type) { return MatInit{r, c, type, 1, 1.0}; }
  static MatInit eye(int r, int c, int type) { return MatInit{r, c, type, 2, 1.0}; }
  void fill(const MatInit& z) {
    for (int y = 0; y < rows; ++y) {
      std::memset(data + (size_t)y * step, 0, (size_t)cols * esz());
      if (z.kind == 0) continue;
      for (int x = 0; x < cols; ++x) {
        if (z.kind == 2 && x != y) continue;
        if (tp == CV_32F) at<float>(y, x) = (float)z.scale; else at<uchar>(y, x) = (uchar)z.scale;
      }
    }
  }
  Mat(const MatInit& z) { alloc(z.rows, z.cols, z.type); fill(z); }
  // like cv::Mat = MatExpr: a destination of the right shape is filled in place (the reference relies on this
  // to zero a rowRange view of the output descriptors, src/ORBextractor.cc:1037)
  Mat& operator=(const MatInit& z) {
    if (z.rows != rows || z.cols != cols || z.type != tp || !data) alloc(z.rows, z.cols, z.type);
    fill(z);
    return *this;
  }
  // N x C one-channel  <->  N x (C / k) k-channel, as Frame::UndistortKeyPoints uses it (src/Frame.cc:421-423)
  Mat reshape(int newcn) const {
    Mat m(*this);
    const int scalars = cols * cn;
    if (scalars % newcn) throw std::logic_error("reshape");
    m.cols = scalars / newcn; m.cn = newcn;
    return m;
  }
  Mat t() const {
    if (tp != CV_32F) throw std::logic_error("t(): CV_32F only");
    Mat m(cols, rows, CV_32F);
    for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) m.at<float>(x, y) = at<float>(y, x);
    return m;
  }
  // cv::Mat::dot: products accumulated in double (dotProd_32f), returned as double
  double dot(const Mat& o) const {
    if (tp != CV_32F || o.tp != CV_32F || total() != o.total()) throw std::logic_error("dot");
    double s = 0;
    if (rows == o.rows) { for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) s += (double)at<float>(y, x) * (double)o.at<float>(y, x); }
    else { for (size_t i = 0; i < total(); ++i) s += (double)(rows == 1 ? at<float>(0, (int)i) : at<float>((int)i, 0)) * (double)(o.rows == 1 ? o.at<float>(0, (int)i) : o.at<float>((int)i, 0)); }
    return s;
  }
  void convertTo(Mat& dst, int type) const {
    Mat out(rows, cols, type);
    for (int y = 0; y < rows; ++y)
      for (int x = 0; x < cols; ++x) {
        const double v = tp == CV_32F ? (double)at<float>(y, x) : (double)at<uchar>(y, x);
        if (type == CV_32F) out.at<float>(y, x) = (float)v; else out.at<uchar>(y, x) = (uchar)cvRound(v);
      }
    dst = out;
  }
};

// ---- the float matrix algebra the reference writes with MatExpr.  Small products follow OpenCV's gemm for
// CV_32F (float products, float sums left to right; pinned against cv2.gemm in tests/test_oracle_cv2.py). ----
static inline void cvshim_need32(const Mat& a) { if (a.type() != CV_32F) throw std::logic_error("CV_32F only"); }
static inline Mat operator*(const Mat& a, const Mat& b) {
  cvshim_need32(a); cvshim_need32(b);
  if (a.cols != b.rows) throw std::logic_error("gemm shape");
  Mat m(a.rows, b.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < b.cols; ++x) {
      float s = a.at<float>(y, 0) * b.at<float>(0, x);
      for (int k = 1; k < a.cols; ++k) s = s + a.at<float>(y, k) * b.at<float>(k, x);
      m.at<float>(y, x) = s;
    }
  return m;
}
template <typename F> static inline Mat cvshim_zip(const Mat& a, const Mat& b, F f) {
  cvshim_need32(a); cvshim_need32(b);
  if (a.rows != b.rows || a.cols != b.cols) throw std::logic_error("shape");
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = f(a.at<float>(y, x), b.at<float>(y, x));
  return m;
}
static inline Mat operator+(const Mat& a, const Mat& b) { return cvshim_zip(a, b, [](float p, float q) { return p + q; }); }
static inline Mat operator-(const Mat& a, const Mat& b) { return cvshim_zip(a, b, [](float p, float q) { return p - q; }); }
static inline Mat operator-(const Mat& a, const MatInit& z) { return a - Mat(z); }
static inline Mat operator*(const Mat& a, double s) {
  cvshim_need32(a);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = (float)((double)a.at<float>(y, x) * s);
  return m;
}
static inline Mat operator*(double s, const Mat& a) { return a * s; }
static inline Mat operator/(const Mat& a, double s) {
  cvshim_need32(a);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = (float)((double)a.at<float>(y, x) / s);
  return m;
}
static inline Mat operator-(const Mat& a) { return a * -1.0; }
static inline MatInit operator*(double s, const MatInit& z) { MatInit r = z; r.scale *= s; return r; }
static inline MatInit operator*(const MatInit& z, double s) { return s * z; }

enum { NORM_L1 = 2, NORM_L2 = 4 };
// cv::norm: CV_32F sums are taken in double
static inline double norm(const Mat& a, int type = NORM_L2) {
  cvshim_need32(a);
  double s = 0;
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) { const double v = a.at<float>(y, x); s += type == NORM_L1 ? std::fabs(v) : v * v; }
  return type == NORM_L1 ? s : std::sqrt(s);
}
static inline double norm(const Mat& a, const Mat& b, int type = NORM_L2) {
  cvshim_need32(a); cvshim_need32(b);
  double s = 0;
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) { const double v = (double)a.at<float>(y, x) - (double)b.at<float>(y, x); s += type == NORM_L1 ? std::fabs(v) : v * v; }
  return type == NORM_L1 ? s : std::sqrt(s);
}

// cv::Mat_<float>(r, c) << a, b, c ...
template <typename T> class Mat_ : public Mat {
 public:
  int fillpos = 0;
  Mat_(int r, int c) : Mat(r, c, CV_32F) { static_assert(sizeof(T) == 4, "float only"); }
  Mat_& operator<<(T v) { put(v); return *this; }
  Mat_& operator,(T v) { put(v); return *this; }
  void put(T v) { at<T>(fillpos / cols, fillpos % cols) = v; ++fillpos; }
};

class _InputArray {
 public:
  const Mat* m;
  _InputArray(const Mat& mm) : m(&mm) {}
  bool empty() const { return m->empty(); }
  Mat getMat() const { return *m; }
};
class _OutputArray {
 public:
  Mat* m;
  _OutputArray(Mat& mm) : m(&mm) {}
  void create(int r, int c, int t) const { m->create(r, c, t); }
  void release() const { m->release(); }
  Mat getMat() const { return *m; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

inline void resize(const Mat& src, Mat& dst, Size sz, double, double, int) {
  dst.create(sz.height, sz.width, CV_8UC1);
  orbo::resize_linear_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}
// src may alias the interior of dst (the reference borders each level in place); border pixels
// only ever read interior pixels, which are left untouched.
inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int) {
  dst.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
  assert(top == left && bottom == right && top == bottom);
  const uchar* interior = dst.data + (size_t)top * dst.step + left;
  if (interior != src.data)
    for (int y = 0; y < src.rows; ++y) std::memcpy(dst.data + (size_t)(y + top) * dst.step + left, src.data + (size_t)y * src.step, src.cols);
  for (int y = -top; y < src.rows + bottom; ++y) {
    int sy = orbo::reflect101(y, src.rows);
    for (int x = -left; x < src.cols + right; ++x) {
      if (y >= 0 && y < src.rows && x >= 0 && x < src.cols) { x = src.cols - 1; continue; }
      dst.data[(size_t)(y + top) * dst.step + x + left] = src.data[(size_t)sy * src.step + orbo::reflect101(x, src.cols)];
    }
  }
}
inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int threshold, bool nms) {
  assert(nms);
  (void)nms;
  std::vector<orbo::FastPt> v;
  orbo::fast9_16(img.data, img.cols, img.rows, (int)img.step, threshold, v);
  kps.clear();
  for (const orbo::FastPt& f : v) kps.push_back(KeyPoint((float)f.x, (float)f.y, 7.f, -1, (float)f.score));
}
inline void GaussianBlur(const Mat& src, Mat& dst, Size k, double sx, double sy, int) {
  assert(k.width == 7 && k.height == 7 && sx == 2 && sy == 2);
  (void)k; (void)sx; (void)sy;
  Mat tmp = src.clone();
  dst.create(src.rows, src.cols, CV_8UC1);
  orbo::gauss7_sigma2(tmp.data, tmp.cols, tmp.rows, (int)tmp.step, dst.data, (int)dst.step);
}
static inline float fastAtan2(float y, float x) { return orbo::fast_atan2_deg(y, x); }
// cv::undistortPoints(src, dst, K, distCoeffs, R = empty, P) for N x 1 two-channel CV_32F points, as
// src/Frame.cc:422,448 call it (P = K): the restatement in cv_prims.hpp, pinned against cv2.undistortPoints.
inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& D, const Mat& R, const Mat& P) {
  if (!R.empty() || src.type() != CV_32F || src.channels() != 2 || src.cols != 1) throw std::logic_error("undistortPoints: shim supports Nx1 CV_32FC2, R empty");
  if (P.at<float>(0, 0) != K.at<float>(0, 0) || P.at<float>(1, 1) != K.at<float>(1, 1) || P.at<float>(0, 2) != K.at<float>(0, 2) || P.at<float>(1, 2) != K.at<float>(1, 2))
    throw std::logic_error("undistortPoints: shim supports P == K");
  orbo::Camera c;
  c.fx = K.at<float>(0, 0); c.fy = K.at<float>(1, 1); c.cx = K.at<float>(0, 2); c.cy = K.at<float>(1, 2);
  for (int i = 0; i < 5; ++i) c.d[i] = i < (int)D.total() ? D.at<float>(i) : 0.f;
  Mat out = src.clone();
  for (int i = 0; i < src.rows; ++i) {
    const float* s = src.ptr<float>(i);
    float* o = out.ptr<float>(i);
    orbo::undistort_point(c, s[0], s[1], &o[0], &o[1]);
  }
  dst = out;
}

struct KeyPointsFilter {  // only named by the reference's disabled ComputeKeyPointsOld path
  static void retainBest(std::vector<KeyPoint>&, int) { throw std::logic_error("dead path"); }
};

// cv::FileStorage / cv::FileNode are only named by DBoW2's YAML save()/load() (Thirdparty/DBoW2/DBoW2/
// TemplatedVocabulary.h:1456-1640), virtual members that must compile but are never called by oracle/_ref: the
// text-file loader the reference uses (:1337-1420) does not touch them.  Every use throws.
struct FileNode {
  FileNode operator[](const char*) const { throw std::logic_error("cv::FileStorage is not part of the shim"); }
  FileNode operator[](const std::string&) const { throw std::logic_error("cv::FileStorage is not part of the shim"); }
  FileNode operator[](int) const { throw std::logic_error("cv::FileStorage is not part of the shim"); }
  size_t size() const { return 0; }
  operator int() const { throw std::logic_error("cv::FileStorage is not part of the shim"); }
  operator double() const { throw std::logic_error("cv::FileStorage is not part of the shim"); }
  operator std::string() const { throw std::logic_error("cv::FileStorage is not part of the shim"); }
};
template <typename T> static inline void operator>>(const FileNode&, T&) { throw std::logic_error("cv::FileStorage is not part of the shim"); }
struct FileStorage {
  enum { READ = 0, WRITE = 1 };
  FileStorage() {}
  FileStorage(const char*, int) {}
  FileStorage(const std::string&, int) {}
  bool isOpened() const { return false; }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
};
template <typename T> static inline FileStorage& operator<<(FileStorage& f, const T&) { throw std::logic_error("cv::FileStorage is not part of the shim"); return f; }

}  // namespace cv
