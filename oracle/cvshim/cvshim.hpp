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

struct MatZeros { int rows, cols; };  // stand-in for the MatExpr that Mat::zeros returns

class Mat {
 public:
  int rows = 0, cols = 0;
  size_t step = 0;
  uchar* data = nullptr;
  std::shared_ptr<std::vector<uchar>> store;

  Mat() {}
  Mat(Size s, int) { alloc(s.height, s.width); }
  Mat(int r, int c, int) { alloc(r, c); }
  Mat(int r, int c, int, void* d, size_t st) : rows(r), cols(c), step(st), data((uchar*)d) {}
  void alloc(int r, int c) {
    rows = r; cols = c; step = (size_t)c;
    store = std::make_shared<std::vector<uchar>>((size_t)r * c);
    data = store->data();
  }
  void create(int r, int c, int) { if (r != rows || c != cols || !data) alloc(r, c); }
  void release() { rows = cols = 0; step = 0; data = nullptr; store.reset(); }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return CV_8UC1; }
  size_t step1() const { return step; }
  Size size() const { return Size(cols, rows); }
  Mat operator()(const Rect& r) const {
    if (r.x < 0 || r.y < 0 || r.x + r.width > cols || r.y + r.height > rows) throw std::out_of_range("Mat roi");
    Mat m(*this); m.rows = r.height; m.cols = r.width; m.data = data + (size_t)r.y * step + r.x; return m;
  }
  Mat rowRange(int a, int b) const {
    if (a < 0 || a > b || b > rows) throw std::out_of_range("rowRange");
    Mat m(*this); m.rows = b - a; m.data = data + (size_t)a * step; return m;
  }
  Mat colRange(int a, int b) const {
    if (a < 0 || a > b || b > cols) throw std::out_of_range("colRange");
    Mat m(*this); m.cols = b - a; m.data = data + a; return m;
  }
  Mat clone() const {
    Mat m; m.alloc(rows, cols);
    for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, cols);
    return m;
  }
  template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
  template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }
  template <typename T> T& at(int i) { return ((T*)data)[i]; }                     // continuous row / column vectors
  template <typename T> const T& at(int i) const { return ((const T*)data)[i]; }
  size_t total() const { return (size_t)rows * cols; }
  uchar* ptr(int y = 0) { return data + (size_t)y * step; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
  static MatZeros zeros(int r, int c, int) { return MatZeros{r, c}; }
  Mat(const MatZeros& z) { alloc(z.rows, z.cols); }
  // like cv::Mat = MatExpr: a destination of the right size is filled in place (the reference
  // relies on this to zero a rowRange view of the output descriptors, src/ORBextractor.cc:1037)
  Mat& operator=(const MatZeros& z) {
    if (z.rows != rows || z.cols != cols || !data) alloc(z.rows, z.cols);
    else for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step, 0, cols);
    return *this;
  }
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
