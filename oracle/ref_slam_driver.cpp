// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// Driver for oracle/_ref/orbref_slam: the reference's own src/Frame.cc, src/ORBmatcher.cc, src/ORBextractor.cc,
// src/KeyFrame.cc, src/MapPoint.cc, src/Map.cc, src/KeyFrameDatabase.cc and Thirdparty/DBoW2, compiled where they lie
// under /root/reference against oracle/cvshim (cvshim/refstubs/Converter.h shadows the one header that needs Eigen
// and g2o).  It builds real ORB_SLAM2::Frame / KeyFrame / MapPoint objects from images handed over in a file and runs
// the reference's own member functions, so that the restatements in orb_oracle.hpp (and through them the CUDA
// kernels) are checked against the reference itself and not only against a re-reading of it.
//
//   orbref_slam <command> <in.bin> <out.bin>
//
// Files are flat little-endian records written / read by oracle/pyoracle.py (ref_slam_*): int32 and float32 scalars,
// then arrays.  Hazards:
//   H1  quad-tree ties: operator new is a bump allocator (addresses grow with creation order), as in ref_driver.cpp;
//       it is thread safe because the stereo constructor runs both extractors in threads.
//   H4  Frame::mb is read by ComputeStereoMatches before the constructor sets it (src/Frame.cc:90 vs :114).  The
//       Frame is constructed by placement new into storage pre-filled with the float bf/fx, so the uninitialised
//       member reads the value DESIGN.md defines for it.
#include <sys/mman.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <list>
#include <map>
#include <mutex>
#include <set>
#include <sstream>
#include <thread>
#include <opencv2/core/core.hpp>
// the driver reads a few protected members of the reference's classes (MapPoint::mfMinDistance / mfMaxDistance);
// the class layouts are unchanged, only the access checks are lifted for this translation unit
#define protected public
#define private public
#include "Frame.h"
#include "KeyFrame.h"
#include "KeyFrameDatabase.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBVocabulary.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"
#undef protected
#undef private

static char* g_base = nullptr;
static std::atomic<size_t> g_off(0);
static size_t g_cap = 0;

static void* bump_alloc(size_t n) {
  n = (n + 15) & ~size_t(15);
  const size_t o = g_off.fetch_add(n);
  if (o + n > g_cap) { std::fprintf(stderr, "orbref_slam: arena exhausted\n"); std::abort(); }
  return g_base + o;
}
static inline bool in_arena(void* p) { return g_base && (char*)p >= g_base && (char*)p < g_base + g_cap; }
void* operator new(size_t n) { return g_base ? bump_alloc(n) : std::malloc(n ? n : 1); }
void* operator new[](size_t n) { return g_base ? bump_alloc(n) : std::malloc(n ? n : 1); }
void operator delete(void* p) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete[](void* p) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete(void* p, size_t) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete[](void* p, size_t) noexcept { if (p && !in_arena(p)) std::free(p); }

using namespace ORB_SLAM2;

struct Reader {
  std::vector<unsigned char> buf; size_t off = 0;
  bool load(const char* path) {
    FILE* f = std::fopen(path, "rb"); if (!f) return false;
    std::fseek(f, 0, SEEK_END); const long n = std::ftell(f); std::fseek(f, 0, SEEK_SET);
    buf.resize((size_t)n); const bool ok = std::fread(buf.data(), 1, (size_t)n, f) == (size_t)n; std::fclose(f); return ok;
  }
  template <typename T> T get() { T v; std::memcpy(&v, &buf[off], sizeof(T)); off += sizeof(T); return v; }
  const unsigned char* bytes(size_t n) { const unsigned char* p = &buf[off]; off += n; return p; }
};
struct Writer {
  FILE* f = nullptr;
  template <typename T> void put(T v) { std::fwrite(&v, sizeof(T), 1, f); }
  void bytes(const void* p, size_t n) { if (n) std::fwrite(p, 1, n, f); }
};

struct Calib { float fx, fy, cx, cy, d[5], bf, thDepth; };
static Calib read_calib(Reader& r) { Calib c; c.fx = r.get<float>(); c.fy = r.get<float>(); c.cx = r.get<float>(); c.cy = r.get<float>();
  for (int i = 0; i < 5; ++i) c.d[i] = r.get<float>(); c.bf = r.get<float>(); c.thDepth = r.get<float>(); return c; }
static cv::Mat make_K(const Calib& c) {
  cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
  K.at<float>(0, 0) = c.fx; K.at<float>(1, 1) = c.fy; K.at<float>(0, 2) = c.cx; K.at<float>(1, 2) = c.cy;
  return K;
}
static cv::Mat make_D(const Calib& c) {                       // as Tracking.cc:70-82 builds it: 4 x 1, 5 x 1 when k3 != 0
  cv::Mat D(c.d[4] != 0.f ? 5 : 4, 1, CV_32F);
  for (int i = 0; i < D.rows; ++i) D.at<float>(i) = c.d[i];
  return D;
}

// placement storage for a Frame, every 32-bit word pre-set to the float `fill` (hazard H4)
static Frame* frame_storage(float fill) {
  const size_t words = (sizeof(Frame) + 3) / 4;
  float* s = (float*)std::malloc(words * 4 + 64);
  for (size_t i = 0; i < words; ++i) s[i] = fill;
  return (Frame*)s;
}

static void write_frame(Writer& w, const Frame& F) {
  static_assert(sizeof(cv::KeyPoint) == 28, "KeyPoint layout");
  w.put<int>(F.N);
  w.bytes(F.mvKeys.data(), 28 * (size_t)F.N);
  w.bytes(F.mvKeysUn.data(), 28 * (size_t)F.N);
  for (int i = 0; i < F.N; ++i) w.bytes(F.mDescriptors.ptr(i), 32);
  w.bytes(F.mvuRight.data(), 4 * (size_t)F.N);
  w.bytes(F.mvDepth.data(), 4 * (size_t)F.N);
}

#include "ref_slam_cmds.inc"
