// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// Driver for oracle/_ref/orbref: the reference's own src/ORBextractor.cc, compiled in place from
// /root/reference against cvshim/, run as a separate process on images handed over in a file.
//
//   orbref extract <in.bin> <out.bin> nfeatures scale nlevels iniTh minTh [reps] [malloc]
//     in.bin : int32 w, h, n ; n tightly packed w*h uint8 images
//     out.bin: per image: int32 count ; count x 28-byte cv::KeyPoint ; count x 32 descriptor bytes
//     stdout : seconds spent inside ORBextractor::operator() (all images, all reps)
//
// Hazard H1: the reference sorts quad-tree nodes by (size, heap address).  By default this
// driver replaces operator new with a bump allocator that is rewound between images, so that
// addresses increase with creation order — the canonical order the restatement and the CUDA
// path reproduce.  With the trailing word "malloc" the system allocator is used instead (for
// timing, and to show that only the tie order depends on it).
#include <sys/mman.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <new>
#include <string>

#include "ORBextractor.h"

static char* g_base = nullptr;
static size_t g_off = 0, g_cap = 0;
static bool g_bump = false;

static void* bump_alloc(size_t n) {
  n = (n + 15) & ~size_t(15);
  if (g_off + n > g_cap) { std::fprintf(stderr, "orbref: arena exhausted\n"); std::abort(); }
  void* p = g_base + g_off;
  g_off += n;
  return p;
}
static inline bool in_arena(void* p) { return g_base && (char*)p >= g_base && (char*)p < g_base + g_cap; }

void* operator new(size_t n) { return g_bump ? bump_alloc(n) : std::malloc(n ? n : 1); }
void* operator new[](size_t n) { return g_bump ? bump_alloc(n) : std::malloc(n ? n : 1); }
void operator delete(void* p) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete[](void* p) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete(void* p, size_t) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete[](void* p, size_t) noexcept { if (p && !in_arena(p)) std::free(p); }

int main(int argc, char** argv) {
  if (argc < 9 || std::string(argv[1]) != "extract") {
    std::fprintf(stderr, "usage: orbref extract in.bin out.bin nfeatures scale nlevels iniTh minTh [reps] [malloc]\n");
    return 2;
  }
  const int nf = std::atoi(argv[4]); const float sf = (float)std::atof(argv[5]);
  const int nl = std::atoi(argv[6]), ini = std::atoi(argv[7]), mn = std::atoi(argv[8]);
  const int reps = argc > 9 ? std::atoi(argv[9]) : 1;
  const bool use_malloc = argc > 10 && std::string(argv[10]) == "malloc";
  FILE* fi = std::fopen(argv[2], "rb");
  if (!fi) return 3;
  int hdr[3];
  if (std::fread(hdr, 4, 3, fi) != 3) return 3;
  const int w = hdr[0], h = hdr[1], n = hdr[2];
  uchar* imgs = (uchar*)std::malloc((size_t)w * h * n);
  if (std::fread(imgs, (size_t)w * h, n, fi) != (size_t)n) return 3;
  std::fclose(fi);
  FILE* fo = std::fopen(argv[3], "wb");
  if (!fo) return 3;
  if (!use_malloc) {
    g_cap = size_t(16) << 30;
    g_base = (char*)mmap(nullptr, g_cap, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (g_base == MAP_FAILED) return 4;
  }
  double secs = 0;
  for (int r = 0; r < reps; ++r)
    for (int i = 0; i < n; ++i) {
      g_off = 0;
      g_bump = !use_malloc;
      int count = 0;
      {
        ORB_SLAM2::ORBextractor ext(nf, sf, nl, ini, mn);
        cv::Mat img(h, w, CV_8UC1, imgs + (size_t)i * w * h, (size_t)w);
        std::vector<cv::KeyPoint> kps;
        cv::Mat desc;
        auto t0 = std::chrono::steady_clock::now();
        ext(img, cv::Mat(), kps, desc);
        secs += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        count = (int)kps.size();
        if (r == reps - 1) {
          std::fwrite(&count, 4, 1, fo);
          static_assert(sizeof(cv::KeyPoint) == 28, "KeyPoint layout");
          if (count) {
            std::fwrite(kps.data(), 28, count, fo);
            for (int k = 0; k < count; ++k) std::fwrite(desc.ptr(k), 1, 32, fo);
          }
        }
      }
      g_bump = false;  // everything allocated for this image is dead; the arena is rewound above
    }
  std::fclose(fo);
  std::printf("%.6f\n", secs);
  return 0;
}
