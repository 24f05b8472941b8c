// Host side of the orbx C ABI (include/orbx.h): handle, geometry tables, launch orchestration.
//
// The arithmetic in make_tables()/configure() follows the reference's constructor and level
// set-up to the letter, because quotas, level sizes and resize weights feed bit-exact kernels:
//   scale/sigma tables, quotas, umax   src/ORBextractor.cc:410-470
//   level sizes                         src/ORBextractor.cc:1111-1112
//   FAST cell grid                      src/ORBextractor.cc:773-787
//   quad-tree roots                     src/ORBextractor.cc:543-563
//   cv::resize tap weights              OpenCV resize.cpp (INTER_LINEAR, 8-bit): see DESIGN.md
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <type_traits>
#include <string>
#include <vector>

#include "orbx_internal.h"

using namespace orbx;

namespace {

thread_local std::string g_err;
std::atomic<unsigned long long> g_launches{0};

orbx_status fail(orbx_status s, const std::string& msg) { g_err = msg; return s; }

#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return fail(ORBX_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));     \
  } while (0)

const int8_t kPattern[1024] = {
#include "brief_pattern.inc"
};

inline int cv_round(double v) { return (int)std::lrint(v); }
inline int cv_round(float v) { return (int)std::lrintf(v); }
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Totals {
  size_t pyr_bytes = 0;
  int cells = 0, pool = 0, lvlkp = 0, taps = 0, win_w = 0, win_h = 0, node_cap = 0;
};

// Makes `dev` current for the scope.  cudaSetDevice is called even when the runtime already reports `dev`: on a host thread
// that has made no runtime call yet it is what binds the primary context to the thread, and configure() goes to the driver
// API directly (cuTensorMapEncodeTiled), which fails with CUDA_ERROR_INVALID_CONTEXT on a thread without one - the
// reference's pattern of constructing the extractors on one thread and calling them on two others (src/Frame.cc:78-81).
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); cudaSetDevice(dev); if (prev == dev) prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

namespace orbx {
orbx_status set_error(orbx_status s, const char* msg) { return fail(s, msg); }
void count_launches(int n) { g_launches += n; }
}  // namespace orbx

struct orbx_extractor {
  orbx_params P{};
  int device = 0, max_w = 0, max_h = 0, max_batch = 0, nlevels = 0;
  float scale[kMaxLevels], inv[kMaxLevels], sigma2[kMaxLevels], invsigma2[kMaxLevels];
  int quota[kMaxLevels];
  int umax[kHalfPatch + 1];
  int kp_cap = 0;
  // geometry of the image size currently configured
  int cur_w = 0, cur_h = 0;
  LevelGeom hg[kMaxLevels];
  Totals cur, capy;
  std::vector<ResizeTap> htaps;
  DeviceLayout L{};
  TmaMaps desc_maps{};       // per-level tensor maps with the K4 tile box
  TmaMaps blur_src_maps{};   // K4a: source box of the whole-plane blur (raw pyramid)
  TmaMaps mom_maps{};        // K4b: raw box for the orientation moments
  TmaMaps tap_maps{};        // K4b: blurred box for the descriptor taps (blur arena)
  bool fused_describe = false;   // ORBX_DESCRIBE=fused at create: the blur-on-demand kernel instead of K4a + K4b (kept for A/B measurements)
  TmaMaps resize_maps{};     // per-level tensor maps with the K1 source-footprint box
  TmaMaps fast_maps{};       // per-level tensor maps with the K2 cell-window box
  unsigned tiled_mask = 0;   // levels whose resize runs through the shared-memory tiled kernel
  // device arenas
  LevelGeom* d_geom = nullptr; ResizeTap* d_taps = nullptr; int* d_umax = nullptr; char4* d_pattern = nullptr;
  uint8_t* d_in = nullptr;
  float* d_uright = nullptr; float* d_depth = nullptr; int* d_sad = nullptr; int* d_nmatch = nullptr;
  uint4* d_strec = nullptr; int* d_sttab = nullptr; int st_tab_stride = 0;   // K5a scratch (sorted right keypoints, bin starts)
  int slots_valid = 0;
  // optional per-stage timing: a ring of event sets, 5 boundaries per extract pass, 3 per stereo pass
  bool profiling = false;
  static constexpr int kRing = 256;
  cudaEvent_t ev_ext[kRing][6] = {}; int n_ext = 0;
  cudaEvent_t ev_st[kRing][3] = {}; int n_st = 0;
};

namespace {

void make_tables(orbx_extractor* e) {
  const int nl = e->nlevels;
  const double sf = (double)e->P.scale_factor;   // the reference keeps scaleFactor as double (ORBextractor.h:101)
  e->scale[0] = 1.0f; e->sigma2[0] = 1.0f;
  for (int i = 1; i < nl; ++i) {
    e->scale[i] = (float)(e->scale[i - 1] * sf);
    e->sigma2[i] = e->scale[i] * e->scale[i];
  }
  for (int i = 0; i < nl; ++i) { e->inv[i] = 1.0f / e->scale[i]; e->invsigma2[i] = 1.0f / e->sigma2[i]; }
  const float factor = (float)(1.0f / sf);
  float want = e->P.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
  int sum = 0;
  for (int l = 0; l < nl - 1; ++l) {
    e->quota[l] = cv_round(want);
    sum += e->quota[l];
    want *= factor;
  }
  e->quota[nl - 1] = std::max(e->P.nfeatures - sum, 0);
  const int vmax = (int)std::floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
  const int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
  const double r2 = kHalfPatch * kHalfPatch;
  for (int v = 0; v <= vmax; ++v) e->umax[v] = cv_round(std::sqrt(r2 - v * v));
  for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
    while (e->umax[v0] == e->umax[v0 + 1]) ++v0;
    e->umax[v] = v0;
    ++v0;
  }
}

void resize_taps(int dn, int sn, ResizeTap* out) {
  const double inv = (double)dn / sn, sc = 1.0 / inv;
  for (int d = 0; d < dn; ++d) {
    float f = (float)((d + 0.5) * sc - 0.5);
    int s = (int)std::floor(f);
    f -= s;
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= sn - 1) { s = sn - 1; f = 0.f; }
    out[d].ofs = s;
    out[d].a0 = (short)cv_round((1.f - f) * 2048.f);
    out[d].a1 = (short)cv_round(f * 2048.f);
  }
}

// Level geometry for a w x h image.  nslots fixes the per-level plane arrays.
orbx_status geometry(const orbx_extractor* e, int w, int h, int nslots, LevelGeom* g, Totals* t,
                     std::vector<ResizeTap>* taps) {
  *t = Totals();
  if (taps) taps->clear();
  size_t pyr_off = 0;
  for (int l = 0; l < e->nlevels; ++l) {
    LevelGeom& q = g[l];
    std::memset(&q, 0, sizeof(q));
    q.w = cv_round((float)w * e->inv[l]);
    q.h = cv_round((float)h * e->inv[l]);
    if (q.w < 2 * kFastBorder + 30 || q.h < 2 * kFastBorder + 30)
      return fail(ORBX_ERR_GEOMETRY, "pyramid level " + std::to_string(l) + " is smaller than 62 px");
    q.pitch = (int)align_up(q.w + 2 * kEdge, 16);
    q.rows = q.h + 2 * kEdge;
    q.plane_stride = align_up((size_t)q.pitch * q.rows, 256);
    q.plane_off = pyr_off;
    pyr_off += q.plane_stride * nslots;
    if (l > 0 && g[l - 1].w == 2 * q.w && g[l - 1].h == 2 * q.h)
      return fail(ORBX_ERR_GEOMETRY, "level " + std::to_string(l) + " is an exact 2x decimation: cv::resize(INTER_LINEAR) silently switches to INTER_AREA there, which this library does not reproduce");
    if (l > 0) {
      // tap tables in BORDERED coordinates: entry bx holds the taps of ROI column reflect101(bx-19),
      // pad columns get zero weights (-> pixel 0); 4-entry (32-byte) aligned for 128-bit loads
      q.xtab_off = t->taps; q.ytab_off = t->taps + q.pitch;
      t->taps += q.pitch + (int)align_up(q.rows, 4);
      if (taps) {
        taps->resize(t->taps);
        std::vector<ResizeTap> tx(q.w), ty(q.h);
        resize_taps(q.w, g[l - 1].w, tx.data());
        resize_taps(q.h, g[l - 1].h, ty.data());
        // pad entries: zero weights (-> pixel 0) at the last valid source offset, so that they do not
        // widen the source footprint of a tile
        ResizeTap padx = tx[reflect101(q.w + kEdge - 1, q.w)], pady = ty[reflect101(q.h + kEdge - 1, q.h)];
        padx.a0 = padx.a1 = 0; pady.a0 = pady.a1 = 0;
        for (int bx = 0; bx < q.pitch; ++bx)
          (*taps)[q.xtab_off + bx] = bx < q.w + 2 * kEdge ? tx[reflect101(bx - kEdge, q.w)] : padx;
        for (int by = 0; by < (int)align_up(q.rows, 4); ++by)
          (*taps)[q.ytab_off + by] = by < q.rows ? ty[reflect101(by - kEdge, q.h)] : pady;
      }
      // source footprint of every output tile of the tiled kernel (it used to be found by shared-memory
      // atomics in every CTA): the TMA box can be requested before the CTA has read a single tap
      const int tiles_x = (q.pitch + kResizeTileW - 1) / kResizeTileW, tiles_y = (q.rows + kResizeTileH - 1) / kResizeTileH;
      q.xtile_off = t->taps; q.ytile_off = t->taps + tiles_x;
      t->taps += (int)align_up(tiles_x + tiles_y, 4);
      if (taps) {
        taps->resize(t->taps);
        for (int i = 0; i < tiles_x; ++i) {
          int lo = 1 << 30;
          for (int c = i * kResizeTileW; c < std::min((i + 1) * kResizeTileW, q.pitch); ++c) lo = std::min(lo, (*taps)[q.xtab_off + c].ofs);
          ResizeTap e; e.ofs = lo; e.a0 = 0; e.a1 = 0;
          (*taps)[q.xtile_off + i] = e;
        }
        for (int i = 0; i < tiles_y; ++i) {
          int lo = 1 << 30, hi = -1;
          for (int r = i * kResizeTileH; r < std::min((i + 1) * kResizeTileH, q.rows); ++r) {
            lo = std::min(lo, (*taps)[q.ytab_off + r].ofs); hi = std::max(hi, (*taps)[q.ytab_off + r].ofs + 1);
          }
          ResizeTap e; e.ofs = lo; e.a0 = (short)(hi - lo + 1); e.a1 = 0;
          (*taps)[q.ytile_off + i] = e;
        }
      }
    }
    // FAST cells
    q.maxBX = q.w - kEdge + 3; q.maxBY = q.h - kEdge + 3;
    const float width = (float)(q.maxBX - kFastBorder), height = (float)(q.maxBY - kFastBorder);
    q.nCols = (int)(width / 30.f); q.nRows = (int)(height / 30.f);
    q.wCell = (int)std::ceil(width / q.nCols); q.hCell = (int)std::ceil(height / q.nRows);
    q.cell_base = t->cells;
    t->cells += q.nCols * q.nRows;
    t->win_w = std::max(t->win_w, q.wCell + 6); t->win_h = std::max(t->win_h, q.hCell + 6);
    // candidate pool bound: kept corners are strict 3x3 maxima, at most ceil(iw/2)*ceil(ih/2) per cell
    long long cap = 0;
    for (int i = 0; i < q.nRows; ++i)
      for (int j = 0; j < q.nCols; ++j) {
        const int x0 = kFastBorder + j * q.wCell, y0 = kFastBorder + i * q.hCell;
        if (y0 >= q.maxBY - 3 || x0 >= q.maxBX - 6) continue;
        const int ww = std::min(x0 + q.wCell + 6, q.maxBX) - x0, hh = std::min(y0 + q.hCell + 6, q.maxBY) - y0;
        if (ww < 7 || hh < 7) continue;
        cap += (long long)((ww - 6 + 1) / 2) * ((hh - 6 + 1) / 2);
      }
    if (cap >= (1 << 24)) return fail(ORBX_ERR_GEOMETRY, "level too large for 24-bit candidate indices");
    q.cand_cap = (int)align_up((size_t)cap + 1, 4);
    q.cand_off = t->pool;
    t->pool += q.cand_cap;
    // quad-tree
    q.quota = e->quota[l];
    q.treeW = q.maxBX - kFastBorder; q.treeH = q.maxBY - kFastBorder;
    q.nIni = (int)std::round((float)q.treeW / q.treeH);
    if (q.nIni < 1 || q.nIni > kMaxRoots)
      return fail(ORBX_ERR_GEOMETRY, "aspect ratio outside [0.5, 16.5): round(W/H) must be 1..16");
    q.hX = (float)q.treeW / q.nIni;
    for (int i = 0; i < q.nIni; ++i) { q.root_x0[i] = (int)(q.hX * (float)i); q.root_x1[i] = (int)(q.hX * (float)(i + 1)); }
    q.node_cap = std::max(q.quota + 3, 4 * q.nIni) + 1;
    t->node_cap = std::max(t->node_cap, q.node_cap);
    q.lvl_kp_off = t->lvlkp;
    t->lvlkp += std::max(q.quota + 3, 4 * kMaxRoots);
    q.scale = e->scale[l]; q.inv_scale = e->inv[l];
    q.patch_size = (int)(31 * e->scale[l]);
  }
  t->pyr_bytes = pyr_off;
  t->node_cap = (int)align_up(t->node_cap, 32);
  return ORBX_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

orbx_status encode_level_maps(const orbx_extractor* e, const LevelGeom* g, int box_w, int box_h, TmaMaps* out, uint8_t* base = nullptr) {
  if (!base) base = e->L.pyr;
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    CU(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    if (q != cudaDriverEntryPointSuccess || !p) return fail(ORBX_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
    fn = (EncodeTiledFn)p;
  }
  for (int l = 0; l < e->nlevels; ++l) {
    const cuuint64_t dims[3] = {(cuuint64_t)g[l].pitch, (cuuint64_t)g[l].rows, (cuuint64_t)e->max_batch};
    const cuuint64_t strides[2] = {(cuuint64_t)g[l].pitch, (cuuint64_t)g[l].plane_stride};
    const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    const CUresult r = fn(&out->m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base + g[l].plane_off, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ORBX_ERR_CUDA, "cuTensorMapEncodeTiled failed for level " + std::to_string(l) + " (CUresult " + std::to_string((int)r) + ")");
  }
  return ORBX_OK;
}

orbx_status configure(orbx_extractor* e, int w, int h, cudaStream_t st) {
  if (w == e->cur_w && h == e->cur_h) return ORBX_OK;
  if (w > e->max_w || h > e->max_h) return fail(ORBX_ERR_CAPACITY, "image larger than max_w x max_h given at create");
  LevelGeom g[kMaxLevels];
  Totals t;
  std::vector<ResizeTap> taps;
  orbx_status s = geometry(e, w, h, e->max_batch, g, &t, &taps);
  if (s != ORBX_OK) return s;
  if (t.pyr_bytes > e->capy.pyr_bytes || t.cells > e->capy.cells || t.pool > e->capy.pool || t.taps > e->capy.taps ||
      t.lvlkp > e->capy.lvlkp || t.node_cap > e->capy.node_cap || t.win_w > e->capy.win_w || t.win_h > e->capy.win_h)
    return fail(ORBX_ERR_CAPACITY, "image geometry needs more device memory than reserved at create");
  TmaMaps maps, rmaps, fmaps, bmaps, mmaps, tmaps;
  s = encode_level_maps(e, g, kBlurBoxW, kBlurBoxH, &bmaps);
  if (s != ORBX_OK) return s;
  s = encode_level_maps(e, g, kMomBoxW, kMomBoxH, &mmaps);
  if (s != ORBX_OK) return s;
  s = encode_level_maps(e, g, kTapBoxW, kTapBoxH, &tmaps, e->L.blur);
  if (s != ORBX_OK) return s;
  s = encode_level_maps(e, g, fast_window_pitch(t.win_w), t.win_h, &fmaps);
  if (s != ORBX_OK) return s;
  s = encode_level_maps(e, g, kDescTileW, kDescTileH, &maps);
  if (s != ORBX_OK) return s;
  s = encode_level_maps(e, g, kResizeBoxW, kResizeBoxH, &rmaps);
  if (s != ORBX_OK) return s;
  // the previous geometry may still be in use by kernels queued on another stream
  CU(cudaDeviceSynchronize());
  e->desc_maps = maps;
  e->resize_maps = rmaps;
  e->fast_maps = fmaps;
  e->blur_src_maps = bmaps; e->mom_maps = mmaps; e->tap_maps = tmaps;
  e->tiled_mask = 0;
  for (int l = 1; l < e->nlevels; ++l)
    if (resize_tiled_fits(g[l], taps.data())) e->tiled_mask |= 1u << l;
  std::memcpy(e->hg, g, sizeof(g));
  e->cur = t;
  e->htaps = taps;
  CU(cudaMemcpyAsync(e->d_geom, e->hg, sizeof(LevelGeom) * e->nlevels, cudaMemcpyHostToDevice, st));
  if (!taps.empty()) CU(cudaMemcpyAsync(e->d_taps, e->htaps.data(), sizeof(ResizeTap) * taps.size(), cudaMemcpyHostToDevice, st));
  CU(cudaStreamSynchronize(st));
  e->cur_w = w; e->cur_h = h;
  e->L.cells_per_slot = t.cells; e->L.pool_per_slot = t.pool; e->L.lvlkp_per_slot = t.lvlkp;   // all <= reserved capacity
  e->slots_valid = 0;
  return ORBX_OK;
}

template <typename T>
orbx_status dev_alloc(T** p, size_t count) {
  CU(cudaMalloc((void**)p, std::max<size_t>(count, 1) * sizeof(T)));
  return ORBX_OK;
}

}  // namespace

extern "C" {

orbx_status orbx_extractor_create(const orbx_params* params, int max_w, int max_h, int max_batch, int device,
                                  orbx_extractor** out) {
  if (!params || !out) return fail(ORBX_ERR_BAD_ARG, "null argument");
  *out = nullptr;
  if (params->nlevels < 1 || params->nlevels > kMaxLevels) return fail(ORBX_ERR_BAD_ARG, "nlevels must be 1..8");
  if (params->nfeatures < 1 || params->nfeatures > 30000) return fail(ORBX_ERR_BAD_ARG, "nfeatures must be 1..30000");
  if (!(params->scale_factor > 1.0f) || params->scale_factor > 4.0f) return fail(ORBX_ERR_BAD_ARG, "scale_factor must be in (1, 4]");
  if (params->ini_th_fast < 1 || params->ini_th_fast > 254 || params->min_th_fast < 1 || params->min_th_fast > 254)
    return fail(ORBX_ERR_BAD_ARG, "FAST thresholds must be in [1, 254]");
  if (max_w < 1 || max_h < 1 || max_w > kMaxDim + 1 || max_h > kMaxDim + 1) return fail(ORBX_ERR_BAD_ARG, "max_w/max_h must be 1..4096");
  if (max_batch < 1) return fail(ORBX_ERR_BAD_ARG, "max_batch must be >= 1");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(ORBX_ERR_CUDA, "no CUDA device: orbx has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(ORBX_ERR_BAD_ARG, "device index out of range");
  DeviceGuard guard(device);

  orbx_extractor* e = new orbx_extractor();
  e->P = *params; e->device = device; e->max_w = max_w; e->max_h = max_h; e->max_batch = max_batch;
  e->nlevels = params->nlevels;
  make_tables(e);
  {
    static const int expect[kHalfPatch + 1] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
    if (std::memcmp(e->umax, expect, sizeof(expect)) != 0) { delete e; return fail(ORBX_ERR_STATE, "umax table differs from the constant compiled into k_describe"); }
  }
  LevelGeom g[kMaxLevels];
  orbx_status s = geometry(e, max_w, max_h, max_batch, g, &e->capy, nullptr);
  if (s != ORBX_OK) { delete e; return s; }
  if (e->capy.node_cap > 3072) { delete e; return fail(ORBX_ERR_CAPACITY, "per-level feature quota too large for the quad-tree kernel (node_cap > 3072)"); }
  e->capy.win_w = e->capy.win_h = 66;   // a FAST cell is at most 60 px wide/high (+6 rim) for any smaller image
  e->kp_cap = e->capy.lvlkp;
  if (e->kp_cap > 65535) { delete e; return fail(ORBX_ERR_CAPACITY, "more than 65535 keypoints per image"); }

  DeviceLayout& L = e->L;
  L.nlevels = e->nlevels; L.nslots = max_batch;
  L.cells_per_slot = e->capy.cells; L.pool_per_slot = e->capy.pool; L.lvlkp_per_slot = e->capy.lvlkp; L.kp_cap = e->kp_cap;
  L.ini_th = params->ini_th_fast; L.min_th = params->min_th_fast;
  const size_t nb = (size_t)max_batch;
#define ALLOC(ptr, count) do { s = dev_alloc(&(ptr), (count)); if (s != ORBX_OK) { orbx_extractor_destroy(e); return s; } } while (0)
  ALLOC(e->d_geom, kMaxLevels);
  ALLOC(e->d_taps, (size_t)e->capy.taps);
  ALLOC(e->d_umax, kHalfPatch + 1);
  ALLOC(e->d_pattern, 256);
  ALLOC(e->d_in, nb * max_w * max_h);
  ALLOC(L.pyr, e->capy.pyr_bytes);
  {
    const char* mode = std::getenv("ORBX_DESCRIBE");
    e->fused_describe = mode && std::strcmp(mode, "fused") == 0;
  }
  ALLOC(L.blur, e->fused_describe ? 1 : e->capy.pyr_bytes);
  ALLOC(L.celltab, nb * e->capy.cells);
  ALLOC(L.poolcnt, nb * kMaxLevels);
  ALLOC(L.pool, nb * e->capy.pool);
  ALLOC(L.okeys, nb * e->capy.pool);
  ALLOC(L.knode, nb * e->capy.pool);
  ALLOC(L.lvlkp, nb * e->capy.lvlkp);
  ALLOC(L.lvlcnt, nb * kMaxLevels);
  ALLOC(L.nkeys, nb * kMaxLevels);
  ALLOC(L.kps, nb * e->kp_cap);
  ALLOC(L.kpk, nb * e->kp_cap);
  ALLOC(L.desc, nb * e->kp_cap * 32);
  ALLOC(L.nkp, nb);
  ALLOC(L.err, nb);
  ALLOC(e->d_uright, nb * e->kp_cap);
  ALLOC(e->d_depth, nb * e->kp_cap);
  ALLOC(e->d_sad, nb * e->kp_cap);
  ALLOC(e->d_nmatch, nb);
  e->st_tab_stride = e->nlevels * max_h + 2;
  ALLOC(e->d_strec, nb * e->kp_cap);
  ALLOC(e->d_sttab, nb * e->st_tab_stride);
#undef ALLOC
  L.geom = e->d_geom; L.taps = e->d_taps;
  cudaError_t ce = cudaMemcpy(e->d_umax, e->umax, sizeof(e->umax), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(e->d_pattern, kPattern, 1024, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemset(L.nkp, 0, nb * sizeof(int));
  if (ce == cudaSuccess) ce = cudaMemset(L.pyr, 0, e->capy.pyr_bytes);
  if (ce != cudaSuccess) { orbx_extractor_destroy(e); return fail(ORBX_ERR_CUDA, cudaGetErrorString(ce)); }
  *out = e;
  return ORBX_OK;
}

void orbx_extractor_destroy(orbx_extractor* e) {
  if (!e) return;
  DeviceGuard guard(e->device);
  cudaDeviceSynchronize();
  DeviceLayout& L = e->L;
  void* ptrs[] = {e->d_geom, e->d_taps, e->d_umax, e->d_pattern, e->d_in, L.pyr, L.blur, L.celltab, L.poolcnt, L.pool, L.okeys,
                  L.knode, L.lvlkp, L.lvlcnt, L.nkeys, L.kps, L.kpk, L.desc, L.nkp, L.err, e->d_uright, e->d_depth, e->d_sad,
                  e->d_nmatch, e->d_strec, e->d_sttab};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (e->ev_ext[0][0])
    for (int i = 0; i < orbx_extractor::kRing; ++i) {
      for (int j = 0; j < 6; ++j) cudaEventDestroy(e->ev_ext[i][j]);
      for (int j = 0; j < 3; ++j) cudaEventDestroy(e->ev_st[i][j]);
    }
  delete e;
}

orbx_status orbx_get_tables(const orbx_extractor* e, int* nlevels, float* scale, float* inv_scale, float* sigma2,
                            float* inv_sigma2, int* quota) {
  if (!e) return fail(ORBX_ERR_BAD_ARG, "null extractor");
  if (nlevels) *nlevels = e->nlevels;
  for (int i = 0; i < e->nlevels; ++i) {
    if (scale) scale[i] = e->scale[i];
    if (inv_scale) inv_scale[i] = e->inv[i];
    if (sigma2) sigma2[i] = e->sigma2[i];
    if (inv_sigma2) inv_sigma2[i] = e->invsigma2[i];
    if (quota) quota[i] = e->quota[i];
  }
  return ORBX_OK;
}

orbx_status orbx_compute_tables(const orbx_params* params, int* nlevels, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                                int* quota) {
  if (!params) return fail(ORBX_ERR_BAD_ARG, "null params");
  if (params->nlevels < 1 || params->nlevels > kMaxLevels || !(params->scale_factor > 1.0f) || params->nfeatures < 1)
    return fail(ORBX_ERR_BAD_ARG, "nlevels must be 1..8, scale_factor > 1, nfeatures >= 1");
  orbx_extractor tmp;                       // host tables only: no device is touched
  tmp.P = *params; tmp.nlevels = params->nlevels;
  make_tables(&tmp);
  return orbx_get_tables(&tmp, nlevels, scale, inv_scale, sigma2, inv_sigma2, quota);
}


int orbx_max_keypoints(const orbx_extractor* e) { return e ? e->kp_cap : 0; }

orbx_status orbx_extract_batch(orbx_extractor* e, const uint8_t* imgs, int n, int w, int h, int stride,
                               size_t image_stride, int src_on_device, int first_slot, void* stream) {
  if (!e || !imgs) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n < 1 || w < 1 || h < 1 || stride < w) return fail(ORBX_ERR_BAD_ARG, "bad image size or stride");
  if (first_slot < 0 || first_slot + n > e->max_batch) return fail(ORBX_ERR_CAPACITY, "slots exceed max_batch");
  cudaStream_t st = (cudaStream_t)stream;
  DeviceGuard guard(e->device);
  orbx_status s = configure(e, w, h, st);
  if (s != ORBX_OK) return s;
  const uint8_t* src = imgs;
  int src_pitch = stride;
  size_t src_istride = image_stride;
  if (!src_on_device) {
    uint8_t* dst = e->d_in + (size_t)first_slot * w * h;
    if (stride == w && image_stride == (size_t)w * h) {
      CU(cudaMemcpyAsync(dst, imgs, (size_t)n * w * h, cudaMemcpyHostToDevice, st));
    } else {
      for (int i = 0; i < n; ++i)
        CU(cudaMemcpy2DAsync(dst + (size_t)i * w * h, w, imgs + (size_t)i * image_stride, stride, w, h, cudaMemcpyHostToDevice, st));
    }
    src = dst; src_pitch = w; src_istride = (size_t)w * h;
  }
  CU(cudaMemsetAsync(e->L.poolcnt + (size_t)first_slot * kMaxLevels, 0, (size_t)n * kMaxLevels * sizeof(int), st));
  int launches = 0;
  cudaEvent_t* ev = (e->profiling && e->n_ext < orbx_extractor::kRing) ? e->ev_ext[e->n_ext++] : nullptr;
  if (ev) cudaEventRecord(ev[0], st);
  launches += launch_pyramid(e->L, e->hg, src, src_pitch, src_istride, first_slot, n, e->tiled_mask, &e->resize_maps, st);
  if (ev) cudaEventRecord(ev[1], st);
  launches += launch_fast(e->L, e->hg, e->fast_maps, first_slot, n, e->cur.win_w, e->cur.win_h, st);
  if (ev) cudaEventRecord(ev[2], st);
  launches += launch_octree(e->L, first_slot, n, e->cur.node_cap, st);
  if (ev) cudaEventRecord(ev[3], st);
  if (!e->fused_describe) launches += launch_blur(e->L, e->hg, e->blur_src_maps, first_slot, n, st);
  if (ev) cudaEventRecord(ev[4], st);
  if (e->fused_describe) launches += launch_describe(e->L, e->desc_maps, first_slot, n, e->d_umax, e->d_pattern, st);
  else launches += launch_describe_planes(e->L, e->mom_maps, e->tap_maps, first_slot, n, e->d_pattern, st);
  if (ev) cudaEventRecord(ev[5], st);
  g_launches += launches;
  CU(cudaGetLastError());
  e->slots_valid = std::max(e->slots_valid, first_slot + n);
  return ORBX_OK;
}

orbx_status orbx_fetch_results(orbx_extractor* e, int first_slot, int n, orbx_keypoint* kps, uint8_t* desc, int* counts,
                               int cap, int dst_on_device, void* stream) {
  if (!e) return fail(ORBX_ERR_BAD_ARG, "null extractor");
  if (n < 1 || first_slot < 0 || first_slot + n > e->max_batch) return fail(ORBX_ERR_CAPACITY, "slots exceed max_batch");
  if ((kps || desc) && cap < e->kp_cap) return fail(ORBX_ERR_CAPACITY, "cap smaller than orbx_max_keypoints()");
  cudaStream_t st = (cudaStream_t)stream;
  DeviceGuard guard(e->device);
  const cudaMemcpyKind kind = dst_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  const size_t kc = (size_t)e->kp_cap;
  if (counts) CU(cudaMemcpyAsync(counts, e->L.nkp + first_slot, (size_t)n * sizeof(int), kind, st));
  if (kps) CU(cudaMemcpy2DAsync(kps, (size_t)cap * sizeof(orbx_keypoint), e->L.kps + first_slot * kc, kc * sizeof(orbx_keypoint),
                                kc * sizeof(orbx_keypoint), n, kind, st));
  if (desc) CU(cudaMemcpy2DAsync(desc, (size_t)cap * 32, e->L.desc + first_slot * kc * 32, kc * 32, kc * 32, n, kind, st));
  return ORBX_OK;
}

orbx_status orbx_fetch_results_packed(orbx_extractor* e, int first_slot, int n, uint64_t* packed, uint8_t* desc, int* level_counts, int cap,
                                      int dst_on_device, void* stream) {
  if (!e) return fail(ORBX_ERR_BAD_ARG, "null extractor");
  if (n < 1 || first_slot < 0 || first_slot + n > e->max_batch) return fail(ORBX_ERR_CAPACITY, "slots exceed max_batch");
  if ((packed || desc) && cap < e->kp_cap) return fail(ORBX_ERR_CAPACITY, "cap smaller than orbx_max_keypoints()");
  cudaStream_t st = (cudaStream_t)stream;
  DeviceGuard guard(e->device);
  const cudaMemcpyKind kind = dst_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  const size_t kc = (size_t)e->kp_cap;
  if (level_counts) CU(cudaMemcpyAsync(level_counts, e->L.lvlcnt + (size_t)first_slot * kMaxLevels, (size_t)n * kMaxLevels * sizeof(int), kind, st));
  if (packed) CU(cudaMemcpy2DAsync(packed, (size_t)cap * 8, e->L.kpk + first_slot * kc, kc * 8, kc * 8, n, kind, st));
  if (desc) CU(cudaMemcpy2DAsync(desc, (size_t)cap * 32, e->L.desc + first_slot * kc * 32, kc * 32, kc * 32, n, kind, st));
  return ORBX_OK;
}

orbx_status orbx_unpack_keypoints(const orbx_params* params, const uint64_t* packed, const int* level_counts, int n_frames, int cap,
                                  orbx_keypoint* kps, int kps_cap, int* counts) {
  if (!params || !packed || !level_counts || !kps) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (params->nlevels < 1 || params->nlevels > kMaxLevels || !(params->scale_factor > 1.0f) || params->nfeatures < 1 || n_frames < 1 || cap < 1 || kps_cap < 1)
    return fail(ORBX_ERR_BAD_ARG, "bad sizes");
  orbx_extractor tmp;                       // host tables only (scale factors exactly as the kernels use them)
  tmp.P = *params; tmp.nlevels = params->nlevels;
  make_tables(&tmp);
  for (int f = 0; f < n_frames; ++f) {
    const uint64_t* src = packed + (size_t)f * cap;
    orbx_keypoint* dst = kps + (size_t)f * kps_cap;
    int j = 0;
    for (int l = 0; l < tmp.nlevels; ++l) {
      const float scale = tmp.scale[l], size = (float)(int)(31 * tmp.scale[l]);      // src/ORBextractor.cc:845, :1095-1101
      const int cnt = level_counts[(size_t)f * kMaxLevels + l];
      for (int i = 0; i < cnt && j < cap; ++i, ++j) {
        if (j >= kps_cap) return fail(ORBX_ERR_CAPACITY, "keypoint buffer too small");
        const uint32_t key = (uint32_t)src[j], ab = (uint32_t)(src[j] >> 32);
        orbx_keypoint kp;
        const float px = (float)unpack_x(key), py = (float)unpack_y(key);
        kp.x = l ? px * scale : px;
        kp.y = l ? py * scale : py;
        kp.size = size;
        std::memcpy(&kp.angle, &ab, 4);
        kp.response = (float)unpack_r(key);
        kp.octave = l; kp.class_id = -1;
        dst[j] = kp;
      }
    }
    if (counts) counts[f] = j;
  }
  return ORBX_OK;
}

orbx_status orbx_extract(orbx_extractor* e, const uint8_t* img, int w, int h, int stride, orbx_keypoint* kps, uint8_t* desc,
                         int cap, int* n, void* stream) {
  if (!e || !n) return fail(ORBX_ERR_BAD_ARG, "null argument");
  *n = 0;
  if (!img || w <= 0 || h <= 0) return ORBX_OK;   // empty image: silent return, src/ORBextractor.cc:1046
  if (!kps || !desc) return fail(ORBX_ERR_BAD_ARG, "null output buffer");
  cudaStream_t st = (cudaStream_t)stream;
  DeviceGuard guard(e->device);
  orbx_status s = orbx_extract_batch(e, img, 1, w, h, stride, (size_t)stride * h, 0, 0, stream);
  if (s != ORBX_OK) return s;
  int count = 0;
  CU(cudaMemcpyAsync(&count, e->L.nkp, sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (count > cap) return fail(ORBX_ERR_CAPACITY, "keypoint buffer too small: need " + std::to_string(count));
  if (count) {
    CU(cudaMemcpyAsync(kps, e->L.kps, (size_t)count * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(desc, e->L.desc, (size_t)count * 32, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  *n = count;
  return ORBX_OK;
}

orbx_status orbx_slot_results(const orbx_extractor* e, int first_slot, const orbx_keypoint** kps_dev, const uint8_t** desc_dev,
                              const int** counts_dev, int* cap) {
  if (!e) return fail(ORBX_ERR_BAD_ARG, "null extractor");
  if (first_slot < 0 || first_slot >= e->max_batch) return fail(ORBX_ERR_BAD_ARG, "slot out of range");
  const size_t kc = (size_t)e->kp_cap;
  if (kps_dev) *kps_dev = e->L.kps + first_slot * kc;
  if (desc_dev) *desc_dev = e->L.desc + first_slot * kc * 32;
  if (counts_dev) *counts_dev = e->L.nkp + first_slot;
  if (cap) *cap = e->kp_cap;
  return ORBX_OK;
}

orbx_status orbx_pyramid_level(const orbx_extractor* e, int slot, int level, const uint8_t** dev_ptr, int* w, int* h, int* stride) {
  if (!e) return fail(ORBX_ERR_BAD_ARG, "null extractor");
  if (e->cur_w == 0) return fail(ORBX_ERR_STATE, "no image extracted yet");
  if (slot < 0 || slot >= e->max_batch || level < 0 || level >= e->nlevels) return fail(ORBX_ERR_BAD_ARG, "slot/level out of range");
  const LevelGeom& g = e->hg[level];
  if (dev_ptr) *dev_ptr = e->L.pyr + g.plane_off + (size_t)slot * g.plane_stride + (size_t)kEdge * g.pitch + kEdge;
  if (w) *w = g.w;
  if (h) *h = g.h;
  if (stride) *stride = g.pitch;
  return ORBX_OK;
}

orbx_status orbx_download_level(orbx_extractor* e, int slot, int level, uint8_t* dst) {
  if (!e || !dst) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (e->cur_w == 0) return fail(ORBX_ERR_STATE, "no image extracted yet");
  if (slot < 0 || slot >= e->max_batch || level < 0 || level >= e->nlevels) return fail(ORBX_ERR_BAD_ARG, "slot/level out of range");
  DeviceGuard guard(e->device);
  const LevelGeom& g = e->hg[level];
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy2D(dst, g.w + 2 * kEdge, e->L.pyr + g.plane_off + (size_t)slot * g.plane_stride, g.pitch, g.w + 2 * kEdge, g.rows,
                  cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

orbx_status orbx_debug_candidates(orbx_extractor* e, int slot, int level, uint32_t* out, int cap, int* n) {
  if (!e || !out || !n) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (e->cur_w == 0) return fail(ORBX_ERR_STATE, "no image extracted yet");
  if (slot < 0 || slot >= e->max_batch || level < 0 || level >= e->nlevels) return fail(ORBX_ERR_BAD_ARG, "slot/level out of range");
  DeviceGuard guard(e->device);
  CU(cudaDeviceSynchronize());
  int cnt = 0;
  CU(cudaMemcpy(&cnt, e->L.nkeys + slot * kMaxLevels + level, sizeof(int), cudaMemcpyDeviceToHost));
  *n = cnt;
  const int m = std::min(cnt, cap);
  if (m > 0) CU(cudaMemcpy(out, e->L.okeys + (size_t)slot * e->L.pool_per_slot + e->hg[level].cand_off, (size_t)m * 4, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

orbx_status orbx_debug_level_keys(orbx_extractor* e, int slot, int level, uint32_t* out, int cap, int* n) {
  if (!e || !out || !n) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (e->cur_w == 0) return fail(ORBX_ERR_STATE, "no image extracted yet");
  if (slot < 0 || slot >= e->max_batch || level < 0 || level >= e->nlevels) return fail(ORBX_ERR_BAD_ARG, "slot/level out of range");
  DeviceGuard guard(e->device);
  CU(cudaDeviceSynchronize());
  int cnt = 0;
  CU(cudaMemcpy(&cnt, e->L.lvlcnt + slot * kMaxLevels + level, sizeof(int), cudaMemcpyDeviceToHost));
  *n = cnt;
  const int m = std::min(cnt, cap);
  if (m > 0) CU(cudaMemcpy(out, e->L.lvlkp + (size_t)slot * e->L.lvlkp_per_slot + e->hg[level].lvl_kp_off, (size_t)m * 4, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

orbx_status orbx_stereo_match_batch(orbx_extractor* eL, int slotL0, orbx_extractor* eR, int slotR0, int n, float bf, float fx,
                                    float* u_right, float* depth, int* n_matches, int cap, int dst_on_device, void* stream) {
  if (!eL || !eR) return fail(ORBX_ERR_BAD_ARG, "null extractor");
  if (n < 1 || slotL0 < 0 || slotR0 < 0 || slotL0 + n > eL->max_batch || slotR0 + n > eR->max_batch)
    return fail(ORBX_ERR_CAPACITY, "slots exceed max_batch");
  if (eL->cur_w == 0 || eR->cur_w == 0 || slotL0 + n > eL->slots_valid || slotR0 + n > eR->slots_valid)
    return fail(ORBX_ERR_STATE, "stereo matching needs a preceding extract on those slots");
  if (eL->cur_w != eR->cur_w || eL->cur_h != eR->cur_h || eL->nlevels != eR->nlevels || eL->device != eR->device)
    return fail(ORBX_ERR_BAD_ARG, "left and right extractor differ in image size, levels or device");
  if (!(bf > 0.f) || !(fx > 0.f)) return fail(ORBX_ERR_BAD_ARG, "bf and fx must be positive");
  if ((u_right || depth) && cap < eL->kp_cap) return fail(ORBX_ERR_CAPACITY, "cap smaller than orbx_max_keypoints()");
  cudaStream_t st = (cudaStream_t)stream;
  DeviceGuard guard(eL->device);
  StereoArgs a;
  a.L = eL->L; a.R = eR->L; a.slotL0 = slotL0; a.slotR0 = slotR0; a.n = n;
  a.bf = bf;
  const float mb = bf / fx;     // Frame::mb (hazard H4)
  a.maxD = bf / mb;             // src/Frame.cc:498
  const bool direct = dst_on_device && u_right && depth && n_matches;
  a.uRight = direct ? u_right : eL->d_uright;
  a.depth = direct ? depth : eL->d_depth;
  a.nMatches = direct ? n_matches : eL->d_nmatch;
  a.sad = eL->d_sad;
  a.rec = eR->d_strec; a.tab = eR->d_sttab; a.tab_stride = eR->st_tab_stride;   // indexed by the right handle's kp_cap
  a.cap = direct ? cap : eL->kp_cap;
  cudaEvent_t* ev = (eL->profiling && eL->n_st < orbx_extractor::kRing) ? eL->ev_st[eL->n_st++] : nullptr;
  if (ev) cudaEventRecord(ev[0], st);
  g_launches += launch_stereo(a, st, ev ? ev[1] : nullptr);
  if (ev) cudaEventRecord(ev[2], st);
  CU(cudaGetLastError());
  if (!direct) {
    const cudaMemcpyKind kind = dst_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    const size_t kc = (size_t)eL->kp_cap;
    if (u_right) CU(cudaMemcpy2DAsync(u_right, (size_t)cap * 4, eL->d_uright, kc * 4, kc * 4, n, kind, st));
    if (depth) CU(cudaMemcpy2DAsync(depth, (size_t)cap * 4, eL->d_depth, kc * 4, kc * 4, n, kind, st));
    if (n_matches) CU(cudaMemcpyAsync(n_matches, eL->d_nmatch, (size_t)n * sizeof(int), kind, st));
  }
  return ORBX_OK;
}

int orbx_hamming256(const uint8_t* a, const uint8_t* b) {
  int d = 0;
  for (int i = 0; i < 4; ++i) {
    unsigned long long u, v;
    std::memcpy(&u, a + 8 * i, 8);
    std::memcpy(&v, b + 8 * i, 8);
    d += __builtin_popcountll(u ^ v);
  }
  return d;
}

}  // extern "C"

// Staging for the matcher entry points called with HOST arrays (on_device == 0): every input is packed into one pinned host
// buffer and reaches the device in ONE copy, scratch and outputs come out of a device buffer with a pinned mirror, results
// return through that mirror.  Both buffers belong to the calling thread, grow to the largest call it has made and are
// then reused: in steady state a call allocates nothing and issues one H2D copy, the kernels, and one small D2H copy per
// output array.  (A SLAM thread calls these members once per frame on a few hundred KB; per-array cudaMallocAsync +
// pageable copies cost more than the kernels.)  With device arrays (on_device == 1) inputs and outputs are used in place
// and only scratch comes from the buffer.
namespace {
struct Arena {                       // a list of blocks: pointers handed out stay valid when more space is needed mid-call
  struct Block { uint8_t* h = nullptr; uint8_t* d = nullptr; size_t cap = 0, used = 0, sent = 0; };
  std::vector<Block> blocks;
  int device = -1;
  void release() {
    for (Block& b : blocks) { if (b.h) cudaFreeHost(b.h); if (b.d) cudaFree(b.d); }
    blocks.clear();
  }
  ~Arena() { release(); }
  // bytes from the current block (a new one when it is full); returns false on allocation failure
  bool take(size_t bytes, uint8_t** h, uint8_t** d, size_t* block_index) {
    bytes = (bytes + 255) & ~(size_t)255;
    if (blocks.empty() || blocks.back().used + bytes > blocks.back().cap) {
      Block b;
      size_t total = 0;
      for (const Block& x : blocks) total += x.cap;
      b.cap = std::max<size_t>(std::max<size_t>(bytes, total), (size_t)1 << 20);      // at least doubles the arena
      if (cudaMallocHost((void**)&b.h, b.cap) != cudaSuccess) return false;
      if (cudaMalloc((void**)&b.d, b.cap) != cudaSuccess) { cudaFreeHost(b.h); return false; }
      blocks.push_back(b);
    }
    Block& b = blocks.back();
    *h = b.h + b.used; *d = b.d + b.used; *block_index = blocks.size() - 1;
    b.used += bytes;
    return true;
  }
  // after a call: one block as large as everything the call needed, so that the next call of this size finds room at once
  void reset() {
    if (blocks.size() > 1) {
      size_t total = 0;
      for (const Block& x : blocks) total += x.cap;
      release();
      Block b; b.cap = total;
      if (cudaMallocHost((void**)&b.h, b.cap) == cudaSuccess) {
        if (cudaMalloc((void**)&b.d, b.cap) == cudaSuccess) blocks.push_back(b); else cudaFreeHost(b.h);
      }
    }
    for (Block& b : blocks) b.used = b.sent = 0;
  }
};
thread_local Arena g_in_arena, g_out_arena;
// With device arrays a call returns before its kernels have run; the next call of this thread hands out the same scratch.
// An event recorded at the end of every call and waited for at the start of the next orders the two, whatever their streams.
thread_local cudaEvent_t g_arena_event = nullptr;

struct Staged {
  cudaStream_t st;
  bool trace = false;                                                    // ORBX_TRACE_CALLS=1: host-side timing of the call on stderr (experiments)
  std::chrono::steady_clock::time_point t0;
  double since() const { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count(); }
  explicit Staged(cudaStream_t s) : st(s) {
    if (getenv("ORBX_TRACE_CALLS")) { trace = true; t0 = std::chrono::steady_clock::now(); }
    int dev = 0;
    cudaGetDevice(&dev);
    if (g_in_arena.device != dev && g_arena_event) { cudaEventDestroy(g_arena_event); g_arena_event = nullptr; }
    for (Arena* a : {&g_in_arena, &g_out_arena}) {
      if (a->device != dev) { a->release(); a->device = dev; }
      a->reset();
    }
    if (g_arena_event) cudaStreamWaitEvent(st, g_arena_event, 0);
  }
  ~Staged() {
    if (!g_arena_event) cudaEventCreateWithFlags(&g_arena_event, cudaEventDisableTiming);
    if (g_arena_event) cudaEventRecord(g_arena_event, st);
    if (trace) std::fprintf(stderr, "orbx call: end %.1f us\n", since());
  }
  // device scratch / output space
  cudaError_t alloc(void** p, size_t bytes) {
    uint8_t *h, *d; size_t bi;
    if (!g_out_arena.take(std::max<size_t>(bytes, 16), &h, &d, &bi)) return cudaErrorMemoryAllocation;
    *p = d;
    return cudaSuccess;
  }
  template <typename T>
  cudaError_t in(const T* host, size_t count, int on_device, const T** out) {
    if (on_device) { *out = host; return cudaSuccess; }
    uint8_t *h, *d; size_t bi;
    if (!g_in_arena.take(std::max<size_t>(count * sizeof(T), 16), &h, &d, &bi)) return cudaErrorMemoryAllocation;
    std::memcpy(h, host, count * sizeof(T));
    *out = reinterpret_cast<const T*>(d);
    return cudaSuccess;
  }
  // the inputs staged since the last upload go to the device: one copy (per arena block).  Called before every launch.
  cudaError_t upload() {
    if (trace) std::fprintf(stderr, "orbx call: staged %.1f us", since());
    for (Arena::Block& b : g_in_arena.blocks)
      if (b.used > b.sent) {
        const cudaError_t e = cudaMemcpyAsync(b.d + b.sent, b.h + b.sent, b.used - b.sent, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return e;
        b.sent = b.used;
      }
    return cudaSuccess;
  }
  // device -> caller's host array: through the pinned mirror when `dev` lies in one of this thread's buffers (the copy is
  // issued by sync(), merged with the call's other downloads from the same block: results that sit side by side in the
  // arena come back in ONE copy; the bytes reach `user` in sync()), else a plain asynchronous copy
  struct Pending { void* user; const uint8_t* mirror; size_t bytes; const Arena::Block* block; size_t off; };
  std::vector<Pending> pending;
  cudaError_t d2h(void* user, const void* dev, size_t bytes) {
    for (Arena* a : {&g_out_arena, &g_in_arena})
      for (const Arena::Block& b : a->blocks)
        if ((const uint8_t*)dev >= b.d && (const uint8_t*)dev + bytes <= b.d + b.cap) {
          const size_t off = (size_t)((const uint8_t*)dev - b.d);
          pending.push_back({user, b.h + off, bytes, &b, off});
          return cudaSuccess;
        }
    return cudaMemcpyAsync(user, dev, bytes, cudaMemcpyDeviceToHost, st);
  }
  cudaError_t sync() {
    // one copy per arena block that has downloads pending: the span from the first to the last wanted byte, unless the
    // span is mostly bytes nobody asked for
    for (size_t i = 0; i < pending.size(); ++i) {
      const Arena::Block* b = pending[i].block;
      bool first = true;
      for (size_t k = 0; k < i; ++k) first = first && pending[k].block != b;
      if (!first) continue;
      size_t lo = (size_t)-1, hi = 0, want = 0;
      for (const Pending& q : pending)
        if (q.block == b) { lo = std::min(lo, q.off); hi = std::max(hi, q.off + q.bytes); want += q.bytes; }
      if (hi - lo <= 2 * want + (64u << 10)) {
        const cudaError_t e = cudaMemcpyAsync(b->h + lo, b->d + lo, hi - lo, cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return e;
      } else {
        for (const Pending& q : pending)
          if (q.block == b) {
            const cudaError_t e = cudaMemcpyAsync(b->h + q.off, b->d + q.off, q.bytes, cudaMemcpyDeviceToHost, st);
            if (e != cudaSuccess) return e;
          }
      }
    }
    if (trace) std::fprintf(stderr, ", issued %.1f us", since());
    const cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    if (trace) std::fprintf(stderr, ", synchronised %.1f us, ", since());
    for (const Pending& p : pending) std::memcpy(p.user, p.mirror, p.bytes);
    pending.clear();
    return cudaSuccess;
  }
};
}  // namespace

extern "C" {

orbx_status orbx_search_init_batch(const orbx_keypoint* kps1, const uint8_t* desc1, const int* n1, const orbx_keypoint* kps2,
                                   const uint8_t* desc2, const int* n2, int n_pairs, int cap, const float* grid_bounds, float* prev_xy,
                                   float nnratio, int window, int check_orientation, int* matches12, int* n_matches,
                                   int on_device, void* stream) {
  if (!kps1 || !desc1 || !n1 || !kps2 || !desc2 || !n2 || !grid_bounds || !prev_xy || !matches12 || !n_matches)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_pairs < 1 || cap < 1 || cap > 65535 || window < 0) return fail(ORBX_ERR_BAD_ARG, "bad sizes");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t np = (size_t)n_pairs, pc = np * cap;
  InitMatchArgs a;
  CU(sg.in(kps1, pc, on_device, &a.k1)); CU(sg.in(desc1, pc * 32, on_device, &a.d1)); CU(sg.in(n1, np, on_device, &a.n1));
  CU(sg.in(kps2, pc, on_device, &a.k2)); CU(sg.in(desc2, pc * 32, on_device, &a.d2)); CU(sg.in(n2, np, on_device, &a.n2));
  const float* prev_in = nullptr;
  CU(sg.in((const float*)prev_xy, pc * 2, on_device, &prev_in));
  a.prev = const_cast<float*>(prev_in);
  if (on_device) { a.m12 = matches12; a.nm = n_matches; }
  else { CU(sg.alloc((void**)&a.m12, pc * 4)); CU(sg.alloc((void**)&a.nm, np * 4)); }
  CU(sg.alloc((void**)&a.scratch, pc * 8 * sizeof(int)));   // one-warp form: 4 int arrays per pair; 1024-thread form: eight list entries per query
  a.n_pairs = n_pairs; a.cap = cap;
  CU(sg.in(grid_bounds, 6, on_device, &a.gb));                  // Frame 2's bounds and inverse cell sizes, src/Frame.cc:207-213
  a.ratio = nnratio; a.window = (float)window; a.checkOri = check_orientation;
  CU(sg.upload());
  g_launches += launch_search_init(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(matches12, a.m12, pc * 4));
    CU(sg.d2h(n_matches, a.nm, np * 4));
    CU(sg.d2h(prev_xy, a.prev, pc * 8));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_knn_ratio_batch(const uint8_t* descA, const float* angA, const int* nA, const uint8_t* descB, const float* angB,
                                 const int* nB, int n_pairs, int cap, int th, float nnratio, int check_orientation, int strict_lt,
                                 int* matches12, int* n_matches, int on_device, void* stream) {
  if (!descA || !angA || !nA || !descB || !angB || !nB || !matches12 || !n_matches) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_pairs < 1 || cap < 1 || cap > 6000) return fail(ORBX_ERR_BAD_ARG, "cap must be 1..6000 (B descriptors are held in shared memory)");
  // a 256-bit distance is at most 256 and "no candidate" is 256 too: th >= 256 would accept a match with no feature
  // (the reference's bestIdx2 = -1 write, src/ORBmatcher.cc:596-600, unreachable there because TH_LOW is 50)
  if (th < 0 || th > 255) return fail(ORBX_ERR_BAD_ARG, "th must be 0..255");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t np = (size_t)n_pairs, pc = np * cap;
  KnnArgs a;
  CU(sg.in(descA, pc * 32, on_device, &a.dA)); CU(sg.in(angA, pc, on_device, &a.angA)); CU(sg.in(nA, np, on_device, &a.nA));
  CU(sg.in(descB, pc * 32, on_device, &a.dB)); CU(sg.in(angB, pc, on_device, &a.angB)); CU(sg.in(nB, np, on_device, &a.nB));
  if (on_device) { a.m12 = matches12; a.nm = n_matches; }
  else { CU(sg.alloc((void**)&a.m12, pc * 4)); CU(sg.alloc((void**)&a.nm, np * 4)); }
  CU(sg.alloc((void**)&a.topk, pc * 8 * sizeof(uint32_t)));
  CU(sg.alloc((void**)&a.done, np * sizeof(int))); CU(cudaMemsetAsync(a.done, 0, np * sizeof(int), st));
  a.n_pairs = n_pairs; a.cap = cap; a.th = th; a.ratio = nnratio; a.checkOri = check_orientation; a.strict = strict_lt;
  CU(sg.upload());
  g_launches += launch_knn_ratio(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(matches12, a.m12, pc * 4));
    CU(sg.d2h(n_matches, a.nm, np * 4));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_search_by_bow_batch(const uint8_t* descA, const float* angA, const uint8_t* validA, const int* nodeA, const int* nA,
                                     const uint8_t* descB, const float* angB, const uint8_t* validB, const int* nodeB, const int* nB,
                                     int n_pairs, int cap, int th, float nnratio, int check_orientation, int strict_lt, int* matches12,
                                     int* n_matches, int on_device, void* stream) {
  if (!descA || !angA || !validA || !nodeA || !nA || !descB || !angB || !validB || !nodeB || !nB || !matches12 || !n_matches)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_pairs < 1 || cap < 1 || cap > 4096) return fail(ORBX_ERR_BAD_ARG, "cap must be 1..4096 (both sort arrays and the B descriptors are held in shared memory)");
  if (th < 0 || th > 255) return fail(ORBX_ERR_BAD_ARG, "th must be 0..255");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t np = (size_t)n_pairs, pc = np * cap;
  BowArgs a;
  CU(sg.in(descA, pc * 32, on_device, &a.dA)); CU(sg.in(angA, pc, on_device, &a.angA)); CU(sg.in(validA, pc, on_device, &a.validA));
  CU(sg.in(nodeA, pc, on_device, &a.nodeA)); CU(sg.in(nA, np, on_device, &a.nA));
  CU(sg.in(descB, pc * 32, on_device, &a.dB)); CU(sg.in(angB, pc, on_device, &a.angB)); CU(sg.in(validB, pc, on_device, &a.validB));
  CU(sg.in(nodeB, pc, on_device, &a.nodeB)); CU(sg.in(nB, np, on_device, &a.nB));
  if (on_device) { a.m12 = matches12; a.nm = n_matches; }
  else { CU(sg.alloc((void**)&a.m12, pc * 4)); CU(sg.alloc((void**)&a.nm, np * 4)); }
  a.n_pairs = n_pairs; a.cap = cap; a.th = th; a.ratio = nnratio; a.checkOri = check_orientation; a.strict = strict_lt;
  CU(sg.alloc((void**)&a.spec, pc * sizeof(uint4)));
  CU(sg.upload());
  g_launches += launch_search_by_bow(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(matches12, a.m12, pc * 4));
    CU(sg.d2h(n_matches, a.nm, np * 4));
    CU(sg.sync());
  }
  return ORBX_OK;
}

struct orbx_vocabulary {
  int device = 0, n_nodes = 0, n_words = 0, L = 0, weighting = 0, norm = 0;
  uint8_t* desc = nullptr; int* child_first = nullptr; int* child_count = nullptr; int* children = nullptr; int* word = nullptr;
  double* weight = nullptr; double* wweight = nullptr;
};

void orbx_vocabulary_destroy(orbx_vocabulary* v) {
  if (!v) return;
  DeviceGuard g(v->device);
  void* ptrs[] = {v->desc, v->child_first, v->child_count, v->children, v->word, v->weight, v->wweight};
  for (void* p : ptrs) if (p) cudaFree(p);
  delete v;
}

orbx_status orbx_vocabulary_create(const uint8_t* node_desc, const int* child_first, const int* child_count, const int* children,
                                   int n_children, const int* node_word, const double* node_weight, int n_nodes, int L, int weighting,
                                   int norm, int device, orbx_vocabulary** out) {
  if (!node_desc || !child_first || !child_count || !children || !node_word || !node_weight || !out) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_nodes < 2 || n_children < 1 || L < 1 || L > 10 || weighting < 0 || weighting > 3 || norm < 0 || norm > 2)
    return fail(ORBX_ERR_BAD_ARG, "bad vocabulary sizes (L 1..10, weighting 0..3, norm 0..2)");
  // the tree must be walkable: children ranges inside the list, child ids inside the node table, the root not a leaf,
  // every leaf carrying a word id; the depth bound keeps a malformed (cyclic) table from hanging the kernel
  int n_words = 0;
  for (int i = 0; i < n_nodes; ++i) {
    if (child_count[i] < 0 || child_first[i] < 0 || (long long)child_first[i] + child_count[i] > n_children) return fail(ORBX_ERR_BAD_ARG, "children range outside the list");
    if (child_count[i] == 0) { if (node_word[i] < 0) return fail(ORBX_ERR_BAD_ARG, "leaf without a word id"); n_words = std::max(n_words, node_word[i] + 1); }
  }
  if (child_count[0] == 0) return fail(ORBX_ERR_BAD_ARG, "the root has no children");
  for (int c = 0; c < n_children; ++c)
    if (children[c] <= 0 || children[c] >= n_nodes) return fail(ORBX_ERR_BAD_ARG, "child id outside the node table");
  {
    std::vector<int> depth(n_nodes, -1);
    std::vector<int> stack{0};
    depth[0] = 0;
    while (!stack.empty()) {
      const int u = stack.back(); stack.pop_back();
      for (int c = 0; c < child_count[u]; ++c) {
        const int w = children[child_first[u] + c];
        if (depth[w] >= 0) return fail(ORBX_ERR_BAD_ARG, "a node has two parents");
        depth[w] = depth[u] + 1;
        if (depth[w] > 16) return fail(ORBX_ERR_BAD_ARG, "tree deeper than 16 levels");
        stack.push_back(w);
      }
    }
  }
  std::vector<double> wweight(n_words, 0.0);
  for (int i = 0; i < n_nodes; ++i) if (child_count[i] == 0) wweight[node_word[i]] = node_weight[i];
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(ORBX_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(ORBX_ERR_BAD_ARG, "bad device index");
  DeviceGuard g(device);
  orbx_vocabulary* v = new orbx_vocabulary;
  v->device = device; v->n_nodes = n_nodes; v->n_words = n_words; v->L = L; v->weighting = weighting; v->norm = norm;
  auto up = [&](auto** dst, const auto* src, size_t count) -> bool {
    using T = std::remove_pointer_t<std::remove_pointer_t<decltype(dst)>>;
    if (cudaMalloc((void**)dst, std::max<size_t>(count, 1) * sizeof(T)) != cudaSuccess) return false;
    return cudaMemcpy(*dst, src, count * sizeof(T), cudaMemcpyHostToDevice) == cudaSuccess;
  };
  const bool ok = up(&v->desc, node_desc, (size_t)n_nodes * 32) && up(&v->child_first, child_first, (size_t)n_nodes) &&
                  up(&v->child_count, child_count, (size_t)n_nodes) && up(&v->children, children, (size_t)n_children) &&
                  up(&v->word, node_word, (size_t)n_nodes) && up(&v->weight, node_weight, (size_t)n_nodes) &&
                  up(&v->wweight, wweight.data(), (size_t)n_words);
  if (!ok) { orbx_vocabulary_destroy(v); return fail(ORBX_ERR_CUDA, "vocabulary upload failed"); }
  *out = v;
  return ORBX_OK;
}

orbx_status orbx_bow_transform_batch(const orbx_vocabulary* voc, const uint8_t* desc, const int* n, int n_frames, int cap, int levelsup,
                                     int* word_of, int* node_of, int* bow_word, double* bow_val, int* bow_count, int on_device,
                                     void* stream) {
  if (!voc || !desc || !n || !word_of || !node_of || !bow_word || !bow_val || !bow_count) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_frames < 1 || cap < 1 || cap > 8192 || levelsup < 0) return fail(ORBX_ERR_BAD_ARG, "bad sizes (cap <= 8192)");
  DeviceGuard g(voc->device);
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t nf = (size_t)n_frames, fc = nf * cap;
  BowTrArgs a;
  a.v_desc = voc->desc; a.v_child_first = voc->child_first; a.v_child_count = voc->child_count; a.v_children = voc->children;
  a.v_word = voc->word; a.v_weight = voc->weight; a.v_wweight = voc->wweight;
  a.L = voc->L; a.levelsup = levelsup; a.weighting = voc->weighting; a.norm = voc->norm;
  CU(sg.in(desc, fc * 32, on_device, &a.desc)); CU(sg.in(n, nf, on_device, &a.n));
  if (on_device) { a.word_of = word_of; a.node_of = node_of; a.bow_word = bow_word; a.bow_val = bow_val; a.bow_count = bow_count; }
  else {
    CU(sg.alloc((void**)&a.word_of, fc * 4)); CU(sg.alloc((void**)&a.node_of, fc * 4)); CU(sg.alloc((void**)&a.bow_word, fc * 4));
    CU(sg.alloc((void**)&a.bow_val, fc * 8)); CU(sg.alloc((void**)&a.bow_count, nf * 4));
  }
  a.n_frames = n_frames; a.cap = cap;
  CU(sg.upload());
  g_launches += launch_bow_transform(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(word_of, a.word_of, fc * 4)); CU(sg.d2h(node_of, a.node_of, fc * 4));
    CU(sg.d2h(bow_word, a.bow_word, fc * 4)); CU(sg.d2h(bow_val, a.bow_val, fc * 8));
    CU(sg.d2h(bow_count, a.bow_count, nf * 4));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_search_triangulation_batch(const orbx_keypoint* kps1, const uint8_t* desc1, const uint8_t* has_mp1, const float* u_right1,
                                            const int* node1, const int* n1, const orbx_keypoint* kps2, const uint8_t* desc2,
                                            const uint8_t* has_mp2, const float* u_right2, const int* node2, const int* n2, int n_pairs,
                                            int cap, const float* F12, const float* pose2, const float* Cw, const float* cam2,
                                            const float* scale_factors2, const float* level_sigma2_2, int nlevels, int only_stereo,
                                            int check_orientation, int* matches12, int* n_matches, int on_device, void* stream) {
  if (!kps1 || !desc1 || !has_mp1 || !u_right1 || !node1 || !n1 || !kps2 || !desc2 || !has_mp2 || !u_right2 || !node2 || !n2 || !F12 ||
      !pose2 || !Cw || !cam2 || !scale_factors2 || !level_sigma2_2 || !matches12 || !n_matches)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_pairs < 1 || cap < 1 || cap > 4096 || nlevels < 1 || nlevels > kMaxLevels) return fail(ORBX_ERR_BAD_ARG, "bad sizes (cap <= 4096, nlevels <= 8)");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t np = (size_t)n_pairs, pc = np * cap;
  TriArgs a;
  CU(sg.in(kps1, pc, on_device, &a.k1)); CU(sg.in(desc1, pc * 32, on_device, &a.d1)); CU(sg.in(has_mp1, pc, on_device, &a.hasMp1));
  CU(sg.in(u_right1, pc, on_device, &a.uR1)); CU(sg.in(node1, pc, on_device, &a.node1)); CU(sg.in(n1, np, on_device, &a.n1));
  CU(sg.in(kps2, pc, on_device, &a.k2)); CU(sg.in(desc2, pc * 32, on_device, &a.d2)); CU(sg.in(has_mp2, pc, on_device, &a.hasMp2));
  CU(sg.in(u_right2, pc, on_device, &a.uR2)); CU(sg.in(node2, pc, on_device, &a.node2)); CU(sg.in(n2, np, on_device, &a.n2));
  CU(sg.in(F12, np * 9, on_device, &a.F12)); CU(sg.in(pose2, np * 12, on_device, &a.pose2)); CU(sg.in(Cw, np * 3, on_device, &a.Cw));
  CU(sg.in(cam2, np * 4, on_device, &a.cam2));
  if (on_device) {
    CU(sg.d2h(a.scale2, scale_factors2, nlevels * sizeof(float)));
    CU(sg.d2h(a.sigma2_2, level_sigma2_2, nlevels * sizeof(float)));
    CU(sg.sync());
  } else {
    std::memcpy(a.scale2, scale_factors2, nlevels * sizeof(float)); std::memcpy(a.sigma2_2, level_sigma2_2, nlevels * sizeof(float));
  }
  for (int l = nlevels; l < kMaxLevels; ++l) { a.scale2[l] = a.scale2[nlevels - 1]; a.sigma2_2[l] = a.sigma2_2[nlevels - 1]; }
  if (on_device) { a.m12 = matches12; a.nm = n_matches; }
  else { CU(sg.alloc((void**)&a.m12, pc * 4)); CU(sg.alloc((void**)&a.nm, np * 4)); }
  a.n_pairs = n_pairs; a.cap = cap; a.onlyStereo = only_stereo; a.checkOri = check_orientation;
  CU(sg.alloc((void**)&a.hist, np * 32 * sizeof(int))); CU(sg.alloc((void**)&a.bins, pc));
  CU(cudaMemsetAsync(a.hist, 0, np * 32 * sizeof(int), st)); CU(cudaMemsetAsync(a.nm, 0, np * sizeof(int), st));
  CU(sg.upload());
  g_launches += launch_search_triangulation(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(matches12, a.m12, pc * 4));
    CU(sg.d2h(n_matches, a.nm, np * 4));
    CU(sg.sync());
  }
  return ORBX_OK;
}

namespace {
// scale tables arrive on the host or on the device; the kernels take them by value
cudaError_t fetch_levels(float* dst, const float* src, int nlevels, int on_device, cudaStream_t st) {
  for (int l = 0; l < kMaxLevels; ++l) dst[l] = 1.f;
  if (!on_device) { std::memcpy(dst, src, nlevels * sizeof(float)); return cudaSuccess; }
  cudaError_t e = cudaMemcpyAsync(dst, src, nlevels * sizeof(float), cudaMemcpyDeviceToHost, st);
  return e != cudaSuccess ? e : cudaStreamSynchronize(st);
}
}  // namespace

}  // extern "C"

// The CSR feature grid of the projection searches: the caller's (K10's output), or none - the 1024-thread kernels build it in
// shared memory from the keypoints.  When the small kernels will run (large capacities), host callers get it built here with
// Frame::PosInGrid's rule (src/Frame.cc:382-392); device callers must pass it then.
template <class SG>
static orbx_status stage_grid(SG& sg, const int* cell_start, const int* cell_items, const orbx_keypoint* kps, const int* n, size_t np, int cap, const float* grid_bounds,
                              int on_device, bool needs_grid, const int** d_cs, const int** d_items) {
  const size_t cells = np * (ORBX_GRID_COLS * ORBX_GRID_ROWS + 1), kc = np * cap;
  if ((cell_start == nullptr) != (cell_items == nullptr)) return fail(ORBX_ERR_BAD_ARG, "cell_start and cell_items go together");
  if (cell_start) {
    if (sg.in(cell_start, cells, on_device, d_cs) != cudaSuccess || sg.in(cell_items, kc, on_device, d_items) != cudaSuccess) return fail(ORBX_ERR_CUDA, "staging the grid");
    return ORBX_OK;
  }
  *d_cs = nullptr; *d_items = nullptr;
  if (!needs_grid) return ORBX_OK;
  if (on_device) return fail(ORBX_ERR_BAD_ARG, "cell_start / cell_items are required for device arrays of this capacity");
  std::vector<int> cs(cells, 0), items(kc, 0);
  const int nc = ORBX_GRID_COLS * ORBX_GRID_ROWS;
  for (size_t p = 0; p < np; ++p) {
    int* c0 = cs.data() + p * (nc + 1);
    const orbx_keypoint* k = kps + p * cap;
    const int np_ = std::min(n[p], cap);
    std::vector<int> cell(np_, -1);
    for (int i = 0; i < np_; ++i) {
      const int gx = (int)roundf((k[i].x - grid_bounds[0]) * grid_bounds[4]), gy = (int)roundf((k[i].y - grid_bounds[2]) * grid_bounds[5]);
      if (gx < 0 || gx >= ORBX_GRID_COLS || gy < 0 || gy >= ORBX_GRID_ROWS) continue;
      cell[i] = gx * ORBX_GRID_ROWS + gy; ++c0[cell[i] + 1];
    }
    for (int c = 1; c <= nc; ++c) c0[c] += c0[c - 1];
    std::vector<int> cur(c0, c0 + nc);
    for (int i = 0; i < np_; ++i) if (cell[i] >= 0) items[p * cap + cur[cell[i]]++] = i;
  }
  if (sg.in(cs.data(), cells, 0, d_cs) != cudaSuccess || sg.in(items.data(), kc, 0, d_items) != cudaSuccess) return fail(ORBX_ERR_CUDA, "staging the grid");
  return ORBX_OK;
}

extern "C" {

orbx_status orbx_fuse_search_batch(const orbx_keypoint* kps_un, const uint8_t* desc, const float* u_right, const int* n, int n_kfs, int cap,
                                   const float* grid_bounds, const int* cell_start, const int* cell_items, const float* kf_bounds,
                                   const float* pose, const float* Ow, float fx, float fy, float cx, float cy, float bf,
                                   const float* scale_factors, const float* inv_level_sigma2, int nlevels, float log_scale_factor,
                                   const uint8_t* mp_valid, const float* mp_xw, const float* mp_normal, const float* mp_min_dist,
                                   const float* mp_max_dist, const uint8_t* mp_desc, const int* m, int mcap, float th, int sim3_variant,
                                   int* best_idx, int on_device, void* stream) {
  if (!kps_un || !desc || !u_right || !n || !grid_bounds || !kf_bounds || !pose || !Ow || !scale_factors ||
      !inv_level_sigma2 || !mp_valid || !mp_xw || !mp_normal || !mp_min_dist || !mp_max_dist || !mp_desc || !m || !best_idx)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_kfs < 1 || n_kfs > 65535 || cap < 1 || mcap < 1 || nlevels < 1 || nlevels > kMaxLevels || !(log_scale_factor > 0.f))
    return fail(ORBX_ERR_BAD_ARG, "bad sizes (n_kfs <= 65535, nlevels <= 8, log_scale_factor > 0)");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t np = (size_t)n_kfs, kc = np * cap, mc = np * mcap, cells = np * (ORBX_GRID_COLS * ORBX_GRID_ROWS + 1);
  ProjSearchArgs a;
  CU(sg.in(kps_un, kc, on_device, &a.kps)); CU(sg.in(desc, kc * 32, on_device, &a.desc)); CU(sg.in(u_right, kc, on_device, &a.u_right));
  CU(sg.in(grid_bounds, 6, on_device, &a.grid_bounds));
  if (stage_grid(sg, cell_start, cell_items, kps_un, n, np, cap, grid_bounds, on_device, project_search_needs_grid(cap, true), &a.cell_start, &a.cell_items) != ORBX_OK) return ORBX_ERR_BAD_ARG;
  CU(sg.in(n, np, on_device, &a.n));
  CU(sg.in(kf_bounds, np * 4, on_device, &a.kf_bounds));
  // xf per problem: R1 t1 | R2 t2 (unused) | Ow
  if (on_device) {
    float* d_xf = nullptr;
    CU(sg.alloc((void**)&d_xf, np * 27 * sizeof(float)));
    CU(cudaMemsetAsync(d_xf, 0, np * 27 * sizeof(float), st));
    CU(cudaMemcpy2DAsync(d_xf, 27 * sizeof(float), pose, 12 * sizeof(float), 12 * sizeof(float), np, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpy2DAsync(d_xf + 24, 27 * sizeof(float), Ow, 3 * sizeof(float), 3 * sizeof(float), np, cudaMemcpyDeviceToDevice, st));
    a.xf = d_xf;
  } else {                                                               // put together on the host: it rides in the call's one upload
    std::vector<float> hx(np * 27, 0.f);
    for (size_t q = 0; q < np; ++q) { std::memcpy(&hx[q * 27], pose + q * 12, 12 * sizeof(float)); std::memcpy(&hx[q * 27 + 24], Ow + q * 3, 3 * sizeof(float)); }
    CU(sg.in(hx.data(), np * 27, 0, &a.xf));
  }
  CU(sg.in(mp_valid, mc, on_device, &a.valid)); CU(sg.in(mp_xw, mc * 3, on_device, &a.xw)); CU(sg.in(mp_normal, mc * 3, on_device, &a.normal));
  CU(sg.in(mp_min_dist, mc, on_device, &a.dmin)); CU(sg.in(mp_max_dist, mc, on_device, &a.dmax)); CU(sg.in(mp_desc, mc * 32, on_device, &a.mp_desc));
  CU(sg.in(m, np, on_device, &a.m));
  if (on_device) a.best = best_idx; else CU(sg.alloc((void**)&a.best, mc * 4));
  CU(fetch_levels(a.scale, scale_factors, nlevels, on_device, st)); CU(fetch_levels(a.inv_sigma2, inv_level_sigma2, nlevels, on_device, st));
  a.n_problems = n_kfs; a.cap = cap; a.mcap = mcap; a.nlevels = nlevels; a.mode = sim3_variant ? (2 | 4 | 16) : (2 | 4 | 8); a.th_dist = 50;   // TH_LOW; :1019 divides 1.0 by z in double, :862 in float
  a.fx = fx; a.fy = fy; a.cx = cx; a.cy = cy; a.bf = bf; a.log_scale = log_scale_factor; a.th = th;
  CU(sg.upload());
  g_launches += launch_project_search(a, st);
  CU(cudaGetLastError());
  if (!on_device) { CU(sg.d2h(best_idx, a.best, mc * 4)); CU(sg.sync()); }
  return ORBX_OK;
}

orbx_status orbx_search_by_sim3_batch(const orbx_keypoint* kps1, const uint8_t* desc1, const int* n1, const int* cell_start1, const int* cell_items1,
                                      const float* kf_bounds1, const float* pose1, const orbx_keypoint* kps2, const uint8_t* desc2, const int* n2,
                                      const int* cell_start2, const int* cell_items2, const float* kf_bounds2, const float* pose2, int n_pairs,
                                      int cap, const float* grid_bounds, float fx, float fy, float cx, float cy, const float* scale_factors,
                                      int nlevels, float log_scale_factor, const float* xf, const int* mp1_index, const float* mp1_xw,
                                      const float* mp1_min_dist, const float* mp1_max_dist, const uint8_t* mp1_desc, const int* m1,
                                      const int* mp2_index, const float* mp2_xw, const float* mp2_min_dist, const float* mp2_max_dist,
                                      const uint8_t* mp2_desc, const int* m2, int mcap, float th, int* matches12, int* n_found, int on_device,
                                      void* stream) {
  if (!kps1 || !desc1 || !n1 || !kf_bounds1 || !pose1 || !kps2 || !desc2 || !n2 ||
      !kf_bounds2 || !pose2 || !grid_bounds || !scale_factors || !xf || !mp1_index || !mp1_xw || !mp1_min_dist || !mp1_max_dist || !mp1_desc || !m1 ||
      !mp2_index || !mp2_xw || !mp2_min_dist || !mp2_max_dist || !mp2_desc || !m2 || !matches12 || !n_found)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_pairs < 1 || n_pairs > 65535 || cap < 1 || mcap < 1 || nlevels < 1 || nlevels > kMaxLevels || !(log_scale_factor > 0.f))
    return fail(ORBX_ERR_BAD_ARG, "bad sizes (n_pairs <= 65535, nlevels <= 8, log_scale_factor > 0)");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t np = (size_t)n_pairs, kc = np * cap, mc = np * mcap, cells = np * (ORBX_GRID_COLS * ORBX_GRID_ROWS + 1);
  const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  const float* d_gb; CU(sg.in(grid_bounds, 6, on_device, &d_gb));
  int* d_vn[2]; int* d_best[2]; const int* d_idx[2]; const int* d_m[2];
  const int* d_n1; CU(sg.in(n1, np, on_device, &d_n1));
  CU(sg.alloc((void**)&d_vn[0], (2 * kc + np) * 4)); d_vn[1] = d_vn[0] + kc;
  int* d_ticket = d_vn[0] + 2 * kc;                                      // -1 per pair: shares the 0xFF fill of the two vnMatch tables
  CU(cudaMemsetAsync(d_vn[0], 0xFF, (2 * kc + np) * 4, st));
  // pass 0: map points of keyframe 1 into keyframe 2 (R1w, t1w then sR21, t21); pass 1: the other way (R2w, t2w then sR12, t12).
  // Everything both passes read is staged first and goes up in one copy; then the five launches.
  ProjSearchArgs pa[2];
  for (int pass = 0; pass < 2; ++pass) {
    ProjSearchArgs& a = pa[pass];
    const bool fwd = pass == 0;
    CU(sg.in(fwd ? kps2 : kps1, kc, on_device, &a.kps)); CU(sg.in(fwd ? desc2 : desc1, kc * 32, on_device, &a.desc));
    a.u_right = nullptr; a.grid_bounds = d_gb;
    if (stage_grid(sg, fwd ? cell_start2 : cell_start1, fwd ? cell_items2 : cell_items1, fwd ? kps2 : kps1, fwd ? n2 : n1, np, cap, grid_bounds, on_device,
                   project_search_needs_grid(cap, false), &a.cell_start, &a.cell_items) != ORBX_OK) return ORBX_ERR_BAD_ARG;
    CU(sg.in(fwd ? n2 : n1, np, on_device, &a.n));
    CU(sg.in(fwd ? kf_bounds2 : kf_bounds1, np * 4, on_device, &a.kf_bounds));
    if (on_device) {
      float* d_xf = nullptr;
      CU(sg.alloc((void**)&d_xf, np * 27 * sizeof(float)));
      CU(cudaMemsetAsync(d_xf, 0, np * 27 * sizeof(float), st));
      CU(cudaMemcpy2DAsync(d_xf, 27 * sizeof(float), fwd ? pose1 : pose2, 12 * sizeof(float), 12 * sizeof(float), np, kind, st));
      CU(cudaMemcpy2DAsync(d_xf + 12, 27 * sizeof(float), xf + (fwd ? 12 : 0), 24 * sizeof(float), 12 * sizeof(float), np, kind, st));
      a.xf = d_xf;
    } else {                                                             // the 27 floats per pair put together on the host
      std::vector<float> hx(np * 27, 0.f);
      for (size_t q = 0; q < np; ++q) {
        std::memcpy(&hx[q * 27], (fwd ? pose1 : pose2) + q * 12, 12 * sizeof(float));
        std::memcpy(&hx[q * 27 + 12], xf + q * 24 + (fwd ? 12 : 0), 12 * sizeof(float));
      }
      CU(sg.in(hx.data(), np * 27, 0, &a.xf));
    }
    a.valid = nullptr; a.normal = nullptr;                               // every listed map point is valid; no viewing-angle test in this member
    CU(sg.in(fwd ? mp1_xw : mp2_xw, mc * 3, on_device, &a.xw)); CU(sg.in(fwd ? mp1_min_dist : mp2_min_dist, mc, on_device, &a.dmin));
    CU(sg.in(fwd ? mp1_max_dist : mp2_max_dist, mc, on_device, &a.dmax)); CU(sg.in(fwd ? mp1_desc : mp2_desc, mc * 32, on_device, &a.mp_desc));
    CU(sg.in(fwd ? m1 : m2, np, on_device, &d_m[pass])); a.m = d_m[pass];
    CU(sg.in(fwd ? mp1_index : mp2_index, mc, on_device, &d_idx[pass]));
    CU(sg.alloc((void**)&d_best[pass], mc * 4)); a.best = d_best[pass];
    if (pass == 0) CU(fetch_levels(a.scale, scale_factors, nlevels, on_device, st));
    else std::memcpy(a.scale, pa[0].scale, sizeof(a.scale));
    for (int l = 0; l < kMaxLevels; ++l) a.inv_sigma2[l] = 1.f;
    a.n_problems = n_pairs; a.cap = cap; a.mcap = mcap; a.nlevels = nlevels; a.mode = 1 | 16; a.th_dist = 100;   // TH_HIGH
    a.fx = fx; a.fy = fy; a.cx = cx; a.cy = cy; a.bf = 0.f; a.log_scale = log_scale_factor; a.th = th;
  }
  Sim3AgreeArgs g;
  g.match1 = d_vn[0]; g.match2 = d_vn[1]; g.n1 = d_n1; g.cap = cap;
  if (on_device) { g.matches12 = matches12; g.nfound = n_found; }
  else { CU(sg.alloc((void**)&g.matches12, kc * 4)); CU(sg.alloc((void**)&g.nfound, np * 4)); }
  CU(sg.upload());
  const int fused = launch_sim3_lat(pa[0], pa[1], d_idx[0], d_idx[1], d_vn[0], d_vn[1], d_ticket, d_n1, g.matches12, g.nfound, st);
  g_launches += fused;
  if (!fused) {
    CU(cudaMemsetAsync(g.nfound, 0, np * 4, st));
    for (int pass = 0; pass < 2; ++pass) {
      g_launches += launch_project_search(pa[pass], st);
      g_launches += launch_scatter_matches(d_best[pass], d_idx[pass], d_m[pass], mcap, cap, n_pairs, d_vn[pass], st);
    }
    g_launches += launch_sim3_agree(g, n_pairs, st);
  }
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(matches12, g.matches12, kc * 4));
    CU(sg.d2h(n_found, g.nfound, np * 4));
    CU(sg.sync());
  }
  (void)n2;
  return ORBX_OK;
}

orbx_status orbx_search_projection_ordered_batch(const orbx_keypoint* kps_un, const uint8_t* desc, const int* n, int n_problems, int cap,
                                                 const float* grid_bounds, const int* cell_start, const int* cell_items,
                                                 const float* target_bounds, const float* pose, const float* Ow, float fx, float fy, float cx,
                                                 float cy, const float* scale_factors, int nlevels, float log_scale_factor,
                                                 const uint8_t* occupied, const uint8_t* mp_valid, const float* mp_xw, const float* mp_normal,
                                                 const float* mp_min_dist, const float* mp_max_dist, const uint8_t* mp_desc,
                                                 const float* mp_src_angle, const int* m, int mcap, float th, int max_dist, int relocalisation,
                                                 int check_orientation, int* assigned, int* n_matches, int on_device, void* stream) {
  if (!kps_un || !desc || !n || !grid_bounds || !target_bounds || !pose || !Ow || !scale_factors || !occupied ||
      !mp_valid || !mp_xw || !mp_normal || !mp_min_dist || !mp_max_dist || !mp_desc || !mp_src_angle || !m || !assigned || !n_matches)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_problems < 1 || cap < 1 || mcap < 1 || nlevels < 1 || nlevels > kMaxLevels || !(log_scale_factor > 0.f) || max_dist < 0 || max_dist > 255)
    return fail(ORBX_ERR_BAD_ARG, "bad sizes (nlevels <= 8, log_scale_factor > 0, max_dist 0..255)");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t np = (size_t)n_problems, kc = np * cap, mc = np * mcap, cells = np * (ORBX_GRID_COLS * ORBX_GRID_ROWS + 1);
  const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  ProjOrderedArgs a;
  CU(sg.in(kps_un, kc, on_device, &a.kps)); CU(sg.in(desc, kc * 32, on_device, &a.desc)); CU(sg.in(n, np, on_device, &a.n));
  CU(sg.in(grid_bounds, 6, on_device, &a.grid_bounds));
  if (stage_grid(sg, cell_start, cell_items, kps_un, n, np, cap, grid_bounds, on_device, project_ordered_needs_grid(cap, mcap), &a.cell_start, &a.cell_items) != ORBX_OK) return ORBX_ERR_BAD_ARG;
  CU(sg.in(target_bounds, np * 4, on_device, &a.tgt_bounds));
  if (on_device) {
    float* d_xf = nullptr;
    CU(sg.alloc((void**)&d_xf, np * 15 * sizeof(float)));
    CU(cudaMemcpy2DAsync(d_xf, 15 * sizeof(float), pose, 12 * sizeof(float), 12 * sizeof(float), np, kind, st));
    CU(cudaMemcpy2DAsync(d_xf + 12, 15 * sizeof(float), Ow, 3 * sizeof(float), 3 * sizeof(float), np, kind, st));
    a.xf = d_xf;
  } else {                                                               // put together on the host: it rides in the call's one upload
    std::vector<float> hx(np * 15, 0.f);
    for (size_t q = 0; q < np; ++q) { std::memcpy(&hx[q * 15], pose + q * 12, 12 * sizeof(float)); std::memcpy(&hx[q * 15 + 12], Ow + q * 3, 3 * sizeof(float)); }
    CU(sg.in(hx.data(), np * 15, 0, &a.xf));
  }
  CU(sg.in(occupied, kc, on_device, &a.occupied));
  CU(sg.in(mp_valid, mc, on_device, &a.valid)); CU(sg.in(mp_xw, mc * 3, on_device, &a.xw)); CU(sg.in(mp_normal, mc * 3, on_device, &a.normal));
  CU(sg.in(mp_min_dist, mc, on_device, &a.dmin)); CU(sg.in(mp_max_dist, mc, on_device, &a.dmax)); CU(sg.in(mp_desc, mc * 32, on_device, &a.mp_desc));
  CU(sg.in(mp_src_angle, mc, on_device, &a.src_angle)); CU(sg.in(m, np, on_device, &a.m));
  if (on_device) { a.assigned = assigned; a.nm = n_matches; }
  else { CU(sg.alloc((void**)&a.assigned, kc * 4)); CU(sg.alloc((void**)&a.nm, np * 4)); }
  CU(sg.alloc((void**)&a.spec, 2 * mc * sizeof(uint4)));
  CU(fetch_levels(a.scale, scale_factors, nlevels, on_device, st));
  a.n_problems = n_problems; a.cap = cap; a.mcap = mcap; a.nlevels = nlevels; a.reloc = relocalisation ? 1 : 0; a.th_dist = max_dist;
  a.checkOri = check_orientation ? 1 : 0; a.fx = fx; a.fy = fy; a.cx = cx; a.cy = cy; a.log_scale = log_scale_factor; a.th = th;
  CU(sg.upload());
  g_launches += launch_project_match_ordered(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(assigned, a.assigned, kc * 4));
    CU(sg.d2h(n_matches, a.nm, np * 4));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_distinctive_descriptors_batch(const uint8_t* desc, const int* counts, int n_points, int kcap, int* best_idx,
                                               uint8_t* best_desc, int on_device, void* stream) {
  if (!desc || !counts || !best_idx) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_points < 1 || kcap < 1 || kcap > 1024) return fail(ORBX_ERR_BAD_ARG, "kcap must be 1..1024 (a map point's descriptors are held in shared memory)");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t np = (size_t)n_points;
  DistinctArgs a;
  CU(sg.in(desc, np * kcap * 32, on_device, &a.desc)); CU(sg.in(counts, np, on_device, &a.counts));
  a.best_desc = nullptr;
  if (on_device) { a.best_idx = best_idx; a.best_desc = best_desc; }
  else {
    CU(sg.alloc((void**)&a.best_idx, np * 4));
    if (best_desc) { CU(sg.alloc((void**)&a.best_desc, np * 32)); CU(cudaMemcpyAsync(a.best_desc, best_desc, np * 32, cudaMemcpyHostToDevice, st)); }
  }
  a.n_points = n_points; a.kcap = kcap;
  CU(sg.upload());
  g_launches += launch_distinctive(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(best_idx, a.best_idx, np * 4));
    if (best_desc) CU(sg.d2h(best_desc, a.best_desc, np * 32));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_undistort_grid_batch(const orbx_keypoint* kps, const int* n, int n_frames, int cap, const orbx_camera* camera,
                                      int cols, int rows, orbx_keypoint* kps_un, float* bounds, int* cell_start, int* cell_items,
                                      int on_device, void* stream) {
  if (!kps || !n || !camera || !kps_un || !bounds || !cell_start || !cell_items) return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_frames < 1 || cap < 1 || cols < 1 || rows < 1) return fail(ORBX_ERR_BAD_ARG, "bad sizes");
  if (!(camera->fx != 0.f) || !(camera->fy != 0.f)) return fail(ORBX_ERR_BAD_ARG, "fx and fy must be non-zero");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t nf = (size_t)n_frames, fc = nf * cap, cells = nf * (ORBX_GRID_COLS * ORBX_GRID_ROWS + 1);
  FrameGridArgs a;
  CU(sg.in(kps, fc, on_device, &a.k)); CU(sg.in(n, nf, on_device, &a.n));
  if (on_device) { a.kun = kps_un; a.bounds = bounds; a.cell_start = cell_start; a.items = cell_items; }
  else {
    CU(sg.alloc((void**)&a.kun, fc * sizeof(orbx_keypoint))); CU(sg.alloc((void**)&a.bounds, 6 * sizeof(float)));
    CU(sg.alloc((void**)&a.cell_start, cells * sizeof(int))); CU(sg.alloc((void**)&a.items, fc * sizeof(int)));
  }
  a.n_frames = n_frames; a.cap = cap; a.cols = cols; a.rows = rows; a.cam = *camera;
  CU(sg.upload());
  g_launches += launch_undistort_grid(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(kps_un, a.kun, fc * sizeof(orbx_keypoint)));
    CU(sg.d2h(bounds, a.bounds, 6 * sizeof(float)));
    CU(sg.d2h(cell_start, a.cell_start, cells * sizeof(int)));
    CU(sg.d2h(cell_items, a.items, fc * sizeof(int)));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_features_in_area_batch(const orbx_keypoint* kps_un, const int* n, int n_frames, int cap, const float* bounds,
                                        const int* cell_start, const int* cell_items, const int* q_frame, const float* q_x,
                                        const float* q_y, const float* q_r, const int* q_min_level, const int* q_max_level,
                                        int n_queries, int qcap, int* out_idx, int* out_cnt, int on_device, void* stream) {
  if (!kps_un || !n || !bounds || !cell_start || !cell_items || !q_frame || !q_x || !q_y || !q_r || !q_min_level || !q_max_level ||
      !out_idx || !out_cnt)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_frames < 1 || cap < 1 || n_queries < 1 || qcap < 1) return fail(ORBX_ERR_BAD_ARG, "bad sizes");
  if (!on_device)
    for (int q = 0; q < n_queries; ++q)
      if (q_frame[q] < 0 || q_frame[q] >= n_frames) return fail(ORBX_ERR_BAD_ARG, "q_frame out of range");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t nf = (size_t)n_frames, fc = nf * cap, cells = nf * (ORBX_GRID_COLS * ORBX_GRID_ROWS + 1), nq = (size_t)n_queries;
  AreaArgs a;
  CU(sg.in(kps_un, fc, on_device, &a.kun)); CU(sg.in(bounds, 6, on_device, &a.bounds));
  CU(sg.in(cell_start, cells, on_device, &a.cell_start)); CU(sg.in(cell_items, fc, on_device, &a.items));
  CU(sg.in(q_frame, nq, on_device, &a.q_frame)); CU(sg.in(q_x, nq, on_device, &a.q_x)); CU(sg.in(q_y, nq, on_device, &a.q_y));
  CU(sg.in(q_r, nq, on_device, &a.q_r)); CU(sg.in(q_min_level, nq, on_device, &a.q_min_level)); CU(sg.in(q_max_level, nq, on_device, &a.q_max_level));
  if (on_device) { a.out_idx = out_idx; a.out_cnt = out_cnt; }
  else { CU(sg.alloc((void**)&a.out_idx, nq * qcap * sizeof(int))); CU(sg.alloc((void**)&a.out_cnt, nq * sizeof(int))); }
  a.cap = cap; a.n_queries = n_queries; a.qcap = qcap;
  CU(sg.upload());
  g_launches += launch_features_in_area(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(out_idx, a.out_idx, nq * qcap * sizeof(int)));
    CU(sg.d2h(out_cnt, a.out_cnt, nq * sizeof(int)));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_search_projection_batch(const orbx_keypoint* kps, const uint8_t* desc, const float* u_right, const int* n,
                                         int n_frames, int cap, const float* grid_bounds, const float* scale_factors, int nlevels,
                                         const uint8_t* mp_in_view, const float* mp_proj_x, const float* mp_proj_y,
                                         const float* mp_proj_xr, const int* mp_level, const float* mp_view_cos,
                                         const uint8_t* mp_desc, const int* mp_obs, const int* m, int mcap, float th,
                                         float nnratio, int* frame_obs, int* assigned, int* n_matches, int on_device,
                                         void* stream) {
  if (!kps || !desc || !u_right || !n || !grid_bounds || !scale_factors || !mp_in_view || !mp_proj_x || !mp_proj_y || !mp_proj_xr || !mp_level ||
      !mp_view_cos || !mp_desc || !mp_obs || !m || !frame_obs || !assigned || !n_matches)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_frames < 1 || cap < 1 || cap > 8192 || mcap < 1 || nlevels < 1 || nlevels > kMaxLevels)
    return fail(ORBX_ERR_BAD_ARG, "bad sizes (cap <= 8192, nlevels <= 8)");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t nf = (size_t)n_frames, fc = nf * cap, fm = nf * mcap;
  ProjArgs a;
  CU(sg.in(kps, fc, on_device, &a.k)); CU(sg.in(desc, fc * 32, on_device, &a.d)); CU(sg.in(u_right, fc, on_device, &a.uRight));
  CU(sg.in(n, nf, on_device, &a.n));
  CU(sg.in(mp_in_view, fm, on_device, &a.inView)); CU(sg.in(mp_proj_x, fm, on_device, &a.projX)); CU(sg.in(mp_proj_y, fm, on_device, &a.projY));
  CU(sg.in(mp_proj_xr, fm, on_device, &a.projXR)); CU(sg.in(mp_level, fm, on_device, &a.level)); CU(sg.in(mp_view_cos, fm, on_device, &a.viewCos));
  CU(sg.in(mp_desc, fm * 32, on_device, &a.mpDesc)); CU(sg.in(mp_obs, fm, on_device, &a.mpObs)); CU(sg.in(m, nf, on_device, &a.m));
  const int* fo_in = nullptr;
  CU(sg.in((const int*)frame_obs, fc, on_device, &fo_in));
  a.frameObs = const_cast<int*>(fo_in);
  if (on_device) { a.assigned = assigned; a.nm = n_matches; }
  else { CU(sg.alloc((void**)&a.assigned, fc * 4)); CU(sg.alloc((void**)&a.nm, nf * 4)); }
  a.n_pairs = n_frames; a.cap = cap; a.mcap = mcap; a.nlevels = nlevels;
  CU(sg.in(grid_bounds, 6, on_device, &a.gb)); a.th = th; a.ratio = nnratio;
  CU(sg.alloc((void**)&a.spec, fm * sizeof(uint4)));
  for (int i = 0; i < kMaxLevels; ++i) a.scale[i] = 1.f;
  if (on_device) CU(sg.d2h(a.scale, scale_factors, sizeof(float) * nlevels));
  else std::memcpy(a.scale, scale_factors, sizeof(float) * nlevels);
  if (on_device) CU(sg.sync());
  CU(sg.upload());
  g_launches += launch_search_projection(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(assigned, a.assigned, fc * 4));
    CU(sg.d2h(n_matches, a.nm, nf * 4));
    CU(sg.d2h(frame_obs, a.frameObs, fc * 4));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_search_projection_frame_batch(const orbx_keypoint* kps, const uint8_t* desc, const float* u_right, const int* n,
                                               int n_frames, int cap, const float* grid_bounds, const float* scale_factors, int nlevels,
                                               const float* pose, const int* mode, float fx, float fy, float cx, float cy, float bf,
                                               const uint8_t* last_valid, const float* last_xw, const int* last_octave,
                                               const float* last_angle, const uint8_t* mp_desc, const int* mp_obs, const int* n_last,
                                               int lcap, float th, int check_orientation, int* frame_obs, int* assigned,
                                               int* n_matches, int on_device, void* stream) {
  if (!kps || !desc || !u_right || !n || !grid_bounds || !scale_factors || !pose || !mode || !last_valid || !last_xw || !last_octave || !last_angle ||
      !mp_desc || !mp_obs || !n_last || !frame_obs || !assigned || !n_matches)
    return fail(ORBX_ERR_BAD_ARG, "null argument");
  if (n_frames < 1 || cap < 1 || cap > 8192 || lcap < 1 || nlevels < 1 || nlevels > kMaxLevels)
    return fail(ORBX_ERR_BAD_ARG, "bad sizes (cap <= 8192, nlevels <= 8)");
  cudaStream_t st = (cudaStream_t)stream;
  Staged sg(st);
  const size_t nf = (size_t)n_frames, fc = nf * cap, fl = nf * lcap;
  ProjFrameArgs a;
  CU(sg.in(kps, fc, on_device, &a.k)); CU(sg.in(desc, fc * 32, on_device, &a.d)); CU(sg.in(u_right, fc, on_device, &a.uRight));
  CU(sg.in(n, nf, on_device, &a.n)); CU(sg.in(pose, nf * 12, on_device, &a.pose)); CU(sg.in(mode, nf, on_device, &a.mode));
  CU(sg.in(last_valid, fl, on_device, &a.lastValid)); CU(sg.in(last_xw, fl * 3, on_device, &a.lastXw));
  CU(sg.in(last_octave, fl, on_device, &a.lastOctave)); CU(sg.in(last_angle, fl, on_device, &a.lastAngle));
  CU(sg.in(mp_desc, fl * 32, on_device, &a.mpDesc)); CU(sg.in(mp_obs, fl, on_device, &a.mpObs)); CU(sg.in(n_last, nf, on_device, &a.nLast));
  const int* fo_in = nullptr;
  CU(sg.in((const int*)frame_obs, fc, on_device, &fo_in));
  a.frameObs = const_cast<int*>(fo_in);
  if (on_device) { a.assigned = assigned; a.nm = n_matches; }
  else { CU(sg.alloc((void**)&a.assigned, fc * 4)); CU(sg.alloc((void**)&a.nm, nf * 4)); }
  CU(sg.alloc((void**)&a.events, fl * 4));
  CU(sg.alloc((void**)&a.spec, fl * sizeof(uint4)));
  a.n_pairs = n_frames; a.cap = cap; a.lcap = lcap; a.nlevels = nlevels; a.checkOri = check_orientation;
  CU(sg.in(grid_bounds, 6, on_device, &a.gb));
  a.fx = fx; a.fy = fy; a.cx = cx; a.cy = cy; a.bf = bf; a.th = th;
  for (int i = 0; i < kMaxLevels; ++i) a.scale[i] = 1.f;
  if (on_device) { CU(sg.d2h(a.scale, scale_factors, sizeof(float) * nlevels)); CU(sg.sync()); }
  else std::memcpy(a.scale, scale_factors, sizeof(float) * nlevels);
  CU(sg.upload());
  g_launches += launch_search_projection_frame(a, st);
  CU(cudaGetLastError());
  if (!on_device) {
    CU(sg.d2h(assigned, a.assigned, fc * 4));
    CU(sg.d2h(n_matches, a.nm, nf * 4));
    CU(sg.d2h(frame_obs, a.frameObs, fc * 4));
    CU(sg.sync());
  }
  return ORBX_OK;
}

orbx_status orbx_set_profiling(orbx_extractor* e, int enable) {
  if (!e) return fail(ORBX_ERR_BAD_ARG, "null extractor");
  DeviceGuard guard(e->device);
  if (enable && !e->ev_ext[0][0]) {
    for (int i = 0; i < orbx_extractor::kRing; ++i) {
      for (int j = 0; j < 6; ++j) CU(cudaEventCreate(&e->ev_ext[i][j]));
      for (int j = 0; j < 3; ++j) CU(cudaEventCreate(&e->ev_st[i][j]));
    }
  }
  e->profiling = enable != 0;
  e->n_ext = e->n_st = 0;
  return ORBX_OK;
}

orbx_status orbx_get_profile(orbx_extractor* e, float* ms, int* calls) {
  if (!e || !ms || !calls) return fail(ORBX_ERR_BAD_ARG, "null argument");
  DeviceGuard guard(e->device);
  CU(cudaDeviceSynchronize());
  for (int s = 0; s < ORBX_N_STAGES; ++s) { ms[s] = 0.f; calls[s] = 0; }
  static const int kExtStage[5] = {0, 1, 2, 6, 3};   // event intervals in launch order: pyramid, FAST, quad-tree, blur, describe
  for (int i = 0; i < e->n_ext; ++i)
    for (int s = 0; s < 5; ++s) {
      float t = 0.f;
      CU(cudaEventElapsedTime(&t, e->ev_ext[i][s], e->ev_ext[i][s + 1]));
      ms[kExtStage[s]] += t; calls[kExtStage[s]] += 1;
    }
  for (int i = 0; i < e->n_st; ++i)
    for (int s = 0; s < 2; ++s) {
      float t = 0.f;
      CU(cudaEventElapsedTime(&t, e->ev_st[i][s], e->ev_st[i][s + 1]));
      ms[4 + s] += t; calls[4 + s] += 1;
    }
  e->n_ext = e->n_st = 0;
  return ORBX_OK;
}

orbx_status orbx_stream_synchronize(void* stream) {
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return ORBX_OK;
}

orbx_status orbx_stream_create(int device, void** stream) {
  if (!stream) return fail(ORBX_ERR_BAD_ARG, "null argument");
  *stream = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(ORBX_ERR_CUDA, "no CUDA device: orbx has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(ORBX_ERR_BAD_ARG, "device index out of range");
  DeviceGuard guard(device);
  cudaStream_t st;
  CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  *stream = (void*)st;
  return ORBX_OK;
}

void orbx_stream_destroy(void* stream) {
  if (stream) cudaStreamDestroy((cudaStream_t)stream);
}

orbx_status orbx_host_alloc(size_t bytes, int write_combined, void** out) {
  if (!out || bytes == 0) return fail(ORBX_ERR_BAD_ARG, "null argument or zero size");
  *out = nullptr;
  CU(cudaHostAlloc(out, bytes, write_combined ? cudaHostAllocWriteCombined : cudaHostAllocDefault));
  return ORBX_OK;
}

void orbx_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

const char* orbx_last_error_string(void) { return g_err.c_str(); }
unsigned long long orbx_kernel_launch_count(void) { return g_launches.load(); }
const char* orbx_version(void) { return "orbx 0.1 (sm_100a)"; }

}  // extern "C"
