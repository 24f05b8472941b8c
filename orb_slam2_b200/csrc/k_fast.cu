// K2 — FAST-9/16 corner detection per 30-px cell with the reference's threshold fallback.
//
// Replaces the per-cell cv::FAST calls of ORBextractor::ComputeKeyPointsOctTree
// (src/ORBextractor.cc:789-829): FAST(window, iniThFAST, nms=true) and, iff that returns no
// keypoint, FAST(window, minThFAST, nms=true).
//
// Formulation (DESIGN.md "K2"): for every tested pixel the arc score
//     S(p) = max over the 16 arcs of 9 contiguous ring pixels of max(min d_k, min -d_k),
//     d_k = I(p) - I(ring_k)
// is threshold independent: p is a corner at t  <=>  S > t, and cv::FAST's response is S-1.
// OpenCV's 3x3 non-max suppression (neighbours that are not corners count as 0, pixels in the
// 3-px window rim are never corners) reduces to "S(p) > t and S(p) > S(q) for the 8 neighbours
// q inside the cell", also threshold independent.  So one S map per cell serves both thresholds.
//
// One warp per cell.  The window is staged in shared memory, each lane owns a column.  S is
// evaluated branch-free with the DPX 3-input packed min/max (VIMNMX3.S16x2): the bright and the
// dark polarity ride in the two 16-bit halves of one register, built by a single IMAD per ring
// pixel: P_k = r_k*0xFFFF + ((256-v)<<16 | (v+256))  ->  lo = d_k+256, hi = -d_k+256.
// Kept corners are appended to the level's candidate pool in pixel row-major order (ballot +
// popc); the cell table records {pool offset, count} so K3 can restore the reference's
// cell-row-major candidate order.
#include "orbx_internal.h"

namespace orbx {

constexpr int kFastWarps = 4;

__device__ __forceinline__ uint32_t vmin3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t vmax3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_s16x2(a, b, c); }

// Ring pixel k of the pixel at c, packed for both polarities: lo = d_k + 256, hi = -d_k + 256.
#define ORBX_RING(k, ptr, off) P[k] = (ptr)[off] * 0xFFFFu + base

// Arc score of the pixel at c (shared-memory window, row pitch wp).  Returns S in [-255, 255].
// P[0], P[4], P[8], P[12] (the compass points) must already be filled in.
__device__ __forceinline__ int arc_score_rest(const uint8_t* c, const int wp, const uint32_t base, uint32_t (&P)[16]) {
  const uint8_t* r3 = c + 3 * wp;
  const uint8_t* r2 = c + 2 * wp;
  const uint8_t* r1 = c + wp;
  const uint8_t* q1 = c - wp;
  const uint8_t* q2 = c - 2 * wp;
  const uint8_t* q3 = c - 3 * wp;
  ORBX_RING(1, r3, 1);   ORBX_RING(2, r2, 2);    ORBX_RING(3, r1, 3);
  ORBX_RING(5, q1, 3);   ORBX_RING(6, q2, 2);    ORBX_RING(7, q3, 1);
  ORBX_RING(9, q3, -1);  ORBX_RING(10, q2, -2);  ORBX_RING(11, q1, -3);
  ORBX_RING(13, r1, -3); ORBX_RING(14, r2, -2);  ORBX_RING(15, r3, -1);
  uint32_t m3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) m3[k] = vmin3(P[k], P[(k + 1) & 15], P[(k + 2) & 15]);
  uint32_t m9[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) m9[k] = vmin3(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
  uint32_t a = vmax3(m9[0], m9[1], m9[2]), b = vmax3(m9[3], m9[4], m9[5]), d = vmax3(m9[6], m9[7], m9[8]);
  uint32_t e = vmax3(m9[9], m9[10], m9[11]), f = vmax3(m9[12], m9[13], m9[14]);
  uint32_t g = vmax3(a, b, d), h = vmax3(e, f, m9[15]);
  uint32_t s = vmax3(g, h, h);
  return max((int)(s & 0xFFFF), (int)(s >> 16)) - 256;
}
// Compass pre-test: any 9-arc of the 16-ring holds two neighbouring compass points, so
// S <= max over the 4 neighbouring compass pairs of min(d) (per polarity).  Fills P[0,4,8,12].
__device__ __forceinline__ int compass_bound(const uint8_t* c, const int wp, const uint32_t base, uint32_t (&P)[16]) {
  ORBX_RING(0, c + 3 * wp, 0); ORBX_RING(4, c, 3); ORBX_RING(8, c - 3 * wp, 0); ORBX_RING(12, c, -3);
  const uint32_t a = vmin3(P[0], P[4], P[4]), b = vmin3(P[4], P[8], P[8]), d = vmin3(P[8], P[12], P[12]), e = vmin3(P[12], P[0], P[0]);
  const uint32_t m = vmax3(vmax3(a, b, d), e, e);
  return max((int)(m & 0xFFFF), (int)(m >> 16)) - 256;
}
__device__ __forceinline__ int arc_score(const uint8_t* c, const int wp) {
  const int v = c[0];
  const uint32_t base = (uint32_t)(v + 256) | ((uint32_t)(256 - v) << 16);
  uint32_t P[16];
  compass_bound(c, wp, base, P);
  return arc_score_rest(c, wp, base, P);
}

// Shared memory per warp (all pitches WP bytes):
//   win   [win_rows][WP]      staged window; pixels keep their global 4-byte phase (aligned word loads)
//   smap  [win_rows-4][WP]    sparse S map with a 1-px zero rim: holds S only where S > minTh
//   queue [(win_rows-6)*(WP-12)] uint16 (yy<<6 | xx): compass survivors, then (in place) the hot list
//   rows  [2][64][2] uint32   kept-corner bit masks per tested row, for minTh and for iniTh
template <int WP>
__global__ void __launch_bounds__(kFastWarps * 32) k_fast_cells(DeviceLayout L, int first_slot, int n, int win_rows) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int win_bytes = WP * win_rows, s_bytes = WP * (win_rows - 4), q_bytes = 2 * (win_rows - 6) * (WP - 12);
  uint8_t* win = smem + (size_t)warp * (win_bytes + s_bytes + q_bytes + 1024);
  uint8_t* smap = win + win_bytes;
  uint16_t* queue = reinterpret_cast<uint16_t*>(smap + s_bytes);
  uint32_t* rows = reinterpret_cast<uint32_t*>(smap + s_bytes + q_bytes);

  const long long gcell = (long long)blockIdx.x * kFastWarps + warp;
  if (gcell >= (long long)n * L.cells_per_slot) return;
  const int img = (int)(gcell / L.cells_per_slot);
  const int c = (int)(gcell - (long long)img * L.cells_per_slot);
  const int slot = first_slot + img;
  int level = 0;
  for (int l = 1; l < L.nlevels; ++l)
    if (c >= L.geom[l].cell_base) level = l;
  const LevelGeom& g = L.geom[level];
  const int ci = c - g.cell_base, ci_row = ci / g.nCols, ci_col = ci - ci_row * g.nCols;
  const int x0 = kFastBorder + ci_col * g.wCell, y0 = kFastBorder + ci_row * g.hCell;
  const int ww = min(x0 + g.wCell + 6, g.maxBX) - x0, hh = min(y0 + g.hCell + 6, g.maxBY) - y0;
  uint2* cell_out = L.celltab + (size_t)slot * L.cells_per_slot + c;
  if (y0 >= g.maxBY - 3 || x0 >= g.maxBX - 6 || ww < 7 || hh < 7) {  // src/ORBextractor.cc:794,802
    if (lane == 0) *cell_out = make_uint2(0u, 0u);
    return;
  }
  const int iw = ww - 6, ih = hh - 6;
  const int min_th = L.min_th, ini_th = L.ini_th;

  // stage the window with aligned word loads (plane base is 256-B aligned, pitch a multiple of 16)
  const int phase = (kEdge + x0) & 3;
  {
    const uint8_t* src = L.pyr + g.plane_off + (size_t)slot * g.plane_stride + (size_t)(kEdge + y0) * g.pitch + (kEdge + x0 - phase);
    const int wpr = (ww + phase + 3) >> 2;          // words per row
    const int total = wpr * hh;
    for (int i = lane; i < total; i += 32) {
      const int r = i / wpr, w = i - r * wpr;
      reinterpret_cast<uint32_t*>(win + r * WP)[w] = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)r * g.pitch) + w);
    }
  }
  for (int i = lane; i < ((ih + 2) * WP) >> 2; i += 32) reinterpret_cast<uint32_t*>(smap)[i] = 0u;
  for (int i = lane; i < 256; i += 32) rows[i] = 0u;
  __syncwarp();

  // stage 1: compass pre-test for every tested pixel (lane = column, up to two 32-column chunks).
  // A pixel whose bound is <= minTh cannot be a corner at either threshold.
  unsigned long long cand[2] = {0ull, 0ull};
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    const int xx = ch * 32 + lane;
    if (xx < iw) {
      const uint8_t* cptr = win + 3 * WP + phase + xx + 3;
      unsigned long long m = 0ull;
      for (int yy = 0; yy < ih; ++yy) {
        const int v = cptr[0];
        const uint32_t base = (uint32_t)(v + 256) | ((uint32_t)(256 - v) << 16);
        uint32_t P[16];
        m |= (unsigned long long)(compass_bound(cptr, WP, base, P) > min_th) << yy;
        cptr += WP;
      }
      cand[ch] = m;
    }
  }
  // survivors -> one warp-wide queue, so that the expensive stage below is load balanced
  int ncand;
  {
    const int mine = __popcll(cand[0]) + __popcll(cand[1]);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    ncand = __shfl_sync(0xffffffffu, incl, 31);
    int pos = incl - mine;
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
      unsigned long long m = cand[ch];
      while (m) {
        const int yy = __ffsll((long long)m) - 1;
        m &= m - 1;
        queue[pos++] = (uint16_t)((yy << 6) | (ch * 32 + lane));
      }
    }
  }
  __syncwarp();

  // stage 2: full 16-point arc score, one survivor per lane per iteration; the ones with S > minTh
  // ("hot") go to the S map and, compacted in place, to the front of the queue
  int nhot = 0;
  for (int i = 0; i < ncand; i += 32) {
    const bool valid = i + lane < ncand;
    const int e = valid ? queue[i + lane] : 0;
    const int yy = e >> 6, xx = e & 63;
    const int sc = arc_score(win + (yy + 3) * WP + phase + xx + 3, WP);
    const bool is_hot = valid && sc > min_th;
    const uint32_t bal = __ballot_sync(0xffffffffu, is_hot);
    if (is_hot) {
      queue[nhot + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)e;
      smap[(yy + 1) * WP + xx + 1] = (uint8_t)sc;
    }
    nhot += __popc(bal);
  }
  __syncwarp();

  // stage 3: strict 3x3 maxima among the hot pixels (neighbours that are not hot read as 0)
  int cnt_ini = 0, cnt_min = 0;
  for (int i = 0; i < nhot; i += 32) {
    const bool valid = i + lane < nhot;
    const int e = valid ? queue[i + lane] : 0;
    const int yy = e >> 6, xx = e & 63;
    const uint8_t* p = smap + (yy + 1) * WP + xx + 1;
    const int sc = p[0];
    const int m1 = __vimax3_s32(p[-WP - 1], p[-WP], p[-WP + 1]), m2 = __vimax3_s32(p[-1], p[1], p[WP - 1]);
    const int m3 = __vimax3_s32(p[WP], p[WP + 1], m1);
    const bool kept = valid && sc > max(m2, m3);
    const bool kept_ini = kept && sc > ini_th;
    if (kept) atomicOr(&rows[yy * 2 + (xx >> 5)], 1u << (xx & 31));
    if (kept_ini) atomicOr(&rows[128 + yy * 2 + (xx >> 5)], 1u << (xx & 31));
    cnt_min += __popc(__ballot_sync(0xffffffffu, kept));
    cnt_ini += __popc(__ballot_sync(0xffffffffu, kept_ini));
  }
  __syncwarp();
  const bool use_ini = cnt_ini > 0;                    // src/ORBextractor.cc:812: retry at minTh iff empty
  const int total = use_ini ? cnt_ini : cnt_min;
  int base = 0;
  if (lane == 0) {
    if (total > 0) base = atomicAdd(L.poolcnt + slot * kMaxLevels + level, total);
    *cell_out = make_uint2((uint32_t)base, (uint32_t)total);
  }
  if (total == 0) return;
  base = __shfl_sync(0xffffffffu, base, 0);

  // emit in pixel row-major order: lane r owns tested rows r and r+32
  const uint32_t* km = rows + (use_ini ? 128 : 0);
  uint32_t* pool = L.pool + (size_t)slot * L.pool_per_slot + g.cand_off;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int yy = half * 32 + lane;
    const uint32_t w0 = yy < ih ? km[yy * 2] : 0u, w1 = yy < ih ? km[yy * 2 + 1] : 0u;
    const int mine = __popc(w0) + __popc(w1);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    int pos = base + incl - mine;
    base += __shfl_sync(0xffffffffu, incl, 31);
    unsigned long long m = ((unsigned long long)w1 << 32) | w0;
    while (m) {
      const int xx = __ffsll((long long)m) - 1;
      m &= m - 1;
      const int resp = (int)smap[(yy + 1) * WP + xx + 1] - 1;
      if (pos < g.cand_cap) pool[pos] = pack_xyr(x0 + 3 + xx - kFastBorder, y0 + 3 + yy - kFastBorder, resp);
      ++pos;
    }
  }
}

template <int WP>
static void launch_fast_t(const DeviceLayout& L, int first_slot, int n, int win_h, unsigned grid, cudaStream_t st) {
  const size_t smem = (size_t)kFastWarps * (WP * win_h + WP * (win_h - 4) + 2 * (win_h - 6) * (WP - 12) + 1024);
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_fast_cells<WP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_fast_cells<WP><<<grid, kFastWarps * 32, smem, st>>>(L, first_slot, n, win_h);
}

int launch_fast(const DeviceLayout& L, const LevelGeom* hg, int first_slot, int n, int win_w, int win_h, cudaStream_t st) {
  (void)hg;
  const long long cells = (long long)n * L.cells_per_slot;
  const unsigned grid = (unsigned)((cells + kFastWarps - 1) / kFastWarps);
  // a staged row (phase + width, rounded up to words) must fit the smem pitch
  if (win_w + 6 <= 48) launch_fast_t<48>(L, first_slot, n, win_h, grid, st);
  else launch_fast_t<72>(L, first_slot, n, win_h, grid, st);
  return 1;
}

}  // namespace orbx
