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

// smem per warp: window [win_rows][WP] (pixels keep their global 4-byte phase so rows are staged
// with aligned 32-bit loads) + S map [win_rows-6][WP].
template <int WP, bool WIDE>
__global__ void __launch_bounds__(kFastWarps * 32) k_fast_cells(DeviceLayout L, int first_slot, int n, int win_rows) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int win_bytes = WP * win_rows, s_bytes = WP * (win_rows - 6);
  uint8_t* win = smem + (size_t)warp * (win_bytes + s_bytes);
  uint8_t* smap = win + win_bytes;

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
  __syncwarp();

  unsigned long long k_ini[2] = {0ull, 0ull}, k_min[2] = {0ull, 0ull};
  if (!WIDE) {
    // Narrow cells (iw <= 32): lane = column, rows streamed top to bottom.  The 3x3 maximum test of
    // row yy-1 runs as soon as row yy is scored, with the neighbour columns read by warp shuffles,
    // so no S map is written; only kept corners store their score (for the response).
    // Pixels that fail the compass pre-test at minTh cannot be corners: they score 0, and a whole
    // warp-row of them skips the 16-point evaluation.
    const bool act = lane < iw;
    const uint8_t* cptr = win + 3 * WP + phase + lane + 3;
    int a = 0, b = 0;
    for (int yy = 0; yy <= ih; ++yy) {
      int cur = 0;
      if (yy < ih) {
        const int v = cptr[0];
        const uint32_t base = (uint32_t)(v + 256) | ((uint32_t)(256 - v) << 16);
        uint32_t P[16];
        const bool need = act && compass_bound(cptr, WP, base, P) > L.min_th;
        if (__any_sync(0xffffffffu, need)) {
          const int sc = arc_score_rest(cptr, WP, base, P);
          cur = need ? max(sc, 0) : 0;
        }
        cptr += WP;
      }
      const int t3 = max(max(a, b), cur);
      int lft = __shfl_up_sync(0xffffffffu, t3, 1), rgt = __shfl_down_sync(0xffffffffu, t3, 1);
      if (lane == 0) lft = 0;
      if (lane == 31) rgt = 0;
      const int nb = max(max(lft, rgt), max(a, cur));
      if (b > nb && b > L.min_th) {
        k_min[0] |= 1ull << (yy - 1);
        if (b > L.ini_th) k_ini[0] |= 1ull << (yy - 1);
        smap[(yy - 1) * WP + lane] = (uint8_t)b;
      }
      a = b; b = cur;
    }
  } else {
  // pass 1: S for every tested pixel; per lane (= column) a bit mask of rows with S > minTh
  unsigned long long hot[2] = {0ull, 0ull};
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    const int xx = ch * 32 + lane;
    if (xx < iw) {
      const uint8_t* cptr = win + 3 * WP + phase + xx + 3;
      uint8_t* sptr = smap + xx;
      unsigned long long m = 0ull;
      for (int yy = 0; yy < ih; ++yy) {
        const int sc = arc_score(cptr, WP);
        sptr[0] = (uint8_t)max(sc, 0);
        m |= (unsigned long long)(sc > L.min_th) << yy;
        cptr += WP; sptr += WP;
      }
      hot[ch] = m;
    }
  }
  __syncwarp();

  // pass 2 (sparse): strict 3x3 maxima among the hot pixels
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    const int xx = ch * 32 + lane;
    unsigned long long m = hot[ch];
    while (m) {
      const int yy = __ffsll((long long)m) - 1;
      m &= m - 1;
      const uint8_t* p = smap + yy * WP + xx;
      const int sc = p[0];
      const bool up = yy > 0, dn = yy < ih - 1, lf = xx > 0, rt = xx < iw - 1;
      bool mx = true;
      if (lf) mx = mx && p[-1] < sc;
      if (rt) mx = mx && p[1] < sc;
      if (up) { mx = mx && p[-WP] < sc; if (lf) mx = mx && p[-WP - 1] < sc; if (rt) mx = mx && p[-WP + 1] < sc; }
      if (dn) { mx = mx && p[WP] < sc; if (lf) mx = mx && p[WP - 1] < sc; if (rt) mx = mx && p[WP + 1] < sc; }
      if (mx) {
        k_min[ch] |= 1ull << yy;
        if (sc > L.ini_th) k_ini[ch] |= 1ull << yy;
      }
    }
  }
  }
  __syncwarp();
  const int cnt_ini = __reduce_add_sync(0xffffffffu, __popcll(k_ini[0]) + __popcll(k_ini[1]));
  const int cnt_min = __reduce_add_sync(0xffffffffu, __popcll(k_min[0]) + __popcll(k_min[1]));
  const bool use_ini = cnt_ini > 0;                    // src/ORBextractor.cc:812: retry at minTh iff empty
  const int total = use_ini ? cnt_ini : cnt_min;
  int base = 0;
  if (lane == 0) {
    if (total > 0) base = atomicAdd(L.poolcnt + slot * kMaxLevels + level, total);
    *cell_out = make_uint2((uint32_t)base, (uint32_t)total);
  }
  if (total == 0) return;
  base = __shfl_sync(0xffffffffu, base, 0);
  const unsigned long long keep0 = use_ini ? k_ini[0] : k_min[0], keep1 = use_ini ? k_ini[1] : k_min[1];
  const unsigned long long any = keep0 | keep1;
  unsigned long long rows = ((unsigned long long)__reduce_or_sync(0xffffffffu, (unsigned)(any >> 32)) << 32) |
                            __reduce_or_sync(0xffffffffu, (unsigned)any);

  // pass 3: emit in pixel row-major order (rows that hold a corner only)
  uint32_t* pool = L.pool + (size_t)slot * L.pool_per_slot + g.cand_off;
  int run = 0;
  while (rows) {
    const int yy = __ffsll((long long)rows) - 1;
    rows &= rows - 1;
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
      const bool keep = ((ch ? keep1 : keep0) >> yy) & 1ull;
      const uint32_t bal = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const int xx = ch * 32 + lane;
        const int pos = base + run + __popc(bal & ((1u << lane) - 1u));
        const int resp = (int)smap[yy * WP + xx] - 1;
        if (pos < g.cand_cap) pool[pos] = pack_xyr(x0 + 3 + xx - kFastBorder, y0 + 3 + yy - kFastBorder, resp);
      }
      run += __popc(bal);
    }
  }
}

template <int WP, bool WIDE>
static void launch_fast_t(const DeviceLayout& L, int first_slot, int n, int win_h, unsigned grid, cudaStream_t st) {
  const size_t smem = (size_t)kFastWarps * WP * (2 * win_h - 6);
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_fast_cells<WP, WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_fast_cells<WP, WIDE><<<grid, kFastWarps * 32, smem, st>>>(L, first_slot, n, win_h);
}

int launch_fast(const DeviceLayout& L, const LevelGeom* hg, int first_slot, int n, int win_w, int win_h, cudaStream_t st) {
  (void)hg;
  const long long cells = (long long)n * L.cells_per_slot;
  const unsigned grid = (unsigned)((cells + kFastWarps - 1) / kFastWarps);
  // a staged row (phase + width, rounded up to words) must fit the pitch; cells of at most 32
  // tested columns take the shuffle-based path
  if (win_w - 6 <= 32) launch_fast_t<48, false>(L, first_slot, n, win_h, grid, st);
  else if (win_w + 6 <= 48) launch_fast_t<48, true>(L, first_slot, n, win_h, grid, st);
  else launch_fast_t<72, true>(L, first_slot, n, win_h, grid, st);
  return 1;
}

}  // namespace orbx
