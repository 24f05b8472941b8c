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
#include "tma_utils.cuh"

namespace orbx {

constexpr int kFastWarps = 3;     // 3 x 8 KB of shared memory per CTA -> 9 CTAs = 27 warps per SM (4 warps: 6 CTAs = 24; measured 2.5 % slower)

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
// S <= max over the 4 neighbouring compass pairs of min(d) (per polarity).  Fills P[0,4,8,12] and
// returns the bound of both polarities packed as (bound + 256) in each half.
__device__ __forceinline__ uint32_t pack_base(const int v) {     // lo = v + 256, hi = 256 - v, one IMAD
  return (uint32_t)v * 0xFFFF0001u + 0x01000100u;
}
__device__ __forceinline__ uint32_t compass_packed(const uint8_t* c, const int wp, uint32_t (&P)[16]) {
  const uint32_t base = pack_base(c[0]);
  ORBX_RING(0, c + 3 * wp, 0); ORBX_RING(4, c, 3); ORBX_RING(8, c - 3 * wp, 0); ORBX_RING(12, c, -3);
  // the four neighbouring pairs are all (one of N/S) x (one of E/W):  max over them of min(a, b) = min(max(N,S), max(E,W))
  return vmin3(vmax3(P[0], P[8], P[8]), vmax3(P[4], P[12], P[12]), 0x7fff7fffu);
}
// Stage-1 form of the same bound with the threshold folded in: `bias` = 0x8000 - 257 - minTh added to both
// halves of the base commutes with min / max (unsigned halves, no carry: every half stays within
// 0x8000 +- 512), and bit 15 of a half is then set iff that polarity's bound exceeds minTh.
__device__ __forceinline__ uint32_t compass_pass_bits(const uint8_t* c, const int wp, const uint32_t base_bias) {
  const uint32_t base = (uint32_t)c[0] * 0xFFFF0001u + base_bias;
  const uint32_t n = c[3 * wp] * 0xFFFFu + base, e = c[3] * 0xFFFFu + base, s = c[-3 * wp] * 0xFFFFu + base, w = c[-3] * 0xFFFFu + base;
  return __vimin3_u16x2(__vimax3_u16x2(n, s, s), __vimax3_u16x2(e, w, w), 0xffffffffu) & 0x80008000u;
}
__device__ __forceinline__ int arc_score(const uint8_t* c, const int wp) {
  uint32_t P[16];
  compass_packed(c, wp, P);
  return arc_score_rest(c, wp, pack_base(c[0]), P);
}

// Shared memory per warp:
//   win   [win_rows][WP]      window box fetched by TMA from the 16-byte boundary below the cell's first
//                             column (WP = 64 or 96); pixel (x, y) of the window sits at [y][phase16 + x]
//   smap  [win_rows-4][SP]    sparse S map with a 1-px zero rim: holds S only where S > minTh
//   queue [(win_rows-6)*iw_max] uint16 (yy<<6 | xx): compass survivors, then (in place) the hot list
//   rows  2 x [win_rows-6][2] uint32  kept-corner bit masks per tested row: strict 3x3 maxima above minTh / above iniTh
//   bar   one mbarrier
// A warp walks kCellsPerWarp consecutive cells; the window of the next cell is requested as soon
// as the scoring stage has finished with the current one, so its latency hides behind the
// maxima / emission stages.
constexpr int kCellsPerWarp = 4;

// the fields of LevelGeom this kernel reads, per level, as a kernel parameter (constant bank instead of global loads)
struct FastGeom { int wCell[kMaxLevels], hCell[kMaxLevels], maxBX[kMaxLevels], maxBY[kMaxLevels], nCols[kMaxLevels], cell_base[kMaxLevels], cand_off[kMaxLevels], cand_cap[kMaxLevels]; };

// NARROW: no cell of the launch has more than 32 tested columns (one column chunk everywhere): the per-row masks are
// not needed and the S map shrinks to a 40-byte pitch - 7.2 KB instead of 8 KB per warp, 10 CTAs per SM instead of 9
template <int WP, bool NARROW>
__global__ void __launch_bounds__(kFastWarps * 32, NARROW ? 10 : 9) k_fast_cells(DeviceLayout L, const __grid_constant__ TmaMaps maps, const __grid_constant__ FastGeom G, int first_slot,
                                                                int n, int win_rows, int iw_max) {
  constexpr int SP = NARROW ? 40 : WP - 16;   // S-map pitch: 40 (up to 32 tested columns), 48 (up to 36) or 80
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int win_bytes = WP * win_rows, s_bytes = SP * (win_rows - 4), q_bytes = (2 * (win_rows - 6) * iw_max + 7) & ~7;
  const int m_words = NARROW ? 0 : 4 * (win_rows - 6);      // two mask sets (kept above minTh / above iniTh), two words per tested row
  const int per_warp = (win_bytes + s_bytes + q_bytes + 4 * m_words + 16 + 127) & ~127;
  uint8_t* win = smem + (size_t)warp * per_warp;
  uint8_t* smap = win + win_bytes;
  uint16_t* queue = reinterpret_cast<uint16_t*>(smap + s_bytes);
  uint32_t* rows = reinterpret_cast<uint32_t*>(smap + s_bytes + q_bytes);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smap + s_bytes + q_bytes + 4 * m_words);
  const int min_th = L.min_th, ini_th = L.ini_th;

  const long long total_cells = (long long)n * L.cells_per_slot;
  const long long cell0 = ((long long)blockIdx.x * kFastWarps + warp) * kCellsPerWarp;
  if (cell0 >= total_cells) return;
  const int ncell = (int)min((long long)kCellsPerWarp, total_cells - cell0);

  struct Cell { int slot, c, level, col, row, x0, y0, ww, hh; bool skip; };
  auto finish = [&](Cell& k) {                // window of cell (col, row) of its level
    const int l = k.level, wCell = G.wCell[l], hCell = G.hCell[l], maxBX = G.maxBX[l], maxBY = G.maxBY[l];
    k.x0 = kFastBorder + k.col * wCell; k.y0 = kFastBorder + k.row * hCell;
    k.ww = min(k.x0 + wCell + 6, maxBX) - k.x0; k.hh = min(k.y0 + hCell + 6, maxBY) - k.y0;
    k.skip = k.y0 >= maxBY - 3 || k.x0 >= maxBX - 6 || k.ww < 7 || k.hh < 7;   // src/ORBextractor.cc:794,802
  };
  auto locate = [&](long long gcell) {
    Cell k;
    const int img = (int)(gcell / L.cells_per_slot);
    k.c = (int)(gcell - (long long)img * L.cells_per_slot);
    k.slot = first_slot + img;
    k.level = 0;
    for (int l = 1; l < L.nlevels; ++l)
      if (k.c >= G.cell_base[l]) k.level = l;
    const int ci = k.c - G.cell_base[k.level], nCols = G.nCols[k.level];
    k.row = ci / nCols; k.col = ci - k.row * nCols;
    finish(k);
    return k;
  };
  auto advance = [&](const Cell& p) {         // the cell after p in (image, level, row, column) order
    Cell k = p;
    ++k.c; ++k.col;
    if (k.c == L.cells_per_slot) { k.c = 0; ++k.slot; k.level = 0; k.col = 0; k.row = 0; }
    else if (k.col == G.nCols[k.level]) {
      k.col = 0; ++k.row;
      if (k.level + 1 < L.nlevels && k.c >= G.cell_base[k.level + 1]) {
        k.row = 0;
        do ++k.level; while (k.level + 1 < L.nlevels && k.c >= G.cell_base[k.level + 1]);
      }
    }
    finish(k);
    return k;
  };
  auto request = [&](const Cell& k) {         // lane 0: one TMA box load for the cell's window
    mbar_expect_tx(bar, (uint32_t)win_bytes);
    tma_load_3d(win, &maps.m[k.level], (kEdge + k.x0) & ~15, kEdge + k.y0, k.slot, bar);
  };

  if (lane == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  __syncwarp();
  Cell nxt = locate(cell0);
  int waited = 0;                             // TMA requests consumed so far (skipped cells make none) -> mbarrier parity
  if (!nxt.skip && lane == 0) request(nxt);

  for (int kc = 0; kc < ncell; ++kc) {
    const Cell cur = nxt;
    if (kc + 1 < ncell) nxt = advance(cur);
    uint2* cell_out = L.celltab + (size_t)cur.slot * L.cells_per_slot + cur.c;
    if (cur.skip) {
      if (lane == 0) *cell_out = make_uint2(0u, 0u);
      if (kc + 1 < ncell && !nxt.skip && lane == 0) request(nxt);
      continue;
    }
    const int x0 = cur.x0, y0 = cur.y0, iw = cur.ww - 6, ih = cur.hh - 6, slot = cur.slot, level = cur.level;
    const int phase = (kEdge + x0) & 15;

    if (NARROW) { for (int i = lane; i < ((ih + 2) * SP) >> 3; i += 32) reinterpret_cast<uint2*>(smap)[i] = make_uint2(0u, 0u); }   // SP % 8 == 0
    else { for (int i = lane; i < ((ih + 2) * SP) >> 4; i += 32) reinterpret_cast<uint4*>(smap)[i] = make_uint4(0u, 0u, 0u, 0u); }   // SP % 16 == 0
    mbar_wait(bar, waited & 1);
    ++waited;

    // stage 1: compass pre-test for every tested pixel (lane = column, up to two 32-column chunks);
    // a pixel whose bound is <= minTh cannot be a corner at either threshold.  The ballot of row yy is
    // parked in lane yy; after the sweep every lane appends its row's survivors to one warp-wide queue
    // (prefix sum of the per-row counts), row-major, so that stage 2 is load balanced and reads
    // neighbouring columns.  The threshold rides in the packed base (compass_pass_bits).
    int ncand = 0;
    const uint32_t kpass = 0x01000100u + (uint32_t)(0x8000 - 257 - min_th) * 0x00010001u;
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
      if (ch * 32 < iw) {                     // warp-uniform
        const int xx = ch * 32 + lane;
        const uint8_t* cptr = win + 3 * WP + phase + min(xx, iw - 1) + 3;
        const uint32_t colmask = iw - ch * 32 >= 32 ? 0xffffffffu : (1u << (iw - ch * 32)) - 1u;
        uint32_t rm0 = 0u, rm1 = 0u;
        const int h0 = min(ih, 32);
        for (int yy = 0; yy < h0; ++yy) {
          const uint32_t bal = __ballot_sync(0xffffffffu, compass_pass_bits(cptr, WP, kpass) != 0u);
          if (lane == yy) rm0 = bal;
          cptr += WP;
        }
        for (int yy = 32; yy < ih; ++yy) {
          const uint32_t bal = __ballot_sync(0xffffffffu, compass_pass_bits(cptr, WP, kpass) != 0u);
          if (lane == yy - 32) rm1 = bal;
          cptr += WP;
        }
        rm0 &= colmask; rm1 &= colmask;
        const int mine = __popc(rm0) | (__popc(rm1) << 16);      // rows 0-31 first, then rows 32+
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        const int tot = __shfl_sync(0xffffffffu, incl, 31), excl = incl - mine;
        uint16_t* q0 = queue + ncand + (excl & 0xffff);
        uint16_t* q1 = queue + ncand + (tot & 0xffff) + (excl >> 16);
        ncand += (tot & 0xffff) + (tot >> 16);
        const int tag = (lane << 6) | (ch * 32);
        while (rm0) { const int b = __ffs((int)rm0) - 1; rm0 &= rm0 - 1u; *q0++ = (uint16_t)(tag + b); }
        while (rm1) { const int b = __ffs((int)rm1) - 1; rm1 &= rm1 - 1u; *q1++ = (uint16_t)(tag + (32 << 6) + b); }
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
        smap[(yy + 1) * SP + xx + 1] = (uint8_t)sc;
      }
      nhot += __popc(bal);
    }
    __syncwarp();
    // the window is dead: fetch the next cell's while the maxima / emission stages run
    if (kc + 1 < ncell && !nxt.skip && lane == 0) { fence_proxy_async(); request(nxt); }

    const uint32_t lt = (1u << lane) - 1u;
    if (NARROW || iw <= 32) {                            // warp-uniform; the usual case
      // One column chunk: the queue, and with it the hot list, is in pixel row-major order, which is the order
      // the candidates must leave in.  Stage 3: strict 3x3 maxima among the hot pixels (neighbours that are not
      // hot read as 0); the kept ones are compacted in place at the front of the queue, bit 15 = also above iniTh.
      int nkept = 0, cnt_ini = 0;
      for (int i = 0; i < nhot; i += 32) {
        const bool valid = i + lane < nhot;
        const int e = valid ? queue[i + lane] : 0;
        const int yy = e >> 6, xx = e & 63;
        const uint8_t* p = smap + (yy + 1) * SP + xx + 1;
        const int sc = p[0];
        const int m1 = __vimax3_s32(p[-SP - 1], p[-SP], p[-SP + 1]), m2 = __vimax3_s32(p[-1], p[1], p[SP - 1]);
        const int m3 = __vimax3_s32(p[SP], p[SP + 1], m1);
        const bool kept = valid && sc > max(m2, m3);
        const bool kept_ini = kept && sc > ini_th;
        const uint32_t bk = __ballot_sync(0xffffffffu, kept), bi = __ballot_sync(0xffffffffu, kept_ini);
        if (kept) queue[nkept + __popc(bk & lt)] = (uint16_t)(e | (kept_ini ? 0x8000 : 0));
        nkept += __popc(bk);
        cnt_ini += __popc(bi);
      }
      __syncwarp();
      const bool use_ini = cnt_ini > 0;                  // src/ORBextractor.cc:812: retry at minTh iff empty
      const int total = use_ini ? cnt_ini : nkept;
      int base = 0;
      if (lane == 0) {
        if (total > 0) base = atomicAdd(L.poolcnt + slot * kMaxLevels + level, total);
        *cell_out = make_uint2((uint32_t)base, (uint32_t)total);
      }
      if (total > 0) {                                   // warp-uniform
        base = __shfl_sync(0xffffffffu, base, 0);
        uint32_t* pool = L.pool + (size_t)slot * L.pool_per_slot + G.cand_off[level];
        const int cand_cap = G.cand_cap[level];
        for (int i = 0; i < nkept; i += 32) {
          const bool valid = i + lane < nkept;
          const int e = valid ? queue[i + lane] : 0;
          const bool sel = valid && (!use_ini || (e & 0x8000));
          const uint32_t bs = __ballot_sync(0xffffffffu, sel);
          if (sel) {
            const int yy = (e >> 6) & 63, xx = e & 63;
            const int resp = (int)smap[(yy + 1) * SP + xx + 1] - 1;
            const int pos = base + __popc(bs & lt);
            if (pos < cand_cap) pool[pos] = pack_xyr(x0 + 3 + xx - kFastBorder, y0 + 3 + yy - kFastBorder, resp);
          }
          base += __popc(bs);
        }
      }
    } else if (!NARROW) {
      // Two column chunks (cells wider than 32 tested columns): the hot list is chunk-major, so the kept corners
      // go through per-row bit masks and leave row by row.
      for (int i = lane; i < m_words; i += 32) rows[i] = 0u;
      __syncwarp();
      // stage 3: strict 3x3 maxima among the hot pixels (neighbours that are not hot read as 0)
      int cnt_ini = 0, cnt_min = 0;
      for (int i = 0; i < nhot; i += 32) {
        const bool valid = i + lane < nhot;
        const int e = valid ? queue[i + lane] : 0;
        const int yy = e >> 6, xx = e & 63;
        const uint8_t* p = smap + (yy + 1) * SP + xx + 1;
        const int sc = p[0];
        const int m1 = __vimax3_s32(p[-SP - 1], p[-SP], p[-SP + 1]), m2 = __vimax3_s32(p[-1], p[1], p[SP - 1]);
        const int m3 = __vimax3_s32(p[SP], p[SP + 1], m1);
        const bool kept = valid && sc > max(m2, m3);
        const bool kept_ini = kept && sc > ini_th;
        if (kept) atomicOr(&rows[yy * 2 + (xx >> 5)], 1u << (xx & 31));
        if (kept_ini) atomicOr(&rows[(m_words >> 1) + yy * 2 + (xx >> 5)], 1u << (xx & 31));
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
      if (total > 0) {                                     // warp-uniform
        base = __shfl_sync(0xffffffffu, base, 0);
        // emit in pixel row-major order: lane r owns tested rows r and r+32; when iniTh found corners only
        // those are emitted (second mask set)
        uint32_t* pool = L.pool + (size_t)slot * L.pool_per_slot + G.cand_off[level];
        const int cand_cap = G.cand_cap[level];
        const uint32_t* km = rows + (use_ini ? (m_words >> 1) : 0);
  #pragma unroll
        for (int half = 0; half < 2; ++half) {
          if (half * 32 >= ih) break;                        // warp-uniform: most cells have at most 32 tested rows
          const int yy = half * 32 + lane;
          unsigned long long m = yy < ih ? (((unsigned long long)km[yy * 2 + 1] << 32) | km[yy * 2]) : 0ull;
          const uint8_t* srow = smap + (yy + 1) * SP + 1;
          const int mine = __popcll(m);
          int incl = mine;
  #pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
          }
          int pos = base + incl - mine;
          base += __shfl_sync(0xffffffffu, incl, 31);
          while (m) {
            const int xx = __ffsll((long long)m) - 1;
            m &= m - 1;
            const int resp = (int)srow[xx] - 1;
            if (pos < cand_cap) pool[pos] = pack_xyr(x0 + 3 + xx - kFastBorder, y0 + 3 + yy - kFastBorder, resp);
            ++pos;
          }
        }
      }
    }
    __syncwarp();                                        // smap / rows / queue are reused by the next cell
  }
}

template <int WP, bool NARROW>
static void launch_fast_t(const DeviceLayout& L, const TmaMaps& maps, const FastGeom& G, int first_slot, int n, int win_w, int win_h, cudaStream_t st) {
  constexpr int SP = NARROW ? 40 : WP - 16;
  const int iw_max = win_w - 6;
  const size_t per_warp = ((size_t)WP * win_h + SP * (win_h - 4) + ((2 * (win_h - 6) * iw_max + 7) & ~7) + (NARROW ? 0 : 16 * (win_h - 6)) + 16 + 127) & ~(size_t)127;
  const size_t smem = kFastWarps * per_warp;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_fast_cells<WP, NARROW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const long long cells = (long long)n * L.cells_per_slot;
  const long long per_cta = (long long)kFastWarps * kCellsPerWarp;
  k_fast_cells<WP, NARROW><<<(unsigned)((cells + per_cta - 1) / per_cta), kFastWarps * 32, smem, st>>>(L, maps, G, first_slot, n, win_h, iw_max);
}

int fast_window_pitch(int win_w) { return win_w + 15 <= 64 ? 64 : 96; }

int launch_fast(const DeviceLayout& L, const LevelGeom* hg, const TmaMaps& maps, int first_slot, int n, int win_w, int win_h, cudaStream_t st) {
  FastGeom G{};
  for (int l = 0; l < L.nlevels; ++l) {
    G.wCell[l] = hg[l].wCell; G.hCell[l] = hg[l].hCell; G.maxBX[l] = hg[l].maxBX; G.maxBY[l] = hg[l].maxBY;
    G.nCols[l] = hg[l].nCols; G.cell_base[l] = hg[l].cell_base; G.cand_off[l] = hg[l].cand_off; G.cand_cap[l] = hg[l].cand_cap;
  }
  // the TMA box starts at the 16-byte boundary below the window: phase (<= 15) + width must fit the pitch
  if (win_w - 6 <= 32) launch_fast_t<64, true>(L, maps, G, first_slot, n, win_w, win_h, st);
  else if (fast_window_pitch(win_w) == 64) launch_fast_t<64, false>(L, maps, G, first_slot, n, win_w, win_h, st);
  else launch_fast_t<96, false>(L, maps, G, first_slot, n, win_w, win_h, st);
  return 1;
}

}  // namespace orbx
