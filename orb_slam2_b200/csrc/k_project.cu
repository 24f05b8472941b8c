// K16 — the back end's projection searches, batched:
//   ORBmatcher::Fuse(pKF, vpMapPoints, th)      src/ORBmatcher.cc:825-975   (the search; the map mutation stays with the caller)
//   ORBmatcher::SearchBySim3(pKF1, pKF2, ...)   src/ORBmatcher.cc:1102-1326 (both passes + the agreement check)
// with KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:543-588), KeyFrame::IsInImage (:610-613) and
// MapPoint::PredictScale (src/MapPoint.cc:385-400).
//
// A map point's search does not depend on what the other map points found (Fuse's Replace / AddObservation only ever
// change the skip status of points that are already in the keyframe, SearchBySim3 writes nothing until the agreement
// check), so one thread handles one map point: transform by one or two steps in gemm arithmetic (float products, sums
// left to right), project, range / viewing-angle tests with cv::norm's and Mat::dot's double sums, predicted level
// = ceil(logf(dmax / dist) / logScale) with glibc's logf restated in double (tools/sweep_logf.c: exhaustive, 0
// mismatches), then the keyframe's CSR grid (orbx_undistort_grid_batch) is walked in the reference's cell order and
// the first least descriptor distance among levels [pred-1, pred] wins.
#include <cstdlib>

#include "orbx_internal.h"

namespace orbx {

constexpr int kPGridCols = 64, kPGridRows = 48, kPCells = kPGridCols * kPGridRows;

__constant__ double c_logf_tab[16][2] = {
    {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
    {0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2},  {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
    {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3},
    {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
    {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
    {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4},
    {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3},
    {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}};

// glibc 2.39 logf for positive normal x (sysdeps/ieee754/flt-32/e_logf.c)
__device__ __forceinline__ float libm_logf(float x) {
  const uint32_t ix = __float_as_uint(x);
  if (ix == 0x3f800000u) return 0.f;
  const uint32_t tmp = ix - 0x3f330000u;
  const int i = (int)((tmp >> 19) & 15u), k = (int)tmp >> 23;
  const double z = (double)__uint_as_float(ix - (tmp & 0xff800000u));
  const double r = __dsub_rn(__dmul_rn(z, c_logf_tab[i][0]), 1.0);
  const double y0 = __dadd_rn(c_logf_tab[i][1], __dmul_rn((double)k, 0x1.62e42fefa39efp-1));
  const double r2 = __dmul_rn(r, r);
  double y = __dadd_rn(__dmul_rn(0x1.5575b0be00b6ap-2, r), -0x1.ffffef20a4123p-2);
  y = __dadd_rn(__dmul_rn(-0x1.00ea348b88334p-2, r2), y);
  y = __dadd_rn(__dmul_rn(y, r2), __dadd_rn(y0, r));
  return (float)y;
}

__device__ __forceinline__ void gemm3(const float* R, const float* t, const float* x, float* o) {
#pragma unroll
  for (int r = 0; r < 3; ++r)
    o[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(R[3 * r], x[0]), __fmul_rn(R[3 * r + 1], x[1])), __fmul_rn(R[3 * r + 2], x[2])), t[r]);
}

__device__ __forceinline__ int ham8p(const uint32_t* q, const uint32_t* __restrict__ g) {
  const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(g)), b1 = __ldg(reinterpret_cast<const uint4*>(g) + 1);
  return __popc(q[0] ^ b0.x) + __popc(q[1] ^ b0.y) + __popc(q[2] ^ b0.z) + __popc(q[3] ^ b0.w) + __popc(q[4] ^ b1.x) +
         __popc(q[5] ^ b1.y) + __popc(q[6] ^ b1.z) + __popc(q[7] ^ b1.w);
}

// one thread per (problem, map point)
__global__ void __launch_bounds__(128) k_project_search(ProjSearchArgs a) {
  const int p = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = min(a.m[p], a.mcap);
  if (i >= m) return;
  const size_t mo = (size_t)p * a.mcap + i, ko = (size_t)p * a.cap;
  int* out = a.best + mo;
  if (!a.valid[mo]) { *out = -1; return; }
  const float* xf = a.xf + (size_t)p * 27;                      // R1 (9) t1 (3) R2 (9) t2 (3) Ow (3)
  const float* kfb = a.kf_bounds + (size_t)p * 4;               // KeyFrame::mnMinX, mnMaxX, mnMinY, mnMaxY
  const float* gb = a.grid_bounds;                              // Frame's mnMinX, mnMaxX, mnMinY, mnMaxY, invW, invH
  const float xw[3] = {a.xw[mo * 3], a.xw[mo * 3 + 1], a.xw[mo * 3 + 2]};
  float pc[3], q[3];
  gemm3(xf, xf + 9, xw, pc);
  if (a.mode & 1) gemm3(xf + 12, xf + 21, pc, q); else { q[0] = pc[0]; q[1] = pc[1]; q[2] = pc[2]; }
  *out = -1;
  if (q[2] < 0.0f) return;
  const float invz = (a.mode & 16) ? (float)__ddiv_rn(1.0, (double)q[2]) : __fdiv_rn(1.0f, q[2]);
  const float x = __fmul_rn(q[0], invz), y = __fmul_rn(q[1], invz);
  const float u = __fadd_rn(__fmul_rn(a.fx, x), a.cx), v = __fadd_rn(__fmul_rn(a.fy, y), a.cy);
  if (!(u >= kfb[0] && u < kfb[1] && v >= kfb[2] && v < kfb[3])) return;
  const float ur = __fsub_rn(u, __fmul_rn(a.bf, invz));
  float PO[3] = {q[0], q[1], q[2]};
  if (a.mode & 2) { PO[0] = __fsub_rn(xw[0], xf[24]); PO[1] = __fsub_rn(xw[1], xf[25]); PO[2] = __fsub_rn(xw[2], xf[26]); }
  const double s2 = __dadd_rn(__dadd_rn(__dmul_rn((double)PO[0], (double)PO[0]), __dmul_rn((double)PO[1], (double)PO[1])), __dmul_rn((double)PO[2], (double)PO[2]));
  const float dist3D = (float)__dsqrt_rn(s2);
  const float dmin = a.dmin[mo], dmax = a.dmax[mo];
  if (dist3D < __fmul_rn(0.8f, dmin) || dist3D > __fmul_rn(1.2f, dmax)) return;
  if (a.mode & 4) {
    const float* nn = a.normal + mo * 3;
    const double dot = __dadd_rn(__dadd_rn(__dmul_rn((double)PO[0], (double)nn[0]), __dmul_rn((double)PO[1], (double)nn[1])), __dmul_rn((double)PO[2], (double)nn[2]));
    if (dot < __dmul_rn(0.5, (double)dist3D)) return;
  }
  int level = (int)ceilf(__fdiv_rn(libm_logf(__fdiv_rn(dmax, dist3D)), a.log_scale));
  level = level < 0 ? 0 : (level >= a.nlevels ? a.nlevels - 1 : level);
  const float radius = __fmul_rn(a.th, a.scale[level]);
  const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, kfb[0]), radius), gb[4])));
  const int cx1 = min(kPGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, kfb[0]), radius), gb[4])));
  const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, kfb[2]), radius), gb[5])));
  const int cy1 = min(kPGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, kfb[2]), radius), gb[5])));
  if (cx0 >= kPGridCols || cx1 < 0 || cy0 >= kPGridRows || cy1 < 0) return;
  uint32_t qd[8];
#pragma unroll
  for (int w = 0; w < 8; ++w) qd[w] = __ldg(reinterpret_cast<const uint32_t*>(a.mp_desc + mo * 32) + w);
  const orbx_keypoint* kps = a.kps + ko;
  const int* cs = a.cell_start + (size_t)p * (kPCells + 1);
  const int* items = a.cell_items + ko;
  int bestDist = INT_MAX, bestIdx = -1;
  for (int ix = cx0; ix <= cx1; ++ix) {
    const int e_hi = cs[ix * kPGridRows + cy1 + 1];
    for (int e = cs[ix * kPGridRows + cy0]; e < e_hi; ++e) {
      const int idx = items[e];
      const orbx_keypoint kp = kps[idx];
      if (!(fabsf(__fsub_rn(kp.x, u)) < radius && fabsf(__fsub_rn(kp.y, v)) < radius)) continue;
      if (kp.octave < level - 1 || kp.octave > level) continue;
      if (a.mode & 8) {
        const float ex = __fsub_rn(u, kp.x), ey = __fsub_rn(v, kp.y);
        const float kr = a.u_right[ko + idx];
        if (kr >= 0.f) {
          const float er = __fsub_rn(ur, kr);
          const float e2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(er, er));
          if ((double)__fmul_rn(e2, a.inv_sigma2[kp.octave]) > 7.8) continue;
        } else {
          const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
          if ((double)__fmul_rn(e2, a.inv_sigma2[kp.octave]) > 5.99) continue;
        }
      }
      const int dist = ham8p(qd, reinterpret_cast<const uint32_t*>(a.desc + (ko + idx) * 32));
      if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
  }
  if (bestDist <= a.th_dist) *out = bestIdx;
}

// SearchBySim3's agreement check (src/ORBmatcher.cc:1303-1323): vnMatch1 / vnMatch2 are indexed by keypoint
__global__ void __launch_bounds__(128) k_sim3_agree(Sim3AgreeArgs a) {
  const int p = blockIdx.y, i1 = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t o = (size_t)p * a.cap;
  int res = -1;
  if (i1 < a.cap) {
    const int idx2 = a.match1[o + i1];
    if (i1 < a.n1[p] && idx2 >= 0 && idx2 < a.cap && a.match2[o + idx2] == i1) res = idx2;
    a.matches12[o + i1] = res;
  }
  const int c = __syncthreads_count(res >= 0);
  if (threadIdx.x == 0 && c) atomicAdd(a.nfound + p, c);
}

// best-per-map-point (indexed by map point) -> vnMatch (indexed by keypoint)
__global__ void __launch_bounds__(128) k_scatter_matches(const int* __restrict__ best, const int* __restrict__ mp_index, const int* __restrict__ m,
                                                         int mcap, int cap, int* __restrict__ vn) {
  const int p = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= min(m[p], mcap)) return;
  const int kpi = mp_index[(size_t)p * mcap + i];
  if (kpi >= 0 && kpi < cap) vn[(size_t)p * cap + kpi] = best[(size_t)p * mcap + i];
}

// K17: the two SearchByProjection members that claim keypoints as they go (hazard H6), one warp per problem:
//   SearchByProjection(pKF, Scw, vpPoints, vpMatched, th)              src/ORBmatcher.cc:290-403   (reloc = 0)
//   SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)  src/ORBmatcher.cc:1472-1599 (reloc = 1)
// Map points are taken in list order; every lane evaluates the (cheap) projection of the current one, the lanes then share
// the candidates of the window, cell column by cell column.  The winner is the least (distance, cell column, position in
// the column's CSR run), i.e. the first least distance in the reference's enumeration order, among free keypoints.
__device__ __forceinline__ int rot_bin30(float a1, float a2) {
  float rot = __fsub_rn(a1, a2);
  if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
  int bin = (int)roundf(__fmul_rn(rot, 1.0f / 30));
  return bin == 30 ? 0 : bin;
}

// The projection and candidate scan of ONE map point of K17 (warp-wide): the least (distance << 8 | cell column, CSR position)
// among the keypoints that are free in `occ`, ~0 when the point is skipped or has no candidate.  Shared by the one-warp
// kernel and by both phases of the latency form.
__device__ __forceinline__ void ordered_scan(const ProjOrderedArgs& a, int p, int iMP, const uint32_t* occ, int lane, unsigned& whi, unsigned& wlo) {
  whi = wlo = 0xffffffffu;
  const size_t ko = (size_t)p * a.cap, mo = (size_t)p * a.mcap;
  const orbx_keypoint* kps = a.kps + ko;
  const int* cs = a.cell_start + (size_t)p * (kPCells + 1);
  const int* items = a.cell_items + ko;
  const float* xf = a.xf + (size_t)p * 15;
  const float* kb = a.tgt_bounds + (size_t)p * 4;
  const float* gb = a.grid_bounds;
    if (!a.valid[mo + iMP]) return;
    const float xw[3] = {a.xw[(mo + iMP) * 3], a.xw[(mo + iMP) * 3 + 1], a.xw[(mo + iMP) * 3 + 2]};
    float pc[3];
    gemm3(xf, xf + 9, xw, pc);
    float u, v;
    if (!a.reloc) {
      if (pc[2] < 0.0f) return;
      const float invz = __fdiv_rn(1.0f, pc[2]);
      u = __fadd_rn(__fmul_rn(a.fx, __fmul_rn(pc[0], invz)), a.cx); v = __fadd_rn(__fmul_rn(a.fy, __fmul_rn(pc[1], invz)), a.cy);
      if (!(u >= kb[0] && u < kb[1] && v >= kb[2] && v < kb[3])) return;
    } else {
      const float invzc = (float)__ddiv_rn(1.0, (double)pc[2]);
      u = __fadd_rn(__fmul_rn(__fmul_rn(a.fx, pc[0]), invzc), a.cx); v = __fadd_rn(__fmul_rn(__fmul_rn(a.fy, pc[1]), invzc), a.cy);
      if (u < kb[0] || u > kb[1] || v < kb[2] || v > kb[3]) return;
    }
    const float PO[3] = {__fsub_rn(xw[0], xf[12]), __fsub_rn(xw[1], xf[13]), __fsub_rn(xw[2], xf[14])};
    const double s2 = __dadd_rn(__dadd_rn(__dmul_rn((double)PO[0], (double)PO[0]), __dmul_rn((double)PO[1], (double)PO[1])), __dmul_rn((double)PO[2], (double)PO[2]));
    const float dist3D = (float)__dsqrt_rn(s2);
    const float dmin = a.dmin[mo + iMP], dmax = a.dmax[mo + iMP];
    if (dist3D < __fmul_rn(0.8f, dmin) || dist3D > __fmul_rn(1.2f, dmax)) return;
    if (!a.reloc) {
      const float* nn = a.normal + (mo + iMP) * 3;
      const double dot = __dadd_rn(__dadd_rn(__dmul_rn((double)PO[0], (double)nn[0]), __dmul_rn((double)PO[1], (double)nn[1])), __dmul_rn((double)PO[2], (double)nn[2]));
      if (dot < __dmul_rn(0.5, (double)dist3D)) return;
    }
    int level = (int)ceilf(__fdiv_rn(libm_logf(__fdiv_rn(dmax, dist3D)), a.log_scale));
    level = level < 0 ? 0 : (level >= a.nlevels ? a.nlevels - 1 : level);
    const int lvl_hi = a.reloc ? level + 1 : level;
    const float radius = __fmul_rn(a.th, a.scale[level]);
    const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, kb[0]), radius), gb[4])));
    const int cx1 = min(kPGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, kb[0]), radius), gb[4])));
    const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, kb[2]), radius), gb[5])));
    const int cy1 = min(kPGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, kb[2]), radius), gb[5])));
    if (cx0 >= kPGridCols || cx1 < 0 || cy0 >= kPGridRows || cy1 < 0) return;
    uint32_t qd[8];
#pragma unroll
    for (int w = 0; w < 8; ++w) qd[w] = __ldg(reinterpret_cast<const uint32_t*>(a.mp_desc + (mo + iMP) * 32) + w);
    unsigned hi = 0xffffffffu, lo = 0xffffffffu;                // (distance << 8 | cell column), CSR position
    for (int ix = cx0; ix <= cx1; ++ix) {
      const int e_hi = cs[ix * kPGridRows + cy1 + 1];
      for (int e = cs[ix * kPGridRows + cy0] + lane; e < e_hi; e += 32) {
        const int idx = items[e];
        if ((occ[idx >> 5] >> (idx & 31)) & 1u) continue;
        const orbx_keypoint kp = kps[idx];
        if (!(fabsf(__fsub_rn(kp.x, u)) < radius && fabsf(__fsub_rn(kp.y, v)) < radius)) continue;
        if (kp.octave < level - 1 || kp.octave > lvl_hi) continue;
        const unsigned h = ((unsigned)ham8p(qd, reinterpret_cast<const uint32_t*>(a.desc + (ko + idx) * 32)) << 8) | (unsigned)ix;
        if (h < hi || (h == hi && (unsigned)e < lo)) { hi = h; lo = (unsigned)e; }
      }
    }
    whi = __reduce_min_sync(0xffffffffu, hi);
    if (whi == 0xffffffffu) return;
    wlo = __reduce_min_sync(0xffffffffu, hi == whi ? lo : 0xffffffffu);
}

__global__ void __launch_bounds__(32) k_project_match_ordered(ProjOrderedArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint32_t* occ = reinterpret_cast<uint32_t*>(smem);            // bit i: keypoint i holds a map point
  __shared__ int s_hist[30];
  const int p = blockIdx.x, lane = threadIdx.x;
  const int n = min(a.n[p], a.cap), m = min(a.m[p], a.mcap);
  const size_t ko = (size_t)p * a.cap, mo = (size_t)p * a.mcap;
  const orbx_keypoint* kps = a.kps + ko;
  const int* items = a.cell_items + ko;
  int* assigned = a.assigned + ko;
  for (int w = lane; w < (a.cap + 31) / 32; w += 32) {
    uint32_t bits = 0u;
    for (int b = 0; b < 32; ++b) { const int i = w * 32 + b; if (i < n && a.occupied[ko + i]) bits |= 1u << b; }
    occ[w] = bits;
  }
  for (int i = lane; i < a.cap; i += 32) assigned[i] = -1;
  if (lane < 30) s_hist[lane] = 0;
  __syncwarp();
  int nm = 0;
  for (int iMP = 0; iMP < m; ++iMP) {
    unsigned whi, wlo;
    ordered_scan(a, p, iMP, occ, lane, whi, wlo);
    if (whi == 0xffffffffu || (int)(whi >> 8) > a.th_dist) continue;      // no candidate (distance starts at 256) or too far
    const int bestIdx = items[wlo];
    if (lane == 0) {
      occ[bestIdx >> 5] |= 1u << (bestIdx & 31);
      assigned[bestIdx] = iMP;
      if (a.reloc && a.checkOri) atomicAdd(&s_hist[rot_bin30(a.src_angle[mo + iMP], kps[bestIdx].angle)], 1);
    }
    ++nm;
    __syncwarp();
  }
  __syncwarp();
  if (a.reloc && a.checkOri) {                                  // rotation consistency, src/ORBmatcher.cc:1577-1596
    int m1 = 0, m2 = 0, m3 = 0, i1 = -1, i2 = -1, i3 = -1;      // ORBmatcher::ComputeThreeMaxima :1601-1642
    for (int i = 0; i < 30; ++i) {
      const int sc = s_hist[i];
      if (sc > m1) { m3 = m2; m2 = m1; m1 = sc; i3 = i2; i2 = i1; i1 = i; }
      else if (sc > m2) { m3 = m2; m2 = sc; i3 = i2; i2 = i; }
      else if (sc > m3) { m3 = sc; i3 = i; }
    }
    if ((float)m2 < __fmul_rn(0.1f, (float)m1)) { i2 = -1; i3 = -1; }
    else if ((float)m3 < __fmul_rn(0.1f, (float)m1)) i3 = -1;
    int drop = 0;
    for (int i = lane; i < n; i += 32) {
      const int mp = assigned[i];
      if (mp >= 0) {
        const int b = rot_bin30(a.src_angle[mo + mp], kps[i].angle);
        if (b != i1 && b != i2 && b != i3) { assigned[i] = -1; ++drop; }
      }
    }
    nm -= __reduce_add_sync(0xffffffffu, drop);
  }
  if (lane == 0) a.nm[p] = nm;
}

// K17, latency form (one or a few problems per call: relocalisation and loop closing ask about one frame / keyframe at a
// time).  A CTA of 8 warps per problem.  Phase A: the winner of every map point against the keypoints that are occupied ON
// ENTRY, map points strided over the warps.  Phase B: warp 0 walks the map points in list order (hazard H6); a phase-A
// winner stands unless its keypoint was taken meanwhile by an earlier map point of this call - only then the point is
// scanned again against the live occupancy (taking away a keypoint that is not a point's winner cannot change its winner).
constexpr int kOrdParWarps = 8;
__global__ void __launch_bounds__(kOrdParWarps * 32) k_project_match_ordered_par(ProjOrderedArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint32_t* occ = reinterpret_cast<uint32_t*>(smem);            // bit i: keypoint i holds a map point
  __shared__ int s_hist[30];
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = min(a.n[p], a.cap), m = min(a.m[p], a.mcap);
  const size_t ko = (size_t)p * a.cap, mo = (size_t)p * a.mcap;
  const orbx_keypoint* kps = a.kps + ko;
  const int* items = a.cell_items + ko;
  int* assigned = a.assigned + ko;
  uint4* spec = a.spec + mo;
  for (int w = tid; w < (a.cap + 31) / 32; w += kOrdParWarps * 32) {
    uint32_t bits = 0u;
    for (int b = 0; b < 32; ++b) { const int i = w * 32 + b; if (i < n && a.occupied[ko + i]) bits |= 1u << b; }
    occ[w] = bits;
  }
  for (int i = tid; i < a.cap; i += kOrdParWarps * 32) assigned[i] = -1;
  if (tid < 30) s_hist[tid] = 0;
  __syncthreads();
  // phase A (occ is only read)
  for (int iMP = warp; iMP < m; iMP += kOrdParWarps) {
    unsigned whi, wlo;
    ordered_scan(a, p, iMP, occ, lane, whi, wlo);
    if (lane == 0) {
      int idx = 0; float ang = 0.f;
      if (whi != 0xffffffffu) { idx = items[wlo]; ang = kps[idx].angle; }
      spec[iMP] = make_uint4(whi, wlo, (unsigned)idx, __float_as_uint(ang));
    }
  }
  __syncthreads();
  if (warp != 0) return;
  // phase B
  int nm = 0;
  for (int base = 0; base < m; base += 32) {
    const int mine = base + lane;
    const uint4 my4 = mine < m ? spec[mine] : make_uint4(0xffffffffu, 0u, 0u, 0u);
    const float mysrc = (mine < m && a.reloc && a.checkOri) ? a.src_angle[mo + mine] : 0.f;
    uint32_t vm = __ballot_sync(0xffffffffu, my4.x != 0xffffffffu);
    while (vm) {
      const int j = __ffs((int)vm) - 1;
      vm &= vm - 1u;
      const int iMP = base + j;
      unsigned whi = __shfl_sync(0xffffffffu, my4.x, j);
      int bestIdx = (int)__shfl_sync(0xffffffffu, my4.z, j);
      float ang = __uint_as_float(__shfl_sync(0xffffffffu, my4.w, j));
      const float src = __shfl_sync(0xffffffffu, mysrc, j);
      if ((occ[bestIdx >> 5] >> (bestIdx & 31)) & 1u) {            // taken meanwhile: scan again (warp-uniform, rare)
        unsigned wlo;
        ordered_scan(a, p, iMP, occ, lane, whi, wlo);
        if (whi == 0xffffffffu) continue;
        bestIdx = items[wlo];
        ang = kps[bestIdx].angle;
      }
      if ((int)(whi >> 8) > a.th_dist) continue;
      if (lane == 0) {
        occ[bestIdx >> 5] |= 1u << (bestIdx & 31);
        assigned[bestIdx] = iMP;
        if (a.reloc && a.checkOri) s_hist[rot_bin30(src, ang)] += 1;
      }
      ++nm;
      __syncwarp();
    }
  }
  __syncwarp();
  if (a.reloc && a.checkOri) {                                  // rotation consistency, src/ORBmatcher.cc:1577-1596
    int m1 = 0, m2 = 0, m3 = 0, i1 = -1, i2 = -1, i3 = -1;      // ORBmatcher::ComputeThreeMaxima :1601-1642
    for (int i = 0; i < 30; ++i) {
      const int sc = s_hist[i];
      if (sc > m1) { m3 = m2; m2 = m1; m1 = sc; i3 = i2; i2 = i1; i1 = i; }
      else if (sc > m2) { m3 = m2; m2 = sc; i3 = i2; i2 = i; }
      else if (sc > m3) { m3 = sc; i3 = i; }
    }
    if ((float)m2 < __fmul_rn(0.1f, (float)m1)) { i2 = -1; i3 = -1; }
    else if ((float)m3 < __fmul_rn(0.1f, (float)m1)) i3 = -1;
    int drop = 0;
    for (int i = lane; i < n; i += 32) {
      const int mp = assigned[i];
      if (mp >= 0) {
        const int b = rot_bin30(a.src_angle[mo + mp], kps[i].angle);
        if (b != i1 && b != i2 && b != i3) { assigned[i] = -1; ++drop; }
      }
    }
    nm -= __reduce_add_sync(0xffffffffu, drop);
  }
  if (lane == 0) a.nm[p] = nm;
}

int launch_project_match_ordered(const ProjOrderedArgs& a, cudaStream_t st) {
  const size_t smem = (size_t)((a.cap + 31) / 32) * 4;
  if (a.n_problems <= 148 && !getenv("ORBX_PROJ_SERIAL")) k_project_match_ordered_par<<<a.n_problems, kOrdParWarps * 32, smem, st>>>(a);
  else k_project_match_ordered<<<a.n_problems, 32, smem, st>>>(a);
  return 1;
}

int launch_project_search(const ProjSearchArgs& a, cudaStream_t st) {
  k_project_search<<<dim3((unsigned)((a.mcap + 127) / 128), (unsigned)a.n_problems), 128, 0, st>>>(a);
  return 1;
}
int launch_scatter_matches(const int* best, const int* mp_index, const int* m, int mcap, int cap, int n_problems, int* vn, cudaStream_t st) {
  k_scatter_matches<<<dim3((unsigned)((mcap + 127) / 128), (unsigned)n_problems), 128, 0, st>>>(best, mp_index, m, mcap, cap, vn);
  return 1;
}
int launch_sim3_agree(const Sim3AgreeArgs& a, int n_problems, cudaStream_t st) {
  k_sim3_agree<<<dim3((unsigned)((a.cap + 127) / 128), (unsigned)n_problems), 128, 0, st>>>(a);
  return 1;
}

}  // namespace orbx
