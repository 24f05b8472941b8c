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
#include "ordered.cuh"

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
  if (a.valid && !a.valid[mo]) { *out = -1; return; }
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

// K16, latency form (one keyframe per call: LoopClosing::ComputeSim3 asks about one candidate pair, SearchInNeighbors fuses
// into one neighbour at a time).  A CTA of 1024 threads per problem; the target's CSR grid, positions, octaves, descriptors
// (and uRight for Fuse's chi-square gate) staged in shared memory.  Map points go through in blocks of 1024: a THREAD per
// map point works out its projection and search window (the double-precision part), the ones that have a window are listed,
// then a WARP per listed map point shares the window's candidates (for_each_in_columns) and keeps the least
// (distance, CSR position) - the reference's first least distance in its enumeration order.
constexpr int kPsLatThreads = 1024;

__host__ __device__ inline size_t ps_lat_bytes(int cap, bool with_ur) {
  size_t b = (size_t)(kPCells + 1 + 3) / 4 * 16 + (size_t)cap * (32 + 8 + 4) + (with_ur ? (size_t)cap * 4 : 0);
  b = (b + 15) & ~(size_t)15;
  return b + (size_t)kPsLatThreads * (20 + 2);              // the block's windows {u, v, ur, radius, packed} and its list
}

__device__ __forceinline__ void project_search_problem(const ProjSearchArgs& a, const int p, uint8_t* smem) {
  __shared__ int s_bcnt[kPsLatThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nth = blockDim.x, nwarps = nth >> 5;
  const int m = min(a.m[p], a.mcap);
  const size_t ko = (size_t)p * a.cap, mo = (size_t)p * a.mcap;
  const bool gate = (a.mode & 8) != 0;
  int* cs = reinterpret_cast<int*>(smem);                       // cell starts (a copy of the caller's, or built here: then shifted by one)
  uint32_t* sdesc = reinterpret_cast<uint32_t*>(smem + (size_t)(kPCells + 1 + 3) / 4 * 16);     // [cap][8] by keypoint
  float2* sxy = reinterpret_cast<float2*>(sdesc + (size_t)a.cap * 8);                           // [cap] in CSR order
  uint32_t* smeta = reinterpret_cast<uint32_t*>(sxy + a.cap);                                   // [cap] octave << 24 | keypoint, CSR order
  float* sur = reinterpret_cast<float*>(smeta + a.cap);                                         // [cap] uRight by keypoint (Fuse only)
  uint8_t* tail = smem + ((((size_t)(kPCells + 1 + 3) / 4 * 16 + (size_t)a.cap * (32 + 8 + 4) + (gate ? (size_t)a.cap * 4 : 0)) + 15) & ~(size_t)15);
  float4* swin = reinterpret_cast<float4*>(tail);                                               // [1024] u, v, ur, radius
  uint32_t* spack = reinterpret_cast<uint32_t*>(swin + kPsLatThreads);                          // [1024] level << 24 | cx0 << 18 | cx1 << 12 | cy0 << 6 | cy1
  uint16_t* slist = reinterpret_cast<uint16_t*>(spack + kPsLatThreads);                         // [1024] block-relative indices that have a window
  {
    const orbx_keypoint* kps = a.kps + ko;
    const float* gbs = a.grid_bounds;
    if (a.cell_start) {                                         // the caller's CSR grid (K10's output)
      const int* gcs = a.cell_start + (size_t)p * (kPCells + 1);
      const int* items = a.cell_items + ko;
      for (int i = tid; i <= kPCells; i += nth) cs[i] = gcs[i];
      const int n_items = min(gcs[kPCells], a.cap);
      for (int e = tid; e < n_items; e += nth) {
        const int idx = items[e];
        const orbx_keypoint kp = kps[idx];
        const int gx = (int)roundf(__fmul_rn(__fsub_rn(kp.x, gbs[0]), gbs[4])), gy = (int)roundf(__fmul_rn(__fsub_rn(kp.y, gbs[2]), gbs[5]));
        sxy[e] = make_float2(kp.x, kp.y);
        smeta[e] = ((uint32_t)kp.octave << 28) | ((uint32_t)(gx * kPGridRows + gy) << 16) | (uint32_t)idx;
      }
    } else {                                                    // no grid given: built here from all `cap` keypoint slots the caller filled (n)
      grid_sort_cta(kps, a.n ? min(a.n[p], a.cap) : a.cap, gbs[0], gbs[2], gbs[4], gbs[5], cs, sxy, smeta);
      cs += 1;
    }
    if (gate) for (int i = tid; i < a.cap; i += nth) sur[i] = a.u_right[ko + i];
    const uint4* src = reinterpret_cast<const uint4*>(a.desc + ko * 32);
    uint4* dst = reinterpret_cast<uint4*>(sdesc);
    for (int i = tid; i < a.cap * 2; i += nth) dst[i] = __ldg(src + i);
  }
  const float* xf = a.xf + (size_t)p * 27;                      // R1 (9) t1 (3) R2 (9) t2 (3) Ow (3)
  const float* kfb = a.kf_bounds + (size_t)p * 4;               // KeyFrame::mnMinX, mnMaxX, mnMinY, mnMaxY
  const float* gb = a.grid_bounds;                              // Frame's mnMinX, mnMaxX, mnMinY, mnMaxY, invW, invH
  const uint32_t* mpd = reinterpret_cast<const uint32_t*>(a.mp_desc + mo * 32);
  for (int base = 0; base < m; base += nth) {
    // a thread per map point of the block: projection, tests, window (the arithmetic of k_project_search, statement by statement)
    const int i = base + tid;
    bool ok = false;
    if (i < m && (!a.valid || a.valid[mo + i])) {
      const float xw[3] = {a.xw[(mo + i) * 3], a.xw[(mo + i) * 3 + 1], a.xw[(mo + i) * 3 + 2]};
      const float dmin = a.dmin[mo + i], dmax = a.dmax[mo + i];
      float nn[3] = {0.f, 0.f, 0.f};
      if (a.mode & 4) { nn[0] = a.normal[(mo + i) * 3]; nn[1] = a.normal[(mo + i) * 3 + 1]; nn[2] = a.normal[(mo + i) * 3 + 2]; }
      float pc[3], q[3];
      gemm3(xf, xf + 9, xw, pc);
      if (a.mode & 1) gemm3(xf + 12, xf + 21, pc, q); else { q[0] = pc[0]; q[1] = pc[1]; q[2] = pc[2]; }
      do {
        if (q[2] < 0.0f) break;
        const float invz = (a.mode & 16) ? (float)__ddiv_rn(1.0, (double)q[2]) : __fdiv_rn(1.0f, q[2]);
        const float x = __fmul_rn(q[0], invz), y = __fmul_rn(q[1], invz);
        const float u = __fadd_rn(__fmul_rn(a.fx, x), a.cx), v = __fadd_rn(__fmul_rn(a.fy, y), a.cy);
        if (!(u >= kfb[0] && u < kfb[1] && v >= kfb[2] && v < kfb[3])) break;
        const float ur = __fsub_rn(u, __fmul_rn(a.bf, invz));
        float PO[3] = {q[0], q[1], q[2]};
        if (a.mode & 2) { PO[0] = __fsub_rn(xw[0], xf[24]); PO[1] = __fsub_rn(xw[1], xf[25]); PO[2] = __fsub_rn(xw[2], xf[26]); }
        const double s2 = __dadd_rn(__dadd_rn(__dmul_rn((double)PO[0], (double)PO[0]), __dmul_rn((double)PO[1], (double)PO[1])), __dmul_rn((double)PO[2], (double)PO[2]));
        const float dist3D = (float)__dsqrt_rn(s2);
        if (dist3D < __fmul_rn(0.8f, dmin) || dist3D > __fmul_rn(1.2f, dmax)) break;
        if (a.mode & 4) {
          const double dot = __dadd_rn(__dadd_rn(__dmul_rn((double)PO[0], (double)nn[0]), __dmul_rn((double)PO[1], (double)nn[1])), __dmul_rn((double)PO[2], (double)nn[2]));
          if (dot < __dmul_rn(0.5, (double)dist3D)) break;
        }
        int level = (int)ceilf(__fdiv_rn(libm_logf(__fdiv_rn(dmax, dist3D)), a.log_scale));
        level = level < 0 ? 0 : (level >= a.nlevels ? a.nlevels - 1 : level);
        const float radius = __fmul_rn(a.th, a.scale[level]);
        const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, kfb[0]), radius), gb[4])));
        const int cx1 = min(kPGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, kfb[0]), radius), gb[4])));
        const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, kfb[2]), radius), gb[5])));
        const int cy1 = min(kPGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, kfb[2]), radius), gb[5])));
        if (cx0 >= kPGridCols || cx1 < 0 || cy0 >= kPGridRows || cy1 < 0) break;
        swin[tid] = make_float4(u, v, ur, radius);
        spack[tid] = ((uint32_t)level << 24) | ((uint32_t)cx0 << 18) | ((uint32_t)cx1 << 12) | ((uint32_t)cy0 << 6) | (uint32_t)cy1;
        ok = true;
      } while (false);
    }
    if (i < m && !ok) a.best[mo + i] = -1;
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) s_bcnt[warp] = __popc(bal);
    __syncthreads();                                            // (also: the staging above is complete)
    int before = 0, nlist = 0;
    for (int w = 0; w < nwarps; ++w) { const int c = s_bcnt[w]; before += w < warp ? c : 0; nlist += c; }
    if (ok) slist[before + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)tid;
    __syncthreads();
    // a warp per listed map point, the next one's descriptor in flight
    auto dword = [&](int c) { return (c < nlist && lane < 8) ? __ldg(mpd + (size_t)(base + slist[c]) * 8 + lane) : 0u; };
    uint32_t ndw = dword(warp);
    for (int c = warp; c < nlist; c += nwarps) {
      const uint32_t dw = ndw;
      ndw = dword(c + nwarps);
      const int r = slist[c];
      const float4 win = swin[r];
      const uint32_t pk = spack[r];
      const float u = win.x, v = win.y, ur = win.z, radius = win.w;
      const int level = (int)(pk >> 24), cx0 = (int)((pk >> 18) & 63u), cx1 = (int)((pk >> 12) & 63u), cy0 = (int)((pk >> 6) & 63u), cy1 = (int)(pk & 63u);
      uint32_t qd[8];
#pragma unroll
      for (int w = 0; w < 8; ++w) qd[w] = __shfl_sync(0xffffffffu, dw, w);
      unsigned bd = 0xffffffffu, be = 0xffffffffu;             // least (distance, cell << 16 | keypoint)
      for_each_in_columns(cx0, cx1, lane,
        [&](int ix, int& e_lo, int& e_hi) { e_lo = cs[ix * kPGridRows + cy0]; e_hi = cs[ix * kPGridRows + cy1 + 1]; },
        [&](int e, int) {
          const uint32_t meta = smeta[e];
          const int idx = (int)(meta & 0xFFFFu), oct = (int)(meta >> 28);
          const uint32_t ck = meta & 0x0FFFFFFFu;               // cell << 16 | keypoint: the reference's enumeration order
          const float2 pt = sxy[e];
          if (!(fabsf(__fsub_rn(pt.x, u)) < radius && fabsf(__fsub_rn(pt.y, v)) < radius)) return;
          if (oct < level - 1 || oct > level) return;
          if (gate) {
            const float ex = __fsub_rn(u, pt.x), ey = __fsub_rn(v, pt.y);
            const float kr = sur[idx];
            if (kr >= 0.f) {
              const float er = __fsub_rn(ur, kr);
              const float e2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(er, er));
              if ((double)__fmul_rn(e2, a.inv_sigma2[oct]) > 7.8) return;
            } else {
              const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
              if ((double)__fmul_rn(e2, a.inv_sigma2[oct]) > 5.99) return;
            }
          }
          const uint32_t* g = sdesc + (size_t)idx * 8;
          const uint4 b0 = *reinterpret_cast<const uint4*>(g), b1 = *reinterpret_cast<const uint4*>(g + 4);
          const unsigned d = __popc(qd[0] ^ b0.x) + __popc(qd[1] ^ b0.y) + __popc(qd[2] ^ b0.z) + __popc(qd[3] ^ b0.w) + __popc(qd[4] ^ b1.x) +
                             __popc(qd[5] ^ b1.y) + __popc(qd[6] ^ b1.z) + __popc(qd[7] ^ b1.w);
          if (d < bd || (d == bd && ck < be)) { bd = d; be = ck; }
        });
      const unsigned wd = __reduce_min_sync(0xffffffffu, bd);
      const unsigned we = __reduce_min_sync(0xffffffffu, bd == wd ? be : 0xffffffffu);
      if (lane == 0) a.best[mo + base + r] = (wd != 0xffffffffu && (int)wd <= a.th_dist) ? (int)(we & 0xFFFFu) : -1;
    }
    __syncthreads();                                            // the block's windows are free again
  }
}

__global__ void __launch_bounds__(kPsLatThreads) k_project_search_lat(ProjSearchArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  project_search_problem(a, blockIdx.x, smem);
}

// SearchBySim3 in ONE launch: CTA 2p searches keyframe pair p's map points 1 -> 2, CTA 2p + 1 the other way; each scatters its
// winners to the per-keypoint table vnMatch (src/ORBmatcher.cc:1215,1295), and the CTA of a pair that finishes second does the
// agreement check (:1303-1323).  ticket[p] starts at -1 (it shares the 0xFF memset of the vnMatch tables).
struct Sim3LatArgs { ProjSearchArgs pass[2]; const int* mp_index[2]; int* vn[2]; int* ticket; const int* n1; int* matches12; int* nfound; };

__global__ void __launch_bounds__(kPsLatThreads) k_sim3_lat(const __grid_constant__ Sim3LatArgs A) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ int s_last, s_cnt;
  const int p = blockIdx.x >> 1, dir = blockIdx.x & 1, tid = threadIdx.x;
  const ProjSearchArgs& a = A.pass[dir];
  project_search_problem(a, p, smem);                          // ends with a block barrier: a.best of this problem is complete (written by this CTA)
  const int m = min(a.m[p], a.mcap), cap = a.cap;
  int* vn = A.vn[dir] + (size_t)p * cap;
  for (int i = tid; i < m; i += blockDim.x) {
    const int kpi = A.mp_index[dir][(size_t)p * a.mcap + i];
    if (kpi >= 0 && kpi < cap) vn[kpi] = a.best[(size_t)p * a.mcap + i];
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) { s_last = atomicAdd(A.ticket + p, 1) == 0; s_cnt = 0; }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int* v1 = A.vn[0] + (size_t)p * cap; const int* v2 = A.vn[1] + (size_t)p * cap;
  int found = 0;
  for (int i1 = tid; i1 < cap; i1 += blockDim.x) {
    const int idx2 = __ldcg(v1 + i1);
    int res = -1;
    if (i1 < A.n1[p] && idx2 >= 0 && idx2 < cap && __ldcg(v2 + idx2) == i1) { res = idx2; ++found; }
    A.matches12[(size_t)p * cap + i1] = res;
  }
  if (found) atomicAdd(&s_cnt, found);
  __syncthreads();
  if (tid == 0) A.nfound[p] = s_cnt;
}

int launch_sim3_lat(const ProjSearchArgs& a0, const ProjSearchArgs& a1, const int* idx0, const int* idx1, int* vn0, int* vn1, int* ticket, const int* n1,
                    int* matches12, int* nfound, cudaStream_t st) {
  const size_t lat = ps_lat_bytes(a0.cap, false);
  if (!(lat <= 200 * 1024 && a0.cap <= 65536) || getenv("ORBX_PROJ_SERIAL")) return 0;      // the caller takes the five-launch route
  Sim3LatArgs A;
  A.pass[0] = a0; A.pass[1] = a1; A.mp_index[0] = idx0; A.mp_index[1] = idx1; A.vn[0] = vn0; A.vn[1] = vn1; A.ticket = ticket; A.n1 = n1;
  A.matches12 = matches12; A.nfound = nfound;
  cudaFuncSetAttribute(k_sim3_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lat);
  k_sim3_lat<<<a0.n_problems * 2, kPsLatThreads, lat, st>>>(A);
  return 1;
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

// The projection of ONE map point of K17 and its search window (every lane computes the same): false when the point is skipped.
struct OrdWin { float u, v, radius; int level, lvl_hi, cx0, cx1, cy0, cy1; };
struct OrdMP { int valid; float xw[3], dmin, dmax, nn[3]; uint32_t qd[8]; };

// every field of map point iMP, requested together: one memory round trip per map point, not one per test (iMP >= m: an invalid point)
template <bool DESC = true>
__device__ __forceinline__ OrdMP ordered_fetch(const ProjOrderedArgs& a, int p, int iMP, int m) {
  const size_t mo = (size_t)p * a.mcap;
  OrdMP q;
  const bool ok = iMP < m;
  q.valid = ok ? a.valid[mo + iMP] : 0;
#pragma unroll
  for (int c = 0; c < 3; ++c) q.xw[c] = ok ? a.xw[(mo + iMP) * 3 + c] : 0.f;
  q.dmin = ok ? a.dmin[mo + iMP] : 0.f; q.dmax = ok ? a.dmax[mo + iMP] : 0.f;
#pragma unroll
  for (int c = 0; c < 3; ++c) q.nn[c] = (ok && !a.reloc) ? a.normal[(mo + iMP) * 3 + c] : 0.f;
#pragma unroll
  for (int c = 0; c < 8; ++c) q.qd[c] = (DESC && ok) ? __ldg(reinterpret_cast<const uint32_t*>(a.mp_desc + (mo + iMP) * 32) + c) : 0u;
  return q;
}

__device__ __forceinline__ bool ordered_window(const ProjOrderedArgs& a, int p, const OrdMP& q, OrdWin& w) {
  const float* xf = a.xf + (size_t)p * 15;
  const float* kb = a.tgt_bounds + (size_t)p * 4;
  const float* gb = a.grid_bounds;
  const float* xw = q.xw; const float* nn = q.nn;
  const float dmin = q.dmin, dmax = q.dmax;
  if (!q.valid) return false;
  float pc[3];
  gemm3(xf, xf + 9, xw, pc);
  float u, v;
  if (!a.reloc) {
    if (pc[2] < 0.0f) return false;
    const float invz = __fdiv_rn(1.0f, pc[2]);
    u = __fadd_rn(__fmul_rn(a.fx, __fmul_rn(pc[0], invz)), a.cx); v = __fadd_rn(__fmul_rn(a.fy, __fmul_rn(pc[1], invz)), a.cy);
    if (!(u >= kb[0] && u < kb[1] && v >= kb[2] && v < kb[3])) return false;
  } else {
    const float invzc = (float)__ddiv_rn(1.0, (double)pc[2]);
    u = __fadd_rn(__fmul_rn(__fmul_rn(a.fx, pc[0]), invzc), a.cx); v = __fadd_rn(__fmul_rn(__fmul_rn(a.fy, pc[1]), invzc), a.cy);
    if (u < kb[0] || u > kb[1] || v < kb[2] || v > kb[3]) return false;
  }
  const float PO[3] = {__fsub_rn(xw[0], xf[12]), __fsub_rn(xw[1], xf[13]), __fsub_rn(xw[2], xf[14])};
  const double s2 = __dadd_rn(__dadd_rn(__dmul_rn((double)PO[0], (double)PO[0]), __dmul_rn((double)PO[1], (double)PO[1])), __dmul_rn((double)PO[2], (double)PO[2]));
  const float dist3D = (float)__dsqrt_rn(s2);
  if (dist3D < __fmul_rn(0.8f, dmin) || dist3D > __fmul_rn(1.2f, dmax)) return false;
  if (!a.reloc) {
    const double dot = __dadd_rn(__dadd_rn(__dmul_rn((double)PO[0], (double)nn[0]), __dmul_rn((double)PO[1], (double)nn[1])), __dmul_rn((double)PO[2], (double)nn[2]));
    if (dot < __dmul_rn(0.5, (double)dist3D)) return false;
  }
  int level = (int)ceilf(__fdiv_rn(libm_logf(__fdiv_rn(dmax, dist3D)), a.log_scale));
  level = level < 0 ? 0 : (level >= a.nlevels ? a.nlevels - 1 : level);
  w.level = level; w.lvl_hi = a.reloc ? level + 1 : level;
  const float radius = __fmul_rn(a.th, a.scale[level]);
  w.u = u; w.v = v; w.radius = radius;
  w.cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, kb[0]), radius), gb[4])));
  w.cx1 = min(kPGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, kb[0]), radius), gb[4])));
  w.cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, kb[2]), radius), gb[5])));
  w.cy1 = min(kPGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, kb[2]), radius), gb[5])));
  return !(w.cx0 >= kPGridCols || w.cx1 < 0 || w.cy0 >= kPGridRows || w.cy1 < 0);
}

// The candidate scan of ONE map point of K17 (warp-wide): the least (distance << 8 | cell column, CSR position)
// among the keypoints that are free in `occ`, ~0 when the point is skipped or has no candidate.
__device__ __forceinline__ void ordered_scan(const ProjOrderedArgs& a, int p, int iMP, const uint32_t* occ, int lane, unsigned& whi, unsigned& wlo) {
  whi = wlo = 0xffffffffu;
  OrdWin w;
  const OrdMP mp = ordered_fetch(a, p, iMP, iMP + 1);
  if (!ordered_window(a, p, mp, w)) return;
  const size_t ko = (size_t)p * a.cap;
  const orbx_keypoint* kps = a.kps + ko;
  const int* cs = a.cell_start + (size_t)p * (kPCells + 1);
  const int* items = a.cell_items + ko;
  unsigned hi = 0xffffffffu, lo = 0xffffffffu;                // (distance << 8 | cell column), CSR position
  for (int ix = w.cx0; ix <= w.cx1; ++ix) {
    const int e_hi = cs[ix * kPGridRows + w.cy1 + 1];
    for (int e = cs[ix * kPGridRows + w.cy0] + lane; e < e_hi; e += 32) {
      const int idx = items[e];
      if ((occ[idx >> 5] >> (idx & 31)) & 1u) continue;
      const orbx_keypoint kp = kps[idx];
      if (!(fabsf(__fsub_rn(kp.x, w.u)) < w.radius && fabsf(__fsub_rn(kp.y, w.v)) < w.radius)) continue;
      if (kp.octave < w.level - 1 || kp.octave > w.lvl_hi) continue;
      const unsigned h = ((unsigned)ham8p(mp.qd, reinterpret_cast<const uint32_t*>(a.desc + (ko + idx) * 32)) << 8) | (unsigned)ix;
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
// time): a CTA of 32 warps per problem, and - as in K8's latency form - nothing in the loops leaves the SM.  The target's
// CSR grid arrives as it is (cell starts; positions / octave / index in CSR order), descriptors and angles by keypoint
// index, the occupancy as a bitset, the output `assigned` stays in shared memory until the end.
//   Phase A: the winner of every map point against the keypoints that are occupied ON ENTRY, map points strided over the warps.
//   Phase B: warp 0 walks the map points in list order (hazard H6); a phase-A winner stands unless its keypoint was taken
//   meanwhile by an earlier map point of this call - only then the point is scanned again against the live occupancy
//   (taking away a keypoint that is not a point's winner cannot change its winner).
constexpr int kOrdParThreads = 1024;

struct OrdStage { int* cs; float2* sxy; uint32_t* smeta; uint32_t* sdesc; float* sang; int* sasg; int* stag; uint32_t* occ; };

__host__ __device__ inline size_t ord_stage_bytes(int cap) {
  size_t b = (size_t)(kPCells + 1 + 3) / 4 * 16;            // cell starts, padded to 16 bytes
  b += (size_t)cap * (8 + 4 + 32 + 4 + 4 + 4) + (size_t)((cap + 31) / 32) * 4;
  return (b + 15) & ~(size_t)15;
}

// The candidate list of a map point with search window `win` ({u, v, radius, packed level / cell window}, ordered_pack) and
// descriptor qd against S.occ as it is now (warp-wide, every lane returns the same):
// {l0 | complete << 31, l1, l2, l3}, l = distance << 16 | keypoint in ascending (distance, cell column, CSR position) = the
// reference's first-least over its enumeration; kNoKey = none.
__device__ __forceinline__ uint32_t ordered_pack(const OrdWin& w) {
  return 0x80000000u | ((uint32_t)w.level << 24) | ((uint32_t)w.cx0 << 18) | ((uint32_t)w.cx1 << 12) | ((uint32_t)w.cy0 << 6) | (uint32_t)w.cy1;
}

__device__ __forceinline__ uint4 ordered_scan_staged(const ProjOrderedArgs& a, const uint4 win, const uint32_t (&qd)[8], const OrdStage& S, int lane) {
  const uint4 none = make_uint4(0xffffffffu, kNoKey, kNoKey, kNoKey);
  if (!(win.w >> 31)) return none;
  OrdWin w;
  w.u = __uint_as_float(win.x); w.v = __uint_as_float(win.y); w.radius = __uint_as_float(win.z);
  w.level = (int)((win.w >> 24) & 15u); w.lvl_hi = a.reloc ? w.level + 1 : w.level;
  w.cx0 = (int)((win.w >> 18) & 63u); w.cx1 = (int)((win.w >> 12) & 63u); w.cy0 = (int)((win.w >> 6) & 63u); w.cy1 = (int)(win.w & 63u);
  unsigned long long k1 = ~0ull, k2 = ~0ull;                  // distance << 32 | cell << 16 | keypoint
  int cnt = 0;
  for_each_in_columns(w.cx0, w.cx1, lane,
    [&](int ix, int& e_lo, int& e_hi) { e_lo = S.cs[ix * kPGridRows + w.cy0]; e_hi = S.cs[ix * kPGridRows + w.cy1 + 1]; },
    [&](int e, int) {
      const uint32_t meta = S.smeta[e];
      const int idx = (int)(meta & 0xFFFFu), oct = (int)(meta >> 28);
      if ((S.occ[idx >> 5] >> (idx & 31)) & 1u) return;
      const float2 pt = S.sxy[e];
      if (!(fabsf(__fsub_rn(pt.x, w.u)) < w.radius && fabsf(__fsub_rn(pt.y, w.v)) < w.radius)) return;
      if (oct < w.level - 1 || oct > w.lvl_hi) return;
      const uint32_t* g = S.sdesc + (size_t)idx * 8;
      const uint4 b0 = *reinterpret_cast<const uint4*>(g), b1 = *reinterpret_cast<const uint4*>(g + 4);
      const int d = __popc(qd[0] ^ b0.x) + __popc(qd[1] ^ b0.y) + __popc(qd[2] ^ b0.z) + __popc(qd[3] ^ b0.w) + __popc(qd[4] ^ b1.x) +
                    __popc(qd[5] ^ b1.y) + __popc(qd[6] ^ b1.z) + __popc(qd[7] ^ b1.w);
      const unsigned long long key = ((unsigned long long)(unsigned)d << 32) | (meta & 0x0FFFFFFFu);   // (distance, cell, index): the reference's order
      ++cnt;
      if (key < k1) { k2 = k1; k1 = key; }
      else if (key < k2) k2 = key;
    });
  unsigned hi[4], lo[4];
  bool complete;
  warp_top4(k1, k2, cnt, hi, lo, complete);
  uint32_t l[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) l[k] = hi[k] == 0xffffffffu ? kNoKey : ((hi[k] << 16) | (lo[k] & 0xFFFFu));
  return make_uint4(l[0] | (complete ? 0x80000000u : 0u), l[1], l[2], l[3]);
}

__global__ void __launch_bounds__(kOrdParThreads) k_project_match_ordered_par(ProjOrderedArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ int s_hist[30];
  __shared__ int s_nm;
  __shared__ int s_bcnt[kOrdParThreads / 32];
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nth = blockDim.x, nwarps = nth >> 5;
  const int n = min(a.n[p], a.cap), m = min(a.m[p], a.mcap);
  const size_t ko = (size_t)p * a.cap, mo = (size_t)p * a.mcap;
  const orbx_keypoint* kps = a.kps + ko;
  int* assigned = a.assigned + ko;
  uint4* spec = a.spec + mo;
  OrdStage S;
  ORBX_CLK(t0);
  S.cs = reinterpret_cast<int*>(smem);
  S.sdesc = reinterpret_cast<uint32_t*>(smem + (size_t)(kPCells + 1 + 3) / 4 * 16);      // 16-byte aligned for any cap
  S.sxy = reinterpret_cast<float2*>(S.sdesc + (size_t)a.cap * 8);
  S.smeta = reinterpret_cast<uint32_t*>(S.sxy + a.cap);
  S.sang = reinterpret_cast<float*>(S.smeta + a.cap);
  S.sasg = reinterpret_cast<int*>(S.sang + a.cap);
  S.stag = S.sasg + a.cap;
  S.occ = reinterpret_cast<uint32_t*>(S.stag + a.cap);
  uint16_t* slist = reinterpret_cast<uint16_t*>(smem + ord_stage_bytes(a.cap));   // [mcap] the map points that have a search window, in list order
  const float* gbs = a.grid_bounds;
  if (a.cell_start) {                                           // the caller's CSR grid (K10's output): copied in its order
    const int* cs = a.cell_start + (size_t)p * (kPCells + 1);
    const int* items = a.cell_items + ko;
    for (int i = tid; i <= kPCells; i += nth) S.cs[i] = cs[i];
    const int n_items = min(cs[kPCells], a.cap);
    for (int e = tid; e < n_items; e += nth) {
      const int idx = items[e];
      const orbx_keypoint kp = kps[idx];
      const int gx = (int)roundf(__fmul_rn(__fsub_rn(kp.x, gbs[0]), gbs[4])), gy = (int)roundf(__fmul_rn(__fsub_rn(kp.y, gbs[2]), gbs[5]));
      S.sxy[e] = make_float2(kp.x, kp.y);
      S.smeta[e] = ((uint32_t)kp.octave << 28) | ((uint32_t)(gx * kPGridRows + gy) << 16) | (uint32_t)idx;
    }
  } else {                                                      // no grid given: built here (cell = PosInGrid of the keypoint)
    grid_sort_cta(kps, n, gbs[0], gbs[2], gbs[4], gbs[5], S.cs, S.sxy, S.smeta);
    S.cs += 1;
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(a.desc + ko * 32);
    uint4* dst = reinterpret_cast<uint4*>(S.sdesc);
    for (int i = tid; i < n * 2; i += nth) dst[i] = __ldg(src + i);
    for (int i = tid; i < a.cap; i += nth) { S.sasg[i] = -1; S.stag[i] = 0; S.sang[i] = i < n ? kps[i].angle : 0.f; }
    for (int base = warp * 32; base < a.cap; base += nth) {
      const int i = base + lane;
      const uint32_t b = __ballot_sync(0xffffffffu, i < n && a.occupied[ko + i]);
      if (lane == 0) S.occ[base >> 5] = b;
    }
    if (tid < 30) s_hist[tid] = 0;
  }
  // pre-pass, one THREAD per map point: the projection and the search window (double-precision square root, logarithm and
  // divisions - a few thousand cycles of latency that a warp per map point would pay 32 lanes wide, once per map point in
  // turn), left in global scratch as 16 bytes; the map points that have a window are listed in order in shared memory
  uint4* wins = a.spec + (size_t)a.n_problems * a.mcap + mo;
  int nlist = 0;
  for (int base = 0; base < m; base += nth) {
    const int i = base + tid;
    OrdWin w;
    const bool ok = i < m && ordered_window(a, p, ordered_fetch<false>(a, p, i, m), w);
    if (ok) wins[i] = make_uint4(__float_as_uint(w.u), __float_as_uint(w.v), __float_as_uint(w.radius), ordered_pack(w));
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) s_bcnt[warp] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
    for (int q = 0; q < nwarps; ++q) { const int c = s_bcnt[q]; before += q < warp ? c : 0; total += c; }
    if (ok) slist[nlist + before + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)i;
    nlist += total;
    __syncthreads();
  }
  ORBX_CLK(t1);
  // phase A (occ is only read): a warp per listed map point, the next one's window and descriptor in flight
  const uint32_t* mpd = reinterpret_cast<const uint32_t*>(a.mp_desc + mo * 32);
  auto fetch = [&](int c, uint4& win, uint32_t& dw) {
    const bool ok = c < nlist;
    const int i = ok ? slist[c] : 0;
    win = ok ? wins[i] : make_uint4(0u, 0u, 0u, 0u);
    dw = (ok && lane < 8) ? __ldg(mpd + (size_t)i * 8 + lane) : 0u;
  };
  {
    uint4 nwin; uint32_t ndw;
    fetch(warp, nwin, ndw);
    for (int c = warp; c < nlist; c += nwarps) {
      const uint4 win = nwin; const uint32_t dw = ndw;
      fetch(c + nwarps, nwin, ndw);
      uint32_t qd[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) qd[q] = __shfl_sync(0xffffffffu, dw, q);
      const uint4 l = ordered_scan_staged(a, win, qd, S, lane);
      if (lane == 0) spec[c] = l;                                        // by list position
    }
  }
  __syncthreads();
  ORBX_CLK(t2);
  // phase B (see ordered.cuh): one key per map point - the first of its list no earlier map point has taken; every match claims
  if (warp == 0) {
    int nm = 0;
    int epoch = 0;
    ORBX_STAT(int dbg_serial = 0; int dbg_iters = 0;)
    const uint4 nokeys = make_uint4(0xffffffffu, kNoKey, kNoKey, kNoKey);
    const bool ori = a.reloc && a.checkOri;
    uint4 nxt = lane < nlist ? spec[lane] : nokeys;
    float nsrc = (lane < nlist && ori) ? a.src_angle[mo + slist[lane]] : 0.f;
    for (int base = 0; base < nlist; base += 32) {
      const int c = base + lane;
      const int iMP = c < nlist ? slist[c] : 0;
      const uint4 my4 = nxt;
      const float src = nsrc;
      nxt = c + 32 < nlist ? spec[c + 32] : nokeys;                       // the next round's inputs are on their way while this one is resolved
      nsrc = (c + 32 < nlist && ori) ? a.src_angle[mo + slist[c + 32]] : 0.f;
      const bool complete = my4.x >> 31;
      const uint32_t key[4] = {my4.x & 0x7FFFFFFFu, my4.y, my4.z, my4.w};
      uint32_t todo = __ballot_sync(0xffffffffu, key[0] != kNoKey);      // lanes whose map point is still to be decided
      while (todo) {
        const bool live = (todo >> lane) & 1u;
        unsigned hitP = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int idx = (int)(key[k] & 0xFFFFu);
          if (key[k] != kNoKey && ((S.occ[idx >> 5] >> (idx & 31)) & 1u)) hitP |= 1u << k;
        }
        bool acc = false, starved = false;
        int bestIdx = -1;
        ORBX_STAT(dbg_iters +=) order_round(key, hitP, S.stag, epoch, lane, [&](unsigned hit) {
          uint32_t k1 = kNoKey;
#pragma unroll
          for (int k = 3; k >= 0; --k)
            if (key[k] != kNoKey && !((hit >> k) & 1u)) k1 = key[k];
          starved = live && !complete && k1 == kNoKey;                   // the list ran out of free keys
          acc = live && k1 != kNoKey && (int)(k1 >> 16) <= a.th_dist;
          bestIdx = acc ? (int)(k1 & 0xFFFFu) : -1;
          return bestIdx;
        });
        // lanes below the first starved one are final (a lane depends on earlier lanes only): they commit together
        const uint32_t sm = __ballot_sync(0xffffffffu, starved);
        const int jstar = sm ? __ffs((int)sm) - 1 : 32;
        const bool commit = acc && lane < jstar;
        if (commit) {
          S.sasg[bestIdx] = iMP;                                         // claims are exclusive here: one writer per keypoint
          atomicOr(&S.occ[bestIdx >> 5], 1u << (bestIdx & 31));
          if (ori) atomicAdd(&s_hist[rot_bin30(src, S.sang[bestIdx])], 1);
        }
        nm += __popc(__ballot_sync(0xffffffffu, commit));
        __syncwarp();
        if (!sm) break;
        // the starved map point is scanned again against the live occupancy (exact), then the lanes after it go round again
        ORBX_STAT(++dbg_serial;)
        const int ij = __shfl_sync(0xffffffffu, iMP, jstar);
        const uint32_t dwj = lane < 8 ? __ldg(mpd + (size_t)ij * 8 + lane) : 0u;
        uint32_t qd[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) qd[q] = __shfl_sync(0xffffffffu, dwj, q);
        const uint4 l = ordered_scan_staged(a, wins[ij], qd, S, lane);
        const uint32_t k1 = l.x & 0x7FFFFFFFu;
        if (k1 != kNoKey && (int)(k1 >> 16) <= a.th_dist) {
          const int bi = (int)(k1 & 0xFFFFu);
          const float sj = __shfl_sync(0xffffffffu, src, jstar);
          if (lane == 0) {
            S.occ[bi >> 5] |= 1u << (bi & 31);
            S.sasg[bi] = ij;
            if (ori) s_hist[rot_bin30(sj, S.sang[bi])] += 1;
          }
          ++nm;
        }
        __syncwarp();
        todo &= ~((2u << jstar) - 1u);
      }
    }
    if (lane == 0) s_nm = nm;
    ORBX_STAT(if (lane == 0 && blockIdx.x == 0) printf("K17 phase B: %d re-scans, %d fixpoint iterations, nm %d of %d listed of %d\n", dbg_serial, dbg_iters, nm, nlist, m);)
  }
  __syncthreads();
  ORBX_CLK(t3);
  if (a.reloc && a.checkOri) {                                  // rotation consistency, src/ORBmatcher.cc:1577-1596, whole CTA
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
    for (int i = tid; i < n; i += nth) {
      const int mp = S.sasg[i];
      if (mp >= 0) {
        const int b = rot_bin30(a.src_angle[mo + mp], S.sang[i]);
        if (b != i1 && b != i2 && b != i3) { S.sasg[i] = -1; ++drop; }
      }
    }
    if (drop) atomicSub(&s_nm, drop);
    __syncthreads();
  }
  for (int i = tid; i < a.cap; i += nth) assigned[i] = S.sasg[i];
  if (tid == 0) a.nm[p] = s_nm;
  ORBX_CLK(t4);
  ORBX_CLK_PRINT("k_project_match_ordered_par stage / A / B / rotation + flush", t0, t1, t2, t3, t4);
}

int launch_project_match_ordered(const ProjOrderedArgs& a, cudaStream_t st) {
  const size_t smem = (size_t)((a.cap + 31) / 32) * 4, par = ord_stage_bytes(a.cap) + (size_t)a.mcap * 2;
  if (par <= 200 * 1024 && a.cap <= 65536 && a.mcap <= 65536 && !getenv("ORBX_PROJ_SERIAL")) {
    cudaFuncSetAttribute(k_project_match_ordered_par, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)par);
    k_project_match_ordered_par<<<a.n_problems, kOrdParThreads, par, st>>>(a);
  } else {
    k_project_match_ordered<<<a.n_problems, 32, smem, st>>>(a);
  }
  return 1;
}

// do the launchers below take the small forms, which walk the caller's CSR grid in global memory (the 1024-thread forms can build it)?
bool project_search_needs_grid(int cap, bool gate) { return !(ps_lat_bytes(cap, gate) <= 200 * 1024 && cap <= 65536) || getenv("ORBX_PROJ_SERIAL"); }
bool project_ordered_needs_grid(int cap, int mcap) {
  return !(ord_stage_bytes(cap) + (size_t)mcap * 2 <= 200 * 1024 && cap <= 65536 && mcap <= 65536) || getenv("ORBX_PROJ_SERIAL");
}

int launch_project_search(const ProjSearchArgs& a, cudaStream_t st) {
  const size_t lat = ps_lat_bytes(a.cap, (a.mode & 8) != 0);
  if (lat <= 200 * 1024 && a.cap <= 65536 && !getenv("ORBX_PROJ_SERIAL")) {
    cudaFuncSetAttribute(k_project_search_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lat);
    k_project_search_lat<<<a.n_problems, kPsLatThreads, lat, st>>>(a);
  } else {
    k_project_search<<<dim3((unsigned)((a.mcap + 127) / 128), (unsigned)a.n_problems), 128, 0, st>>>(a);
  }
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
