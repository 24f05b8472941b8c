// K5 — stereo matching for rectified pairs, batched over frames.
//
// Replaces Frame::ComputeStereoMatches (src/Frame.cc:466-640).
//   k_stereo_rows  : one CTA per pair sorts the right keypoints by (octave, row) into 16-byte records and a
//                    table of bin starts (the reference's per-row table vRowIndices, :476-493, turned inside out).
//   k_stereo_match : one warp per left keypoint, no block-level step.
//     (i)  descriptor search.  The reference scans the left keypoint's row of vRowIndices in ascending iR
//          with a strict "<" (:522-548).  Equivalent: among all iR with floor(yR-r) <= (int)vL <= ceil(yR+r),
//          |octave difference| <= 1 and uL-maxD <= uR <= uL, take the minimum of (distance, iR)
//          lexicographically.  The candidates are the three contiguous runs (one per octave) of rows that can
//          reach the keypoint's row; one record per lane, 4x __popcll, one REDUX min over the packed key.
//     (ii) 11x11 SAD over 11 shifts on the pyramid level of the left keypoint (:552-592), integer
//          arithmetic (the reference's float L1 norm of integer-valued patches is exact),
//          first minimum wins; parabola refinement in the reference's float operation order
//          with un-fused intrinsics (:597-622).
//   k_stereo_median: one CTA per frame; the median SAD (element size/2 of the sorted list, :627)
//     by bisection over the value range, then the 1.5*1.4*median cut (:628-639).
#include "orbx_internal.h"

namespace orbx {

constexpr int kStWarps = 2;     // warps are independent; small CTAs keep the tail short (8 / 4 / 2 / 1 warps: 0.384 / 0.362 / 0.355 / 0.402 ms per 256 pairs)

__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint4* __restrict__ b) {
  const uint4 b0 = __ldg(b), b1 = __ldg(b + 1);
  const unsigned long long x0 = ((unsigned long long)(a0.y ^ b0.y) << 32) | (a0.x ^ b0.x);
  const unsigned long long x1 = ((unsigned long long)(a0.w ^ b0.w) << 32) | (a0.z ^ b0.z);
  const unsigned long long x2 = ((unsigned long long)(a1.y ^ b1.y) << 32) | (a1.x ^ b1.x);
  const unsigned long long x3 = ((unsigned long long)(a1.w ^ b1.w) << 32) | (a1.z ^ b1.z);
  return __popcll(x0) + __popcll(x1) + __popcll(x2) + __popcll(x3);
}

constexpr int kLP = 12, kRP = 24;       // row pitch of the left patch / right strip in shared memory (words per row: 3 / 6)

// K5a: the right keypoints of one pair, counting-sorted by (octave, (int)y) into 16-byte records
// {x, floor(y-r) | ceil(y+r) << 16, iR, 0} (r = 2 * scale[octave], src/Frame.cc:483-493) plus the table of bin
// starts T[octave * H + row].  This is the reference's vRowIndices turned inside out: a left keypoint on row v
// of level l reads, for each octave l-1 .. l+1, the one contiguous run of rows that can reach v.  The order
// inside a bin is irrelevant because the winner is the lexicographic minimum of (distance, iR).
constexpr int kRowThreads = 1024;
__global__ void __launch_bounds__(kRowThreads) k_stereo_rows(StereoArgs a) {
  __shared__ int s_part[kRowThreads / 32];
  const int img = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int slotR = a.slotR0 + img;
  const int nR = a.R.nkp[slotR];
  const int H = a.R.geom[0].h, nb = a.R.nlevels * H;
  int* T = a.tab + (size_t)img * a.tab_stride;           // nb + 2 entries; T[b] = start of bin b when the kernel ends
  uint4* rec = a.rec + (size_t)img * a.R.kp_cap;
  const orbx_keypoint* kR = a.R.kps + (size_t)slotR * a.R.kp_cap;
  for (int i = tid; i < nb + 2; i += kRowThreads) T[i] = 0;
  __syncthreads();
  for (int i = tid; i < nR; i += kRowThreads) {
    const int b = kR[i].octave * H + min(max((int)kR[i].y, 0), H - 1);
    atomicAdd(&T[b + 2], 1);
  }
  __syncthreads();
  // inclusive scan of T[1 .. nb+1]: T[b+1] becomes the start of bin b, advanced to its end by the scatter below
  const int len = nb + 1, per = (len + kRowThreads - 1) / kRowThreads;
  const int lo = 1 + tid * per, hi = min(lo + per, 1 + len);
  int sum = 0;
  for (int i = lo; i < hi; ++i) sum += T[i];
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_part[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int v = s_part[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    s_part[lane] = v;
  }
  __syncthreads();
  int run = incl - sum + (warp ? s_part[warp - 1] : 0);
  for (int i = lo; i < hi; ++i) { run += T[i]; T[i] = run; }
  __syncthreads();
  for (int i = tid; i < nR; i += kRowThreads) {
    const orbx_keypoint kp = kR[i];
    const int b = kp.octave * H + min(max((int)kp.y, 0), H - 1);
    const int pos = atomicAdd(&T[b + 1], 1);
    const float r = __fmul_rn(2.0f, a.L.geom[kp.octave].scale);                     // :487
    const int ymax = (int)ceilf(__fadd_rn(kp.y, r)), ymin = (int)floorf(__fsub_rn(kp.y, r));
    rec[pos] = make_uint4(__float_as_uint(kp.x), ((unsigned)ymin & 0xffffu) | ((unsigned)ymax << 16), (unsigned)i, 0u);
  }
}

// K5b: one warp per left keypoint, no block-level step.
__global__ void __launch_bounds__(kStWarps * 32, 24) k_stereo_match(StereoArgs a, int ctas_per_img) {
  __shared__ __align__(16) uint8_t s_patch[kStWarps * 416];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int img = blockIdx.x / ctas_per_img;
  const int iL = (blockIdx.x - img * ctas_per_img) * kStWarps + warp;
  const int slotL = a.slotL0 + img, slotR = a.slotR0 + img;
  if (iL >= a.L.nkp[slotL]) {            // rows are kp_cap long: the entries behind the last keypoint read "no match" as well
    if (lane == 0 && iL < a.L.kp_cap) { a.uRight[(size_t)img * a.cap + iL] = -1.f; a.depth[(size_t)img * a.cap + iL] = -1.f; }
    return;
  }
  const orbx_keypoint* kL = a.L.kps + (size_t)slotL * a.L.kp_cap;
  const uint8_t* dL = a.L.desc + (size_t)slotL * a.L.kp_cap * 32;
  const uint8_t* dR = a.R.desc + (size_t)slotR * a.R.kp_cap * 32;
  const int* T = a.tab + (size_t)img * a.tab_stride;
  const uint4* rec = a.rec + (size_t)img * a.R.kp_cap;
  uint8_t* lp = s_patch + warp * 416;   // 11x11 left patch, pitch 12
  uint8_t* rp = lp + 144;               // 11x21 right strip, pitch 24
  do {                                  // one pass; "continue" = this keypoint has no stereo match
    const size_t oi = (size_t)img * a.cap + iL;
    const size_t si = (size_t)img * a.L.kp_cap + iL;      // the SAD scratch belongs to the left handle: its own pitch, not the caller's
    if (lane == 0) { a.uRight[oi] = -1.f; a.depth[oi] = -1.f; a.sad[si] = -1; }
    const float uL = kL[iL].x, vL = kL[iL].y;
    const int lvl = kL[iL].octave;
    const int row = (int)vL;
    const float minU = __fsub_rn(uL, a.maxD), maxU = uL;
    if (maxU < 0) continue;
    const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(dL + (size_t)iL * 32));
    const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(dL + (size_t)iL * 32) + 1);
    // the three runs (octaves lvl-1, lvl, lvl+1): lanes 0..5 fetch their bounds from the table.
    // floor(y-r) <= row <= ceil(y+r)  =>  |(int)y - row| <= ceil(r) + 1
    const int H = a.R.geom[0].h;
    int tv = 0;
    if (lane < 6) {
      const int o = lvl - 1 + (lane >> 1);
      if (o >= 0 && o < a.R.nlevels) {
        const int reach = (int)ceilf(__fmul_rn(2.0f, a.L.geom[o].scale)) + 1;
        const int rr = (lane & 1) ? min(max(row + reach, 0), H - 1) + 1 : min(max(row - reach, 0), H - 1);
        tv = T[o * H + rr];
      }
    }
    const int s0 = __shfl_sync(0xffffffffu, tv, 0), n0 = __shfl_sync(0xffffffffu, tv, 1) - s0;
    const int s1 = __shfl_sync(0xffffffffu, tv, 2), n1 = __shfl_sync(0xffffffffu, tv, 3) - s1;
    const int s2 = __shfl_sync(0xffffffffu, tv, 4), n2 = __shfl_sync(0xffffffffu, tv, 5) - s2;
    const int total = n0 + n1 + n2;
    unsigned best = 100u << 16;                                                      // TH_HIGH
    float bestU = 0.f;
    for (int j = lane; j < total; j += 32) {
      const int pos = j < n0 ? s0 + j : j < n0 + n1 ? s1 + (j - n0) : s2 + (j - n0 - n1);
      const uint4 rc = __ldg(rec + pos);
      const int ymin = (int)(short)(rc.y & 0xffffu), ymax = (int)rc.y >> 16;
      const float uR = __uint_as_float(rc.x);
      if (row >= ymin && row <= ymax && uR >= minU && uR <= maxU) {
        const unsigned key = ((unsigned)hamming256(a0, a1, reinterpret_cast<const uint4*>(dR + (size_t)rc.z * 32)) << 16) | rc.z;
        if (key < best) { best = key; bestU = uR; }
      }
    }
    const unsigned mykey = best;
    best = __reduce_min_sync(0xffffffffu, best);
    if ((best >> 16) >= 75u) continue;                                               // (TH_HIGH+TH_LOW)/2
    const float uR0 = __shfl_sync(0xffffffffu, bestU, __ffs((int)__ballot_sync(0xffffffffu, mykey == best)) - 1);
    const LevelGeom& gl = a.L.geom[lvl];
    const LevelGeom& gr = a.R.geom[lvl];
    const float inv = gl.inv_scale;
    const float su = roundf(__fmul_rn(uL, inv)), sv = roundf(__fmul_rn(vL, inv)), sr = roundf(__fmul_rn(uR0, inv));
    if (sr < 0.f || __fadd_rn(sr, 11.f) >= (float)gr.w) continue;                    // :573-575
    const int cy = (int)sv, cxl = (int)su, cxr = (int)sr;
    const uint8_t* pl = a.L.pyr + gl.plane_off + (size_t)slotL * gl.plane_stride + (size_t)(kEdge + cy - 5) * gl.pitch + kEdge + cxl - 5;
    const uint8_t* pr = a.R.pyr + gr.plane_off + (size_t)slotR * gr.plane_stride + (size_t)(kEdge + cy - 5) * gr.pitch + kEdge + cxr - 10;
    // SAD of the centre-subtracted 11x11 patches over the 11 shifts (:578-592), in integers.
    // Lanes 0..10 fetch one patch row each as aligned words (11 left bytes, 21 right bytes), byte-align them with
    // funnel shifts and park them in shared memory; then every lane owns 4 of the 121 pixels and evaluates
    // |(L - cl) - (R - cr)| = |(L - cl + cr) - R| with one VABSDIFF per pixel and shift.  Lane s ends up holding
    // the sum of shift s.  (A packed-byte VABSDIFF4 form is not exact here: R + cl - cr leaves a byte for most
    // shifts of a corner patch.)
    int mine = 0x7fffffff;
    {
      __syncwarp();
      if (lane < 11) {
        const uint8_t* al = pl + (size_t)lane * gl.pitch;
        const uint8_t* ar = pr + (size_t)lane * gr.pitch;
        const uint32_t* wl = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(al) & ~(uintptr_t)3);
        const uint32_t* wr = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(ar) & ~(uintptr_t)3);
        const int shl = (int)(reinterpret_cast<uintptr_t>(al) & 3) * 8, shr = (int)(reinterpret_cast<uintptr_t>(ar) & 3) * 8;
        const uint32_t L0 = __ldg(wl), L1 = __ldg(wl + 1), L2 = __ldg(wl + 2), L3 = __ldg(wl + 3);
        uint32_t W[7];
#pragma unroll
        for (int k = 0; k < 6; ++k) W[k] = __ldg(wr + k);
        W[6] = 0u;
        uint32_t* lrow = reinterpret_cast<uint32_t*>(lp) + lane * (kLP / 4);
        uint32_t* rrow = reinterpret_cast<uint32_t*>(rp) + lane * (kRP / 4);
        lrow[0] = __funnelshift_r(L0, L1, shl); lrow[1] = __funnelshift_r(L1, L2, shl); lrow[2] = __funnelshift_r(L2, L3, shl);
#pragma unroll
        for (int k = 0; k < 6; ++k) rrow[k] = __funnelshift_r(W[k], W[k + 1], shr);
      }
      __syncwarp();
      // pixel q of this lane: row 3q + lane/11, column lane%11; lanes 22..24 use their idle fourth slot for the
      // three pixels (2,10), (5,10), (8,10) that this cut leaves out
      const int lr = lane >= 22 ? 2 : lane >= 11 ? 1 : 0, lc = lane - 11 * lr;
      const bool v3 = lr < 2 || lane < 25;
      const int row3 = lr < 2 ? 9 + lr : 2 + 3 * (lane - 22), col3 = lr < 2 ? lc : 10;
      const int cl = lp[5 * kLP + 5];
      int la[4];
      const uint8_t* rq[4];
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        la[q] = (int)lp[(3 * q + lr) * kLP + lc] - cl;
        rq[q] = rp + (3 * q + lr) * kRP + lc + 10;       // right pixel of shift 0 is column +5, centre of the strip at +10: index by inc below
      }
      la[3] = v3 ? (int)lp[row3 * kLP + col3] - cl : 0;
      rq[3] = rp + (v3 ? row3 * kRP + col3 + 10 : 10);
#pragma unroll
      for (int inc = -5; inc <= 5; ++inc) {
        const int cr = rp[5 * kRP + 10 + inc];
        unsigned sum = __sad(la[0] + cr, (int)rq[0][inc - 5], 0u);
        sum = __sad(la[1] + cr, (int)rq[1][inc - 5], sum);
        sum = __sad(la[2] + cr, (int)rq[2][inc - 5], sum);
        if (v3) sum = __sad(la[3] + cr, (int)rq[3][inc - 5], sum);
        sum = __reduce_add_sync(0xffffffffu, sum);
        if (lane == inc + 5) mine = (int)sum;
      }
    }
    // first minimum over the shifts (:587-591): least (sum, shift)
    const unsigned bkey = __reduce_min_sync(0xffffffffu, lane < 11 ? ((unsigned)mine << 4) | (unsigned)lane : 0xffffffffu);
    const int bestSad = (int)(bkey >> 4), bs = (int)(bkey & 15u), bestInc = bs - 5;
    if (bs == 0 || bs == 10) continue;                                               // :594
    const float d1 = (float)__shfl_sync(0xffffffffu, mine, bs - 1), d2 = (float)bestSad, d3 = (float)__shfl_sync(0xffffffffu, mine, bs + 1);
    const float delta = __fdiv_rn(__fsub_rn(d1, d3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
    if (delta < -1.f || delta > 1.f) continue;
    float bestu = __fmul_rn(gl.scale, __fadd_rn(__fadd_rn(sr, (float)bestInc), delta));
    float disp = __fsub_rn(uL, bestu);
    if (disp >= 0.f && disp < a.maxD) {
      if (disp <= 0.f) { disp = 0.01f; bestu = (float)((double)uL - 0.01); }
      if (lane == 0) { a.depth[oi] = __fdiv_rn(a.bf, disp); a.uRight[oi] = bestu; a.sad[si] = bestSad; }
    }
  } while (false);
}

constexpr int kMedThreads = 1024;
// The reference sorts (SAD, iL) and reads element size/2 (src/Frame.cc:626-627).  Only that element's
// SAD value matters: it is the smallest v with #(SAD <= v) > size/2, found by bisection over the
// value range (SAD <= 121*510 < 2^16) with one block-wide count per step.
__global__ void __launch_bounds__(kMedThreads) k_stereo_median(StereoArgs a) {
  __shared__ int s_part[kMedThreads / 32];
  __shared__ int s_tot;
  const int img = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nL = a.L.nkp[a.slotL0 + img];
  const size_t base = (size_t)img * a.cap, sbase = (size_t)img * a.L.kp_cap;
  // up to 4 values per thread in registers (kp_cap <= 4096 for any accepted configuration is not
  // guaranteed, so the tail is re-read from memory in the rare larger case)
  int v[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) { const int i = tid + q * kMedThreads; v[q] = i < nL ? a.sad[sbase + i] : -1; }
  auto block_count = [&](int limit) {      // #(0 <= SAD <= limit) over the whole frame
    int c = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) c += (v[q] >= 0 && v[q] <= limit);
    for (int i = tid + 4 * kMedThreads; i < nL; i += kMedThreads) { const int s = a.sad[sbase + i]; c += (s >= 0 && s <= limit); }
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) s_part[warp] = c;
    __syncthreads();
    if (warp == 0) {
      const int t = __reduce_add_sync(0xffffffffu, s_part[lane]);
      if (lane == 0) s_tot = t;
    }
    __syncthreads();
    const int r = s_tot;
    __syncthreads();
    return r;
  };
  const int cnt = block_count(0x7fffffff);
  if (cnt == 0) { if (tid == 0) a.nMatches[img] = 0; return; }   // the reference reads an empty vector here
  const int k = cnt / 2;
  int lo = 0, hi = 65535;                  // invariant: count(<= hi) > k
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (block_count(mid) > k) hi = mid; else lo = mid + 1;
  }
  const float th = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)lo);
  int kept = 0;
  for (int i = tid; i < nL; i += kMedThreads) {
    const int s = a.sad[sbase + i];
    if (s < 0) continue;
    if ((float)s < th) ++kept;
    else { a.uRight[base + i] = -1.f; a.depth[base + i] = -1.f; }
  }
  kept = __reduce_add_sync(0xffffffffu, kept);
  if (lane == 0) s_part[warp] = kept;
  __syncthreads();
  if (warp == 0) {
    const int t = __reduce_add_sync(0xffffffffu, s_part[lane]);
    if (lane == 0) a.nMatches[img] = t;
  }
}

int launch_stereo(const StereoArgs& a, cudaStream_t st, cudaEvent_t ev_mid) {
  const int ctas_per_img = (a.L.kp_cap + kStWarps - 1) / kStWarps;
  k_stereo_rows<<<a.n, kRowThreads, 0, st>>>(a);
  k_stereo_match<<<a.n * ctas_per_img, kStWarps * 32, 0, st>>>(a, ctas_per_img);
  if (ev_mid) cudaEventRecord(ev_mid, st);
  k_stereo_median<<<a.n, kMedThreads, 0, st>>>(a);
  return 3;
}

}  // namespace orbx
