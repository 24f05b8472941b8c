// Shared by the latency forms of the order-dependent matchers (K8, K9 in k_match.cu; K17 in k_project.cu).
#pragma once
#include <stdint.h>

#include "orbx_internal.h"
#ifdef ORBX_PHASE_CLOCKS
#include <cstdio>
#endif

#ifdef ORBX_PHASE_CLOCKS
// experiments: SM clock at phase boundaries, printed by thread 0 of CTA 0 (build with ORBX_NVCC_EXTRA=-DORBX_PHASE_CLOCKS)
#define ORBX_CLK(v) const long long v = clock64()
#define ORBX_CLK_PRINT(name, ...) do { if (threadIdx.x == 0 && blockIdx.x == 0) { const long long c_[] = {__VA_ARGS__}; printf("%s:", name); \
  for (unsigned i_ = 1; i_ < sizeof(c_) / sizeof(c_[0]); ++i_) printf(" %lld", c_[i_] - c_[i_ - 1]); printf(" cycles\n"); } } while (0)
#define ORBX_STAT(...) __VA_ARGS__
#else
#define ORBX_CLK(v)
#define ORBX_CLK_PRINT(...)
#define ORBX_STAT(...)
#endif

namespace orbx {

// ---- ordered resolution of the latency forms -------------------------------------------------------------------
// Hazard H6 in one sentence: map point j takes the best keypoint(s) among those no EARLIER map point of the call has
// claimed.  A single warp doing that one map point at a time costs ~450 cycles per point (every step is a chain of
// dependent shuffles / shared-memory reads) and ~3000 per point that must be scanned again: 0.2 ms for 400 points, four
// fifths of the kernel.  So:
//   A (all warps): every map point gets the FOUR smallest candidate keys against the frame as it is on entry, sorted, and
//     whether that list is all of its candidates (warp_top4);
//   B (warp 0), 32 map points per round, one per lane: a lane's answer is the first key(s) of its list not claimed by an
//     earlier round or by an earlier lane of this round.  The lanes iterate to the fixpoint of exactly that equation
//     (publish the keypoint each lane would claim, look up who wants my keys, pick again; lane j is final after j + 1
//     iterations at the latest, in practice after two or three), then commit together.  The sequential answer is the
//     only fixpoint, so the result is the reference's, bit for bit.  A round in which a truncated list runs out of
//     unclaimed keys (rare) is redone one map point at a time with full re-scans.
constexpr uint32_t kNoKey = 0x7FFFFFFFu;

// The K smallest 64-bit keys (hi, lo) of a warp whose lanes each kept their two smallest (k1 <= k2) of `cnt` candidates.
// A lane that has given both kept keys and had more candidates ends the list there (its third key is unknown).
// Returns the list length; `complete` = the list holds every candidate of the warp.
template <int K>
__device__ __forceinline__ int warp_topk(unsigned long long k1, unsigned long long k2, int cnt, unsigned (&ohi)[K], unsigned (&olo)[K], bool& complete) {
  int popped = 0, len = 0;
  bool open = true;
#pragma unroll
  for (int r = 0; r < K; ++r) {
    ohi[r] = olo[r] = 0xffffffffu;
    if (open) {
      if (__any_sync(0xffffffffu, k1 == ~0ull && cnt > popped)) open = false;
      else {
        const unsigned hi = (unsigned)(k1 >> 32), lo = (unsigned)k1;
        const unsigned ghi = __reduce_min_sync(0xffffffffu, hi);
        if (ghi == 0xffffffffu) open = false;
        else {
          const unsigned glo = __reduce_min_sync(0xffffffffu, hi == ghi ? lo : 0xffffffffu);
          ohi[r] = ghi; olo[r] = glo; len = r + 1;
          if (hi == ghi && lo == glo) { k1 = k2; k2 = ~0ull; ++popped; }
        }
      }
    }
  }
  complete = !__any_sync(0xffffffffu, cnt > popped);
  return len;
}
__device__ __forceinline__ int warp_top4(unsigned long long k1, unsigned long long k2, int cnt, unsigned (&ohi)[4], unsigned (&olo)[4], bool& complete) {
  return warp_topk<4>(k1, k2, cnt, ohi, olo, complete);
}

// One fixpoint round of phase B (see above).  Every lane brings the keypoints of its list (key[k] & 0xFFFF, kNoKey = none),
// `hitP` = which of them were claimed by earlier rounds, and a functor pick(hit) -> keypoint it would claim (-1: none) that
// also leaves the lane's decision in the caller's variables.  Returns after the iteration in which no lane changed its claim
// (with the number of iterations).  stag[i] = (iteration stamp << 5) | (31 - lowest lane that wants keypoint i), written
// with atomicMax: a newer stamp beats every older entry, so nothing has to be reset; `epoch` counts the iterations of the
// whole kernel (start it at 0 over a zeroed stag).
template <int NK, class Pick>
__device__ __forceinline__ int order_round(const uint32_t (&key)[NK], unsigned hitP, int* stag, int& epoch, int lane, Pick pick) {
  int claim = pick(hitP);
  int it = 0;
  for (; it < 33; ++it) {
    ++epoch;
    if (claim >= 0) atomicMax(&stag[claim], (epoch << 5) | (31 - lane));
    __syncwarp();
    unsigned hitR = 0u;
#pragma unroll
    for (int k = 0; k < NK; ++k) {
      const int t = key[k] != kNoKey ? stag[key[k] & 0xFFFFu] : 0;
      if ((t >> 5) == epoch && 31 - (t & 31) < lane) hitR |= 1u << k;
    }
    const int next = pick(hitP | hitR);
    const bool moved = next != claim;
    claim = next;
    if (!__any_sync(0xffffffffu, moved)) break;                          // (also orders this iteration's reads before the next one's atomics)
  }
  return it + 1;
}

// Calls body(e, ix) once for every entry e of the runs [lo(ix), hi(ix)) of columns cx0 .. cx1, the entries dealt round-robin
// over the lanes of the warp (a lane per column entry would leave most lanes idle and give the low lanes several
// candidates each: one loop trip per column instead of one per 32 candidates).  bounds(ix, lo, hi) gives a column's run.
// body must not contain warp-synchronous operations.
template <class Bounds, class Body>
__device__ __forceinline__ void for_each_in_columns(int cx0, int cx1, int lane, Bounds bounds, Body body) {
  for (int cb = cx0; cb <= cx1; cb += 32) {
    const int ix = cb + lane, nc = min(32, cx1 - cb + 1);
    int e_lo = 0, len = 0;
    if (ix <= cx1) { int e_hi; bounds(ix, e_lo, e_hi); len = max(e_hi - e_lo, 0); }
    int incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31), excl = incl - len;
    for (int fb = 0; fb < total; fb += 32) {
      const int f = fb + lane;
      int col = 0;                                                       // column of entry f = the number of columns that end at or before it
      for (int c = 0; c < nc - 1; ++c) col += (__shfl_sync(0xffffffffu, incl, c) <= f) ? 1 : 0;
      const int e0 = __shfl_sync(0xffffffffu, e_lo, col), x0 = __shfl_sync(0xffffffffu, excl, col);
      if (f < total) body(e0 + (f - x0), cb + col);
    }
  }
}

// Counting sort of a frame's keypoints by 64 x 48 grid cell (Frame::AssignFeaturesToGrid / PosInGrid, src/Frame.cc:230-245,
// 382-392) by the whole CTA, into shared memory: on return soff[c + 1] = first position of cell c (c = gx * 48 + gy) and
// soff[kOrdGridCells + 1] = the number of keypoints in the grid; sxy / smeta hold position and  octave << 28 | cell << 16 | index
// per position.  The order inside a cell is arbitrary - callers rank candidates by (distance, cell, index), which is the
// reference's enumeration order whatever the storage order.  soff needs kOrdGridCells + 2 ints.  Ends with a block barrier.
constexpr int kOrdGridCells = 64 * 48;
__device__ __forceinline__ int grid_sort_cta(const orbx_keypoint* k, int n, float minX, float minY, float invW, float invH, int* soff, float2* sxy,
                                             uint32_t* smeta) {
  __shared__ int s_gw[32];
  __shared__ int s_gt;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nth = blockDim.x;
  for (int i = tid; i <= kOrdGridCells + 1; i += nth) soff[i] = 0;
  __syncthreads();
  for (int i = tid; i < n; i += nth) {
    const orbx_keypoint kp = k[i];
    const int gx = (int)roundf(__fmul_rn(__fsub_rn(kp.x, minX), invW)), gy = (int)roundf(__fmul_rn(__fsub_rn(kp.y, minY), invH));
    if (gx >= 0 && gx < 64 && gy >= 0 && gy < 48) atomicAdd(&soff[gx * 48 + gy + 1], 1);
  }
  __syncthreads();
  const int ch = (kOrdGridCells + nth) / nth, b0 = tid * ch;    // inclusive scan of soff[0 .. kOrdGridCells]
  int sum = 0;
  for (int j = 0; j < ch; ++j) sum += (b0 + j <= kOrdGridCells) ? soff[b0 + j] : 0;
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_gw[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const int own = lane < (nth >> 5) ? s_gw[lane] : 0;
    int v = own;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    s_gw[lane] = v - own;
    if (lane == 31) s_gt = v;
  }
  __syncthreads();
  int run = s_gw[warp] + incl - sum;
  for (int j = 0; j < ch; ++j)
    if (b0 + j <= kOrdGridCells) { run += soff[b0 + j]; soff[b0 + j] = run; }   // soff[c + 1] = end of cell c
  __syncthreads();
  for (int i = tid; i < n; i += nth) {                           // fill each cell from its end downwards: soff[c + 1] = start of cell c
    const orbx_keypoint kp = k[i];
    const int gx = (int)roundf(__fmul_rn(__fsub_rn(kp.x, minX), invW)), gy = (int)roundf(__fmul_rn(__fsub_rn(kp.y, minY), invH));
    if (gx >= 0 && gx < 64 && gy >= 0 && gy < 48) {
      const int c = gx * 48 + gy;
      const int pos = atomicSub(&soff[c + 1], 1) - 1;
      sxy[pos] = make_float2(kp.x, kp.y);
      smeta[pos] = ((uint32_t)kp.octave << 28) | ((uint32_t)c << 16) | (uint32_t)i;
    }
  }
  __syncthreads();
  if (tid == 0) soff[kOrdGridCells + 1] = s_gt;
  __syncthreads();
  return s_gt;
}

}  // namespace orbx
