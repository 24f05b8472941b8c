// Shared by the latency forms of the order-dependent matchers (K8, K9 in k_match.cu; K17 in k_project.cu).
#pragma once
#include <stdint.h>

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

// The four smallest 64-bit keys (hi, lo) of a warp whose lanes each kept their two smallest (k1 <= k2) of `cnt` candidates.
// A lane that has given both kept keys and had more candidates ends the list there (its third key is unknown).
// Returns the list length; `complete` = the list holds every candidate of the warp.
__device__ __forceinline__ int warp_top4(unsigned long long k1, unsigned long long k2, int cnt, unsigned (&ohi)[4], unsigned (&olo)[4], bool& complete) {
  int popped = 0, len = 0;
  bool open = true;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
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

// One fixpoint round of phase B (see above).  Every lane brings the keypoints of its list (key[k] & 0xFFFF, kNoKey = none),
// `hitP` = which of them were claimed by earlier rounds, and a functor pick(hit) -> keypoint it would claim (-1: none) that
// also leaves the lane's decision in the caller's variables.  Returns after the iteration in which no lane changed its claim.
template <class Pick>
__device__ __forceinline__ void order_round(const uint32_t (&key)[4], unsigned hitP, int* stag, int lane, Pick pick) {
  int claim = pick(hitP);
  for (int it = 0; it < 33; ++it) {
    if (claim >= 0) atomicMin(&stag[claim], lane);
    __syncwarp();
    unsigned hitR = 0u;
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (key[k] != kNoKey && stag[key[k] & 0xFFFFu] < lane) hitR |= 1u << k;
    __syncwarp();
    if (claim >= 0) stag[claim] = 255;
    __syncwarp();
    const int next = pick(hitP | hitR);
    const bool moved = next != claim;
    claim = next;
    if (!__any_sync(0xffffffffu, moved)) break;
  }
}

}  // namespace orbx
