// K6 / K7 — ORBmatcher Hamming searches, batched over independent frame pairs.
//
// K6  k_search_init  <- ORBmatcher::SearchForInitialization   src/ORBmatcher.cc:405-520
//                       (+ Frame::GetFeaturesInArea / PosInGrid, src/Frame.cc:327-392)
// K7  k_knn_ratio    <- ORBmatcher::SearchByBoW(KF,KF), one bucket  src/ORBmatcher.cc:522-655
//     three_maxima   <- ORBmatcher::ComputeThreeMaxima        src/ORBmatcher.cc:1601-1642
//
// Both reference routines carry a sequential dependency (hazard H6): a query's candidate set
// depends on what earlier queries matched (vMatchedDistance / vbMatched2).  Queries are
// therefore resolved in index order by one warp (K6) or one CTA (K7) per pair, with the
// per-query candidate scan spread over the lanes/threads; pairs are independent and fill the GPU.
//
// K6 needs no bucket grid: a keypoint of frame 2 is returned by GetFeaturesInArea iff it is on
// level 0, PosInGrid accepted it, and |dx| < r, |dy| < r (the cell range computed with
// floor/ceil always covers the round()-assigned cell of such a keypoint).  The grid only fixes
// the enumeration order (cell x, cell y, index), which matters for ties on the best distance;
// it is carried as a 64-bit key  distance<<32 | cell<<16 | index  reduced with min.
#include <algorithm>
#include <climits>

#include <cstdlib>

#include "orbx_internal.h"
#include "ordered.cuh"

namespace orbx {

constexpr int kHisto = 30;

__device__ __forceinline__ int ham8(const uint32_t* a, const uint32_t* b) {
  int d = 0;
#pragma unroll
  for (int i = 0; i < 8; i += 2)
    d += __popcll(((unsigned long long)(a[i + 1] ^ b[i + 1]) << 32) | (a[i] ^ b[i]));
  return d;
}

__device__ __forceinline__ int rot_bin(float a1, float a2) {
  float rot = __fsub_rn(a1, a2);
  if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
  int bin = (int)roundf(__fmul_rn(rot, 1.0f / kHisto));
  if (bin == kHisto) bin = 0;
  return bin;
}

// hist[0..29] -> the up-to-three bins to keep
__device__ __forceinline__ void three_maxima(const int* hist, int& i1, int& i2, int& i3) {
  int m1 = 0, m2 = 0, m3 = 0;
  i1 = i2 = i3 = -1;
  for (int i = 0; i < kHisto; ++i) {
    const int s = hist[i];
    if (s > m1) { m3 = m2; m2 = m1; m1 = s; i3 = i2; i2 = i1; i1 = i; }
    else if (s > m2) { m3 = m2; m2 = s; i3 = i2; i2 = i; }
    else if (s > m3) { m3 = s; i3 = i; }
  }
  if ((float)m2 < __fmul_rn(0.1f, (float)m1)) { i2 = -1; i3 = -1; }
  else if ((float)m3 < __fmul_rn(0.1f, (float)m1)) { i3 = -1; }
}

// ------------------------------------------------------------------------------------------
// K6: one warp per frame pair.  scratch per pair: 4*cap ints
//   [0,cap)      compact list of frame-2 level-0 keypoints: cell<<16 | index
//   [cap,2cap)   vMatchedDistance per frame-2 keypoint
//   [2cap,3cap)  vnMatches21
//   [3cap,4cap)  rotation bin per frame-1 keypoint (-1 = none)
__global__ void __launch_bounds__(32) k_search_init(InitMatchArgs a) {
  __shared__ int s_hist[kHisto];
  const int p = blockIdx.x, lane = threadIdx.x;
  const int n1 = a.n1[p], n2 = a.n2[p];
  const float minX = a.gb[0], minY = a.gb[2], invW = a.gb[4], invH = a.gb[5];   // Frame::mnMinX, mnMinY, mfGridElement{Width,Height}Inv
  const orbx_keypoint* k1 = a.k1 + (size_t)p * a.cap;
  const orbx_keypoint* k2 = a.k2 + (size_t)p * a.cap;
  const uint32_t* d1 = reinterpret_cast<const uint32_t*>(a.d1 + (size_t)p * a.cap * 32);
  const uint32_t* d2 = reinterpret_cast<const uint32_t*>(a.d2 + (size_t)p * a.cap * 32);
  float* prev = a.prev + (size_t)p * a.cap * 2;
  int* m12 = a.m12 + (size_t)p * a.cap;
  int* clist = a.scratch + (size_t)p * a.cap * 4;
  int* mdist = clist + a.cap;
  int* m21 = mdist + a.cap;
  int* binof = m21 + a.cap;

  for (int i = lane; i < n1; i += 32) { m12[i] = -1; binof[i] = -1; }
  for (int i = lane; i < n2; i += 32) { mdist[i] = INT_MAX; m21[i] = -1; }
  if (lane < kHisto) s_hist[lane] = 0;
  int nc = 0;
  for (int base = 0; base < n2; base += 32) {
    const int i = base + lane;
    bool ok = false;
    int cell = 0;
    if (i < n2) {
      const orbx_keypoint kp = k2[i];
      const int gx = (int)roundf(__fmul_rn(__fsub_rn(kp.x, minX), invW)), gy = (int)roundf(__fmul_rn(__fsub_rn(kp.y, minY), invH));   // PosInGrid, src/Frame.cc:382-392
      ok = kp.octave == 0 && gx >= 0 && gx < 64 && gy >= 0 && gy < 48;
      cell = gx * 48 + gy;
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if (ok) clist[nc + __popc(bal & ((1u << lane) - 1u))] = (cell << 16) | i;
    nc += __popc(bal);
  }
  __syncwarp();

  int nm = 0;
  const float r = a.window;
  for (int i1 = 0; i1 < n1; ++i1) {
    if (k1[i1].octave > 0) continue;
    const float qx = prev[2 * i1], qy = prev[2 * i1 + 1];
    uint32_t q[8];
#pragma unroll
    for (int w = 0; w < 8; ++w) q[w] = __ldg(d1 + (size_t)i1 * 8 + w);
    unsigned long long b1 = ~0ull;
    int b2 = INT_MAX;
    for (int j = lane; j < nc; j += 32) {
      const int ck = clist[j], i2 = ck & 0xFFFF;
      const orbx_keypoint kp = k2[i2];
      if (!(fabsf(__fsub_rn(kp.x, qx)) < r && fabsf(__fsub_rn(kp.y, qy)) < r)) continue;
      const int d = ham8(q, d2 + (size_t)i2 * 8);
      if (mdist[i2] <= d) continue;
      const unsigned long long key = ((unsigned long long)d << 32) | (unsigned)ck;
      if (key < b1) { if (b1 != ~0ull) b2 = (int)(b1 >> 32); b1 = key; }
      else if (d < b2) b2 = d;
    }
    // warp merge: global best key; second best = min(owner's second, everybody else's best)
    unsigned hi = (unsigned)(b1 >> 32), lo = (unsigned)b1;
    const unsigned ghi = __reduce_min_sync(0xffffffffu, hi);
    const unsigned glo = __reduce_min_sync(0xffffffffu, hi == ghi ? lo : 0xffffffffu);
    if (ghi == 0xffffffffu) continue;
    const bool owner = hi == ghi && lo == glo;
    const int contrib = owner ? b2 : (b1 == ~0ull ? INT_MAX : (int)hi);
    const int bestDist2 = (int)__reduce_min_sync(0xffffffffu, (unsigned)contrib);
    const int bestDist = (int)ghi, bi = glo & 0xFFFF;
    if (bestDist <= 50 && (float)bestDist < __fmul_rn((float)bestDist2, a.ratio)) {
      if (lane == 0) {
        if (m21[bi] >= 0) m12[m21[bi]] = -1;
        m12[i1] = bi; m21[bi] = i1; mdist[bi] = bestDist;
        if (a.checkOri) { const int bin = rot_bin(k1[i1].angle, k2[bi].angle); binof[i1] = bin; s_hist[bin] += 1; }
      }
      __syncwarp();
    }
  }
  __syncwarp();
  int k_a = -1, k_b = -1, k_c = -1;
  if (a.checkOri) three_maxima(s_hist, k_a, k_b, k_c);
  for (int i = lane; i < n1; i += 32) {
    int m = m12[i];
    if (a.checkOri && m >= 0) {
      const int b = binof[i];
      if (b != k_a && b != k_b && b != k_c) { m = -1; m12[i] = -1; }
    }
    if (m >= 0) { prev[2 * i] = k2[m].x; prev[2 * i + 1] = k2[m].y; ++nm; }
  }
  nm = __reduce_add_sync(0xffffffffu, nm);
  if (lane == 0) a.nm[p] = nm;
}

// ------------------------------------------------------------------------------------------
// K7: one CTA per pair; B descriptors transposed in shared memory ([word][j], conflict-free).
constexpr int kKnnThreads = 256;
constexpr int kTopK = 8;

// K7 runs in two launches.
//   k_knn_topk    every A descriptor scans ALL B descriptors of its pair (no dependency yet) and keeps its
//                 kTopK nearest as sorted keys  distance<<16 | j.  One thread = one query, B rows are
//                 broadcast from shared memory.  The 256-bit popcount goes through a carry-save adder
//                 tree (8 LOP3) so that only 4 POPC are issued per distance: POPC is the slowest pipe on
//                 this path (0.5 warp instructions / clk / SM, profiles/r1d_pipe_rates.jsonl).
//                 Keys at or above `sentinel` are never kept: a second-best that far away passes the ratio
//                 test against any best <= th, a best that far away fails the threshold (launch_knn_ratio).
//   k_knn_resolve one warp per pair walks the queries in index order (hazard H6: vbMatched2): best and
//                 second-best are the first two list entries whose B feature is still free.  The list is
//                 exact unless fewer than two of its entries are free while it is full; then the warp
//                 rescans all B descriptors for that query.
__device__ __forceinline__ uint32_t lop3_xor3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
__device__ __forceinline__ uint32_t lop3_maj(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
// popcount of the 256-bit xor of q and (b0, b1)
__device__ __forceinline__ int ham256_csa(const uint32_t (&q)[8], const uint4 b0, const uint4 b1) {
  const uint32_t x0 = q[0] ^ b0.x, x1 = q[1] ^ b0.y, x2 = q[2] ^ b0.z, x3 = q[3] ^ b0.w;
  const uint32_t x4 = q[4] ^ b1.x, x5 = q[5] ^ b1.y, x6 = q[6] ^ b1.z, x7 = q[7] ^ b1.w;
  const uint32_t s0 = lop3_xor3(x0, x1, x2), c0 = lop3_maj(x0, x1, x2);
  const uint32_t s1 = lop3_xor3(x3, x4, x5), c1 = lop3_maj(x3, x4, x5);
  const uint32_t s2 = lop3_xor3(s0, s1, x6), c2 = lop3_maj(s0, s1, x6);
  const uint32_t t0 = lop3_xor3(c0, c1, c2), f0 = lop3_maj(c0, c1, c2);
  return (__popc(s2) + __popc(x7)) + 2 * __popc(t0) + 4 * __popc(f0);
}

// FUSED: the query block that finishes a pair's lists last goes on to resolve the pair (its B descriptors are already in this
// CTA's shared memory), so the ordered part of one pair runs beside the distance work of the others instead of after all of it.
template <bool FUSED> __device__ __forceinline__ void knn_resolve_tail(const KnnArgs& a, uint32_t sentinel, int p, uint8_t* smem);

template <bool FUSED>
__global__ void __launch_bounds__(kKnnThreads) k_knn_topk(KnnArgs a, int qblocks, uint32_t sentinel) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint4* Bs = reinterpret_cast<uint4*>(smem);                  // [nB][2]
  __shared__ int s_last;
  const int p = blockIdx.x / qblocks, qb = blockIdx.x - p * qblocks, tid = threadIdx.x;
  const int nA = a.nA[p], nB = a.nB[p];
  if (qb * kKnnThreads >= nA) {
    if (FUSED && nA == 0 && qb == 0 && tid == 0) a.nm[p] = 0;  // nothing to resolve
    return;
  }
  const uint4* dB = reinterpret_cast<const uint4*>(a.dB + (size_t)p * a.cap * 32);
  for (int idx = tid; idx < nB * 2; idx += kKnnThreads) Bs[idx] = __ldg(dB + idx);
  __syncthreads();
  const int i = qb * kKnnThreads + tid;
  if (i < nA) {
    uint32_t q[8];
    {
      const uint4* dA = reinterpret_cast<const uint4*>(a.dA + ((size_t)p * a.cap + i) * 32);
      const uint4 q0 = __ldg(dA), q1 = __ldg(dA + 1);
      q[0] = q0.x; q[1] = q0.y; q[2] = q0.z; q[3] = q0.w; q[4] = q1.x; q[5] = q1.y; q[6] = q1.z; q[7] = q1.w;
    }
    uint32_t top[kTopK];
#pragma unroll
    for (int k = 0; k < kTopK; ++k) top[k] = sentinel;
#pragma unroll 2
    for (int j = 0; j < nB; ++j) {
      const uint32_t key = ((uint32_t)ham256_csa(q, Bs[2 * j], Bs[2 * j + 1]) << 16) + (uint32_t)j;
      if (key < top[kTopK - 1]) {                              // rare: only neighbours closer than the sentinel
        top[kTopK - 1] = key;
#pragma unroll
        for (int k = kTopK - 1; k > 0; --k) {
          const uint32_t lo = min(top[k - 1], top[k]), hi = max(top[k - 1], top[k]);
          top[k - 1] = lo; top[k] = hi;
        }
      }
    }
    uint4* out = reinterpret_cast<uint4*>(a.topk + ((size_t)p * a.cap + i) * kTopK);
    out[0] = make_uint4(top[0], top[1], top[2], top[3]);
    out[1] = make_uint4(top[4], top[5], top[6], top[7]);
  }
  if (!FUSED) return;
  __threadfence();                                             // this block's lists are visible device-wide before its ticket
  __syncthreads();
  if (tid == 0) s_last = atomicAdd(a.done + p, 1) == (nA + kKnnThreads - 1) / kKnnThreads - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  knn_resolve_tail<FUSED>(a, sentinel, p, smem);
}

constexpr int kKnnResolveThreads = 128;
// The ordered part of K7 for pair p, by the whole CTA (any size >= 32).  BSM: the pair's B descriptors are at the start of the
// dynamic shared memory ([nB][2] uint4) - the exact re-scans then cost no global round trip.
template <bool BSM>
__device__ __forceinline__ void knn_resolve_pair(const KnnArgs& a, uint32_t sentinel, int p, uint8_t* smem, bool stage_b) {
  const int uwords = (a.cap + 31) / 32;
  uint4* Bsm = reinterpret_cast<uint4*>(smem);                 // BSM: [nB][2] B descriptors for the exact re-scans
  uint32_t* used = reinterpret_cast<uint32_t*>(smem + (BSM ? (size_t)a.cap * 32 : 0));   // bit j: B feature j already matched (vbMatched2)
  int* claim = reinterpret_cast<int*>(used + uwords);          // [cap] order_round's tags: who wants B feature j in the round being resolved
  __shared__ int s_hist[kHisto];
  __shared__ int s_nm;
  const int tid = threadIdx.x, lane = tid & 31, nthr = blockDim.x;
  const int nA = a.nA[p], nB = a.nB[p];
  const uint32_t* dA = reinterpret_cast<const uint32_t*>(a.dA + (size_t)p * a.cap * 32);
  const uint32_t* dB = reinterpret_cast<const uint32_t*>(a.dB + (size_t)p * a.cap * 32);
  const float* angA = a.angA + (size_t)p * a.cap;
  const float* angB = a.angB + (size_t)p * a.cap;
  const uint4* tk = reinterpret_cast<const uint4*>(a.topk + (size_t)p * a.cap * kTopK);
  int* m12 = a.m12 + (size_t)p * a.cap;
  // the whole CTA sets up (and, at the end, does the rotation check); the ordered part in between belongs to warp 0
  ORBX_CLK(c0);
  for (int w = tid; w < uwords; w += nthr) used[w] = 0u;
  for (int j = tid; j < a.cap; j += nthr) claim[j] = 0;                                // order_round's tags
  if (tid < kHisto) s_hist[tid] = 0;
  if (BSM && stage_b) {
    const uint4* src = reinterpret_cast<const uint4*>(dB);
    for (int i = tid; i < nB * 2; i += nthr) Bsm[i] = __ldg(src + i);
  }
  __syncthreads();
  if (tid < 32) {
  const int far = min((int)(sentinel >> 16), 256);             // stands for "at least the sentinel distance, or none"
  int nm = 0, epoch = 0;
  ORBX_STAT(int dbg_it = 0; int dbg_rs = 0; long long dbg_rc = 0;)
  ORBX_CLK(c1);
  // Queries are taken 32 at a time, one per lane, and decided together (ordered.cuh): a lane's best and second best are the
  // first two entries of its list that no earlier query has matched; the lanes iterate to the fixpoint of that rule, commit
  // together, and a lane whose full list has fewer than two free entries left is re-scanned exactly, alone, before the
  // lanes after it go round again - the outcome is the one of the index-ordered loop of the reference.
  uint4 n0 = make_uint4(sentinel, sentinel, sentinel, sentinel), n1 = n0;
  if (lane < nA) { n0 = __ldcg(tk + 2 * lane); n1 = __ldcg(tk + 2 * lane + 1); }      // written by other CTAs of this launch: L2, not the read-only path
  for (int i0 = 0; i0 < nA; i0 += 32) {
    const int i = i0 + lane;
    const bool active = i < nA;
    const uint32_t raw[kTopK] = {n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, n1.z, n1.w};
    if (i + 32 < nA) { n0 = __ldcg(tk + 2 * (i + 32)); n1 = __ldcg(tk + 2 * (i + 32) + 1); }
    const bool full = raw[kTopK - 1] < sentinel;               // sorted: the last entry is valid iff all are
    uint32_t key[kTopK];
#pragma unroll
    for (int k = 0; k < kTopK; ++k) key[k] = (active && raw[k] < sentinel) ? raw[k] : kNoKey;
    uint32_t todo = __ballot_sync(0xffffffffu, active);
    int mine = -1;                                             // this lane's m12 entry
    while (todo) {
      const bool live = (todo >> lane) & 1u;
      unsigned hitP = 0u;
#pragma unroll
      for (int k = 0; k < kTopK; ++k)
        if (key[k] != kNoKey && ((used[(key[k] & 0xffffu) >> 5] >> (key[k] & 31u)) & 1u)) hitP |= 1u << k;
      bool pass = false, starved = false;
      int bi = -1;
      ORBX_STAT(dbg_it +=) order_round(key, hitP, claim, epoch, lane, [&](unsigned hit) {
        int cnt = 0, best = 256, sec = 256;
        bi = -1;
#pragma unroll
        for (int k = 0; k < kTopK; ++k)
          if (key[k] != kNoKey && !((hit >> k) & 1u)) {
            if (cnt == 0) { best = (int)(key[k] >> 16); bi = (int)(key[k] & 0xffffu); }
            else if (cnt == 1) sec = (int)(key[k] >> 16);
            ++cnt;
          }
        starved = live && cnt < 2 && full;                      // the list cannot tell best and second best any more
        if (cnt == 1 && !full) sec = far;                       // every B below the sentinel distance was on the list
        pass = live && !starved && cnt > 0 && (a.strict ? best < a.th : best <= a.th) && (float)best < __fmul_rn(a.ratio, (float)sec);
        return pass ? bi : -1;
      });
      const uint32_t sm = __ballot_sync(0xffffffffu, starved);
      const int jstar = sm ? __ffs((int)sm) - 1 : 32;
      if (live && lane < jstar) {
        mine = pass ? bi : -1;
        if (pass) { atomicOr(&used[bi >> 5], 1u << (bi & 31)); ++nm; }
      }
      __syncwarp();
      if (!sm) break;
      {                                                         // exact rescan of query i0 + jstar (src/ORBmatcher.cc:574-594)
        ORBX_STAT(++dbg_rs;)
        ORBX_CLK(r0);
        const int iq = i0 + jstar;
        uint32_t qd[8];
#pragma unroll
        for (int w = 0; w < 8; ++w) qd[w] = __ldg(dA + (size_t)iq * 8 + w);
        unsigned b1 = 256u << 16;
        int b2 = 256;
        auto take = [&](int d, int jj) {                        // first least distance wins (ascending jj per lane, lanes merged by key)
          const unsigned kk = ((unsigned)d << 16) | (unsigned)jj;
          if (d < (int)(b1 >> 16)) { b2 = (int)(b1 >> 16); b1 = kk; }
          else if (d < b2) b2 = d;
        };
        if (BSM) {
          // four B features per lane and trip, their loads issued together: the walk is one warp's latency chain
          for (int j0 = lane; j0 < nB; j0 += 128) {
            int d[4]; bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int jj = j0 + 32 * u;
              ok[u] = jj < nB && !((used[jj >> 5] >> (jj & 31)) & 1u);
              const int js = jj < nB ? jj : 0;
              d[u] = ham256_csa(qd, Bsm[2 * js], Bsm[2 * js + 1]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) if (ok[u]) take(d[u], j0 + 32 * u);
          }
        } else {
          for (int jj = lane; jj < nB; jj += 32) {
            if ((used[jj >> 5] >> (jj & 31)) & 1u) continue;
            uint32_t bd[8];
#pragma unroll
            for (int w = 0; w < 8; ++w) bd[w] = __ldg(dB + (size_t)jj * 8 + w);
            take(ham8(qd, bd), jj);
          }
        }
        const unsigned wmin = __reduce_min_sync(0xffffffffu, b1);
        const int rsec = (int)__reduce_min_sync(0xffffffffu, (unsigned)(b1 == wmin ? b2 : (int)(b1 >> 16)));
        const int rbest = (int)(wmin >> 16), rbi = (int)(wmin & 0xffffu);
        const bool rpass = (a.strict ? rbest < a.th : rbest <= a.th) && (float)rbest < __fmul_rn(a.ratio, (float)rsec);
        if (lane == jstar) {
          mine = rpass ? rbi : -1;
          if (rpass) { atomicOr(&used[rbi >> 5], 1u << (rbi & 31)); ++nm; }
        }
        __syncwarp();
        ORBX_CLK(r1);
        ORBX_STAT(dbg_rc += r1 - r0;)
      }
      todo &= ~((2u << jstar) - 1u);
    }
    if (active) m12[i] = mine;                                  // one coalesced store per group
  }
  nm = __reduce_add_sync(0xffffffffu, nm);
  if (lane == 0) s_nm = nm;
  ORBX_CLK(c2);
  ORBX_STAT(if (lane == 0 && p == 0) printf("K7 resolve pair 0: setup %lld, ordered %lld cycles: %d queries, %d fixpoint iterations, %d re-scans in %lld cycles\n", c1 - c0, c2 - c1, nA, dbg_it, dbg_rs, dbg_rc);)
  }   // warp 0
  __syncthreads();
  if (a.checkOri) {                                            // rotation consistency, src/ORBmatcher.cc:607-652, whole CTA
    int mybin[64];
    const bool keep = nA <= 64 * nthr;            // bins remembered in registers for the usual sizes, recomputed otherwise
    int r = 0;
    for (int base = 0; base < nA; base += nthr, ++r) {
      const int i = base + tid;
      const int m = i < nA ? m12[i] : -1;
      const unsigned act = __ballot_sync(0xffffffffu, m >= 0);
      int b = -1;
      if (m >= 0) {                                            // one atomic per distinct bin of the warp
        b = rot_bin(angA[i], angB[m]);
        const unsigned peers = __match_any_sync(act, b);
        if (lane == __ffs((int)peers) - 1) atomicAdd(&s_hist[b], __popc(peers));
      }
      if (keep) mybin[r] = b;
    }
    __syncthreads();
    int k_a, k_b, k_c;
    three_maxima(s_hist, k_a, k_b, k_c);
    int drop = 0;
    r = 0;
    for (int i = tid; i < nA; i += nthr, ++r) {
      int b;
      if (keep) b = mybin[r];
      else { const int m = m12[i]; b = m >= 0 ? rot_bin(angA[i], angB[m]) : -1; }
      if (b >= 0 && b != k_a && b != k_b && b != k_c) { m12[i] = -1; ++drop; }
    }
    if (drop) atomicSub(&s_nm, drop);
    __syncthreads();
  }
  if (tid == 0) a.nm[p] = s_nm;
}

template <bool FUSED> __device__ __forceinline__ void knn_resolve_tail(const KnnArgs& a, uint32_t sentinel, int p, uint8_t* smem) {
  if (FUSED) knn_resolve_pair<true>(a, sentinel, p, smem, false);
}

// stand-alone form (a pair whose B side does not fit the fused kernel's shared memory, or ORBX_KNN_SPLIT for A/B runs)
template <bool BSM>
__global__ void __launch_bounds__(kKnnResolveThreads) k_knn_resolve(KnnArgs a, uint32_t sentinel) {
  extern __shared__ __align__(16) uint8_t smem[];
  knn_resolve_pair<BSM>(a, sentinel, blockIdx.x, smem, true);
}

// ------------------------------------------------------------------------------------------
// K8  k_search_projection <- ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th)
//                            src/ORBmatcher.cc:45-129 (+ RadiusByViewingCos :131-137), the local-map
//                            search of every tracked frame (src/Tracking.cc:1192).
// One warp per frame, map points in index order (an accepted match occupies its keypoint for the
// points that follow, H6).  The frame's keypoints are counting-sorted by their 64x48 grid cell
// (Frame::AssignFeaturesToGrid / PosInGrid, src/Frame.cc:230-245,382-392) in shared memory, so a map
// point only touches the cell columns of its search window, each a contiguous run; best and second
// best are the two smallest keys  distance<<32 | cell<<16 | index  (= the reference's first-wins
// streaming in GetFeaturesInArea order).
constexpr int kGridCells = 64 * 48;

// Counting sort of a frame's keypoints by 64x48 grid cell (Frame::AssignFeaturesToGrid / PosInGrid,
// src/Frame.cc:230-245,382-392), one warp.  On return cell c occupies [soff[c+1], soff[c+2]) of
// sxy / smeta (the last cell ends at the returned count); keypoints outside the grid are dropped,
// as the reference does.  smeta = octave<<28 | cell<<16 | keypoint index.
__device__ __forceinline__ int grid_sort(const orbx_keypoint* k, int n, float minX, float minY, float invW, float invH, float2* sxy, uint32_t* smeta,
                                         int* soff, int lane) {
  for (int i = lane; i <= kGridCells; i += 32) soff[i] = 0;
  __syncwarp();
  for (int i = lane; i < n; i += 32) {
    const int gx = (int)roundf(__fmul_rn(__fsub_rn(k[i].x, minX), invW)), gy = (int)roundf(__fmul_rn(__fsub_rn(k[i].y, minY), invH));
    if (gx >= 0 && gx < 64 && gy >= 0 && gy < 48) atomicAdd(&soff[gx * 48 + gy + 1], 1);
  }
  __syncwarp();
  int carry = 0;                                                       // inclusive scan: soff[c+1] = end of cell c
  for (int base = 0; base <= kGridCells; base += 32) {
    const int i = base + lane;
    int v = i <= kGridCells ? soff[i] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    if (i <= kGridCells) soff[i] = v + carry;
    carry += __shfl_sync(0xffffffffu, v, 31);
  }
  __syncwarp();
  // fill each cell from its end downwards: afterwards soff[c+1] = start of cell c
  for (int i = lane; i < n; i += 32) {
    const orbx_keypoint kp = k[i];
    const int gx = (int)roundf(__fmul_rn(__fsub_rn(kp.x, minX), invW)), gy = (int)roundf(__fmul_rn(__fsub_rn(kp.y, minY), invH));
    if (gx >= 0 && gx < 64 && gy >= 0 && gy < 48) {
      const int c = gx * 48 + gy;
      const int pos = atomicSub(&soff[c + 1], 1) - 1;
      sxy[pos] = make_float2(kp.x, kp.y);
      smeta[pos] = ((uint32_t)kp.octave << 28) | ((uint32_t)c << 16) | (uint32_t)i;
    }
  }
  __syncwarp();
  return carry;
}

// DSM: the frame's descriptors are copied to shared memory first (they fit for the usual 1000-2000 features); a candidate's
// descriptor then costs a shared-memory read instead of a dependent L2 round trip per map point - with ONE frame per call
// (the SLAM loop's pattern) the kernel is a single warp walking the map points in order and those round trips were its time.
// The fields of map point i + 1 are fetched while point i is processed.
template <bool DSM>
__global__ void __launch_bounds__(32) k_search_projection(ProjArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  float2* sxy = reinterpret_cast<float2*>(smem);                       // [cap] sorted by cell
  uint32_t* smeta = reinterpret_cast<uint32_t*>(sxy + a.cap);          // [cap] octave<<28 | cell<<16 | index
  int* soff = reinterpret_cast<int*>(smeta + a.cap);                   // [kGridCells + 2] cell c = [soff[c+1], soff[c+2])
  uint32_t* sdesc = reinterpret_cast<uint32_t*>((reinterpret_cast<uintptr_t>(soff + kGridCells + 2) + 15) & ~(uintptr_t)15);  // DSM: [cap][8]
  const int p = blockIdx.x, lane = threadIdx.x;
  const int n = a.n[p], m = a.m[p];
  const orbx_keypoint* k = a.k + (size_t)p * a.cap;
  const uint32_t* dg = reinterpret_cast<const uint32_t*>(a.d + (size_t)p * a.cap * 32);
  const float* uRight = a.uRight + (size_t)p * a.cap;
  int* frameObs = a.frameObs + (size_t)p * a.cap;
  int* assigned = a.assigned + (size_t)p * a.cap;
  const size_t mo = (size_t)p * a.mcap;

  for (int i = lane; i < n; i += 32) assigned[i] = -1;
  if (DSM) {
    const uint4* src = reinterpret_cast<const uint4*>(dg);
    uint4* dst = reinterpret_cast<uint4*>(sdesc);
    for (int i = lane; i < n * 2; i += 32) dst[i] = __ldg(src + i);
  }
  const uint32_t* d = DSM ? sdesc : dg;
  const float minX = a.gb[0], minY = a.gb[2], invW = a.gb[4], invH = a.gb[5];
  const int n_in_grid = grid_sort(k, n, minX, minY, invW, invH, sxy, smeta, soff, lane);
  const int* cstart = soff + 1;                                         // cell c = [cstart[c], cstart[c+1]); the last cell ends at n_in_grid

  // fields of one map point: lanes 0..7 hold one descriptor word each, every lane the scalars
  struct MP { int inView, level; float viewCos, x, y, xr; uint32_t dw; };
  auto fetch = [&](int i) {
    MP q;
    q.inView = i < m ? a.inView[mo + i] : 0;
    q.level = i < m ? a.level[mo + i] : 0;
    q.viewCos = i < m ? a.viewCos[mo + i] : 0.f;
    q.x = i < m ? a.projX[mo + i] : 0.f; q.y = i < m ? a.projY[mo + i] : 0.f; q.xr = i < m ? a.projXR[mo + i] : 0.f;
    q.dw = (i < m && lane < 8) ? __ldg(reinterpret_cast<const uint32_t*>(a.mpDesc + (mo + i) * 32) + lane) : 0u;
    return q;
  };
  int nm = 0;
  MP nx = fetch(0);
  for (int iMP = 0; iMP < m; ++iMP) {
    const MP c = nx;
    nx = fetch(iMP + 1);
    if (!c.inView) continue;
    const int lvl = min(max(c.level, 0), a.nlevels - 1);
    float r = ((double)c.viewCos > 0.998) ? 2.5f : 4.0f;
    if (a.th != 1.0f) r = __fmul_rn(r, a.th);
    const float R = __fmul_rn(r, a.scale[lvl]);
    const float x = c.x, y = c.y;
    // GetFeaturesInArea's cell window, src/Frame.cc:332-346: (x - mnMinX -+ r) * inv, left to right
    const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, minX), R), invW)));
    if (cx0 >= 64) continue;
    const int cx1 = min(63, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, minX), R), invW)));
    if (cx1 < 0) continue;
    const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, minY), R), invH)));
    if (cy0 >= 48) continue;
    const int cy1 = min(47, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, minY), R), invH)));
    if (cy1 < 0) continue;
    const float xr = c.xr;
    uint32_t q[8];
#pragma unroll
    for (int w = 0; w < 8; ++w) q[w] = __shfl_sync(0xffffffffu, c.dw, w);
    unsigned long long k1 = ~0ull, k2 = ~0ull;
    for (int ix = cx0; ix <= cx1; ++ix) {
      const int c_lo = ix * 48 + cy0, c_hi = ix * 48 + cy1;
      const int e_lo = cstart[c_lo], e_hi = (c_hi + 1 < kGridCells) ? cstart[c_hi + 1] : n_in_grid;
      for (int e = e_lo + lane; e < e_hi; e += 32) {
        const uint32_t meta = smeta[e];
        const int oct = (int)(meta >> 28), idx = (int)(meta & 0xFFFF);
        if (oct < lvl - 1 || oct > lvl) continue;                                     // GetFeaturesInArea level window
        const float2 pt = sxy[e];
        if (!(fabsf(__fsub_rn(pt.x, x)) < R && fabsf(__fsub_rn(pt.y, y)) < R)) continue;
        if (frameObs[idx] > 0) continue;                                              // already holds an observed map point
        const float ur = uRight[idx];
        if (ur > 0.f && fabsf(__fsub_rn(xr, ur)) > R) continue;
        const int dist = ham8(q, d + (size_t)idx * 8);
        const unsigned long long key = ((unsigned long long)dist << 32) | (meta & 0x0FFFFFFFu);
        if (key < k1) { k2 = k1; k1 = key; }
        else if (key < k2) k2 = key;
      }
    }
    // warp merge of the two smallest keys
    const unsigned hi1 = (unsigned)(k1 >> 32), lo1 = (unsigned)k1;
    const unsigned ghi = __reduce_min_sync(0xffffffffu, hi1);
    if (ghi == 0xffffffffu) continue;                                                 // no candidate at all
    const unsigned glo = __reduce_min_sync(0xffffffffu, hi1 == ghi ? lo1 : 0xffffffffu);
    const bool owner = hi1 == ghi && lo1 == glo;
    const unsigned long long mine2 = owner ? k2 : k1;                                 // best key of "the rest" held by this lane
    const unsigned hi2 = (unsigned)(mine2 >> 32), lo2 = (unsigned)mine2;
    const unsigned shi = __reduce_min_sync(0xffffffffu, hi2);
    const unsigned slo = __reduce_min_sync(0xffffffffu, hi2 == shi ? lo2 : 0xffffffffu);
    const int bestDist = (int)ghi, bestIdx = (int)(glo & 0xFFFF);
    if (bestDist > 100) continue;                                                     // TH_HIGH
    const int bestLvl = k[bestIdx].octave;
    const int bestDist2 = shi == 0xffffffffu ? 256 : (int)shi;
    const int bestLvl2 = shi == 0xffffffffu ? -1 : k[slo & 0xFFFF].octave;
    if (bestLvl == bestLvl2 && (float)bestDist > __fmul_rn(a.ratio, (float)bestDist2)) continue;
    if (lane == 0) { assigned[bestIdx] = iMP; frameObs[bestIdx] = a.mpObs[mo + iMP]; }
    ++nm;
    __syncwarp();
  }
  if (lane == 0) a.nm[p] = nm;
}

// ------------------------------------------------------------------------------------------
// Latency forms of K8 / K9: ONE frame per call is the SLAM loop's pattern (src/Tracking.cc:1192, :886), so a call is one
// CTA on one SM and what it costs is that CTA's critical path.  Everything the candidate loops touch is therefore staged
// in shared memory by the whole CTA first (a dependent global load per candidate column was most of the kernel: two L2
// round trips for frameObs / uRight per column and map point), and the cell sort is done by all warps, not by one.
constexpr int kParThreads = 1024;


struct FrameStage {                // shared-memory views of one frame
  float2* sxy;                     // [cap] positions sorted by grid cell
  uint32_t* smeta;                 // [cap] octave<<28 | cell<<16 | keypoint index, same order
  int* soff;                       // [kGridCells + 2] cell c = [soff[c+1], soff[c+2])
  uint32_t* sdesc;                 // [cap][8] descriptors by keypoint index
  float* sur;                      // [cap] uRight by keypoint index
  uint32_t* sobs;                  // [(cap+31)/32] bit i: keypoint i holds a map point with Observations() > 0 (live)
  uint32_t* claim;                 // [(cap+31)/32] bit i: keypoint i was given such a map point by THIS call
  uint8_t* soct;                   // [cap] octave by keypoint index
  int* sasg;                       // [cap] the call's output `assigned`, kept here until the ordered phase is over
  int* stag;                       // [cap] who wants keypoint i in the round being resolved (order_round)
  int n_in_grid;
};

__host__ __device__ inline size_t frame_stage_bytes(int cap) {
  size_t b = (size_t)cap * 12 + (kGridCells + 2) * sizeof(int);
  b = ((b + 15) & ~(size_t)15) + (size_t)cap * 32 + (size_t)cap * 12 + (size_t)((cap + 31) / 32) * 8 + (size_t)cap;
  return (b + 15) & ~(size_t)15;
}

// All threads of the CTA.  Ends with a block barrier.  Keypoints above `max_octave` stay out of the grid (K6 searches level 0 only).
__device__ __forceinline__ void frame_stage(FrameStage& S, uint8_t* smem, int cap, const orbx_keypoint* k, const uint32_t* dg, const float* uRight,
                                            const int* frameObs, int n, float minX, float minY, float invW, float invH, int max_octave = 255) {
  __shared__ int s_wsum[32];
  __shared__ int s_total;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nth = blockDim.x;
  S.sxy = reinterpret_cast<float2*>(smem);
  S.smeta = reinterpret_cast<uint32_t*>(S.sxy + cap);
  S.soff = reinterpret_cast<int*>(S.smeta + cap);
  S.sdesc = reinterpret_cast<uint32_t*>((reinterpret_cast<uintptr_t>(S.soff + kGridCells + 2) + 15) & ~(uintptr_t)15);
  S.sur = reinterpret_cast<float*>(S.sdesc + (size_t)cap * 8);
  S.sasg = reinterpret_cast<int*>(S.sur + cap);
  S.stag = S.sasg + cap;
  S.sobs = reinterpret_cast<uint32_t*>(S.stag + cap);
  S.claim = S.sobs + (cap + 31) / 32;
  S.soct = reinterpret_cast<uint8_t*>(S.claim + (cap + 31) / 32);
  int* soff = S.soff;
  ORBX_CLK(fs0);
  for (int i = tid; i <= kGridCells + 1; i += nth) soff[i] = 0;
  for (int base = warp * 32; base < cap; base += nth) {                 // 32 keypoints per ballot
    const int i = base + lane;
    const uint32_t b = __ballot_sync(0xffffffffu, frameObs != nullptr && i < n && frameObs[i] > 0);
    if (lane == 0) { S.sobs[base >> 5] = b; S.claim[base >> 5] = 0u; }
  }
  for (int i = tid; i < n; i += nth) { S.sasg[i] = -1; S.stag[i] = 0; S.sur[i] = uRight ? uRight[i] : -1.f; }
  {
    const uint4* src = reinterpret_cast<const uint4*>(dg);
    uint4* dst = reinterpret_cast<uint4*>(S.sdesc);
    for (int i = tid; i < n * 2; i += nth) dst[i] = __ldg(src + i);
  }
  __syncthreads();
  ORBX_CLK(fs1);
  for (int i = tid; i < n; i += nth) {                                  // Frame::PosInGrid, src/Frame.cc:382-392
    const orbx_keypoint kp = k[i];
    S.soct[i] = (uint8_t)kp.octave;
    const int gx = (int)roundf(__fmul_rn(__fsub_rn(kp.x, minX), invW)), gy = (int)roundf(__fmul_rn(__fsub_rn(kp.y, minY), invH));
    if (gx >= 0 && gx < 64 && gy >= 0 && gy < 48 && kp.octave <= max_octave) atomicAdd(&soff[gx * 48 + gy + 1], 1);
  }
  __syncthreads();
  ORBX_CLK(fs2);
  // inclusive scan of soff[0 .. kGridCells]: a contiguous chunk per thread, warp scan of the chunk sums, scan of the warp sums
  const int ch = (kGridCells + nth) / nth, b0 = tid * ch;
  int sum = 0;
  for (int j = 0; j < ch; ++j) sum += (b0 + j <= kGridCells) ? soff[b0 + j] : 0;
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_wsum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const int own = lane < (nth >> 5) ? s_wsum[lane] : 0;
    int v = own;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    s_wsum[lane] = v - own;
    if (lane == 31) s_total = v;
  }
  __syncthreads();
  int run = s_wsum[warp] + incl - sum;
  for (int j = 0; j < ch; ++j)
    if (b0 + j <= kGridCells) { run += soff[b0 + j]; soff[b0 + j] = run; }   // soff[c+1] = end of cell c
  __syncthreads();
  ORBX_CLK(fs3);
  for (int i = tid; i < n; i += nth) {                                  // fill each cell from its end downwards: soff[c+1] = start of cell c
    const orbx_keypoint kp = k[i];
    const int gx = (int)roundf(__fmul_rn(__fsub_rn(kp.x, minX), invW)), gy = (int)roundf(__fmul_rn(__fsub_rn(kp.y, minY), invH));
    if (gx >= 0 && gx < 64 && gy >= 0 && gy < 48 && kp.octave <= max_octave) {
      const int c = gx * 48 + gy;
      const int pos = atomicSub(&soff[c + 1], 1) - 1;
      S.sxy[pos] = make_float2(kp.x, kp.y);
      S.smeta[pos] = ((uint32_t)kp.octave << 28) | ((uint32_t)c << 16) | (uint32_t)i;
    }
  }
  __syncthreads();
  ORBX_CLK(fs4);
  ORBX_CLK_PRINT("frame_stage copy / histogram / scan / fill", fs0, fs1, fs2, fs3, fs4);
  S.n_in_grid = s_total;
}

__device__ __forceinline__ bool bit_of(const uint32_t* w, int i) { return (w[i >> 5] >> (i & 31)) & 1u; }

// K8, latency form.
struct ProjMP { int inView, level, obs; float viewCos, x, y, xr; uint32_t dw; };

// Map point c's candidate list against S.sobs as it is now (warp-wide; every lane returns the same):
// {l0 | complete << 31, l1 | (Observations() > 0) << 31, l2, l3}, l = distance << 16 | keypoint in ascending (distance, cell,
// keypoint) = the reference's first-wins streaming over GetFeaturesInArea's order; kNoKey = none.
__device__ __forceinline__ uint4 proj_scan(const ProjArgs& a, const ProjMP& c, const FrameStage& S, float minX, float minY, float invW, float invH, int lane) {
  const uint4 none = make_uint4(0xffffffffu, kNoKey | (c.obs > 0 ? 0x80000000u : 0u), kNoKey, kNoKey);
  if (!c.inView) return none;
  const int lvl = min(max(c.level, 0), a.nlevels - 1);
  float r = ((double)c.viewCos > 0.998) ? 2.5f : 4.0f;
  if (a.th != 1.0f) r = __fmul_rn(r, a.th);
  const float R = __fmul_rn(r, a.scale[lvl]);
  const float x = c.x, y = c.y;
  // GetFeaturesInArea's cell window, src/Frame.cc:332-346: (x - mnMinX -+ r) * inv, left to right
  const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, minX), R), invW)));
  if (cx0 >= 64) return none;
  const int cx1 = min(63, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, minX), R), invW)));
  if (cx1 < 0) return none;
  const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, minY), R), invH)));
  if (cy0 >= 48) return none;
  const int cy1 = min(47, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, minY), R), invH)));
  if (cy1 < 0) return none;
  uint32_t q[8];
#pragma unroll
  for (int w = 0; w < 8; ++w) q[w] = __shfl_sync(0xffffffffu, c.dw, w);
  const int* cstart = S.soff + 1;
  unsigned long long k1 = ~0ull, k2 = ~0ull;
  int cnt = 0;
  for_each_in_columns(cx0, cx1, lane,
    [&](int ix, int& e_lo, int& e_hi) {
      const int c_hi = ix * 48 + cy1;
      e_lo = cstart[ix * 48 + cy0]; e_hi = (c_hi + 1 < kGridCells) ? cstart[c_hi + 1] : S.n_in_grid;
    },
    [&](int e, int) {
      const uint32_t meta = S.smeta[e];
      const int oct = (int)(meta >> 28), idx = (int)(meta & 0xFFFF);
      if (oct < lvl - 1 || oct > lvl) return;                                         // GetFeaturesInArea level window
      const float2 pt = S.sxy[e];
      if (!(fabsf(__fsub_rn(pt.x, x)) < R && fabsf(__fsub_rn(pt.y, y)) < R)) return;
      if (bit_of(S.sobs, idx)) return;                                                // already holds an observed map point
      const float ur = S.sur[idx];
      if (ur > 0.f && fabsf(__fsub_rn(c.xr, ur)) > R) return;
      const unsigned long long key = ((unsigned long long)ham8(q, S.sdesc + (size_t)idx * 8) << 32) | (meta & 0x0FFFFFFFu);
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
  return make_uint4(l[0] | (complete ? 0x80000000u : 0u), l[1] | (c.obs > 0 ? 0x80000000u : 0u), l[2], l[3]);
}

// the reference's acceptance of a (best, second best) pair, src/ORBmatcher.cc:108-124; keys distance << 16 | keypoint
__device__ __forceinline__ bool proj_accept(uint32_t k1, uint32_t k2, const uint8_t* soct, float ratio) {
  if (k1 == kNoKey) return false;
  const int bestDist = (int)(k1 >> 16);
  if (bestDist > 100) return false;                                                   // TH_HIGH
  const int bestLvl = soct[k1 & 0xFFFFu];
  const int bestDist2 = k2 == kNoKey ? 256 : (int)(k2 >> 16);
  const int bestLvl2 = k2 == kNoKey ? -1 : (int)soct[k2 & 0xFFFFu];
  return !(bestLvl == bestLvl2 && (float)bestDist > __fmul_rn(ratio, (float)bestDist2));
}

__global__ void __launch_bounds__(kParThreads) k_search_projection_par(ProjArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int n = a.n[p], m = a.m[p];
  int* frameObs = a.frameObs + (size_t)p * a.cap;
  int* assigned = a.assigned + (size_t)p * a.cap;
  const size_t mo = (size_t)p * a.mcap;
  uint4* spec = a.spec + mo;
  const float minX = a.gb[0], minY = a.gb[2], invW = a.gb[4], invH = a.gb[5];
  FrameStage S;
  ORBX_CLK(t0);
  frame_stage(S, smem, a.cap, a.k + (size_t)p * a.cap, reinterpret_cast<const uint32_t*>(a.d + (size_t)p * a.cap * 32), a.uRight + (size_t)p * a.cap,
              frameObs, n, minX, minY, invW, invH);
  auto fetch = [&](int i) {
    ProjMP q;
    q.inView = i < m ? a.inView[mo + i] : 0;
    q.level = i < m ? a.level[mo + i] : 0;
    q.obs = i < m ? a.mpObs[mo + i] : 0;
    q.viewCos = i < m ? a.viewCos[mo + i] : 0.f;
    q.x = i < m ? a.projX[mo + i] : 0.f; q.y = i < m ? a.projY[mo + i] : 0.f; q.xr = i < m ? a.projXR[mo + i] : 0.f;
    q.dw = (i < m && lane < 8) ? __ldg(reinterpret_cast<const uint32_t*>(a.mpDesc + (mo + i) * 32) + lane) : 0u;
    return q;
  };
  // phase A
  ORBX_CLK(t1);
  {
    ProjMP nx = fetch(warp);
    for (int iMP = warp; iMP < m; iMP += nwarps) {
      const ProjMP c = nx;
      nx = fetch(iMP + nwarps);
      const uint4 l = proj_scan(a, c, S, minX, minY, invW, invH, lane);
      if (lane == 0) spec[iMP] = l;
    }
  }
  __syncthreads();
  ORBX_CLK(t2);
  // phase B
  if (warp == 0) {
    int nm = 0;
    ORBX_STAT(int dbg_serial = 0; int dbg_iters = 0;)
    int epoch = 0;
    const uint4 nokeys = make_uint4(0xffffffffu, kNoKey, kNoKey, kNoKey);
    uint4 nxt = lane < m ? spec[lane] : nokeys;
    for (int base = 0; base < m; base += 32) {
      const int iMP = base + lane;
      const uint4 my4 = nxt;
      nxt = iMP + 32 < m ? spec[iMP + 32] : nokeys;                       // the next round's lists are on their way while this one is resolved
      const bool complete = my4.x >> 31, obs = my4.y >> 31;              // obs: Observations() > 0
      const uint32_t key[4] = {my4.x & 0x7FFFFFFFu, my4.y & 0x7FFFFFFFu, my4.z, my4.w};
      uint32_t todo = __ballot_sync(0xffffffffu, key[0] != kNoKey);      // lanes whose map point is still to be decided
      while (todo) {
        const bool mine = (todo >> lane) & 1u;
        unsigned hitP = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (key[k] != kNoKey && bit_of(S.claim, (int)(key[k] & 0xFFFFu))) hitP |= 1u << k;
        bool acc = false, starved = false;
        int bestIdx = -1;
        const int iters = order_round(key, hitP, S.stag, epoch, lane, [&](unsigned hit) {
          uint32_t k1 = kNoKey, k2 = kNoKey;                             // the first two keys nobody earlier has claimed
#pragma unroll
          for (int k = 3; k >= 0; --k)
            if (key[k] != kNoKey && !((hit >> k) & 1u)) { k2 = k1; k1 = key[k]; }
          starved = mine && !complete && k2 == kNoKey;                   // the list ran out before two free keys were found
          acc = mine && proj_accept(k1, k2, S.soct, a.ratio);
          bestIdx = acc ? (int)(k1 & 0xFFFFu) : -1;
          return (acc && obs) ? bestIdx : -1;
        });
        ORBX_STAT(dbg_iters += iters;)
        (void)iters;
        // lanes below the first starved one are final (a lane depends on earlier lanes only): they commit together
        const uint32_t sm = __ballot_sync(0xffffffffu, starved);
        const int jstar = sm ? __ffs((int)sm) - 1 : 32;
        const bool commit = acc && lane < jstar;
        if (commit) {
          atomicMax(&S.sasg[bestIdx], iMP);                              // a keypoint given away twice keeps the later map point
          if (obs) { atomicOr(&S.claim[bestIdx >> 5], 1u << (bestIdx & 31)); atomicOr(&S.sobs[bestIdx >> 5], 1u << (bestIdx & 31)); }
        }
        nm += __popc(__ballot_sync(0xffffffffu, commit));
        __syncwarp();
        if (!sm) break;
        // the starved map point is scanned again against the live occupancy (exact), then the lanes after it go round again
        ORBX_STAT(++dbg_serial;)
        const ProjMP c = fetch(base + jstar);
        const uint4 l = proj_scan(a, c, S, minX, minY, invW, invH, lane);
        const uint32_t k1 = l.x & 0x7FFFFFFFu;
        if (proj_accept(k1, l.y & 0x7FFFFFFFu, S.soct, a.ratio)) {
          const int bi = (int)(k1 & 0xFFFFu);
          if (lane == 0) {
            S.sasg[bi] = base + jstar;
            if (l.y >> 31) { S.claim[bi >> 5] |= 1u << (bi & 31); S.sobs[bi >> 5] |= 1u << (bi & 31); }
          }
          ++nm;
        }
        __syncwarp();
        todo &= ~((2u << jstar) - 1u);
      }
    }
    if (lane == 0) a.nm[p] = nm;
    ORBX_STAT(if (lane == 0 && blockIdx.x == 0) printf("K8 phase B: %d re-scans, %d fixpoint iterations, nm %d of %d\n", dbg_serial, dbg_iters, nm, m);)
  }
  __syncthreads();
  ORBX_CLK(t3);
  // results leave together: assigned[], and frameObs[i] = Observations() of the map point keypoint i ended up with
  for (int i = tid; i < n; i += blockDim.x) {
    const int v = S.sasg[i];
    assigned[i] = v;
    if (v >= 0) frameObs[i] = a.mpObs[mo + v];
  }
  ORBX_CLK(t4);
  ORBX_CLK_PRINT("k_search_projection_par stage / A / B / flush", t0, t1, t2, t3, t4);
}

// ------------------------------------------------------------------------------------------
// K6, 1024-thread form (the monocular initialiser asks about one frame pair per call; batches take it too).  Frame 2 is staged
// in shared memory in grid-cell order (frame_stage), so a query touches the cells of its window instead of every level-0
// keypoint, and the walk over the queries - order matters: a query skips a frame-2 feature whose recorded match distance is
// not above its own (vMatchedDistance), and may take a feature away from an earlier query - is split in two:
//   A (all warps, no state): every level-0 keypoint of frame 1 gets its eight nearest window candidates, sorted by
//     (distance, cell, index) = the reference's first-wins order, and whether that is all of them;
//   B (warp 0), 32 queries per pass, one per lane: each lane judges its list against the recorded distances as they stand -
//     best and second best are its first two entries whose recorded distance is above their own.  A lane is held back when
//     an earlier lane of the pass is about to record, on one of those two features, a distance at or below the lane's own
//     (claims of one feature by successive lanes have falling distances, so the latest earlier claimer decides); the lanes
//     before the first held-back lane commit together - a feature claimed twice among them goes to the later lane, as in the
//     loop - and the pass resumes from that lane.  A lane whose truncated list has fewer than two survivors is re-scanned
//     exactly, alone.  Every decision is therefore taken against exactly the state the reference's loop would have.
__host__ __device__ inline size_t init_lat_bytes(int cap) { return frame_stage_bytes(cap) + (size_t)cap * (4 + 4 + 4 + 2 + 1) + 16; }

struct InitQ { float x, y; uint32_t dw; };

// frame-1 query (prev position x, y; descriptor q): its candidates in frame 2, level 0, |dx| < r and |dy| < r, whose recorded
// match distance is above their own when `live`; the list format of proj_scan.
constexpr int kInitK = 8;
__device__ __forceinline__ void init_scan(const InitMatchArgs& a, const InitQ& c, const FrameStage& S, const int* smd, bool live, const float* gbv, int lane,
                                          uint32_t (&l)[kInitK], bool& complete) {
#pragma unroll
  for (int k = 0; k < kInitK; ++k) l[k] = kNoKey;
  complete = true;
  const float minX = gbv[0], minY = gbv[2], invW = gbv[4], invH = gbv[5], r = a.window, x = c.x, y = c.y;
  // GetFeaturesInArea's cell window, src/Frame.cc:332-346
  const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, minX), r), invW)));
  if (cx0 >= 64) return;
  const int cx1 = min(63, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, minX), r), invW)));
  if (cx1 < 0) return;
  const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, minY), r), invH)));
  if (cy0 >= 48) return;
  const int cy1 = min(47, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, minY), r), invH)));
  if (cy1 < 0) return;
  uint32_t q[8];
#pragma unroll
  for (int w = 0; w < 8; ++w) q[w] = __shfl_sync(0xffffffffu, c.dw, w);
  const int* cstart = S.soff + 1;
  unsigned long long k1 = ~0ull, k2 = ~0ull;
  int cnt = 0;
  for_each_in_columns(cx0, cx1, lane,
    [&](int ix, int& e_lo, int& e_hi) {
      const int c_hi = ix * 48 + cy1;
      e_lo = cstart[ix * 48 + cy0]; e_hi = (c_hi + 1 < kGridCells) ? cstart[c_hi + 1] : S.n_in_grid;
    },
    [&](int e, int) {
      const uint32_t meta = S.smeta[e];
      if (meta >> 28) return;                                                         // level 0 only (minLevel = maxLevel = 0)
      const int idx = (int)(meta & 0xFFFF);
      const float2 pt = S.sxy[e];
      if (!(fabsf(__fsub_rn(pt.x, x)) < r && fabsf(__fsub_rn(pt.y, y)) < r)) return;
      const int d = ham8(q, S.sdesc + (size_t)idx * 8);
      if (live && smd[idx] <= d) return;                                              // vMatchedDistance[i2] <= dist, src/ORBmatcher.cc:436
      const unsigned long long key = ((unsigned long long)d << 32) | (meta & 0x0FFFFFFFu);
      ++cnt;
      if (key < k1) { k2 = k1; k1 = key; }
      else if (key < k2) k2 = key;
    });
  unsigned hi[kInitK], lo[kInitK];
  warp_topk<kInitK>(k1, k2, cnt, hi, lo, complete);
#pragma unroll
  for (int k = 0; k < kInitK; ++k) l[k] = hi[k] == 0xffffffffu ? kNoKey : ((hi[k] << 16) | (lo[k] & 0xFFFFu));
}

__global__ void __launch_bounds__(kParThreads) k_search_init_lat(InitMatchArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ int s_hist[kHisto];
  __shared__ int s_bcnt[kParThreads / 32];
  __shared__ int s_nm;
  __shared__ float s_gb[6];
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nth = blockDim.x, nwarps = nth >> 5;
  const int n1 = a.n1[p], n2 = a.n2[p];
  const orbx_keypoint* k1 = a.k1 + (size_t)p * a.cap;
  const orbx_keypoint* k2 = a.k2 + (size_t)p * a.cap;
  const uint32_t* d1 = reinterpret_cast<const uint32_t*>(a.d1 + (size_t)p * a.cap * 32);
  float* prev = a.prev + (size_t)p * a.cap * 2;
  int* m12 = a.m12 + (size_t)p * a.cap;
  uint4* spec = reinterpret_cast<uint4*>(a.scratch + (size_t)p * a.cap * 8);           // candidate lists by query-list position, two uint4 each
  uint8_t* tail = smem + frame_stage_bytes(a.cap);
  int* sm12 = reinterpret_cast<int*>(tail);                    // [cap] vnMatches12
  int* sm21 = sm12 + a.cap;                                    // [cap] vnMatches21
  int* smd = sm21 + a.cap;                                     // [cap] vMatchedDistance
  uint16_t* sq = reinterpret_cast<uint16_t*>(smd + a.cap);     // [cap] the level-0 keypoints of frame 1, in order
  uint8_t* sbin = reinterpret_cast<uint8_t*>(sq + a.cap);      // [cap] rotation bin of a frame-1 keypoint's latest match
  ORBX_CLK(t0);
  for (int i = tid; i < a.cap; i += nth) { sm12[i] = -1; sm21[i] = -1; smd[i] = INT_MAX; sbin[i] = 255; }
  if (tid < kHisto) s_hist[tid] = 0;
  if (tid < 6) s_gb[tid] = a.gb[tid];
  int nq = 0;
  for (int base = 0; base < n1; base += nth) {                 // ordered list of the queries
    const int i = base + tid;
    const bool ok = i < n1 && k1[i].octave <= 0;
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) s_bcnt[warp] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < nwarps; ++w) { const int c = s_bcnt[w]; before += w < warp ? c : 0; total += c; }
    if (ok) sq[nq + before + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)i;
    nq += total;
    __syncthreads();
  }
  FrameStage S;
  frame_stage(S, smem, a.cap, k2, reinterpret_cast<const uint32_t*>(a.d2 + (size_t)p * a.cap * 32), nullptr, nullptr, n2, a.gb[0], a.gb[2], a.gb[4], a.gb[5], 0);
  ORBX_CLK(t1);
  auto fetch = [&](int c) {
    InitQ q;
    const bool ok = c < nq;
    const int i = ok ? sq[c] : 0;
    q.x = ok ? prev[2 * i] : 0.f; q.y = ok ? prev[2 * i + 1] : 0.f;
    q.dw = (ok && lane < 8) ? __ldg(d1 + (size_t)i * 8 + lane) : 0u;
    return q;
  };
  // phase A
  {
    InitQ nx = fetch(warp);
    for (int c = warp; c < nq; c += nwarps) {
      const InitQ cur = nx;
      nx = fetch(c + nwarps);
      uint32_t l[kInitK]; bool complete;
      init_scan(a, cur, S, smd, false, s_gb, lane, l, complete);
      if (lane == 0) { spec[2 * c] = make_uint4(l[0] | (complete ? 0x80000000u : 0u), l[1], l[2], l[3]); spec[2 * c + 1] = make_uint4(l[4], l[5], l[6], l[7]); }
    }
  }
  __syncthreads();
  ORBX_CLK(t2);
  // phase B
  if (warp == 0) {
    ORBX_STAT(int dbg_pass = 0; int dbg_rs = 0;)
    const uint4 nokeys = make_uint4(kNoKey, kNoKey, kNoKey, kNoKey);
    uint4 nx0 = lane < nq ? spec[2 * lane] : make_uint4(0xffffffffu, kNoKey, kNoKey, kNoKey), nx1 = lane < nq ? spec[2 * lane + 1] : nokeys;
    uint32_t* tag = reinterpret_cast<uint32_t*>(S.stag);        // [cap] zero between passes: the lanes of the pass that claim frame-2 feature i
    for (int base = 0; base < nq; base += 32) {
      const int c = base + lane;
      const int i1 = c < nq ? sq[c] : 0;
      const uint4 m0 = nx0, m1 = nx1;
      nx0 = c + 32 < nq ? spec[2 * (c + 32)] : make_uint4(0xffffffffu, kNoKey, kNoKey, kNoKey); nx1 = c + 32 < nq ? spec[2 * (c + 32) + 1] : nokeys;
      const float ang1 = (c < nq && a.checkOri) ? k1[i1].angle : 0.f;
      const bool complete = m0.x >> 31;
      const uint32_t key[kInitK] = {m0.x & 0x7FFFFFFFu, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
      uint32_t todo = __ballot_sync(0xffffffffu, key[0] != kNoKey);
      while (todo) {
        const bool live = (todo >> lane) & 1u;
        ORBX_STAT(++dbg_pass;)
        // the first two entries whose recorded distance is above their own
        uint32_t f1 = kNoKey, f2 = kNoKey;
#pragma unroll
        for (int k = kInitK - 1; k >= 0; --k)
          if (key[k] != kNoKey && !(smd[key[k] & 0xFFFFu] <= (int)(key[k] >> 16))) { f2 = f1; f1 = key[k]; }
        const bool starved = live && !complete && f2 == kNoKey;
        const int best = (int)(f1 >> 16), bi = (int)(f1 & 0xFFFFu);
        // bestDist2 stays INT_MAX without a second candidate: the product is then far above any distance
        const bool pass = live && !starved && f1 != kNoKey && best <= 50 && (float)best < __fmul_rn(f2 == kNoKey ? (float)INT_MAX : (float)(int)(f2 >> 16), a.ratio);
        if (pass) atomicOr(&tag[bi], 1u << lane);
        __syncwarp();
        // held back: an earlier lane is about to record, on my best or second best, a distance that takes it off my list
        const uint32_t e1 = (live && f1 != kNoKey) ? tag[f1 & 0xFFFFu] & ((1u << lane) - 1u) : 0u;
        const uint32_t e2 = (live && f2 != kNoKey) ? tag[f2 & 0xFFFFu] & ((1u << lane) - 1u) : 0u;
        const int b1 = __shfl_sync(0xffffffffu, best, e1 ? 31 - __clz((int)e1) : lane);
        const int b2 = __shfl_sync(0xffffffffu, best, e2 ? 31 - __clz((int)e2) : lane);
        const bool held = (e1 && b1 <= (int)(f1 >> 16)) || (e2 && b2 <= (int)(f2 >> 16));
        const uint32_t stop = __ballot_sync(0xffffffffu, live && (held || starved));
        const int first = stop ? __ffs((int)stop) - 1 : 32;
        const uint32_t cm = __ballot_sync(0xffffffffu, pass && lane < first);
        if (pass && lane < first) {                             // src/ORBmatcher.cc:449-466
          if (a.checkOri) { const int bin = rot_bin(ang1, k2[bi].angle); sbin[i1] = (uint8_t)bin; atomicAdd(&s_hist[bin], 1); }
          if (!(tag[bi] & cm & ~((2u << lane) - 1u))) {         // not taken away again by a later lane of this commit
            const int old = sm21[bi];
            if (old >= 0) sm12[old] = -1;
            sm12[i1] = bi; sm21[bi] = i1; smd[bi] = best;
          }
        }
        __syncwarp();
        if (pass) tag[bi] = 0u;
        __syncwarp();
        if (first == 32) break;
        if (__shfl_sync(0xffffffffu, (int)starved, first)) {    // exact re-scan against the recorded distances as they stand
          ORBX_STAT(++dbg_rs;)
          const int cs = base + first, is = __shfl_sync(0xffffffffu, i1, first);
          uint32_t l[kInitK]; bool lc;
          init_scan(a, fetch(cs), S, smd, true, s_gb, lane, l, lc);
          const uint32_t g1 = l[0], g2 = l[1];
          const int rb = (int)(g1 >> 16), rbi = (int)(g1 & 0xFFFFu);
          if (g1 != kNoKey && rb <= 50 && (float)rb < __fmul_rn(g2 == kNoKey ? (float)INT_MAX : (float)(int)(g2 >> 16), a.ratio)) {
            if (lane == 0) {
              const int old = sm21[rbi];
              if (old >= 0) sm12[old] = -1;
              sm12[is] = rbi; sm21[rbi] = is; smd[rbi] = rb;
              if (a.checkOri) { const int bin = rot_bin(k1[is].angle, k2[rbi].angle); sbin[is] = (uint8_t)bin; s_hist[bin] += 1; }
            }
          }
          __syncwarp();
          todo &= ~((2u << first) - 1u);
        } else {
          todo &= ~((1u << first) - 1u);                       // lane `first` judges again, now against the committed state
        }
      }
    }
    ORBX_STAT(if (lane == 0 && blockIdx.x == 0) printf("K6 phase B: %d queries, %d passes, %d re-scans\n", nq, dbg_pass, dbg_rs);)
  }
  __syncthreads();
  ORBX_CLK(t3);
  // rotation consistency (src/ORBmatcher.cc:474-492), vbPrevMatched update (:495-497), write-back: whole CTA
  if (tid == 0) s_nm = 0;
  int k_a = -1, k_b = -1, k_c = -1;
  if (a.checkOri) three_maxima(s_hist, k_a, k_b, k_c);
  __syncthreads();
  int nm = 0;
  for (int i = tid; i < n1; i += nth) {
    int m = sm12[i];
    if (a.checkOri && m >= 0) { const int b = sbin[i]; if (b != k_a && b != k_b && b != k_c) m = -1; }
    m12[i] = m;
    if (m >= 0) { prev[2 * i] = k2[m].x; prev[2 * i + 1] = k2[m].y; ++nm; }
  }
  if (nm) atomicAdd(&s_nm, nm);
  __syncthreads();
  if (tid == 0) a.nm[p] = s_nm;
  ORBX_CLK(t4);
  ORBX_CLK_PRINT("k_search_init_lat list + stage / A / B / rotation + flush", t0, t1, t2, t3, t4);
}

int launch_search_projection(const ProjArgs& a, cudaStream_t st) {
  const size_t base = (size_t)a.cap * 12 + (kGridCells + 2) * sizeof(int);
  const size_t with_desc = ((base + 15) & ~(size_t)15) + (size_t)a.cap * 32;
  // latency mode (a few frames per call): one warp per SM anyway, so spend the shared memory on the descriptors;
  // throughput mode keeps the small footprint (6 CTAs per SM)
  const size_t par = frame_stage_bytes(a.cap);
  if (par <= 200 * 1024 && !getenv("ORBX_PROJ_SERIAL")) {               // also for batches: 57 us of one SM per frame against 850 us of one warp
    cudaFuncSetAttribute(k_search_projection_par, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)par);
    k_search_projection_par<<<a.n_pairs, kParThreads, par, st>>>(a);
  } else if (a.n_pairs <= 148 && with_desc <= 200 * 1024) {
    cudaFuncSetAttribute(k_search_projection<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)with_desc);
    k_search_projection<true><<<a.n_pairs, 32, with_desc, st>>>(a);
  } else {
    if (base > 48 * 1024) cudaFuncSetAttribute(k_search_projection<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)base);
    k_search_projection<false><<<a.n_pairs, 32, base, st>>>(a);
  }
  return 1;
}

// ------------------------------------------------------------------------------------------
// K9  k_search_projection_frame <- ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono)
//                                  src/ORBmatcher.cc:1328-1470, the motion-model search (src/Tracking.cc:886,892).
// Same structure as K8.  Camera-frame coordinates follow OpenCV's small-matrix gemm (float products
// and sums, left to right, + t; pinned against cv2.gemm in tests/test_oracle_cv2.py).  The rotation
// histogram is applied after the ordered pass from a per-match event list.
template <bool DSM>
__global__ void __launch_bounds__(32) k_search_projection_frame(ProjFrameArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  float2* sxy = reinterpret_cast<float2*>(smem);
  uint32_t* smeta = reinterpret_cast<uint32_t*>(sxy + a.cap);
  int* soff = reinterpret_cast<int*>(smeta + a.cap);
  uint32_t* sdesc = reinterpret_cast<uint32_t*>((reinterpret_cast<uintptr_t>(soff + kGridCells + 2) + 15) & ~(uintptr_t)15);   // DSM: [cap][8], see K8
  __shared__ int s_hist[kHisto];
  const int p = blockIdx.x, lane = threadIdx.x;
  const int n = a.n[p], nLast = a.nLast[p], mode = a.mode[p];
  const orbx_keypoint* k = a.k + (size_t)p * a.cap;
  const uint32_t* dg = reinterpret_cast<const uint32_t*>(a.d + (size_t)p * a.cap * 32);
  const float* uRight = a.uRight + (size_t)p * a.cap;
  int* frameObs = a.frameObs + (size_t)p * a.cap;
  int* assigned = a.assigned + (size_t)p * a.cap;
  int* events = a.events + (size_t)p * a.lcap;
  const size_t lo = (size_t)p * a.lcap;
  const float* T = a.pose + (size_t)p * 12;

  for (int i = lane; i < n; i += 32) assigned[i] = -1;
  if (lane < kHisto) s_hist[lane] = 0;
  if (DSM) {
    const uint4* src = reinterpret_cast<const uint4*>(dg);
    uint4* dst = reinterpret_cast<uint4*>(sdesc);
    for (int i = lane; i < n * 2; i += 32) dst[i] = __ldg(src + i);
  }
  const uint32_t* d = DSM ? sdesc : dg;
  const float minX = a.gb[0], maxX = a.gb[1], minY = a.gb[2], maxY = a.gb[3], invW = a.gb[4], invH = a.gb[5];
  const int n_in_grid = grid_sort(k, n, minX, minY, invW, invH, sxy, smeta, soff, lane);
  const int* cstart = soff + 1;
  float Tr[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) Tr[i] = T[i];

  // The last frame's keypoints that carry a map point are a minority (a few hundred of ~2000): their indices come from
  // ballots over 32 validity flags at a time (the next block's flags are loaded one block ahead), and the fields of the
  // next valid one are fetched while the current one is processed.  Lanes 0..7 hold one descriptor word each.
  struct LP { int octave; float X0, X1, X2; uint32_t dw; };
  auto fetch = [&](int i) {
    LP q;
    const bool ok = i >= 0;
    q.octave = ok ? a.lastOctave[lo + i] : 0;
    q.X0 = ok ? a.lastXw[(lo + i) * 3] : 0.f; q.X1 = ok ? a.lastXw[(lo + i) * 3 + 1] : 0.f; q.X2 = ok ? a.lastXw[(lo + i) * 3 + 2] : 0.f;
    q.dw = (ok && lane < 8) ? __ldg(reinterpret_cast<const uint32_t*>(a.mpDesc + (lo + i) * 32) + lane) : 0u;
    return q;
  };
  int vbase = 0;
  uint32_t vmask = __ballot_sync(0xffffffffu, lane < nLast && a.lastValid[lo + lane]);
  uint32_t vnext = __ballot_sync(0xffffffffu, 32 + lane < nLast && a.lastValid[lo + 32 + lane]);
  auto pop = [&]() {                                   // next valid index in ascending order, -1 at the end (warp-uniform)
    while (vmask == 0u) {
      vbase += 32;
      if (vbase >= nLast) return -1;
      vmask = vnext;
      vnext = __ballot_sync(0xffffffffu, vbase + 32 + lane < nLast && a.lastValid[lo + vbase + 32 + lane]);
    }
    const int b = __ffs((int)vmask) - 1;
    vmask &= vmask - 1u;
    return vbase + b;
  };
  int nm = 0;
  int i = pop();
  LP nx = fetch(i);
  while (i >= 0) {
    const LP c = nx;
    const int inext = pop();
    nx = fetch(inext);
    const int icur = i;
    i = inext;
    const float X0 = c.X0, X1 = c.X1, X2 = c.X2;
    float c3[3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
      c3[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(Tr[3 * r], X0), __fmul_rn(Tr[3 * r + 1], X1)), __fmul_rn(Tr[3 * r + 2], X2)), Tr[9 + r]);
    const float invzc = (float)(1.0 / (double)c3[2]);
    if (invzc < 0.f) continue;
    const float u = __fadd_rn(__fmul_rn(__fmul_rn(a.fx, c3[0]), invzc), a.cx), v = __fadd_rn(__fmul_rn(__fmul_rn(a.fy, c3[1]), invzc), a.cy);
    if (u < minX || u > maxX || v < minY || v > maxY) continue;                        // src/ORBmatcher.cc:1373-1376
    if (!(u == u) || !(v == v)) continue;                                              // NaN never passes the reference's window either
    const int oct = min(max(c.octave, 0), a.nlevels - 1);
    const float R = __fmul_rn(a.th, a.scale[oct]);
    const int l_lo = mode == 1 ? oct : (mode == 2 ? 0 : oct - 1), l_hi = mode == 1 ? 99 : (mode == 2 ? oct : oct + 1);
    const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, minX), R), invW)));
    if (cx0 >= 64) continue;
    const int cx1 = min(63, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, minX), R), invW)));
    if (cx1 < 0) continue;
    const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, minY), R), invH)));
    if (cy0 >= 48) continue;
    const int cy1 = min(47, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, minY), R), invH)));
    if (cy1 < 0) continue;
    const float ur_pred = __fsub_rn(u, __fmul_rn(a.bf, invzc));
    uint32_t q[8];
#pragma unroll
    for (int w = 0; w < 8; ++w) q[w] = __shfl_sync(0xffffffffu, c.dw, w);
    unsigned long long best = ~0ull;
    for (int ix = cx0; ix <= cx1; ++ix) {
      const int c_lo = ix * 48 + cy0, c_hi = ix * 48 + cy1;
      const int e_lo = cstart[c_lo], e_hi = (c_hi + 1 < kGridCells) ? cstart[c_hi + 1] : n_in_grid;
      for (int e = e_lo + lane; e < e_hi; e += 32) {
        const uint32_t meta = smeta[e];
        const int oct2 = (int)(meta >> 28), idx = (int)(meta & 0xFFFF);
        if (oct2 < l_lo || oct2 > l_hi) continue;
        const float2 pt = sxy[e];
        if (!(fabsf(__fsub_rn(pt.x, u)) < R && fabsf(__fsub_rn(pt.y, v)) < R)) continue;
        if (frameObs[idx] > 0) continue;
        const float ur = uRight[idx];
        if (ur > 0.f && fabsf(__fsub_rn(ur_pred, ur)) > R) continue;
        const unsigned long long key = ((unsigned long long)ham8(q, d + (size_t)idx * 8) << 32) | (meta & 0x0FFFFFFFu);
        best = min(best, key);
      }
    }
    const unsigned hi1 = (unsigned)(best >> 32), lo1 = (unsigned)best;
    const unsigned ghi = __reduce_min_sync(0xffffffffu, hi1);
    if (ghi > 100u) continue;                                                         // none, or worse than TH_HIGH
    const unsigned glo = __reduce_min_sync(0xffffffffu, hi1 == ghi ? lo1 : 0xffffffffu);
    const int bestIdx = (int)(glo & 0xFFFF);
    if (lane == 0) {
      assigned[bestIdx] = icur;
      frameObs[bestIdx] = a.mpObs[lo + icur];
      int bin = 0;
      if (a.checkOri) { bin = rot_bin(a.lastAngle[lo + icur], k[bestIdx].angle); s_hist[bin] += 1; }
      events[nm] = (bin << 16) | bestIdx;
    }
    ++nm;
    __syncwarp();
  }
  __syncwarp();
  if (a.checkOri) {
    int k_a, k_b, k_c;
    three_maxima(s_hist, k_a, k_b, k_c);
    int drop = 0;
    for (int e = lane; e < nm; e += 32) {
      const int bin = events[e] >> 16, idx = events[e] & 0xFFFF;
      if (bin != k_a && bin != k_b && bin != k_c) { assigned[idx] = -1; ++drop; }
    }
    nm -= __reduce_add_sync(0xffffffffu, drop);
  }
  if (lane == 0) a.nm[p] = nm;
}

// ------------------------------------------------------------------------------------------
// K9, latency form (see K8's): the frame side staged by frame_stage; the last-frame keypoints that carry a map point are
// compacted (in index order) into a list in shared memory; phase A - every entry of the list gets the smallest candidate
// key against the current frame as it is on entry, entries strided over the warps, the next entry's fields in flight;
// phase B - warp 0 walks the list in order, re-scans only the entries whose keypoint was claimed meanwhile, applies
// TH_HIGH, claims, and records the rotation events.
struct ProjLP { int octave, obs; float X0, X1, X2; uint32_t dw; };

// list entry c's candidate list against S.sobs as it is now, in K8's format (proj_scan)
__device__ __forceinline__ uint4 projf_scan(const ProjFrameArgs& a, const ProjLP& c, const float* Tr, int mode, const FrameStage& S,
                                            const float* gbv, int lane) {
  const uint4 none = make_uint4(0xffffffffu, kNoKey | (c.obs > 0 ? 0x80000000u : 0u), kNoKey, kNoKey);
  const float minX = gbv[0], maxX = gbv[1], minY = gbv[2], maxY = gbv[3], invW = gbv[4], invH = gbv[5];
  float c3[3];
#pragma unroll
  for (int r = 0; r < 3; ++r)
    c3[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(Tr[3 * r], c.X0), __fmul_rn(Tr[3 * r + 1], c.X1)), __fmul_rn(Tr[3 * r + 2], c.X2)), Tr[9 + r]);
  const float invzc = (float)(1.0 / (double)c3[2]);
  if (invzc < 0.f) return none;
  const float u = __fadd_rn(__fmul_rn(__fmul_rn(a.fx, c3[0]), invzc), a.cx), v = __fadd_rn(__fmul_rn(__fmul_rn(a.fy, c3[1]), invzc), a.cy);
  if (u < minX || u > maxX || v < minY || v > maxY) return none;                     // src/ORBmatcher.cc:1373-1376
  if (!(u == u) || !(v == v)) return none;                                           // NaN never passes the reference's window either
  const int oct = min(max(c.octave, 0), a.nlevels - 1);
  const float R = __fmul_rn(a.th, a.scale[oct]);
  const int l_lo = mode == 1 ? oct : (mode == 2 ? 0 : oct - 1), l_hi = mode == 1 ? 99 : (mode == 2 ? oct : oct + 1);
  const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, minX), R), invW)));
  if (cx0 >= 64) return none;
  const int cx1 = min(63, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, minX), R), invW)));
  if (cx1 < 0) return none;
  const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, minY), R), invH)));
  if (cy0 >= 48) return none;
  const int cy1 = min(47, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, minY), R), invH)));
  if (cy1 < 0) return none;
  const float ur_pred = __fsub_rn(u, __fmul_rn(a.bf, invzc));
  uint32_t q[8];
#pragma unroll
  for (int w = 0; w < 8; ++w) q[w] = __shfl_sync(0xffffffffu, c.dw, w);
  const int* cstart = S.soff + 1;
  unsigned long long k1 = ~0ull, k2 = ~0ull;
  int cnt = 0;
  for_each_in_columns(cx0, cx1, lane,
    [&](int ix, int& e_lo, int& e_hi) {
      const int c_hi = ix * 48 + cy1;
      e_lo = cstart[ix * 48 + cy0]; e_hi = (c_hi + 1 < kGridCells) ? cstart[c_hi + 1] : S.n_in_grid;
    },
    [&](int e, int) {
      const uint32_t meta = S.smeta[e];
      const int oct2 = (int)(meta >> 28), idx = (int)(meta & 0xFFFF);
      if (oct2 < l_lo || oct2 > l_hi) return;
      const float2 pt = S.sxy[e];
      if (!(fabsf(__fsub_rn(pt.x, u)) < R && fabsf(__fsub_rn(pt.y, v)) < R)) return;
      if (bit_of(S.sobs, idx)) return;
      const float ur = S.sur[idx];
      if (ur > 0.f && fabsf(__fsub_rn(ur_pred, ur)) > R) return;
      const unsigned long long key = ((unsigned long long)ham8(q, S.sdesc + (size_t)idx * 8) << 32) | (meta & 0x0FFFFFFFu);
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
  return make_uint4(l[0] | (complete ? 0x80000000u : 0u), l[1] | (c.obs > 0 ? 0x80000000u : 0u), l[2], l[3]);
}

__global__ void __launch_bounds__(kParThreads) k_search_projection_frame_par(ProjFrameArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ int s_hist[kHisto];
  __shared__ float s_gb[6], s_Tr[12];
  __shared__ int s_bcnt[kParThreads / 32 + 1];
  __shared__ int s_nm;
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nth = blockDim.x, nwarps = nth >> 5;
  const int n = a.n[p], nLast = a.nLast[p], mode = a.mode[p];
  const orbx_keypoint* k = a.k + (size_t)p * a.cap;
  int* frameObs = a.frameObs + (size_t)p * a.cap;
  int* assigned = a.assigned + (size_t)p * a.cap;
  const size_t lo = (size_t)p * a.lcap;
  uint4* spec = a.spec + lo;                                             // indexed by list position
  uint16_t* svalid = reinterpret_cast<uint16_t*>(smem + frame_stage_bytes(a.cap));   // [lcap] last-frame indices with a map point, ascending
  int* sev = reinterpret_cast<int*>(smem + frame_stage_bytes(a.cap) + (((size_t)a.lcap * 2 + 15) & ~(size_t)15));   // [lcap] accepted matches
  if (tid < kHisto) s_hist[tid] = 0;
  if (tid < 6) s_gb[tid] = a.gb[tid];
  if (tid >= 32 && tid < 44) s_Tr[tid - 32] = a.pose[(size_t)p * 12 + tid - 32];
  // ordered compaction of the valid last-frame keypoints: rounds of blockDim.x flags
  ORBX_CLK(t0);
  int nvalid = 0;
  for (int base = 0; base < nLast; base += nth) {
    const int i = base + tid;
    const bool ok = i < nLast && a.lastValid[lo + i];
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) s_bcnt[warp] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < nwarps; ++w) { const int c = s_bcnt[w]; before += w < warp ? c : 0; total += c; }
    if (ok) svalid[nvalid + before + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)i;
    nvalid += total;
    __syncthreads();
  }
  FrameStage S;
  ORBX_CLK(t1);
  frame_stage(S, smem, a.cap, k, reinterpret_cast<const uint32_t*>(a.d + (size_t)p * a.cap * 32), a.uRight + (size_t)p * a.cap, frameObs, n,
              a.gb[0], a.gb[2], a.gb[4], a.gb[5]);
  const float* Tr = s_Tr;                                                // written before frame_stage's barriers
  auto fetch = [&](int c) {                                              // fields of list entry c
    ProjLP q;
    const bool ok = c < nvalid;
    const int i = ok ? svalid[c] : 0;
    q.octave = ok ? a.lastOctave[lo + i] : 0;
    q.obs = ok ? a.mpObs[lo + i] : 0;
    q.X0 = ok ? a.lastXw[(lo + i) * 3] : 0.f; q.X1 = ok ? a.lastXw[(lo + i) * 3 + 1] : 0.f; q.X2 = ok ? a.lastXw[(lo + i) * 3 + 2] : 0.f;
    q.dw = (ok && lane < 8) ? __ldg(reinterpret_cast<const uint32_t*>(a.mpDesc + (lo + i) * 32) + lane) : 0u;
    return q;
  };
  // phase A
  ORBX_CLK(t2);
  {
    ProjLP nx = fetch(warp);
    for (int c = warp; c < nvalid; c += nwarps) {
      const ProjLP cur = nx;
      nx = fetch(c + nwarps);
      const uint4 l = projf_scan(a, cur, Tr, mode, S, s_gb, lane);
      if (lane == 0) spec[c] = l;
    }
  }
  __syncthreads();
  ORBX_CLK(t3);
  // phase B (see "ordered resolution" above): one key per map point here - the first of its list nobody earlier has claimed
  if (warp == 0) {
    int nm = 0;
    ORBX_STAT(int dbg_serial = 0; int dbg_iters = 0;)
    int epoch = 0;
    const uint4 nokeys = make_uint4(0xffffffffu, kNoKey, kNoKey, kNoKey);
    uint4 nxt = lane < nvalid ? spec[lane] : nokeys;
    for (int base = 0; base < nvalid; base += 32) {
      const int mine = base + lane;
      const int myi = mine < nvalid ? svalid[mine] : 0;
      const uint4 my4 = nxt;
      nxt = mine + 32 < nvalid ? spec[mine + 32] : nokeys;
      const bool complete = my4.x >> 31, obs = my4.y >> 31;              // obs: Observations() > 0
      const uint32_t key[4] = {my4.x & 0x7FFFFFFFu, my4.y & 0x7FFFFFFFu, my4.z, my4.w};
      uint32_t todo = __ballot_sync(0xffffffffu, key[0] != kNoKey);      // lanes whose map point is still to be decided
      while (todo) {
        const bool live = (todo >> lane) & 1u;
        unsigned hitP = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (key[k] != kNoKey && bit_of(S.claim, (int)(key[k] & 0xFFFFu))) hitP |= 1u << k;
        bool acc = false, starved = false;
        int bestIdx = -1;
        const int iters = order_round(key, hitP, S.stag, epoch, lane, [&](unsigned hit) {
          uint32_t k1 = kNoKey;
#pragma unroll
          for (int k = 3; k >= 0; --k)
            if (key[k] != kNoKey && !((hit >> k) & 1u)) k1 = key[k];
          starved = live && !complete && k1 == kNoKey;                   // the list ran out of free keys
          acc = live && k1 != kNoKey && (k1 >> 16) <= 100u;              // TH_HIGH
          bestIdx = acc ? (int)(k1 & 0xFFFFu) : -1;
          return (acc && obs) ? bestIdx : -1;
        });
        ORBX_STAT(dbg_iters += iters;)
        (void)iters;
        // lanes below the first starved one are final (a lane depends on earlier lanes only): they commit together
        const uint32_t sm = __ballot_sync(0xffffffffu, starved);
        const int jstar = sm ? __ffs((int)sm) - 1 : 32;
        const bool commit = acc && lane < jstar;
        const uint32_t cm = __ballot_sync(0xffffffffu, commit);
        if (commit) {
          atomicMax(&S.sasg[bestIdx], myi);                              // a keypoint given away twice keeps the later map point
          if (obs) { atomicOr(&S.claim[bestIdx >> 5], 1u << (bestIdx & 31)); atomicOr(&S.sobs[bestIdx >> 5], 1u << (bestIdx & 31)); }
          sev[nm + __popc(cm & ((1u << lane) - 1u))] = (mine << 16) | bestIdx;   // list position | keypoint: the rotation bin is worked out afterwards
        }
        nm += __popc(cm);
        __syncwarp();
        if (!sm) break;
        // the starved entry is scanned again against the live occupancy (exact), then the lanes after it go round again
        ORBX_STAT(++dbg_serial;)
        const ProjLP c = fetch(base + jstar);
        const uint4 l = projf_scan(a, c, Tr, mode, S, s_gb, lane);
        const uint32_t k1 = l.x & 0x7FFFFFFFu;
        if (k1 != kNoKey && (k1 >> 16) <= 100u) {
          const int bi = (int)(k1 & 0xFFFFu), i = __shfl_sync(0xffffffffu, myi, jstar);
          if (lane == 0) {
            S.sasg[bi] = i;
            if (l.y >> 31) { S.claim[bi >> 5] |= 1u << (bi & 31); S.sobs[bi >> 5] |= 1u << (bi & 31); }
            sev[nm] = ((base + jstar) << 16) | bi;
          }
          ++nm;
        }
        __syncwarp();
        todo &= ~((2u << jstar) - 1u);
      }
    }
    if (lane == 0) s_nm = nm;
    ORBX_STAT(if (lane == 0 && blockIdx.x == 0) printf("K9 phase B: %d re-scans, %d fixpoint iterations, nm %d of %d\n", dbg_serial, dbg_iters, nm, nvalid);)
  }
  __syncthreads();
  ORBX_CLK(t4);
  // frameObs[i] = Observations() of the map point keypoint i ended up with (before the rotation check takes matches away)
  for (int i = tid; i < n; i += nth) {
    const int v = S.sasg[i];
    if (v >= 0) frameObs[i] = a.mpObs[lo + v];
  }
  const int nm0 = s_nm;
  if (a.checkOri) {
    // rotation histogram of the accepted matches (counts do not depend on the order), whole CTA: no angle load in the ordered loop
    for (int e = tid; e < nm0; e += nth) {
      const unsigned ev = (unsigned)sev[e];
      const int idx = (int)(ev & 0xFFFF), bin = rot_bin(a.lastAngle[lo + svalid[ev >> 16]], k[idx].angle);
      atomicAdd(&s_hist[bin], 1);
      sev[e] = (bin << 16) | idx;
    }
    __syncthreads();
    int k_a, k_b, k_c;
    three_maxima(s_hist, k_a, k_b, k_c);
    int drop = 0;
    for (int e = tid; e < nm0; e += nth) {
      const int bin = sev[e] >> 16, idx = sev[e] & 0xFFFF;
      if (bin != k_a && bin != k_b && bin != k_c) { S.sasg[idx] = -1; ++drop; }
    }
    if (drop) atomicSub(&s_nm, drop);
    __syncthreads();
  }
  for (int i = tid; i < n; i += nth) assigned[i] = S.sasg[i];
  if (tid == 0) a.nm[p] = s_nm;
  ORBX_CLK(t5);
  ORBX_CLK_PRINT("k_search_projection_frame_par compact / stage / A / B / rotation + flush", t0, t1, t2, t3, t4, t5);
}

int launch_search_projection_frame(const ProjFrameArgs& a, cudaStream_t st) {
  const size_t base = (size_t)a.cap * 12 + (kGridCells + 2) * sizeof(int);
  const size_t with_desc = ((base + 15) & ~(size_t)15) + (size_t)a.cap * 32;
  const size_t par = frame_stage_bytes(a.cap) + (((size_t)a.lcap * 2 + 15) & ~(size_t)15) + (size_t)a.lcap * 4;
  if (par <= 200 * 1024 && a.lcap <= 65535 && !getenv("ORBX_PROJ_SERIAL")) {    // see launch_search_projection
    cudaFuncSetAttribute(k_search_projection_frame_par, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)par);
    k_search_projection_frame_par<<<a.n_pairs, kParThreads, par, st>>>(a);
  } else if (a.n_pairs <= 148 && with_desc <= 200 * 1024) {
    cudaFuncSetAttribute(k_search_projection_frame<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)with_desc);
    k_search_projection_frame<true><<<a.n_pairs, 32, with_desc, st>>>(a);
  } else {
    if (base > 48 * 1024) cudaFuncSetAttribute(k_search_projection_frame<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)base);
    k_search_projection_frame<false><<<a.n_pairs, 32, base, st>>>(a);
  }
  return 1;
}

// ------------------------------------------------------------------------------------------
// K12: ORBmatcher::SearchByBoW with real vocabulary buckets, both forms:
//   (KeyFrame*, KeyFrame*)  src/ORBmatcher.cc:522-655   strict: bestDist1 <  TH_LOW, validB = pMP2 && !isBad()
//   (KeyFrame*, Frame&)     src/ORBmatcher.cc:159-288           bestDist1 <= TH_LOW, validB = all
// The FeatureVector of a side is given as one vocabulary node id per feature.  One CTA per pair:
//   1. both sides are sorted by (node, feature index) with a block bitonic sort - that is the
//      std::map / push_back order the reference iterates in - and the B descriptors are staged in
//      shared memory (bit j of `used` starts as !validB[j], so invalid and matched features are
//      skipped by the same test);
//   2. warp 0 walks the A side in that order (hazard H6: vbMatched2 / vpMapPointMatches), its lanes
//      scan the B run of the same node; the A descriptors are fetched four queries ahead.
constexpr int kBowThreads = 128;

__global__ void __launch_bounds__(kBowThreads) k_search_by_bow(BowArgs a, int capPad) {
  extern __shared__ __align__(16) uint8_t smem[];
  unsigned long long* keyA = reinterpret_cast<unsigned long long*>(smem);          // [capPad] node<<32 | index, sorted
  unsigned long long* keyB = keyA + capPad;                                         // [capPad]
  uint32_t* Bs = reinterpret_cast<uint32_t*>(keyB + capPad);                        // [cap][8] B descriptors
  uint32_t* used = Bs + (size_t)a.cap * 8;                                          // [(cap+31)/32]
  __shared__ int s_hist[kHisto];
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
  const int nA = a.nA[p], nB = a.nB[p];
  const size_t po = (size_t)p * a.cap;
  const uint32_t* dA = reinterpret_cast<const uint32_t*>(a.dA + po * 32);
  const uint32_t* dB = reinterpret_cast<const uint32_t*>(a.dB + po * 32);
  const float* angA = a.angA + po;
  const float* angB = a.angB + po;
  int* m12 = a.m12 + po;

  for (int i = tid; i < capPad; i += kBowThreads) {
    keyA[i] = i < nA ? (((unsigned long long)(uint32_t)a.nodeA[po + i] << 32) | (unsigned)i) : ~0ull;
    keyB[i] = i < nB ? (((unsigned long long)(uint32_t)a.nodeB[po + i] << 32) | (unsigned)i) : ~0ull;
  }
  for (int i = tid; i < nB * 8; i += kBowThreads) Bs[i] = __ldg(dB + i);
  for (int w = tid; w < (a.cap + 31) / 32; w += kBowThreads) {
    uint32_t bits = 0u;
    for (int b = 0; b < 32; ++b) { const int j = w * 32 + b; if (j < nB && !a.validB[po + j]) bits |= 1u << b; }
    used[w] = bits;
  }
  for (int i = tid; i < nA; i += kBowThreads) m12[i] = -1;
  if (tid < kHisto) s_hist[tid] = 0;
  __syncthreads();
  for (int k = 2; k <= capPad; k <<= 1)                                             // bitonic sort, both arrays together
    for (int jj = k >> 1; jj > 0; jj >>= 1) {
      for (int i = tid; i < capPad; i += kBowThreads) {
        const int l = i ^ jj;
        if (l > i) {
          const bool up = (i & k) == 0;
          unsigned long long x = keyA[i], y = keyA[l];
          if ((x > y) == up) { keyA[i] = y; keyA[l] = x; }
          x = keyB[i]; y = keyB[l];
          if ((x > y) == up) { keyB[i] = y; keyB[l] = x; }
        }
      }
      __syncthreads();
    }
  if (tid >= 32) return;

  int nm = 0, b0 = 0, b1 = 0;
  uint32_t cur_node = 0xffffffffu;
  bool have_node = false;
  // A descriptors four queries ahead: lane l holds word (l & 7) of query t + (l >> 3)
  auto fetch = [&](int t) -> uint32_t {
    const int tt = t + (lane >> 3);
    return tt < nA ? __ldg(dA + (size_t)(uint32_t)(keyA[tt] & 0xffffffffu) * 8 + (lane & 7)) : 0u;
  };
  uint32_t qw = fetch(0);
  for (int t0 = 0; t0 < nA; t0 += 4) {
    const uint32_t qcur = qw;
    qw = fetch(t0 + 4);
    for (int s = 0; s < 4 && t0 + s < nA; ++s) {
      const unsigned long long ka = keyA[t0 + s];
      const uint32_t node = (uint32_t)(ka >> 32);
      const int i = (int)(ka & 0xffffffffu);
      if (!have_node || node != cur_node) {                    // B run of this node: [b0, b1) (the lower_bound steps)
        have_node = true; cur_node = node;
        b0 = b1;
        while (true) {
          const int e = b0 + lane;
          const int c = __popc(__ballot_sync(0xffffffffu, e < nB && (uint32_t)(keyB[e] >> 32) < node));
          b0 += c;
          if (c < 32) break;
        }
        b1 = b0;
        while (true) {
          const int e = b1 + lane;
          const int c = __popc(__ballot_sync(0xffffffffu, e < nB && (uint32_t)(keyB[e] >> 32) == node));
          b1 += c;
          if (c < 32) break;
        }
      }
      if (!a.validA[po + i] || b0 == b1) continue;              // warp-uniform
      uint32_t q[8];
#pragma unroll
      for (int w = 0; w < 8; ++w) q[w] = __shfl_sync(0xffffffffu, qcur, s * 8 + w);
      unsigned k1 = 256u << 16;
      int k2 = 256;
      for (int e = b0 + lane; e < b1; e += 32) {
        const int j = (int)(keyB[e] & 0xffffffffu);
        if ((used[j >> 5] >> (j & 31)) & 1u) continue;
        const int d = ham8(q, Bs + (size_t)j * 8);
        const unsigned kk = ((unsigned)d << 16) | (unsigned)j;
        if (d < (int)(k1 >> 16)) { k2 = (int)(k1 >> 16); k1 = kk; }
        else if (d < k2) k2 = d;
      }
      const unsigned wmin = __reduce_min_sync(0xffffffffu, k1);
      const int sec = (int)__reduce_min_sync(0xffffffffu, (unsigned)(k1 == wmin ? k2 : (int)(k1 >> 16)));
      const int best = (int)(wmin >> 16), bi = (int)(wmin & 0xffffu);
      const bool pass = (a.strict ? best < a.th : best <= a.th) && (float)best < __fmul_rn(a.ratio, (float)sec);
      if (pass) {
        if (lane == 0) { m12[i] = bi; used[bi >> 5] |= 1u << (bi & 31); }
        ++nm;
        __syncwarp();
      }
    }
  }
  __syncwarp();
  if (a.checkOri) {
    for (int i = lane; i < nA; i += 32) {
      const int m = m12[i];
      if (m >= 0) atomicAdd(&s_hist[rot_bin(angA[i], angB[m])], 1);
    }
    __syncwarp();
    int k_a, k_b, k_c;
    three_maxima(s_hist, k_a, k_b, k_c);
    int drop = 0;
    for (int i = lane; i < nA; i += 32) {
      const int m = m12[i];
      if (m >= 0) {
        const int b = rot_bin(angA[i], angB[m]);
        if (b != k_a && b != k_b && b != k_c) { m12[i] = -1; ++drop; }
      }
    }
    nm -= __reduce_add_sync(0xffffffffu, drop);
  }
  if (lane == 0) a.nm[p] = nm;
}

// Bitonic sort of one or two shared-memory arrays of `n` (a power of two, >= 32) 64-bit keys by the whole CTA (blockDim.x a
// multiple of 32, n a multiple of 32).  The compare-exchange steps whose partner distance is below 32 stay inside a warp: the
// thread that owns element i keeps it in a register and trades it by shuffle - 45 of the 66 steps of a 2048-key sort, no
// barrier, no shared-memory traffic; only the 21 wide steps go through shared memory with a barrier each.
template <int NARR>
__device__ __forceinline__ void cta_bitonic_sort(unsigned long long* k0, unsigned long long* k1, int n) {
  const int tid = threadIdx.x, nth = blockDim.x;
  auto narrow = [&](int k, int jtop) {                       // steps jj = jtop, jtop / 2, ..., 1 (jtop < 32) of merge size k, in registers
    for (int base = 0; base < n; base += nth) {
      const int i = base + tid;
      if (base + (tid & ~31) >= n) break;                    // warp-uniform
#pragma unroll
      for (int arr = 0; arr < NARR; ++arr) {
        unsigned long long* key = arr ? k1 : k0;
        unsigned long long x = key[i];
        for (int jj = jtop; jj > 0; jj >>= 1) {
          const unsigned long long y = __shfl_xor_sync(0xffffffffu, x, jj);
          const bool keep_min = ((i & jj) == 0) == ((i & k) == 0);
          x = keep_min ? (x < y ? x : y) : (x < y ? y : x);
        }
        key[i] = x;
      }
    }
  };
  for (int k = 2; k <= 32 && k <= n; k <<= 1) narrow(k, k >> 1);     // own elements only: no barrier needed between these
  __syncthreads();
  for (int k = 64; k <= n; k <<= 1) {
    for (int jj = k >> 1; jj >= 32; jj >>= 1) {
      for (int q = tid; q < (n >> 1); q += nth) {             // one compare-exchange PAIR per thread and trip: nobody idles
        const int i = ((q & ~(jj - 1)) << 1) | (q & (jj - 1)), l = i | jj;
        const bool up = (i & k) == 0;
#pragma unroll
        for (int arr = 0; arr < NARR; ++arr) {
          unsigned long long* key = arr ? k1 : k0;
          const unsigned long long x = key[i], y = key[l];
          if ((x > y) == up) { key[i] = y; key[l] = x; }
        }
      }
      __syncthreads();
    }
    narrow(k, 16);
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// K12, latency form (a few pairs per call: TrackReferenceKeyFrame, Relocalization and DetectLoop ask about one keyframe at
// a time).  One CTA of 1024 threads per pair, everything the loops touch in shared memory:
//   1. both sides sorted by (node, index) by the whole CTA; B descriptors, `used` (= !validB on entry) and validA as bits;
//   2. one THREAD per sorted A entry: the B run of its node by binary search, then the four nearest free B features of that
//      run (keys distance << 16 | j; ascending j on equal distances = the reference's first-wins walk) - most runs are a
//      handful of features, a warp per query would idle; the entries that have a candidate are listed in order;
//   3. the ordered phase (hazard H6, ordered.cuh).  Queries of different nodes never compete for a B feature, so the list
//      is cut at node boundaries into one piece per warp and the 32 warps resolve their pieces at the same time, each in
//      rounds of 32 queries;
//   4. rotation histogram and the write-back by the whole CTA.
constexpr int kBowLatThreads = 1024;

__host__ __device__ inline size_t bow_lat_bytes(int cap, int capPad) {
  return (size_t)capPad * 16 + (size_t)cap * 32 + (size_t)((cap + 31) / 32) * 8 + (size_t)capPad * 4 + (size_t)cap * 8 + (size_t)capPad * 2 + 16;
}

__global__ void __launch_bounds__(kBowLatThreads) k_search_by_bow_lat(BowArgs a, int capPad) {
  extern __shared__ __align__(16) uint8_t smem[];
  unsigned long long* keyA = reinterpret_cast<unsigned long long*>(smem);          // [capPad] node<<32 | index, sorted
  unsigned long long* keyB = keyA + capPad;                                         // [capPad]
  uint32_t* Bs = reinterpret_cast<uint32_t*>(keyB + capPad);                        // [cap][8] B descriptors
  uint32_t* used = Bs + (size_t)a.cap * 8;                                          // [(cap+31)/32] B feature invalid or matched
  uint32_t* vA = used + (a.cap + 31) / 32;                                          // [(cap+31)/32] validA
  uint32_t* srun = vA + (a.cap + 31) / 32;                                          // [capPad] by sorted A position: B run rb0 | rb1 << 16
  int* sasg = reinterpret_cast<int*>(srun + capPad);                                // [cap] m12 until the end
  int* stag = sasg + a.cap;                                                         // [cap] order_round's tags, by B feature
  uint16_t* slist = reinterpret_cast<uint16_t*>(stag + a.cap);                      // [capPad] sorted A positions that have a candidate
  __shared__ int s_hist[kHisto];
  __shared__ int s_bcnt[kBowLatThreads / 32];
  __shared__ int s_nm;
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nth = blockDim.x, nwarps = nth >> 5;
  const int nA = a.nA[p], nB = a.nB[p];
  const size_t po = (size_t)p * a.cap;
  const uint32_t* dA = reinterpret_cast<const uint32_t*>(a.dA + po * 32);
  const uint32_t* dB = reinterpret_cast<const uint32_t*>(a.dB + po * 32);
  uint4* spec = a.spec + po;

  ORBX_CLK(t0);
  for (int i = tid; i < capPad; i += nth) {
    keyA[i] = i < nA ? (((unsigned long long)(uint32_t)a.nodeA[po + i] << 32) | (unsigned)i) : ~0ull;
    keyB[i] = i < nB ? (((unsigned long long)(uint32_t)a.nodeB[po + i] << 32) | (unsigned)i) : ~0ull;
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(dB);
    uint4* dst = reinterpret_cast<uint4*>(Bs);
    for (int i = tid; i < nB * 2; i += nth) dst[i] = __ldg(src + i);
  }
  for (int base = warp * 32; base < a.cap; base += nth) {
    const int j = base + lane;
    const uint32_t ub = __ballot_sync(0xffffffffu, j < nB && !a.validB[po + j]);
    const uint32_t vb = __ballot_sync(0xffffffffu, j < nA && a.validA[po + j]);
    if (lane == 0) { used[base >> 5] = ub; vA[base >> 5] = vb; }
  }
  for (int i = tid; i < a.cap; i += nth) { sasg[i] = -1; stag[i] = 0; }
  if (tid < kHisto) s_hist[tid] = 0;
  if (tid == 0) s_nm = 0;
  __syncthreads();
  ORBX_CLK(t1);
  cta_bitonic_sort<2>(keyA, keyB, capPad);
  // 2a. a thread per sorted A entry: the B run of its node; the entries that are valid and have a run are listed in order
  ORBX_CLK(t2);
  int nlist = 0;
  for (int base = 0; base < nA; base += nth) {
    const int t = base + tid;
    bool ok = false;
    if (t < nA) {
      const unsigned long long ka = keyA[t];
      const uint32_t node = (uint32_t)(ka >> 32);
      int lo = 0, hi = nB;                                                          // first B entry with node >= this node
      while (lo < hi) { const int mid = (lo + hi) >> 1; if ((uint32_t)(keyB[mid] >> 32) < node) lo = mid + 1; else hi = mid; }
      const int rb0 = lo;
      hi = nB;                                                                      // first B entry with node > this node
      while (lo < hi) { const int mid = (lo + hi) >> 1; if ((uint32_t)(keyB[mid] >> 32) <= node) lo = mid + 1; else hi = mid; }
      srun[t] = (uint32_t)rb0 | ((uint32_t)lo << 16);
      ok = bit_of(vA, (int)(ka & 0xffffffffu)) && rb0 < lo;
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) s_bcnt[warp] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < nwarps; ++w) { const int c = s_bcnt[w]; before += w < warp ? c : 0; total += c; }
    if (ok) slist[nlist + before + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)t;
    nlist += total;
    __syncthreads();
  }
  // 2b. FOUR threads per LISTED entry, each walking every fourth feature of the run (a walk is a chain of dependent
  // shared-memory reads, ~400 cycles per feature: splitting it four ways is worth more than the merge costs), then the four
  // sorted lists of a quad are merged by shuffles.  Neighbouring entries mostly share a node, so the lanes of a warp read
  // the same few B descriptors.
  for (int c0 = 0; c0 < nlist; c0 += nth / 4) {
    const int c = c0 + (tid >> 2), sub = tid & 3;
    uint32_t l0 = kNoKey, l1 = kNoKey, l2 = kNoKey, l3 = kNoKey;
    int cnt = 0;
    if (c < nlist) {
      const int t = slist[c];
      const int i = (int)(keyA[t] & 0xffffffffu);
      const uint32_t rr = srun[t];
      const uint4 q0 = __ldg(reinterpret_cast<const uint4*>(dA + (size_t)i * 8)), q1 = __ldg(reinterpret_cast<const uint4*>(dA + (size_t)i * 8) + 1);
      const uint32_t q[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
      for (int e = (int)(rr & 0xFFFFu) + sub; e < (int)(rr >> 16); e += 4) {
        const int j = (int)(keyB[e] & 0xffffffffu);
        if (bit_of(used, j)) continue;
        const uint4 b0 = *reinterpret_cast<const uint4*>(Bs + (size_t)j * 8), b1 = *reinterpret_cast<const uint4*>(Bs + (size_t)j * 8 + 4);
        const int d = ham256_csa(q, b0, b1);                                        // 4 POPC per distance instead of 8: POPC is the slow pipe
        uint32_t kk = ((uint32_t)d << 16) | (uint32_t)j;                            // sorted insertion into (l0 <= l1 <= l2 <= l3)
        ++cnt;
        if (kk < l0) { const uint32_t sw = l0; l0 = kk; kk = sw; }
        if (kk < l1) { const uint32_t sw = l1; l1 = kk; kk = sw; }
        if (kk < l2) { const uint32_t sw = l2; l2 = kk; kk = sw; }
        if (kk < l3) l3 = kk;
      }
    }
    // merge within the quad (lanes 4q .. 4q+3): after two exchange steps every lane holds the four smallest keys of the entry
#pragma unroll
    for (int step = 1; step <= 2; step <<= 1) {
      const uint32_t p0 = __shfl_xor_sync(0xffffffffu, l0, step), p1 = __shfl_xor_sync(0xffffffffu, l1, step);
      const uint32_t p2 = __shfl_xor_sync(0xffffffffu, l2, step), p3 = __shfl_xor_sync(0xffffffffu, l3, step);
      cnt += __shfl_xor_sync(0xffffffffu, cnt, step);
      const uint32_t in[4] = {p0, p1, p2, p3};
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint32_t kk = in[u];
        if (kk < l0) { const uint32_t sw = l0; l0 = kk; kk = sw; }
        if (kk < l1) { const uint32_t sw = l1; l1 = kk; kk = sw; }
        if (kk < l2) { const uint32_t sw = l2; l2 = kk; kk = sw; }
        if (kk < l3) l3 = kk;
      }
    }
    if (c < nlist && sub == 0) spec[c] = make_uint4(l0 | (cnt <= 4 ? 0x80000000u : 0u), l1, l2, l3);   // no candidate: every key kNoKey, nothing to decide
  }
  __syncthreads();
  // 3. ordered phase, a piece of the list per warp, cut at node boundaries
  ORBX_CLK(t3);
  {
    auto node_at = [&](int c) { return (uint32_t)(keyA[slist[c]] >> 32); };
    auto cut = [&](int w) {                                                          // first list position of warp w's piece (warp-uniform)
      int c = (int)(((long long)nlist * w) / nwarps);
      if (w <= 0) return 0;
      if (w >= nwarps) return nlist;
      while (c < nlist) {                                                           // forwards to the next node change
        const int cc = c + lane;
        const uint32_t same = __ballot_sync(0xffffffffu, cc < nlist && cc > 0 && node_at(cc) == node_at(cc - 1));
        const int run = __ffs((int)~same) - 1;                                      // lanes before the first change
        if (run < 0) { c += 32; continue; }
        c += run;
        break;
      }
      return min(c, nlist);
    };
    const int c_lo = cut(warp), c_hi = cut(warp + 1);
    auto pass_of = [&](uint32_t k1, uint32_t k2) {
      if (k1 == kNoKey) return false;
      const int best = (int)(k1 >> 16), sec = k2 == kNoKey ? 256 : (int)(k2 >> 16);
      return (a.strict ? best < a.th : best <= a.th) && (float)best < __fmul_rn(a.ratio, (float)sec);
    };
    int nm = 0, epoch = 0;
    for (int base = c_lo; base < c_hi; base += 32) {
      const int c = base + lane;
      const bool in = c < c_hi;
      const int t = in ? slist[c] : 0;
      const uint4 my4 = in ? spec[c] : make_uint4(0xffffffffu, kNoKey, kNoKey, kNoKey);
      const int iA = (int)(keyA[t] & 0xffffffffu);
      const bool complete = my4.x >> 31;
      const uint32_t key[4] = {my4.x & 0x7FFFFFFFu, my4.y, my4.z, my4.w};
      uint32_t todo = __ballot_sync(0xffffffffu, key[0] != kNoKey);
      while (todo) {
        const bool live = (todo >> lane) & 1u;
        unsigned hitP = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (key[k] != kNoKey && bit_of(used, (int)(key[k] & 0xFFFFu))) hitP |= 1u << k;
        bool acc = false, starved = false;
        int bi = -1;
        order_round(key, hitP, stag, epoch, lane, [&](unsigned hit) {
          uint32_t k1 = kNoKey, k2 = kNoKey;                             // the two nearest features nobody earlier has matched
#pragma unroll
          for (int k = 3; k >= 0; --k)
            if (key[k] != kNoKey && !((hit >> k) & 1u)) { k2 = k1; k1 = key[k]; }
          starved = live && !complete && k2 == kNoKey;
          acc = live && pass_of(k1, k2);
          bi = acc ? (int)(k1 & 0xFFFFu) : -1;
          return bi;
        });
        const uint32_t sm = __ballot_sync(0xffffffffu, starved);
        const int jstar = sm ? __ffs((int)sm) - 1 : 32;
        const bool commit = acc && lane < jstar;
        if (commit) { sasg[iA] = bi; atomicOr(&used[bi >> 5], 1u << (bi & 31)); }
        nm += __popc(__ballot_sync(0xffffffffu, commit));
        __syncwarp();
        if (!sm) break;
        // the starved query: its whole run again, against the live `used` (warp-wide, exact)
        const int ts = __shfl_sync(0xffffffffu, t, jstar), is = __shfl_sync(0xffffffffu, iA, jstar);
        const uint32_t rr = srun[ts];
        const uint32_t qw = lane < 8 ? __ldg(dA + (size_t)is * 8 + lane) : 0u;
        uint32_t q[8];
#pragma unroll
        for (int w = 0; w < 8; ++w) q[w] = __shfl_sync(0xffffffffu, qw, w);
        uint32_t k1 = 0xffffffffu, k2 = 0xffffffffu;
        for (int e = (int)(rr & 0xFFFFu) + lane; e < (int)(rr >> 16); e += 32) {
          const int j = (int)(keyB[e] & 0xffffffffu);
          if (bit_of(used, j)) continue;
          const uint32_t kk = ((uint32_t)ham8(q, Bs + (size_t)j * 8) << 16) | (uint32_t)j;
          if (kk < k1) { k2 = k1; k1 = kk; }
          else if (kk < k2) k2 = kk;
        }
        const uint32_t w1 = __reduce_min_sync(0xffffffffu, k1);
        const uint32_t w2 = __reduce_min_sync(0xffffffffu, k1 == w1 ? k2 : k1);
        if (pass_of(w1 == 0xffffffffu ? kNoKey : w1, w2 == 0xffffffffu ? kNoKey : w2)) {
          const int bj = (int)(w1 & 0xFFFFu);
          if (lane == 0) { sasg[is] = bj; atomicOr(&used[bj >> 5], 1u << (bj & 31)); }   // other warps own other bits of the same word
          ++nm;
        }
        __syncwarp();
        todo &= ~((2u << jstar) - 1u);
      }
    }
    if (lane == 0 && nm) atomicAdd(&s_nm, nm);
  }
  __syncthreads();
  ORBX_CLK(t4);
  // 4. rotation consistency (src/ORBmatcher.cc:270-285, :636-652) and the write-back, whole CTA
  const float* angA = a.angA + po;
  const float* angB = a.angB + po;
  if (a.checkOri) {
    int mybin[(4096 + kBowLatThreads - 1) / kBowLatThreads];
    int r = 0;
    for (int base = 0; base < nA; base += nth, ++r) {
      const int i = base + tid;
      const int m = i < nA ? sasg[i] : -1;
      mybin[r] = -1;
      const unsigned act = __ballot_sync(0xffffffffu, m >= 0);
      if (m >= 0) {                                                    // one atomic per distinct bin of the warp
        const int b = rot_bin(angA[i], angB[m]);
        mybin[r] = b;
        const unsigned peers = __match_any_sync(act, b);
        if (lane == __ffs((int)peers) - 1) atomicAdd(&s_hist[b], __popc(peers));
      }
    }
    __syncthreads();
    int k_a, k_b, k_c;
    three_maxima(s_hist, k_a, k_b, k_c);
    int drop = 0;
    r = 0;
    for (int i = tid; i < nA; i += nth, ++r) {
      const int b = mybin[r];
      if (b >= 0 && b != k_a && b != k_b && b != k_c) { sasg[i] = -1; ++drop; }
    }
    if (drop) atomicSub(&s_nm, drop);
    __syncthreads();
  }
  int* m12 = a.m12 + po;
  for (int i = tid; i < nA; i += nth) m12[i] = sasg[i];
  if (tid == 0) a.nm[p] = s_nm;
  ORBX_CLK(t5);
  ORBX_CLK_PRINT("k_search_by_bow_lat stage / sort / lists / ordered / rotation + flush", t0, t1, t2, t3, t4, t5);
  ORBX_STAT(if (tid == 0 && blockIdx.x == 0) printf("  K12: nA %d nB %d listed %d\n", nA, nB, nlist);)
}

int launch_search_by_bow(const BowArgs& a, cudaStream_t st) {
  int capPad = 32;
  while (capPad < a.cap) capPad <<= 1;
  const size_t lat = bow_lat_bytes(a.cap, capPad);
  // the 1024-thread form whenever its shared memory fits: it finishes a pair in a fifteenth of the small-CTA form's time, which
  // is worth more than the two small CTAs an SM could hold instead (the small form stays for cap > ~3300 and for A/B tests)
  if (lat <= 200 * 1024 && !getenv("ORBX_BOW_SERIAL")) {
    cudaFuncSetAttribute(k_search_by_bow_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lat);
    k_search_by_bow_lat<<<a.n_pairs, kBowLatThreads, lat, st>>>(a, capPad);
    return 1;
  }
  const size_t smem = (size_t)capPad * 16 + (size_t)a.cap * 32 + (size_t)((a.cap + 31) / 32) * 4;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_search_by_bow, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_search_by_bow<<<a.n_pairs, kBowThreads, smem, st>>>(a, capPad);
  return 1;
}

// ------------------------------------------------------------------------------------------
// K14: ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-823) with CheckDistEpipolarLine (:140-157),
// batched over keyframe pairs.  The reference never sets vbMatched2, so the features of
// KF1 are independent: one thread per KF1 feature.  One CTA per pair sorts the KF2 side by (node, index) and
// stages its descriptors in shared memory; a thread finds the run of its node by binary search and walks it in
// index order with the reference's acceptance rule (dist <= TH_LOW, dist <= best so far, epipole and epipolar
// checks), i.e. the LAST candidate of least distance that passes the checks.
__global__ void __launch_bounds__(kBowThreads) k_search_triangulation(TriArgs a, int capPad) {
  extern __shared__ __align__(16) uint8_t smem[];
  unsigned long long* keyB = reinterpret_cast<unsigned long long*>(smem);          // [capPad] node<<32 | index, sorted
  uint32_t* Bs = reinterpret_cast<uint32_t*>(keyB + capPad);                        // [cap][8]
  __shared__ int s_hist[kHisto];
  __shared__ int s_nm;
  __shared__ float s_epi[2];
  const int p = blockIdx.x, tid = threadIdx.x;
  const int n1 = a.n1[p], n2 = a.n2[p];
  const size_t po = (size_t)p * a.cap;
  const orbx_keypoint* k1 = a.k1 + po; const orbx_keypoint* k2 = a.k2 + po;
  const uint32_t* d1 = reinterpret_cast<const uint32_t*>(a.d1 + po * 32);
  const uint32_t* d2 = reinterpret_cast<const uint32_t*>(a.d2 + po * 32);
  int* m12 = a.m12 + po;
  for (int i = tid; i < capPad; i += kBowThreads)
    keyB[i] = i < n2 ? (((unsigned long long)(uint32_t)a.node2[po + i] << 32) | (unsigned)i) : ~0ull;
  for (int i = tid; i < n2 * 8; i += kBowThreads) Bs[i] = __ldg(d2 + i);
  if (tid < kHisto) s_hist[tid] = 0;
  if (tid == 0) {
    s_nm = 0;
    const float* T = a.pose2 + (size_t)p * 12; const float* C = a.Cw + (size_t)p * 3; const float* K = a.cam2 + (size_t)p * 4;
    float C2[3];
    for (int r = 0; r < 3; ++r)                                                     // R2w*Cw + t2w, cv::gemm arithmetic (src/ORBmatcher.cc:667)
      C2[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[3 * r], C[0]), __fmul_rn(T[3 * r + 1], C[1])), __fmul_rn(T[3 * r + 2], C[2])), T[9 + r]);
    const float invz = __fdiv_rn(1.0f, C2[2]);
    s_epi[0] = __fadd_rn(__fmul_rn(__fmul_rn(K[0], C2[0]), invz), K[2]);
    s_epi[1] = __fadd_rn(__fmul_rn(__fmul_rn(K[1], C2[1]), invz), K[3]);
  }
  __syncthreads();
  for (int k = 2; k <= capPad; k <<= 1)
    for (int jj = k >> 1; jj > 0; jj >>= 1) {
      for (int i = tid; i < capPad; i += kBowThreads) {
        const int l = i ^ jj;
        if (l > i) {
          const bool up = (i & k) == 0;
          const unsigned long long x = keyB[i], y = keyB[l];
          if ((x > y) == up) { keyB[i] = y; keyB[l] = x; }
        }
      }
      __syncthreads();
    }
  const float* F = a.F12 + (size_t)p * 9;
  const float ex = s_epi[0], ey = s_epi[1];
  int nm = 0;
  for (int i = tid; i < n1; i += kBowThreads) {
    int best2 = -1;
    const bool st1 = a.uR1[po + i] >= 0.f;
    if (!a.hasMp1[po + i] && !(a.onlyStereo && !st1)) {
      const uint32_t node = (uint32_t)a.node1[po + i];
      int lo = 0, hi = n2;                                                          // first entry with node2 >= node
      while (lo < hi) { const int mid = (lo + hi) >> 1; if ((uint32_t)(keyB[mid] >> 32) < node) lo = mid + 1; else hi = mid; }
      uint32_t q[8];
#pragma unroll
      for (int w = 0; w < 8; ++w) q[w] = __ldg(d1 + (size_t)i * 8 + w);
      const orbx_keypoint kp1 = k1[i];
      const float la = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, F[0]), __fmul_rn(kp1.y, F[3])), F[6]);
      const float lb = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, F[1]), __fmul_rn(kp1.y, F[4])), F[7]);
      const float lc = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, F[2]), __fmul_rn(kp1.y, F[5])), F[8]);
      const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
      int bestDist = 50;                                                            // TH_LOW
      for (int e = lo; e < n2 && (uint32_t)(keyB[e] >> 32) == node; ++e) {
        const int j = (int)(keyB[e] & 0xffffffffu);
        const int dist = ham8(q, Bs + (size_t)j * 8);
        if (dist > 50 || dist > bestDist) continue;
        if (a.hasMp2[po + j]) continue;
        const bool st2 = a.uR2[po + j] >= 0.f;
        if (a.onlyStereo && !st2) continue;
        const orbx_keypoint kp2 = k2[j];
        if (!st1 && !st2) {
          const float dx = __fsub_rn(ex, kp2.x), dy = __fsub_rn(ey, kp2.y);
          if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, a.scale2[kp2.octave])) continue;
        }
        if (den == 0.f) continue;
        const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, kp2.x), __fmul_rn(lb, kp2.y)), lc);
        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
        if ((double)dsqr < __dmul_rn(3.84, (double)a.sigma2_2[kp2.octave])) { best2 = j; bestDist = dist; }
      }
    }
    m12[i] = best2;
    if (best2 >= 0) {
      ++nm;
      if (a.checkOri) atomicAdd(&s_hist[rot_bin(k1[i].angle, k2[best2].angle)], 1);
    }
  }
  if (nm) atomicAdd(&s_nm, nm);
  __syncthreads();
  if (a.checkOri) {
    int k_a, k_b, k_c;
    three_maxima(s_hist, k_a, k_b, k_c);
    int drop = 0;
    for (int i = tid; i < n1; i += kBowThreads) {
      const int m = m12[i];                                                         // written by this same thread above
      if (m >= 0) {
        const int b = rot_bin(k1[i].angle, k2[m].angle);
        if (b != k_a && b != k_b && b != k_c) { m12[i] = -1; ++drop; }
      }
    }
    if (drop) atomicSub(&s_nm, drop);
    __syncthreads();
  }
  if (tid == 0) a.nm[p] = s_nm;
}

// K14, latency form (one keyframe pair per call is how LocalMapping::CreateNewMapPoints asks): the same walk, 1024 threads
// instead of 128, the KF2 side sorted by the whole CTA, and everything the per-candidate tests read - position, octave,
// "has a map point", "is stereo" - staged in shared memory beside the descriptors, so that a candidate costs no global
// round trip (the 128-thread form spends its time on three dependent ones per candidate that survives the distance test).
constexpr int kTriLatThreads = 1024;

__host__ __device__ inline size_t tri_lat_bytes(int cap, int capPad) {
  return (size_t)capPad * 16 + (size_t)cap * 32 + (size_t)cap * 8 + (size_t)((cap + 31) / 32) * 8 + (size_t)cap + 16;
}

__global__ void __launch_bounds__(kTriLatThreads) k_search_triangulation_lat(TriArgs a, int capPad, int G) {
  extern __shared__ __align__(16) uint8_t smem[];
  unsigned long long* keyB = reinterpret_cast<unsigned long long*>(smem);          // [capPad] KF2 node<<32 | index, sorted
  unsigned long long* keyA = keyB + capPad;                                         // [capPad] KF1 the same: neighbouring threads then walk the same run (broadcast reads)
  uint32_t* Bs = reinterpret_cast<uint32_t*>(keyA + capPad);                        // [cap][8]
  float2* sxy = reinterpret_cast<float2*>(Bs + (size_t)a.cap * 8);                  // [cap] KF2 keypoint positions
  uint32_t* hasb = reinterpret_cast<uint32_t*>(sxy + a.cap);                        // [(cap+31)/32] KF2 feature has a map point
  uint32_t* stb = hasb + (a.cap + 31) / 32;                                         // [(cap+31)/32] KF2 feature is stereo
  uint8_t* soct = reinterpret_cast<uint8_t*>(stb + (a.cap + 31) / 32);              // [cap] KF2 octave
  __shared__ int s_hist[kHisto];
  __shared__ int s_nm;
  __shared__ float s_epi[2], s_F[9];
  const int p = blockIdx.x / G, g = blockIdx.x - p * G, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nth = blockDim.x;
  const int n1 = a.n1[p], n2 = a.n2[p];
  const size_t po = (size_t)p * a.cap;
  const orbx_keypoint* k1 = a.k1 + po; const orbx_keypoint* k2 = a.k2 + po;
  const uint32_t* d1 = reinterpret_cast<const uint32_t*>(a.d1 + po * 32);
  const uint32_t* d2 = reinterpret_cast<const uint32_t*>(a.d2 + po * 32);
  int* m12 = a.m12 + po;
  ORBX_CLK(t0);
  for (int i = tid; i < capPad; i += nth) {
    keyB[i] = i < n2 ? (((unsigned long long)(uint32_t)a.node2[po + i] << 32) | (unsigned)i) : ~0ull;
    keyA[i] = i < n1 ? (((unsigned long long)(uint32_t)a.node1[po + i] << 32) | (unsigned)i) : ~0ull;
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(d2);
    uint4* dst = reinterpret_cast<uint4*>(Bs);
    for (int i = tid; i < n2 * 2; i += nth) dst[i] = __ldg(src + i);
  }
  for (int j = tid; j < n2; j += nth) { const orbx_keypoint kp = k2[j]; sxy[j] = make_float2(kp.x, kp.y); soct[j] = (uint8_t)kp.octave; }
  for (int base = warp * 32; base < a.cap; base += nth) {
    const int j = base + lane;
    const uint32_t hb = __ballot_sync(0xffffffffu, j < n2 && a.hasMp2[po + j]);
    const uint32_t sb = __ballot_sync(0xffffffffu, j < n2 && a.uR2[po + j] >= 0.f);
    if (lane == 0) { hasb[base >> 5] = hb; stb[base >> 5] = sb; }
  }
  if (tid < kHisto) s_hist[tid] = 0;
  if (tid >= 32 && tid < 41) s_F[tid - 32] = a.F12[(size_t)p * 9 + tid - 32];
  if (tid == 0) {
    s_nm = 0;
    const float* T = a.pose2 + (size_t)p * 12; const float* C = a.Cw + (size_t)p * 3; const float* K = a.cam2 + (size_t)p * 4;
    float C2[3];
    for (int r = 0; r < 3; ++r)                                                     // R2w*Cw + t2w, cv::gemm arithmetic (src/ORBmatcher.cc:667)
      C2[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[3 * r], C[0]), __fmul_rn(T[3 * r + 1], C[1])), __fmul_rn(T[3 * r + 2], C[2])), T[9 + r]);
    const float invz = __fdiv_rn(1.0f, C2[2]);
    s_epi[0] = __fadd_rn(__fmul_rn(__fmul_rn(K[0], C2[0]), invz), K[2]);
    s_epi[1] = __fadd_rn(__fmul_rn(__fmul_rn(K[1], C2[1]), invz), K[3]);
  }
  __syncthreads();
  ORBX_CLK(t1);
  cta_bitonic_sort<2>(keyB, keyA, capPad);
  ORBX_CLK(t2);
  const float ex = s_epi[0], ey = s_epi[1];
  int nm = 0;
  ORBX_STAT(int dbg_cand = 0;)
  // this CTA's share of the sorted KF1 features: a contiguous piece (neighbouring threads keep walking the same runs)
  const int piece = (((n1 + G - 1) / G) + 31) & ~31, t_lo = g * piece, t_hi = min(n1, t_lo + piece);
  // FOUR threads per KF1 feature, each walking every fourth KF2 feature of the run (the walk is a chain of dependent
  // shared-memory reads); the winner - least distance, the LAST such candidate in index order - is order-independent, so the
  // four partial winners are merged by the key  distance << 16 | (65535 - j).
  for (int t0 = t_lo; t0 < t_hi; t0 += nth / 4) {
    const int t = t0 + (tid >> 2), sub = tid & 3;
    uint32_t win = 0xffffffffu;
    int i = 0;
    float ang1 = 0.f;
    if (t < t_hi) {
      const unsigned long long ka = keyA[t];
      i = (int)(ka & 0xffffffffu);
      const uint32_t node = (uint32_t)(ka >> 32);
      // every field of the KF1 feature is requested before the first test
      const bool st1 = a.uR1[po + i] >= 0.f;
      const int has1 = a.hasMp1[po + i];
      const orbx_keypoint kp1 = k1[i];
      ang1 = kp1.angle;
      const uint4 q0 = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i * 8)), q1 = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i * 8) + 1);
      const uint32_t qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
      if (!has1 && !(a.onlyStereo && !st1)) {
        int lo = 0, hi = n2;                                                        // first entry with node2 >= node
        while (lo < hi) { const int mid = (lo + hi) >> 1; if ((uint32_t)(keyB[mid] >> 32) < node) lo = mid + 1; else hi = mid; }
        const float la = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, s_F[0]), __fmul_rn(kp1.y, s_F[3])), s_F[6]);
        const float lb = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, s_F[1]), __fmul_rn(kp1.y, s_F[4])), s_F[7]);
        const float lc = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, s_F[2]), __fmul_rn(kp1.y, s_F[5])), s_F[8]);
        const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
        int bestDist = 50;                                                          // TH_LOW
        for (int e = lo + sub; e < n2 && (uint32_t)(keyB[e] >> 32) == node; e += 4) {
          ORBX_STAT(++dbg_cand;)
          const int j = (int)(keyB[e] & 0xffffffffu);
          const uint4 b0 = *reinterpret_cast<const uint4*>(Bs + (size_t)j * 8), b1 = *reinterpret_cast<const uint4*>(Bs + (size_t)j * 8 + 4);
          const int dist = ham256_csa(qv, b0, b1);                                  // 4 POPC per distance instead of 8: POPC is the slow pipe
          if (dist > 50 || dist > bestDist) continue;
          if (bit_of(hasb, j)) continue;
          const bool st2 = bit_of(stb, j);
          if (a.onlyStereo && !st2) continue;
          const float2 p2 = sxy[j];
          const int o2 = soct[j];
          if (!st1 && !st2) {
            const float dx = __fsub_rn(ex, p2.x), dy = __fsub_rn(ey, p2.y);
            if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, a.scale2[o2])) continue;
          }
          if (den == 0.f) continue;
          const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, p2.x), __fmul_rn(lb, p2.y)), lc);
          const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
          if ((double)dsqr < __dmul_rn(3.84, (double)a.sigma2_2[o2])) { win = ((uint32_t)dist << 16) | (uint32_t)(65535 - j); bestDist = dist; }
        }
      }
    }
    win = min(win, __shfl_xor_sync(0xffffffffu, win, 1));
    win = min(win, __shfl_xor_sync(0xffffffffu, win, 2));
    if (t < t_hi && sub == 0) {
      const int best2 = win == 0xffffffffu ? -1 : 65535 - (int)(win & 0xFFFFu);
      m12[i] = best2;
      if (best2 >= 0) {
        ++nm;
        if (a.checkOri) { const int bin = rot_bin(ang1, k2[best2].angle); a.bins[po + i] = (uint8_t)bin; atomicAdd(&s_hist[bin], 1); }
      }
    }
  }
  ORBX_CLK(t3w);
  if (nm) atomicAdd(&s_nm, nm);
  __syncthreads();
  ORBX_CLK(t3);
  // this CTA's counts join the pair's; the rotation check is k_triangulation_finish's once every CTA of the pair is done
  if (a.checkOri && tid < kHisto && s_hist[tid]) atomicAdd(&a.hist[(size_t)p * 32 + tid], s_hist[tid]);
  if (tid == 0 && s_nm) atomicAdd(&a.nm[p], s_nm);
  ORBX_CLK(t4);
  ORBX_CLK_PRINT("k_search_triangulation_lat stage / sort / walk (thread 0) / walk (all) / totals", t0, t1, t2, t3w, t3, t4);
}

// rotation consistency of K14's 1024-thread form (src/ORBmatcher.cc:792-808): one CTA per pair, after all walk CTAs
__global__ void __launch_bounds__(256) k_triangulation_finish(TriArgs a) {
  __shared__ int s_hist[kHisto];
  __shared__ int s_drop;
  const int p = blockIdx.x, tid = threadIdx.x;
  const size_t po = (size_t)p * a.cap;
  if (tid < kHisto) s_hist[tid] = a.hist[(size_t)p * 32 + tid];
  if (tid == 0) s_drop = 0;
  __syncthreads();
  int k_a, k_b, k_c;
  three_maxima(s_hist, k_a, k_b, k_c);
  int* m12 = a.m12 + po;
  int drop = 0;
  for (int i = tid; i < a.n1[p]; i += 256)
    if (m12[i] >= 0) { const int bn = a.bins[po + i]; if (bn != k_a && bn != k_b && bn != k_c) { m12[i] = -1; ++drop; } }
  if (drop) atomicAdd(&s_drop, drop);
  __syncthreads();
  if (tid == 0 && s_drop) a.nm[p] -= s_drop;
}

int launch_search_triangulation(const TriArgs& a, cudaStream_t st) {
  int capPad = 32;
  while (capPad < a.cap) capPad <<= 1;
  const size_t lat = tri_lat_bytes(a.cap, capPad);
  if (lat <= 200 * 1024 && !getenv("ORBX_BOW_SERIAL")) {
    // KF1 features are independent, so a pair is spread over up to 8 CTAs while the GPU has SMs to spare (each sorts the KF2
    // side for itself: ~20 us, against a walk that is ~90 us on one SM when the runs are long)
    const int G = std::max(1, std::min(8, 148 / std::max(a.n_pairs, 1)));
    cudaFuncSetAttribute(k_search_triangulation_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lat);
    k_search_triangulation_lat<<<a.n_pairs * G, kTriLatThreads, lat, st>>>(a, capPad, G);
    if (a.checkOri) { k_triangulation_finish<<<a.n_pairs, 256, 0, st>>>(a); return 2; }
    return 1;
  }
  const size_t smem = (size_t)capPad * 8 + (size_t)a.cap * 32;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_search_triangulation, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_search_triangulation<<<a.n_pairs, kBowThreads, smem, st>>>(a, capPad);
  return 1;
}

// ------------------------------------------------------------------------------------------
// K13: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-305), batched over map points
// (SURVEY.md 8f rank 3).  One warp per map point: its N observed descriptors sit in shared memory; for
// every row i the lanes take the columns j, the 257-bin histogram of the row gives the element
// [(N-1)/2] of the sorted row without sorting; the least median wins, first index on ties.
constexpr int kDistWarps = 4;

__global__ void __launch_bounds__(kDistWarps * 32) k_distinctive(DistinctArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mp = blockIdx.x * kDistWarps + warp;
  if (mp >= a.n_points) return;
  uint32_t* D = reinterpret_cast<uint32_t*>(smem) + (size_t)warp * (a.kcap * 8 + 288);   // [kcap][8]
  int* hist = reinterpret_cast<int*>(D + (size_t)a.kcap * 8);                             // [288] (257 used)
  const int N = min(a.counts[mp], a.kcap);
  const uint32_t* src = reinterpret_cast<const uint32_t*>(a.desc + (size_t)mp * a.kcap * 32);
  for (int i = lane; i < N * 8; i += 32) D[i] = __ldg(src + i);
  __syncwarp();
  int bestMedian = INT_MAX, bestIdx = N > 0 ? 0 : -1;
  const int rank = (N - 1) / 2;                                  // (int)(0.5 * (N - 1))
  for (int i = 0; i < N; ++i) {
    for (int b = lane; b < 288; b += 32) hist[b] = 0;
    __syncwarp();
    uint32_t q[8];
#pragma unroll
    for (int w = 0; w < 8; ++w) q[w] = D[i * 8 + w];
    for (int j = lane; j < N; j += 32) atomicAdd(&hist[ham8(q, D + (size_t)j * 8)], 1);   // d(i,i) = 0 as in the reference
    __syncwarp();
    int run = 0, median = 256;
    for (int c = 0; c < 9; ++c) {
      int v = hist[c * 32 + lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
      }
      const uint32_t hit = __ballot_sync(0xffffffffu, run + v >= rank + 1);
      if (hit) { median = c * 32 + __ffs((int)hit) - 1; break; }
      run += __shfl_sync(0xffffffffu, v, 31);
    }
    if (median < bestMedian) { bestMedian = median; bestIdx = i; }
    __syncwarp();
  }
  if (lane == 0) a.best_idx[mp] = bestIdx;
  if (a.best_desc && bestIdx >= 0 && lane < 8) reinterpret_cast<uint32_t*>(a.best_desc + (size_t)mp * 32)[lane] = D[bestIdx * 8 + lane];
}

int launch_distinctive(const DistinctArgs& a, cudaStream_t st) {
  const size_t smem = (size_t)kDistWarps * ((size_t)a.kcap * 8 + 288) * 4;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_distinctive, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_distinctive<<<(a.n_points + kDistWarps - 1) / kDistWarps, kDistWarps * 32, smem, st>>>(a);
  return 1;
}

int launch_search_init(const InitMatchArgs& a, cudaStream_t st) {
  const size_t lat = init_lat_bytes(a.cap);
  if (lat <= 200 * 1024 && a.cap <= 65535 && !getenv("ORBX_PROJ_SERIAL")) {
    cudaFuncSetAttribute(k_search_init_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lat);
    k_search_init_lat<<<a.n_pairs, kParThreads, lat, st>>>(a);
  } else {
    k_search_init<<<a.n_pairs, 32, 0, st>>>(a);
  }
  return 1;
}

int launch_knn_ratio(const KnnArgs& a, cudaStream_t st) {
  // Smallest second-best distance that passes the ratio test against ANY acceptable best (<= th); it is also
  // kept above th, so a best at that distance can never be accepted.  Neighbours that far need no tracking.
  int dstar = 257;
  for (int s2 = 0; s2 <= 256; ++s2)
    if ((float)a.th < a.ratio * (float)s2) { dstar = s2; break; }
  dstar = std::min(std::max(dstar, a.th + 1), 257);
  const uint32_t sentinel = (uint32_t)dstar << 16;
  const int qblocks = (a.cap + kKnnThreads - 1) / kKnnThreads;
  const size_t smem = (size_t)a.cap * 32;
  const size_t rs = (size_t)((a.cap + 31) / 32) * 4 + (size_t)a.cap * 4;
  if (rs + smem <= 200 * 1024 && !getenv("ORBX_KNN_SPLIT")) {
    cudaFuncSetAttribute(k_knn_topk<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(rs + smem));
    k_knn_topk<true><<<a.n_pairs * qblocks, kKnnThreads, rs + smem, st>>>(a, qblocks, sentinel);
    return 1;
  }
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_knn_topk<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_knn_topk<false><<<a.n_pairs * qblocks, kKnnThreads, smem, st>>>(a, qblocks, sentinel);
  if (rs + smem <= 200 * 1024) {
    cudaFuncSetAttribute(k_knn_resolve<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(rs + smem));
    k_knn_resolve<true><<<a.n_pairs, kKnnResolveThreads, rs + smem, st>>>(a, sentinel);
  } else {
    if (rs > 48 * 1024) cudaFuncSetAttribute(k_knn_resolve<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs);
    k_knn_resolve<false><<<a.n_pairs, kKnnResolveThreads, rs, st>>>(a, sentinel);
  }
  return 2;
}

// ------------------------------------------------------------------------------------------
// K7 without any allocation (orbx_knn_ratio_noalloc, include/orbx.h).
//
// Who needs this: code that runs where the CUDA allocator must not be touched — a tracking
// thread with a hard per-frame deadline (cudaMalloc/cudaFree take a device-wide lock and can
// stall for milliseconds behind another thread's allocation), a process that has handed all of
// its device memory to a pool and forbids further driver allocations, or a plug-in loaded into a
// host that audits allocations.  The batched entry point above obtains its scratch with
// cudaMallocAsync; this one keeps the scratch in a fixed array that is part of the library image.
//
// The array below is that scratch: one rotation-histogram bin per A feature (int) followed by one
// "already matched" flag per B feature (byte).  It exists once per GPU that loads liborbx.so and
// stays until the process exits; every call on that GPU uses the same bytes.  Consequences:
//   * sizes are capped at kNoAllocMax features per side; larger inputs are rejected
//   * the array still holds the previous call's contents, so the kernel clears the flags it
//     relies on being zero before it reads them (the bins are written before they are read)
//   * two calls that overlap on the GPU (different streams, or different host threads without
//     ordering) race on the array; serialise them
// In host code `g_knn_static` is only a handle for the runtime, not a device address: the kernel
// must be given the address that cudaGetSymbolAddress reports for the current device.
constexpr int kNoAllocMax = 4096;   // features per side; scratch = 4096*4 + 4096 bytes = 20 KiB per GPU
static __device__ __align__(256) unsigned char g_knn_static[kNoAllocMax * sizeof(int) + kNoAllocMax];

__global__ void __launch_bounds__(kKnnThreads) k_knn_ratio_static(const uint32_t* __restrict__ dA, const float* __restrict__ angA, int nA,
                                                                  const uint32_t* __restrict__ dB, const float* __restrict__ angB, int nB,
                                                                  int th, float ratio, int checkOri, int strict, int* m12, int* nm,
                                                                  unsigned char* scratch, int capPad) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint32_t* Bt = reinterpret_cast<uint32_t*>(smem);                 // [8][capPad] transposed B descriptors
  int* binof = reinterpret_cast<int*>(scratch);                     // [kNoAllocMax]
  unsigned char* used = scratch + kNoAllocMax * sizeof(int);        // [kNoAllocMax], must start at 0
  __shared__ unsigned s_key[kKnnThreads / 32];
  __shared__ int s_sec[kKnnThreads / 32];
  __shared__ int s_hist[kHisto];
  __shared__ int s_nm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int idx = tid; idx < nB * 8; idx += kKnnThreads) { const int j = idx >> 3, w = idx & 7; Bt[w * capPad + j] = __ldg(dB + idx); }
  for (int j = tid; j < nB; j += kKnnThreads) used[j] = 0;          // stale flags from the previous call
  if (tid < kHisto) s_hist[tid] = 0;
  if (tid == 0) s_nm = 0;
  __syncthreads();
  for (int i = 0; i < nA; ++i) {
    uint32_t q[8];
#pragma unroll
    for (int w = 0; w < 8; ++w) q[w] = __ldg(dA + (size_t)i * 8 + w);
    unsigned b1 = 256u << 16;
    int b2 = 256;
    for (int j = tid; j < nB; j += kKnnThreads) {
      if (used[j]) continue;
      int d = 0;
#pragma unroll
      for (int w = 0; w < 8; w += 2)
        d += __popcll(((unsigned long long)(q[w + 1] ^ Bt[(w + 1) * capPad + j]) << 32) | (q[w] ^ Bt[w * capPad + j]));
      const unsigned key = ((unsigned)d << 16) | (unsigned)j;
      if (d < (int)(b1 >> 16)) { b2 = (int)(b1 >> 16); b1 = key; }
      else if (d < b2) b2 = d;
    }
    const unsigned wmin = __reduce_min_sync(0xffffffffu, b1);
    const int wsec = (int)__reduce_min_sync(0xffffffffu, (unsigned)(b1 == wmin ? b2 : (int)(b1 >> 16)));
    if (lane == 0) { s_key[warp] = wmin; s_sec[warp] = wsec; }
    __syncthreads();
    if (tid == 0) {
      unsigned g = s_key[0];
      int own = 0;
      for (int w = 1; w < kKnnThreads / 32; ++w) if (s_key[w] < g) { g = s_key[w]; own = w; }
      int sec = s_sec[own];
      for (int w = 0; w < kKnnThreads / 32; ++w) if (w != own) sec = min(sec, (int)(s_key[w] >> 16));
      const int best = (int)(g >> 16), bi = g & 0xFFFF;
      int m = -1, bin = -1;
      const bool pass = strict ? best < th : best <= th;
      if (pass && (float)best < __fmul_rn(ratio, (float)sec)) {
        m = bi; used[bi] = 1; s_nm += 1;
        if (checkOri) { bin = rot_bin(angA[i], angB[bi]); s_hist[bin] += 1; }
      }
      m12[i] = m; binof[i] = bin;
    }
    __syncthreads();
  }
  if (checkOri) {
    int k_a, k_b, k_c;
    three_maxima(s_hist, k_a, k_b, k_c);
    int drop = 0;
    for (int i = tid; i < nA; i += kKnnThreads) {
      const int b = binof[i];
      if (b >= 0 && b != k_a && b != k_b && b != k_c) { m12[i] = -1; ++drop; }
    }
    if (drop) atomicSub(&s_nm, drop);
    __syncthreads();
  }
  if (tid == 0) *nm = s_nm;
}

}  // namespace orbx

// Device address of the static scratch, looked up once per device (with that device current).
static unsigned char* g_knn_static_addr[64] = {};

extern "C" orbx_status orbx_knn_ratio_noalloc(const uint8_t* descA, const float* angA, int nA, const uint8_t* descB,
                                              const float* angB, int nB, int th, float nnratio, int check_orientation,
                                              int strict_lt, int* matches12, int* n_matches, void* stream) {
  using namespace orbx;
  if (!descA || !angA || !descB || !angB || !matches12 || !n_matches) return set_error(ORBX_ERR_BAD_ARG, "null argument");
  if (nA < 0 || nB < 0) return set_error(ORBX_ERR_BAD_ARG, "negative feature count");
  // a best distance of 256 is the "no candidate" value of the scan and must never pass the threshold
  if (th < 0 || th > 255) return set_error(ORBX_ERR_BAD_ARG, "th must be 0..255 (the reference passes TH_LOW = 50)");
  if (nA > kNoAllocMax || nB > kNoAllocMax)
    return set_error(ORBX_ERR_CAPACITY, "orbx_knn_ratio_noalloc handles at most 4096 features per side (fixed static scratch)");
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return set_error(ORBX_ERR_CUDA, "cudaGetDevice failed");
  if (dev < 0 || dev >= 64) return set_error(ORBX_ERR_CAPACITY, "orbx_knn_ratio_noalloc keeps its per-device set-up for device indices 0..63 only");
  const int capPad = (kNoAllocMax + 31) & ~31;
  const size_t smem = (size_t)capPad * 32;
  if (!g_knn_static_addr[dev]) {   // one-time set-up for this device: allocates nothing, but must precede stream capture
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_knn_static);
    if (e != cudaSuccess || !p) return set_error(ORBX_ERR_CUDA, cudaGetErrorString(e));
    e = cudaFuncSetAttribute(k_knn_ratio_static, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_error(ORBX_ERR_CUDA, cudaGetErrorString(e));
    g_knn_static_addr[dev] = static_cast<unsigned char*>(p);
  }
  k_knn_ratio_static<<<1, kKnnThreads, smem, (cudaStream_t)stream>>>(
      reinterpret_cast<const uint32_t*>(descA), angA, nA, reinterpret_cast<const uint32_t*>(descB), angB, nB, th, nnratio,
      check_orientation, strict_lt, matches12, n_matches, g_knn_static_addr[dev], capPad);
  count_launches(1);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(ORBX_ERR_CUDA, cudaGetErrorString(e));
  return ORBX_OK;
}
