// K15 — DBoW2 TemplatedVocabulary::transform for batches of frames.
//
// Replaces, for Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:395-402, levelsup = 4):
//   transform(features, BowVector, FeatureVector, levelsup)  Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1196
//   transform(feature, word, weight, node, levelsup)         :1216-1258   (tree descent, first least distance wins)
//   FORB::distance                                           FORB.cpp:81-101
//   BowVector::addWeight / addIfNotExist / normalize         BowVector.cpp:32-92
//   FeatureVector::addFeature                                FeatureVector.cpp:28-45
//
// The vocabulary lives on the device as flat arrays (orbx_vocabulary_create).  One CTA per frame:
//   1. one thread per feature walks the tree (k children per level, 256-bit Hamming, strict '<' keeps the
//      first child on ties) and records the word, its weight and the node at level L - levelsup;
//   2. (word, feature) keys are bitonic-sorted in shared memory; the first key of each word owns the word and
//      adds the word's weight once per feature IN FEATURE ORDER - the double sum BowVector::addWeight builds;
//   3. the L1 / L2 norm is a sum in ascending word order (std::map iteration): one thread adds the values
//      sequentially, then every value is divided by it.  Doubles use explicit round-to-nearest intrinsics.
// The FeatureVector comes back as one node id per feature, which is what K12 / K14 take.
#include "orbx_internal.h"

namespace orbx {

constexpr int kBowTrThreads = 128;

__device__ __forceinline__ int ham8g(const uint32_t* q, const uint32_t* __restrict__ g) {
  const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(g)), b1 = __ldg(reinterpret_cast<const uint4*>(g) + 1);
  return __popc(q[0] ^ b0.x) + __popc(q[1] ^ b0.y) + __popc(q[2] ^ b0.z) + __popc(q[3] ^ b0.w) + __popc(q[4] ^ b1.x) +
         __popc(q[5] ^ b1.y) + __popc(q[6] ^ b1.z) + __popc(q[7] ^ b1.w);
}

__global__ void __launch_bounds__(kBowTrThreads) k_bow_transform(BowTrArgs a, int capPad) {
  extern __shared__ __align__(16) uint8_t smem[];
  unsigned long long* key = reinterpret_cast<unsigned long long*>(smem);     // [capPad] word<<32 | feature, sorted
  double* val = reinterpret_cast<double*>(key + capPad);                      // [capPad] value of the word owned by slot e
  int* slot = reinterpret_cast<int*>(val + capPad);                           // [capPad] output position of an owner, -1 otherwise
  __shared__ int s_scan[kBowTrThreads];
  __shared__ double s_norm;
  const int f = blockIdx.x, tid = threadIdx.x;
  const int n = min(a.n[f], a.cap);
  const size_t fo = (size_t)f * a.cap;
  const uint32_t* desc = reinterpret_cast<const uint32_t*>(a.desc + fo * 32);
  const uint32_t* vdesc = reinterpret_cast<const uint32_t*>(a.v_desc);
  const int nid_level = a.L - a.levelsup;

  // 1. tree descent, one feature per thread
  for (int i = tid; i < capPad; i += kBowTrThreads) {
    unsigned long long k = ~0ull;
    if (i < n) {
      uint32_t q[8];
#pragma unroll
      for (int w = 0; w < 8; ++w) q[w] = __ldg(desc + (size_t)i * 8 + w);
      int node = 0, level = 0, nid = 0;
      do {
        ++level;
        const int* ch = a.v_children + a.v_child_first[node];
        const int nc = a.v_child_count[node];
        int bestn = __ldg(ch), best = ham8g(q, vdesc + (size_t)bestn * 8);
        for (int c = 1; c < nc; ++c) {
          const int id = __ldg(ch + c);
          const int d = ham8g(q, vdesc + (size_t)id * 8);
          if (d < best) { best = d; bestn = id; }
        }
        node = bestn;
        if (level == nid_level) nid = node;
      } while (a.v_child_count[node] > 0);
      const bool keep = a.v_weight[node] > 0.0;                              // "not stopped", TemplatedVocabulary.h:1161
      a.word_of[fo + i] = keep ? a.v_word[node] : -1;
      a.node_of[fo + i] = keep ? nid : -1;
      if (keep) k = ((unsigned long long)(uint32_t)a.v_word[node] << 32) | (unsigned)i;
    }
    key[i] = k;
  }
  __syncthreads();
  // 2. sort by (word, feature)
  for (int k = 2; k <= capPad; k <<= 1)
    for (int jj = k >> 1; jj > 0; jj >>= 1) {
      for (int i = tid; i < capPad; i += kBowTrThreads) {
        const int l = i ^ jj;
        if (l > i) {
          const bool up = (i & k) == 0;
          const unsigned long long x = key[i], y = key[l];
          if ((x > y) == up) { key[i] = y; key[l] = x; }
        }
      }
      __syncthreads();
    }
  // owners: the first key of each word.  TF_IDF / TF: value = the word's weight added once per feature, in
  // feature order (addWeight); IDF / BINARY: the weight once (addIfNotExist).
  for (int e = tid; e < capPad; e += kBowTrThreads) {
    const unsigned long long k = key[e];
    bool owner = false;
    if (k != ~0ull) {
      const uint32_t w = (uint32_t)(k >> 32);
      owner = e == 0 || (uint32_t)(key[e - 1] >> 32) != w;
      if (owner) {
        const double wt = a.v_wweight[w];
        double v = wt;
        if (a.weighting <= 1)
          for (int t = e + 1; t < capPad && (uint32_t)(key[t] >> 32) == w && key[t] != ~0ull; ++t) v = __dadd_rn(v, wt);
        val[e] = v;
      }
    }
    slot[e] = owner ? 1 : 0;
  }
  __syncthreads();
  // output positions: exclusive scan of the owner flags in slot order.  Thread t holds slots t, t+128, ... so the
  // scan is done over the flags themselves, 128 per sweep.
  int carry = 0;
  for (int base = 0; base < capPad; base += kBowTrThreads) {
    const int e = base + tid;
    const int flag = slot[e];
    s_scan[tid] = flag;
    __syncthreads();
    for (int o = 1; o < kBowTrThreads; o <<= 1) {
      const int t = tid >= o ? s_scan[tid - o] : 0;
      __syncthreads();
      s_scan[tid] += t;
      __syncthreads();
    }
    const int incl = s_scan[tid], total = s_scan[kBowTrThreads - 1];
    slot[e] = flag ? carry + incl - 1 : -1;
    carry += total;
    __syncthreads();
  }
  const int nwords = carry;
  // 3. normalisation (BowVector.cpp:62-92) or, without it, the division by the vector size (:1168-1174)
  if (tid == 0) {
    double nrm = 0.0;
    if (a.norm == 1) { for (int e = 0; e < capPad; ++e) if (slot[e] >= 0) nrm = __dadd_rn(nrm, fabs(val[e])); }
    else if (a.norm == 2) { for (int e = 0; e < capPad; ++e) if (slot[e] >= 0) nrm = __dadd_rn(nrm, __dmul_rn(val[e], val[e])); nrm = __dsqrt_rn(nrm); }
    s_norm = nrm;
    a.bow_count[f] = nwords;
  }
  __syncthreads();
  const double nrm = s_norm;
  for (int e = tid; e < capPad; e += kBowTrThreads) {
    const int o = slot[e];
    if (o < 0) continue;
    double v = val[e];
    if (a.norm == 0) { if (a.weighting <= 1) v = __ddiv_rn(v, (double)nwords); }
    else if (nrm > 0.0) v = __ddiv_rn(v, nrm);
    a.bow_word[fo + o] = (int)(key[e] >> 32);
    a.bow_val[fo + o] = v;
  }
}

int launch_bow_transform(const BowTrArgs& a, cudaStream_t st) {
  int capPad = kBowTrThreads;
  while (capPad < a.cap) capPad <<= 1;
  const size_t smem = (size_t)capPad * (8 + 8 + 4);
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_bow_transform, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_bow_transform<<<a.n_frames, kBowTrThreads, smem, st>>>(a, capPad);
  return 1;
}

}  // namespace orbx
