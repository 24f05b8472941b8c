// K3 — quad-tree keypoint distribution, one CTA per (image, level).
//
// Replaces ORBextractor::DistributeOctTree / ExtractorNode::DivideNode
// (src/ORBextractor.cc:539-763, 481-537).  The reference walks a std::list of nodes, each
// owning a std::vector of keypoints, and splits nodes into 4 until it has N of them; the output
// order is the list order and, per node, the first keypoint with the maximal response.
//
// B200 formulation (DESIGN.md "K3"): nothing in the algorithm depends on the physical key
// vectors — keys inside a node always stay in candidate order — so a key only carries the list
// position of its node (16-bit, global scratch) and nodes are SoA arrays in shared memory indexed
// by list position.  One breadth-first pass is: per-key quadrant + shared-memory child counters,
// a block scan over nodes in list order (children are push_front'ed: new position = C-1-creation
// index; untouched leaves follow in old order), a key remap.  The reference's "careful" phase
// (sort open nodes by (size, address), split from the back, stop at N) is order-determined too:
// rank = #(bigger, or equal and earlier in the list), gains prefix-summed in rank order, first
// rank reaching N found by a block min — so it is parallel as well.  Heap-address ties (hazard
// H1) are resolved as creation order, which equals list position among same-round nodes.
#include "orbx_internal.h"

namespace orbx {

constexpr int kOctThreads = 256;

struct Box { short x0, y0, x1, y1; };

__device__ __forceinline__ int quadrant(uint32_t key, Box b) {
  const int mx = b.x0 + ((b.x1 - b.x0 + 1) >> 1), my = b.y0 + ((b.y1 - b.y0 + 1) >> 1);  // ceil(d/2), :483-484
  const int kx = unpack_x(key), ky = unpack_y(key);
  return (kx < mx) ? ((ky < my) ? 0 : 2) : ((ky < my) ? 1 : 3);                              // :515-525
}
__device__ __forceinline__ Box child_box(Box b, int q) {
  const short mx = (short)(b.x0 + ((b.x1 - b.x0 + 1) >> 1)), my = (short)(b.y0 + ((b.y1 - b.y0 + 1) >> 1));
  Box c;
  c.x0 = (q & 1) ? mx : b.x0;  c.x1 = (q & 1) ? b.x1 : mx;
  c.y0 = (q & 2) ? my : b.y0;  c.y1 = (q & 2) ? b.y1 : my;
  return c;
}

// Exclusive block scan of one value per thread; *total receives the block sum.  sw: int[34].
__device__ __forceinline__ int block_scan_excl(int v, int* sw, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) sw[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const int nw = kOctThreads / 32;
    int w = lane < nw ? sw[lane] : 0, wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    if (lane < nw) sw[lane] = wi - w;
    if (lane == nw - 1) sw[33] = wi;
  }
  __syncthreads();
  const int ex = sw[warp] + incl - v;
  *total = sw[33];
  __syncthreads();
  return ex;
}
// In-place exclusive scan of a shared array of any length; returns the total.
__device__ __forceinline__ int array_scan_excl(int* a, int len, int* sw) {
  int carry = 0;
  for (int base = 0; base < len; base += kOctThreads) {
    const int i = base + threadIdx.x;
    int tot;
    const int ex = block_scan_excl(i < len ? a[i] : 0, sw, &tot);
    if (i < len) a[i] = carry + ex;
    carry += tot;
  }
  __syncthreads();
  return carry;
}
__device__ __forceinline__ int block_sum(int v, int* sw) {
  int tot;
  block_scan_excl(v, sw, &tot);
  return tot;
}

__global__ void __launch_bounds__(kOctThreads) k_octree(DeviceLayout L, int first_slot, int n_img, int NC, int KC) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  Box* box0 = reinterpret_cast<Box*>(smem_raw);
  Box* box1 = box0 + NC;
  int* cnt0 = reinterpret_cast<int*>(box1 + NC);
  int* cnt1 = cnt0 + NC;
  int* cc = cnt1 + NC;        // [4*NC] child counts, then new positions
  int* a1 = cc + 4 * NC;      // [NC]
  int* a2 = a1 + NC;
  int* a3 = a2 + NC;
  int* a4 = a3 + NC;
  int* a5 = a4 + NC;
  int* sw = a5 + NC;          // [34] scan scratch
  int* s_misc = sw + 34;      // [2]
  // the level's ordered keys and their node ids, kept in shared memory when there are at most KC of them (every
  // pass reads both; from global memory those loads were 35 % of the kernel's stall samples)
  uint32_t* s_keys = reinterpret_cast<uint32_t*>(s_misc + 2);
  uint16_t* s_node = reinterpret_cast<uint16_t*>(s_keys + KC);

  const int tid = threadIdx.x;
  const int img = blockIdx.x / L.nlevels, level = blockIdx.x - img * L.nlevels;
  if (img >= n_img) return;
  const int slot = first_slot + img;
  const LevelGeom& g = L.geom[level];
  const int N = g.quota;
  const uint2* tab = L.celltab + (size_t)slot * L.cells_per_slot + g.cell_base;
  const uint32_t* pool = L.pool + (size_t)slot * L.pool_per_slot + g.cand_off;
  uint32_t* gkeys = L.okeys + (size_t)slot * L.pool_per_slot + g.cand_off;   // global copy: debug API, and the working set when it exceeds KC
  uint16_t* gnode = L.knode + (size_t)slot * L.pool_per_slot + g.cand_off;

  // ---- A. candidates in reference order: cells row-major, pixels row-major inside a cell ----
  const int ncells = g.nCols * g.nRows;
  int nkeys = 0;
  for (int base = 0; base < ncells; base += kOctThreads) {
    const int c = base + tid;
    uint2 e = make_uint2(0u, 0u);
    if (c < ncells) e = tab[c];
    int tot;
    const int ex = block_scan_excl((int)e.y, sw, &tot);
    for (uint32_t j = 0; j < e.y; ++j) {
      const uint32_t key = pool[e.x + j];
      const int k = nkeys + ex + (int)j;
      gkeys[k] = key;
      if (k < KC) s_keys[k] = key;
    }
    nkeys += tot;
  }
  if (tid == 0) L.nkeys[slot * kMaxLevels + level] = nkeys;
  const bool in_smem = nkeys <= KC;                  // CTA-uniform
  const uint32_t* okeys = in_smem ? s_keys : gkeys;
  uint16_t* knode = in_smem ? s_node : gnode;

  // ---- B. root nodes (src/ORBextractor.cc:543-585) ----
  Box* box[2] = {box0, box1};
  int* cnt[2] = {cnt0, cnt1};
  int cur = 0;
  for (int i = tid; i < NC; i += kOctThreads) cnt0[i] = 0;
  if (tid < g.nIni) {
    Box b; b.x0 = (short)g.root_x0[tid]; b.x1 = (short)g.root_x1[tid]; b.y0 = 0; b.y1 = (short)g.treeH;
    box0[tid] = b;
  }
  __syncthreads();
  for (int k = tid; k < nkeys; k += kOctThreads) {
    const float kx = (float)unpack_x(okeys[k]);
    int r = (int)__fdiv_rn(kx, g.hX);                           // vpIniNodes[kp.pt.x/hX], :569
    r = min(max(r, 0), g.nIni - 1);
    knode[k] = (uint16_t)r;
    atomicAdd(&cnt0[r], 1);
  }
  __syncthreads();
  int n = g.nIni;         // length of the node arrays (may hold empty roots before the first pass)
  int before = 0;         // lNodes.size() at the top of the reference's loop
  for (int i = 0; i < g.nIni; ++i) before += cnt0[i] > 0;
  int C = 0;              // nodes created by the last pass/round = list positions [0, C)
  bool careful = false;

  // ---- C. breadth-first passes (src/ORBextractor.cc:594-672) ----
  while (true) {
    for (int i = tid; i < 4 * n; i += kOctThreads) cc[i] = 0;
    __syncthreads();
    for (int k = tid; k < nkeys; k += kOctThreads) {
      const int nd = knode[k];
      if (cnt[cur][nd] != 1) atomicAdd(&cc[4 * nd + quadrant(okeys[k], box[cur][nd])], 1);
    }
    __syncthreads();
    int my_open = 0;
    for (int i = tid; i < n; i += kOctThreads) {
      const int c = cnt[cur][i];
      int nch = 0;
      if (c > 1)
        for (int q = 0; q < 4; ++q) { const int v = cc[4 * i + q]; nch += v > 0; my_open += v > 1; }
      a1[i] = nch;
      a2[i] = (c == 1);
    }
    __syncthreads();
    C = array_scan_excl(a1, n, sw);
    const int K = array_scan_excl(a2, n, sw);
    const int n_open = block_sum(my_open, sw);
    const int nxt = cur ^ 1;
    for (int i = tid; i < n; i += kOctThreads) {
      const int c = cnt[cur][i];
      const Box b = box[cur][i];
      if (c == 1) {
        const int p = C + a2[i];
        box[nxt][p] = b; cnt[nxt][p] = 1; cc[4 * i] = p;
      } else if (c > 1) {
        int r = 0;
        for (int q = 0; q < 4; ++q) {
          const int v = cc[4 * i + q];
          if (v > 0) {
            const int p = C - 1 - (a1[i] + r);                  // push_front: newest first
            box[nxt][p] = child_box(b, q); cnt[nxt][p] = v; cc[4 * i + q] = p;
            ++r;
          }
        }
      }
    }
    __syncthreads();
    for (int k = tid; k < nkeys; k += kOctThreads) {
      const int nd = knode[k];
      const int p = (cnt[cur][nd] == 1) ? cc[4 * nd] : cc[4 * nd + quadrant(okeys[k], box[cur][nd])];
      knode[k] = (uint16_t)p;
    }
    __syncthreads();
    cur = nxt;
    n = C + K;
    if (n >= N || n == before) break;                            // :669
    if (n + 3 * n_open > N) { careful = true; break; }           // :673
    before = n;
  }

  // ---- D. careful rounds (src/ORBextractor.cc:676-737) ----
  while (careful) {
    before = n;
    // open nodes: created last round (positions < C) and holding more than one key
    for (int i = tid; i < 4 * C; i += kOctThreads) cc[i] = 0;
    if (tid == 0) { s_misc[0] = 0x7fffffff; s_misc[1] = 0; }
    __syncthreads();
    for (int k = tid; k < nkeys; k += kOctThreads) {
      const int nd = knode[k];
      if (nd < C && cnt[cur][nd] > 1) atomicAdd(&cc[4 * nd + quadrant(okeys[k], box[cur][nd])], 1);
    }
    __syncthreads();
    // rank by (size desc, list position asc) == the reference's sort() walked from the back.
    // One packed key per open node, size << 16 | (0xFFFF - position) (0 for the others): rank = number of larger
    // keys, read two at a time.  Positions are < node_cap <= 3072; a node's size fits 16 bits whenever the level
    // has fewer than 65536 candidates (`packed`, CTA-uniform).  A level with more (a multi-megapixel image of
    // dense texture) ranks on the unpacked sizes with the position as the tie-break: slower, same order.
    const bool packed = nkeys < 65536;
    unsigned* rkey = reinterpret_cast<unsigned*>(a5);            // a5 is free until the survivor scan below
    for (int i = tid; i < ((C + 1) & ~1); i += kOctThreads) {   // padded to an even count
      const int ci = i < C ? cnt[cur][i] : 0;
      rkey[i] = ci > 1 ? (packed ? ((unsigned)ci << 16) | (unsigned)(0xFFFF - i) : (unsigned)ci) : 0u;
    }
    __syncthreads();
    int my_m = 0;
    for (int i = tid; i < C; i += kOctThreads) {
      const int ci = cnt[cur][i];
      int rank = -1;
      if (ci > 1) {
        rank = 0;
        const unsigned mk = rkey[i];
        if (packed) {
          for (int j = 0; j < C; j += 2) {
            const uint2 kj = *reinterpret_cast<const uint2*>(rkey + j);
            rank += (kj.x > mk) + (kj.y > mk);
          }
        } else {
          for (int j = 0; j < C; ++j) rank += (rkey[j] > mk) || (rkey[j] == mk && j < i);
        }
        int nch = 0;
        for (int q = 0; q < 4; ++q) nch += cc[4 * i + q] > 0;
        a1[rank] = nch - 1;   // gain of splitting this node, in rank order
        a3[rank] = nch - 1;
        a2[rank] = i;         // node by rank
        ++my_m;
      }
      a4[i] = rank;           // rank by node
    }
    __syncthreads();
    const int m = block_sum(my_m, sw);
    array_scan_excl(a1, m, sw);                                  // a1[r] = sum of gains before rank r
    for (int r = tid; r < m; r += kOctThreads)
      if (n + a1[r] + a3[r] >= N) atomicMin(&s_misc[0], r);      // first split that reaches N, :730
    __syncthreads();
    const int nsplit = (m == 0) ? 0 : (s_misc[0] == 0x7fffffff ? m : s_misc[0] + 1);
    const int Cn = (nsplit == 0) ? 0 : a1[nsplit - 1] + a3[nsplit - 1] + nsplit;   // children created
    for (int i = tid; i < n; i += kOctThreads) {
      const bool split = i < C && a4[i] >= 0 && a4[i] < nsplit;
      a5[i] = split ? 0 : 1;
    }
    __syncthreads();
    array_scan_excl(a5, n, sw);                                  // rank among surviving old nodes
    const int nxt = cur ^ 1;
    for (int i = tid; i < n; i += kOctThreads) {
      const int rank = (i < C) ? a4[i] : -1;
      const Box b = box[cur][i];
      if (rank >= 0 && rank < nsplit) {
        const int cbase = a1[rank] + rank;                       // children created before this node's
        int r = 0;
        for (int q = 0; q < 4; ++q) {
          const int v = cc[4 * i + q];
          if (v > 0) {
            const int p = Cn - 1 - (cbase + r);
            box[nxt][p] = child_box(b, q); cnt[nxt][p] = v; cc[4 * i + q] = p;
            ++r;
          }
        }
      } else {
        const int p = Cn + a5[i];
        box[nxt][p] = b; cnt[nxt][p] = cnt[cur][i];
        a5[i] = p;
      }
    }
    __syncthreads();
    for (int k = tid; k < nkeys; k += kOctThreads) {
      const int nd = knode[k];
      const int rank = (nd < C) ? a4[nd] : -1;
      const int p = (rank >= 0 && rank < nsplit) ? cc[4 * nd + quadrant(okeys[k], box[cur][nd])] : a5[nd];
      knode[k] = (uint16_t)p;
    }
    __syncthreads();
    cur = nxt;
    n = Cn + (n - nsplit);
    C = Cn;
    if (n >= N || n == before) break;                            // :734
  }

  // ---- E. best keypoint per node, first one on ties (src/ORBextractor.cc:742-760) ----
  for (int i = tid; i < n; i += kOctThreads) a1[i] = 0;
  __syncthreads();
  for (int k = tid; k < nkeys; k += kOctThreads)
    atomicMax(reinterpret_cast<unsigned*>(&a1[knode[k]]), ((unsigned)unpack_r(okeys[k]) << 24) | (0xFFFFFFu - (unsigned)k));
  __syncthreads();
  uint32_t* out = L.lvlkp + (size_t)slot * L.lvlkp_per_slot + g.lvl_kp_off;
  for (int i = tid; i < n; i += kOctThreads) {
    const uint32_t key = okeys[0xFFFFFFu - ((unsigned)a1[i] & 0xFFFFFFu)];
    out[i] = pack_xyr(unpack_x(key) + kFastBorder, unpack_y(key) + kFastBorder, unpack_r(key));  // :842-843
  }
  if (tid == 0) L.lvlcnt[slot * kMaxLevels + level] = n;
}

constexpr int kOctKeyCap = 3072;   // keys per (image, level) held in shared memory (6 bytes each); beyond that the kernel works from global memory
size_t octree_smem_bytes(int NC) { return (size_t)NC * (2 * sizeof(Box) + 2 * 4 + 16 + 5 * 4) + 36 * 4 + (size_t)kOctKeyCap * 6; }

int launch_octree(const DeviceLayout& L, int first_slot, int n, int node_cap, cudaStream_t st) {
  const size_t smem = octree_smem_bytes(node_cap);
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_octree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_octree<<<n * L.nlevels, kOctThreads, smem, st>>>(L, first_slot, n, node_cap, kOctKeyCap);
  return 1;
}

}  // namespace orbx
