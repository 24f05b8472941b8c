// Internal definitions shared by the orbx kernels and the C-ABI host code.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/orbx.h"

namespace orbx {

constexpr int kEdge = 19;          // EDGE_THRESHOLD, src/ORBextractor.cc:74
constexpr int kFastBorder = 16;    // EDGE_THRESHOLD-3, src/ORBextractor.cc:773
constexpr int kHalfPatch = 15;     // HALF_PATCH_SIZE
constexpr int kMaxLevels = 8;
constexpr int kMaxRoots = 16;
constexpr int kMaxDim = 4095;      // x,y are packed in 12 bits

// packed candidate / level keypoint: x | y<<12 | response<<24 (response = FAST score, 1..254)
__host__ __device__ inline uint32_t pack_xyr(int x, int y, int r) { return (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)r << 24); }
__host__ __device__ inline int unpack_x(uint32_t p) { return p & 0xFFF; }
__host__ __device__ inline int unpack_y(uint32_t p) { return (p >> 12) & 0xFFF; }
__host__ __device__ inline int unpack_r(uint32_t p) { return p >> 24; }

struct ResizeTap {  // one destination coordinate of cv::resize INTER_LINEAR
  int ofs;          // source index of the first tap
  short a0, a1;     // 11-bit fixed-point weights of taps ofs, ofs+1
};

struct LevelGeom {
  int w, h;              // level (ROI) size
  int pitch, rows;       // bordered plane: pitch bytes per row (multiple of 16), rows = h + 38
  size_t plane_off;      // byte offset of slot 0's plane of this level inside the pyramid arena
  size_t plane_stride;   // bytes between consecutive slots' planes of this level
  int xtab_off, ytab_off;  // offsets (in ResizeTap) of this level's resize tables
  int xtile_off, ytile_off;  // per 128-column / 32-row output tile of the tiled resize: first source column {ofs}, first source row and row count {ofs, a0}
  // FAST cell grid (src/ORBextractor.cc:773-806)
  int maxBX, maxBY, nCols, nRows, wCell, hCell;
  int cell_base;         // index of this level's first cell in the per-slot cell table
  int cand_off, cand_cap;  // this level's segment of the per-slot candidate pool
  // quad-tree (src/ORBextractor.cc:539-563)
  int quota, nIni, treeW, treeH, node_cap, lvl_kp_off;
  float hX;
  int root_x0[kMaxRoots], root_x1[kMaxRoots];
  float scale, inv_scale;
  int patch_size;
};

struct DeviceLayout {    // per-handle pointers handed to kernels by value
  int nlevels, nslots;
  int cells_per_slot, pool_per_slot, lvlkp_per_slot, kp_cap;
  int ini_th, min_th;
  const LevelGeom* geom;     // [nlevels]
  const ResizeTap* taps;
  uint8_t* pyr;
  uint8_t* blur;             // 7x7 sigma-2 blurred level planes, same layout as pyr (K4a); only the ROI part is written / read
  uint2* celltab;            // [nslots][cells_per_slot] {pool offset, count}
  int* poolcnt;              // [nslots][kMaxLevels]
  uint32_t* pool;            // [nslots][pool_per_slot]
  uint32_t* okeys;           // [nslots][pool_per_slot] candidates in reference order
  uint16_t* knode;           // [nslots][pool_per_slot]
  uint32_t* lvlkp;           // [nslots][lvlkp_per_slot] post-quad-tree keypoints, level coords
  int* lvlcnt;               // [nslots][kMaxLevels]
  int* nkeys;                // [nslots][kMaxLevels] ordered candidate count
  orbx_keypoint* kps;        // [nslots][kp_cap]
  uint2* kpk;                // [nslots][kp_cap] the same keypoints in 8 bytes: {x | y<<12 | response<<24 (level coordinates), angle bits}; level from lvlcnt
  uint8_t* desc;             // [nslots][kp_cap][32]
  int* nkp;                  // [nslots]
  int* err;                  // [nslots]
};

__host__ __device__ inline int reflect101(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i;
}

// one TMA tensor map per pyramid level (passed to kernels as a __grid_constant__ parameter)
struct TmaMaps { CUtensorMap m[kMaxLevels]; };
constexpr int kResizeTileW = 128, kResizeTileH = 32;   // output tile of k_resize_tiled (bordered pixels)
constexpr int kResizeBoxW = 192, kResizeBoxH = 48;  // K1 tiled resize: source footprint box
constexpr int kDescTileW = 64, kDescTileH = 43;   // K4 tile box: TMA needs a 16-byte aligned start, so 43+15 columns -> 64
constexpr int kBlurBoxW = 160, kBlurBoxH = 38;    // K4a source box of a 128 x 32 output tile: 3-px halo, 16-byte aligned start
constexpr int kMomBoxW = 48, kMomBoxH = 31;       // K4b raw box for the orientation moments: 31 + 15 columns -> 48
constexpr int kTapBoxW = 64, kTapBoxH = 37;       // K4b blurred box for the descriptor taps: 37 + 15 columns -> 64

// per-level fields the blur kernel reads, as a kernel parameter
constexpr int kBlurMaxRows = 8 * ((4096 + 3 + 31) / 32);   // tile rows of all levels stacked: at most 8 levels of a 4096-px high image
struct BlurGeom {
  int tiles_x[kMaxLevels], pitch[kMaxLevels], rows[kMaxLevels];
  size_t plane_off[kMaxLevels], plane_stride[kMaxLevels];
  uint16_t row_info[kBlurMaxRows];   // per stacked tile row: level | tile row inside the level << 4
};

// kernel launchers (each returns the number of kernel launches it made)
int launch_pyramid(const DeviceLayout& L, const LevelGeom* hgeom, const uint8_t* src, int src_pitch,
                   size_t src_img_stride, int first_slot, int n, unsigned tiled_mask, const TmaMaps* resize_maps, cudaStream_t st);
bool resize_tiled_fits(const LevelGeom& g, const ResizeTap* taps);
int launch_fast(const DeviceLayout& L, const LevelGeom* hg, const TmaMaps& maps, int first_slot, int n, int win_w, int win_h, cudaStream_t st);
int fast_window_pitch(int win_w);   // TMA box width of the FAST window (64 or 96)
int launch_octree(const DeviceLayout& L, int first_slot, int n, int node_cap, cudaStream_t st);
int launch_describe(const DeviceLayout& L, const TmaMaps& maps, int first_slot, int n, const int* d_umax, const char4* d_pattern, cudaStream_t st);
// K4a + K4b: whole-plane blur, then orientation + descriptors from a raw moments box and a blurred tap box per keypoint
int launch_blur(const DeviceLayout& L, const LevelGeom* hg, const TmaMaps& blur_src_maps, int first_slot, int n, cudaStream_t st);
int launch_describe_planes(const DeviceLayout& L, const TmaMaps& mom_maps, const TmaMaps& tap_maps, int first_slot, int n, const char4* d_pattern, cudaStream_t st);

struct StereoArgs {
  DeviceLayout L, R;
  int slotL0, slotR0, n;
  float bf, maxD;
  float* uRight; float* depth; int* nMatches; int* sad;  // device, [n][cap]
  int cap;
  uint4* rec; int* tab; int tab_stride;   // K5a output: sorted right-keypoint records [n][R.kp_cap], bin starts [n][tab_stride]
};
int launch_stereo(const StereoArgs& a, cudaStream_t st, cudaEvent_t ev_mid);

struct InitMatchArgs {
  const orbx_keypoint* k1; const uint8_t* d1; const int* n1;
  const orbx_keypoint* k2; const uint8_t* d2; const int* n2;
  int n_pairs, cap; const float* gb; float* prev; float ratio; float window; int checkOri;   // gb: device, Frame 2's {mnMinX, mnMaxX, mnMinY, mnMaxY, invW, invH}
  int* m12; int* nm;
  int* scratch;  // per pair: cap ints (matched distance) + cap ints (matches21) + cap ints (grid order)
};
int launch_search_init(const InitMatchArgs& a, cudaStream_t st);

struct KnnArgs {
  const uint8_t* dA; const float* angA; const int* nA;
  const uint8_t* dB; const float* angB; const int* nB;
  int n_pairs, cap, th; float ratio; int checkOri, strict;
  int* m12; int* nm;
  uint32_t* topk;  // per pair: cap x 8 sorted keys distance<<16 | j (nearest B features of every A feature)
  int* done;       // [n_pairs], zero on entry: how many query blocks of the pair have written their lists (the last one resolves the pair)
};
int launch_knn_ratio(const KnnArgs& a, cudaStream_t st);

struct ProjArgs {
  const orbx_keypoint* k; const uint8_t* d; const float* uRight; const int* n;          // frame side, [P][cap]
  const uint8_t* inView; const float* projX; const float* projY; const float* projXR;  // map-point side, [P][mcap]
  const int* level; const float* viewCos; const uint8_t* mpDesc; const int* mpObs; const int* m;
  int n_pairs, cap, mcap, nlevels;
  const float* gb;            // device: the Frame's {mnMinX, mnMaxX, mnMinY, mnMaxY, invW, invH}
  float th, ratio;
  float scale[kMaxLevels];
  int* frameObs; int* assigned; int* nm;
  uint4* spec;                // scratch [P][mcap]: the candidate list (four smallest keys) per map point (latency mode)
};
int launch_search_projection(const ProjArgs& a, cudaStream_t st);

struct ProjFrameArgs {
  const orbx_keypoint* k; const uint8_t* d; const float* uRight; const int* n;          // current frame, [P][cap]
  const uint8_t* lastValid; const float* lastXw; const int* lastOctave; const float* lastAngle;   // last frame, [P][lcap]
  const uint8_t* mpDesc; const int* mpObs; const int* nLast;
  const float* pose;          // [P][12]: Rcw row-major (9) then tcw (3)
  const int* mode;            // [P]: 0 neither, 1 forward, 2 backward
  int n_pairs, cap, lcap, nlevels, checkOri;
  const float* gb;            // device: the current Frame's {mnMinX, mnMaxX, mnMinY, mnMaxY, invW, invH}
  float fx, fy, cx, cy, bf, th;
  float scale[kMaxLevels];
  int* frameObs; int* assigned; int* nm;
  int* events;                // scratch [P][lcap]: bin<<16 | keypoint index per accepted match
  uint4* spec;                // scratch [P][lcap]: the candidate list per last-frame keypoint that carries a map point (latency mode)
};
int launch_search_projection_frame(const ProjFrameArgs& a, cudaStream_t st);

struct BowArgs {
  const uint8_t* dA; const float* angA; const uint8_t* validA; const int* nodeA; const int* nA;
  const uint8_t* dB; const float* angB; const uint8_t* validB; const int* nodeB; const int* nB;
  int n_pairs, cap, th; float ratio; int checkOri, strict;
  int* m12; int* nm;
  uint4* spec;     // scratch [P][cap]: per sorted query its two nearest free B features (distance << 16 | j) and its node's B run
};
int launch_search_by_bow(const BowArgs& a, cudaStream_t st);

struct BowTrArgs {
  // vocabulary (device): node descriptors, children lists, leaf word ids / weights, weight per word id
  const uint8_t* v_desc; const int* v_child_first; const int* v_child_count; const int* v_children; const int* v_word;
  const double* v_weight; const double* v_wweight;
  int L, levelsup, weighting, norm;
  const uint8_t* desc; const int* n; int n_frames, cap;
  int* word_of; int* node_of; int* bow_word; double* bow_val; int* bow_count;
};
int launch_bow_transform(const BowTrArgs& a, cudaStream_t st);

struct ProjSearchArgs {
  const orbx_keypoint* kps; const uint8_t* desc; const float* u_right;                 // target keyframes, [P][cap]
  const float* grid_bounds; const int* cell_start; const int* cell_items; const float* kf_bounds; const float* xf;
  const uint8_t* valid; const float* xw; const float* normal; const float* dmin; const float* dmax; const uint8_t* mp_desc;   // [P][mcap]
  const int* m;
  const int* n;               // [P] keypoints per target (used when cell_start is null: the 1024-thread form then builds the grid itself)
  int n_problems, cap, mcap, nlevels, mode, th_dist;
  float fx, fy, cx, cy, bf, log_scale, th;
  float scale[kMaxLevels], inv_sigma2[kMaxLevels];
  int* best;
};
int launch_project_search(const ProjSearchArgs& a, cudaStream_t st);
bool project_search_needs_grid(int cap, bool gate);          // true: the small form will run, which needs the caller's CSR grid
bool project_ordered_needs_grid(int cap, int mcap);
int launch_scatter_matches(const int* best, const int* mp_index, const int* m, int mcap, int cap, int n_problems, int* vn, cudaStream_t st);
struct Sim3AgreeArgs { const int* match1; const int* match2; const int* n1; int cap; int* matches12; int* nfound; };
int launch_sim3_agree(const Sim3AgreeArgs& a, int n_problems, cudaStream_t st);
// SearchBySim3 in one launch (both passes, scatter, agreement); returns 0 when the 1024-thread form does not fit: use the five launches
int launch_sim3_lat(const ProjSearchArgs& a0, const ProjSearchArgs& a1, const int* idx0, const int* idx1, int* vn0, int* vn1, int* ticket, const int* n1,
                    int* matches12, int* nfound, cudaStream_t st);

struct ProjOrderedArgs {
  const orbx_keypoint* kps; const uint8_t* desc; const int* n;                          // targets, [P][cap]
  const float* grid_bounds; const int* cell_start; const int* cell_items; const float* tgt_bounds; const float* xf;
  const uint8_t* occupied;
  const uint8_t* valid; const float* xw; const float* normal; const float* dmin; const float* dmax; const uint8_t* mp_desc;
  const float* src_angle; const int* m;
  int n_problems, cap, mcap, nlevels, reloc, th_dist, checkOri;
  float fx, fy, cx, cy, log_scale, th;
  float scale[kMaxLevels];
  int* assigned; int* nm;
  uint4* spec;                // scratch [2][P][mcap]: candidate lists (four smallest keys, distance << 16 | keypoint) by list position, then the search windows by map point (latency form)
};
int launch_project_match_ordered(const ProjOrderedArgs& a, cudaStream_t st);

struct TriArgs {
  const orbx_keypoint* k1; const uint8_t* d1; const uint8_t* hasMp1; const float* uR1; const int* node1; const int* n1;
  const orbx_keypoint* k2; const uint8_t* d2; const uint8_t* hasMp2; const float* uR2; const int* node2; const int* n2;
  const float* F12; const float* pose2; const float* Cw; const float* cam2;   // [P][9], [P][12], [P][3], [P][4]
  float scale2[kMaxLevels], sigma2_2[kMaxLevels];
  int n_pairs, cap, onlyStereo, checkOri;
  int* m12; int* nm;
  int* hist;       // scratch [P][32], zero on entry: the pair's rotation histogram, summed over its CTAs (1024-thread form)
  uint8_t* bins;   // scratch [P][cap]: rotation bin of every KF1 feature's match
};
int launch_search_triangulation(const TriArgs& a, cudaStream_t st);

struct DistinctArgs {
  const uint8_t* desc; const int* counts; int n_points, kcap;
  int* best_idx; uint8_t* best_desc;
};
int launch_distinctive(const DistinctArgs& a, cudaStream_t st);

struct FrameGridArgs {
  const orbx_keypoint* k; const int* n; int n_frames, cap, cols, rows;
  orbx_camera cam;
  orbx_keypoint* kun; float* bounds; int* cell_start; int* items;
};
int launch_undistort_grid(const FrameGridArgs& a, cudaStream_t st);

struct AreaArgs {
  const orbx_keypoint* kun; int cap;
  const float* bounds; const int* cell_start; const int* items;
  const int* q_frame; const float* q_x; const float* q_y; const float* q_r; const int* q_min_level; const int* q_max_level;
  int n_queries, qcap;
  int* out_idx; int* out_cnt;
};
int launch_features_in_area(const AreaArgs& a, cudaStream_t st);

// shared by the extern "C" entry points that live outside orbx_api.cu
orbx_status set_error(orbx_status s, const char* msg);
void count_launches(int n);

}  // namespace orbx
