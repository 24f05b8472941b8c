// K10 / K11 — the Frame-constructor steps that follow extraction, batched over frames.
//
// K10 k_undistort_grid   <- Frame::UndistortKeyPoints      src/Frame.cc:404-434  (cv::undistortPoints, P = K)
//                           Frame::ComputeImageBounds      src/Frame.cc:436-464
//                           Frame::AssignFeaturesToGrid    src/Frame.cc:230-245  (+ PosInGrid :382-392)
// K11 k_features_in_area <- Frame::GetFeaturesInArea        src/Frame.cc:327-380
//
// cv::undistortPoints (OpenCV 4.13, the 6-argument overload) is 5 fixed-point iterations of the
// radial/tangential model in double, no early exit; it is restated here with explicit round-to-nearest
// double intrinsics so that nothing is contracted into an FMA (the oracle's restatement is pinned against
// cv2.undistortPoints bit for bit).  The grid is delivered as the reference's mGrid flattened to CSR:
// cell = ix*48 + iy, items inside a cell in keypoint-index order (push_back order).
#include "orbx_internal.h"

namespace orbx {

constexpr int kGridCols = 64, kGridRows = 48, kCells = kGridCols * kGridRows;

struct CamD { double fx, fy, cx, cy, ifx, ify, k[5]; bool distorted; };

__device__ __forceinline__ CamD load_cam(const orbx_camera& c) {
  CamD d;
  d.fx = c.fx; d.fy = c.fy; d.cx = c.cx; d.cy = c.cy;
  d.ifx = __ddiv_rn(1.0, d.fx); d.ify = __ddiv_rn(1.0, d.fy);
  d.k[0] = c.k1; d.k[1] = c.k2; d.k[2] = c.p1; d.k[3] = c.p2; d.k[4] = c.k3;
  d.distorted = c.k1 != 0.0f;                                   // src/Frame.cc:406
  return d;
}

#define DM(a, b) __dmul_rn((a), (b))
#define DA(a, b) __dadd_rn((a), (b))
#define DS(a, b) __dsub_rn((a), (b))

// the reference's call: undistortPoints(mat, mat, mK, mDistCoef, Mat(), mK); k[5..11] of OpenCV's model are 0
__device__ __forceinline__ float2 undistort_point(const CamD& c, float u, float v) {
  double x = DM(DS((double)u, c.cx), c.ifx), y = DM(DS((double)v, c.cy), c.ify);
  const double x0 = x, y0 = y;
#pragma unroll 1
  for (int j = 0; j < 5; ++j) {
    const double r2 = DA(DM(x, x), DM(y, y));
    const double num = DA(1.0, DM(DA(DM(DA(DM(0.0, r2), 0.0), r2), 0.0), r2));
    const double den = DA(1.0, DM(DA(DM(DA(DM(c.k[4], r2), c.k[1]), r2), c.k[0]), r2));
    const double icdist = __ddiv_rn(num, den);
    if (icdist < 0) { x = x0; y = y0; break; }
    const double dX = DA(DA(DA(DM(DM(DM(2.0, c.k[2]), x), y), DM(c.k[3], DA(r2, DM(DM(2.0, x), x)))), DM(0.0, r2)), DM(DM(0.0, r2), r2));
    const double dY = DA(DA(DA(DM(c.k[2], DA(r2, DM(DM(2.0, y), y))), DM(DM(DM(2.0, c.k[3]), x), y)), DM(0.0, r2)), DM(DM(0.0, r2), r2));
    x = DM(DS(x0, dX), icdist);
    y = DM(DS(y0, dY), icdist);
  }
  const double xx = DA(DA(DM(c.fx, x), DM(0.0, y)), c.cx), yy = DA(DA(DM(0.0, x), DM(c.fy, y)), c.cy);
  const double ww = __ddiv_rn(1.0, DA(DA(DM(0.0, x), DM(0.0, y)), 1.0));
  return make_float2((float)DM(xx, ww), (float)DM(yy, ww));
}

__device__ __forceinline__ int grid_cell(float x, float y, const float* b) {
  const int gx = (int)roundf(__fmul_rn(__fsub_rn(x, b[0]), b[4])), gy = (int)roundf(__fmul_rn(__fsub_rn(y, b[2]), b[5]));
  return (gx < 0 || gx >= kGridCols || gy < 0 || gy >= kGridRows) ? -1 : gx * kGridRows + gy;
}

// one CTA per frame
__global__ void __launch_bounds__(256) k_undistort_grid(FrameGridArgs a) {
  __shared__ int s_cnt[kCells + 1];
  __shared__ float s_b[6];
  __shared__ int s_warp[8];
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = min(a.n[f], a.cap);
  const orbx_keypoint* k = a.k + (size_t)f * a.cap;
  orbx_keypoint* kun = a.kun + (size_t)f * a.cap;
  int* cell_start = a.cell_start + (size_t)f * (kCells + 1);
  int* items = a.items + (size_t)f * a.cap;
  const CamD cam = load_cam(a.cam);

  for (int c = tid; c <= kCells; c += 256) s_cnt[c] = 0;
  if (tid == 0) {                                              // ComputeImageBounds
    float b0 = 0.f, b1 = (float)a.cols, b2 = 0.f, b3 = (float)a.rows;
    if (cam.distorted) {
      const float2 c0 = undistort_point(cam, 0.f, 0.f), c1 = undistort_point(cam, (float)a.cols, 0.f);
      const float2 c2 = undistort_point(cam, 0.f, (float)a.rows), c3 = undistort_point(cam, (float)a.cols, (float)a.rows);
      b0 = fminf(c0.x, c2.x); b1 = fmaxf(c1.x, c3.x); b2 = fminf(c0.y, c1.y); b3 = fmaxf(c2.y, c3.y);
    }
    s_b[0] = b0; s_b[1] = b1; s_b[2] = b2; s_b[3] = b3;
    s_b[4] = __fdiv_rn((float)kGridCols, __fsub_rn(b1, b0));   // src/Frame.cc:101-102
    s_b[5] = __fdiv_rn((float)kGridRows, __fsub_rn(b3, b2));
    if (f == 0) for (int i = 0; i < 6; ++i) a.bounds[i] = s_b[i];
  }
  __syncthreads();
  // UndistortKeyPoints + per-cell counts
  for (int i = tid; i < n; i += 256) {
    orbx_keypoint kp = k[i];
    if (cam.distorted) { const float2 p = undistort_point(cam, kp.x, kp.y); kp.x = p.x; kp.y = p.y; }
    kun[i] = kp;
    const int c = grid_cell(kp.x, kp.y, s_b);
    if (c >= 0) atomicAdd(&s_cnt[c], 1);
  }
  __syncthreads();
  // exclusive scan of the 3072 counts: 12 consecutive cells per thread, then warp and block scans
  constexpr int kPer = kCells / 256;
  int local[kPer], sum = 0;
#pragma unroll
  for (int q = 0; q < kPer; ++q) { local[q] = s_cnt[tid * kPer + q]; sum += local[q]; }
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int base = incl - sum;
  for (int w = 0; w < warp; ++w) base += s_warp[w];
  __syncthreads();
#pragma unroll
  for (int q = 0; q < kPer; ++q) { s_cnt[tid * kPer + q] = base; cell_start[tid * kPer + q] = base; base += local[q]; }
  if (tid == 255) { s_cnt[kCells] = base; cell_start[kCells] = base; }
  __syncthreads();
  // fill (any order inside a cell), then restore index order cell by cell: lists are short (N / 3072 on average)
  for (int i = tid; i < n; i += 256) {
    const int c = grid_cell(kun[i].x, kun[i].y, s_b);
    if (c >= 0) items[atomicAdd(&s_cnt[c], 1)] = i;
  }
  __syncthreads();
  for (int c = tid; c < kCells; c += 256) {
    const int lo = cell_start[c], hi = s_cnt[c];                // s_cnt[c] ran up to the end of cell c
    for (int p = lo + 1; p < hi; ++p) {
      const int v = items[p];
      int q = p - 1;
      while (q >= lo && items[q] > v) { items[q + 1] = items[q]; --q; }
      items[q + 1] = v;
    }
  }
}

// one warp per query
__global__ void __launch_bounds__(256) k_features_in_area(AreaArgs a) {
  const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (q >= a.n_queries) return;
  const int f = a.q_frame[q];
  const orbx_keypoint* kun = a.kun + (size_t)f * a.cap;
  const int* cs = a.cell_start + (size_t)f * (kCells + 1);
  const int* items = a.items + (size_t)f * a.cap;
  int* out = a.out_idx + (size_t)q * a.qcap;
  const float x = a.q_x[q], y = a.q_y[q], r = a.q_r[q];
  const int minLevel = a.q_min_level[q], maxLevel = a.q_max_level[q];
  const float* b = a.bounds;
  int cnt = 0;
  const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, b[0]), r), b[4])));
  const int cx1 = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, b[0]), r), b[4])));
  const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, b[2]), r), b[5])));
  const int cy1 = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, b[2]), r), b[5])));
  if (cx0 < kGridCols && cx1 >= 0 && cy0 < kGridRows && cy1 >= 0) {
    const bool chk = minLevel > 0 || maxLevel >= 0;
    for (int ix = cx0; ix <= cx1; ++ix) {
      // cells (ix, cy0..cy1) are contiguous in the CSR and already in (iy, index) order
      const int e_lo = cs[ix * kGridRows + cy0], e_hi = cs[ix * kGridRows + cy1 + 1];
      for (int e0 = e_lo; e0 < e_hi; e0 += 32) {
        const int e = e0 + lane;
        bool take = false;
        int idx = 0;
        if (e < e_hi) {
          idx = items[e];
          const orbx_keypoint kp = kun[idx];
          take = fabsf(__fsub_rn(kp.x, x)) < r && fabsf(__fsub_rn(kp.y, y)) < r;
          if (chk && (kp.octave < minLevel || (maxLevel >= 0 && kp.octave > maxLevel))) take = false;
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, take);
        const int pos = cnt + __popc(bal & ((1u << lane) - 1u));
        if (take && pos < a.qcap) out[pos] = idx;
        cnt += __popc(bal);
      }
    }
  }
  if (lane == 0) a.out_cnt[q] = cnt;
}

int launch_undistort_grid(const FrameGridArgs& a, cudaStream_t st) {
  k_undistort_grid<<<a.n_frames, 256, 0, st>>>(a);
  return 1;
}
int launch_features_in_area(const AreaArgs& a, cudaStream_t st) {
  k_features_in_area<<<(a.n_queries + 7) / 8, 256, 0, st>>>(a);
  return 1;
}

}  // namespace orbx
