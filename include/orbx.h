/*
 * orbx — C ABI of the B200 (sm_100a) ORB feature front end.
 *
 * This is the drop-in boundary for ONE path of ORB-SLAM2: ORBextractor::operator(),
 * Frame::ComputeStereoMatches and the ORBmatcher Hamming searches.  The reference has no FFI
 * layer; its boundary is the C++ class surface that Frame / Tracking call, so every entry point
 * below names the reference member it stands in for (paths relative to the reference tree).
 * include/orbx_cxx.hpp re-creates those C++ classes on top of this ABI; INTEGRATION.md shows
 * the binding a maintainer would add.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ or torch types cross this boundary
 *   - every call returns an orbx_status; nothing throws, asserts or exits
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream)
 *   - an extractor handle owns all device memory it needs (allocated once in create);
 *     extract/match calls allocate nothing
 *   - the ORBmatcher / Frame entry points take their arrays from host or from device memory (`on_device`).  Host arrays are
 *     staged through a pinned buffer and a device scratch buffer owned by the CALLING THREAD: they grow to the largest call
 *     that thread has made and are reused, so in steady state a call allocates nothing, uploads all inputs in one copy
 *     and returns when the results are in the caller's arrays.  With device arrays only kernel scratch comes from that buffer
 *   - a handle is a single-stream object: calls on one handle must not overlap; distinct
 *     handles may be used from different host threads (the reference runs its left and right
 *     extractor on two threads, src/Frame.cc:78-81)
 *   - there is no CPU fallback: without a CUDA device create() fails with ORBX_ERR_CUDA
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum orbx_status {
  ORBX_OK = 0,
  ORBX_ERR_BAD_ARG = 1,   /* null pointer, non-positive size, threshold outside [1,254], ... */
  ORBX_ERR_GEOMETRY = 2,  /* a pyramid level would be narrower than 62 px or round(W/H) == 0 or > 16
                             (the reference divides by zero there, src/ORBextractor.cc:543,784) */
  ORBX_ERR_CAPACITY = 3,  /* batch larger than max_batch, image larger than max_w x max_h,
                             or a caller buffer too small */
  ORBX_ERR_CUDA = 4,      /* CUDA runtime error, see orbx_last_error_string() */
  ORBX_ERR_STATE = 5      /* e.g. stereo matching before any extract on that slot */
} orbx_status;

/* ORBextractor constructor arguments, include/ORBextractor.h:51-52 / src/ORBextractor.cc:410 */
typedef struct orbx_params {
  int nfeatures;
  float scale_factor;
  int nlevels; /* 1..8 */
  int ini_th_fast;
  int min_th_fast;
} orbx_params;

/* Same 28-byte layout as cv::KeyPoint: pt.x, pt.y, size, angle, response, octave, class_id. */
typedef struct orbx_keypoint {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} orbx_keypoint;

typedef struct orbx_extractor orbx_extractor;

/* ---- extractor life cycle -------------------------------------------------------------- */

/* ORBextractor::ORBextractor (src/ORBextractor.cc:410-470).  Scale tables, per-level quotas and
 * the orientation disc are computed on the host with the reference's float/double arithmetic.
 * max_w/max_h bound the image size, max_batch the number of image slots held on the device
 * (a stereo batch of F frames uses 2F slots of one handle, or F slots of two handles). */
orbx_status orbx_extractor_create(const orbx_params* params, int max_w, int max_h, int max_batch,
                                  int device, orbx_extractor** out);
void orbx_extractor_destroy(orbx_extractor* ext);

/* GetLevels / GetScaleFactor(s) / GetInverseScaleFactors / GetScaleSigmaSquares /
 * GetInverseScaleSigmaSquares (include/ORBextractor.h:63-83); any pointer may be NULL.
 * quota = mnFeaturesPerLevel (src/ORBextractor.cc:436-446). */
orbx_status orbx_get_tables(const orbx_extractor* ext, int* nlevels, float* scale, float* inv_scale,
                            float* sigma2, float* inv_sigma2, int* quota);

/* The same tables from the constructor arguments alone, without an extractor handle or a device: the reference's
 * constructor fills them (src/ORBextractor.cc:414-446) and every Frame constructor reads them BEFORE the first
 * extraction (src/Frame.cc:69-75), so a stand-in class must be able to answer at once.  Any output may be NULL. */
orbx_status orbx_compute_tables(const orbx_params* params, int* nlevels, float* scale_factors, float* inv_scale_factors,
                                float* level_sigma2, float* inv_level_sigma2, int* features_per_level);

/* Upper bound on keypoints one image can return (nfeatures plus up to 3 per level, because the
 * quad-tree's last split may overshoot, src/ORBextractor.cc:730). Size caller buffers with it. */
int orbx_max_keypoints(const orbx_extractor* ext);

/* ---- extraction ------------------------------------------------------------------------ */

/* ORBextractor::operator()(image, mask, keypoints, descriptors), src/ORBextractor.cc:1043-1105,
 * for ONE host image, synchronous.  mask is ignored by the reference (include/ORBextractor.h:58)
 * and has no parameter here.  A NULL/empty image returns ORBX_OK with *n = 0 (silent return,
 * src/ORBextractor.cc:1046).  kps/desc are host buffers of `cap` entries (desc: cap x 32 bytes).
 * Results stay resident in slot 0 for orbx_stereo_match / orbx_download_level. */
orbx_status orbx_extract(orbx_extractor* ext, const uint8_t* img, int w, int h, int stride,
                         orbx_keypoint* kps, uint8_t* desc, int cap, int* n, void* stream);

/* The same path for `n` equally sized images in one go (slots first_slot .. first_slot+n-1),
 * asynchronous on `stream`.  imgs points at image 0; image i starts image_stride bytes after
 * image i-1, rows are `stride` bytes apart.  src_on_device: 0 = host memory (pinned for true
 * async; copied with cudaMemcpy2DAsync), 1 = device memory (read in place). */
orbx_status orbx_extract_batch(orbx_extractor* ext, const uint8_t* imgs, int n, int w, int h,
                               int stride, size_t image_stride, int src_on_device, int first_slot,
                               void* stream);

/* Copies results of slots [first_slot, first_slot+n) to caller memory (host or device, per
 * dst_on_device), asynchronous on `stream`: counts[i], kps[i*cap .. ], desc[(i*cap .. )*32].
 * Any of kps/desc/counts may be NULL.  cap must be >= orbx_max_keypoints(). */
orbx_status orbx_fetch_results(orbx_extractor* ext, int first_slot, int n, orbx_keypoint* kps,
                               uint8_t* desc, int* counts, int cap, int dst_on_device, void* stream);

/* orbx_fetch_results with the keypoints in 8 bytes instead of the 28 of cv::KeyPoint, for callers that fetch many frames
 * per call and are bound by the device -> host copy: packed[i*cap + j] = {x | y << 12 | response << 24 in level coordinates,
 * angle bits << 32} of keypoint j of slot first_slot + i, level_counts[i*8 + l] = keypoints of level l (the output is
 * level-major, src/ORBextractor.cc:1073-1103), descriptors as in orbx_fetch_results.  orbx_unpack_keypoints (host code, no
 * device needed) rebuilds the cv::KeyPoint records - pt scaled by mvScaleFactor[level] in float, size, octave, class_id -
 * byte for byte what orbx_fetch_results returns, plus counts[i] = the keypoint count of frame i. */
orbx_status orbx_fetch_results_packed(orbx_extractor* ext, int first_slot, int n, uint64_t* packed, uint8_t* desc,
                                      int* level_counts, int cap, int dst_on_device, void* stream);
orbx_status orbx_unpack_keypoints(const orbx_params* params, const uint64_t* packed, const int* level_counts, int n_frames,
                                  int cap, orbx_keypoint* kps, int kps_cap, int* counts);

/* Where the results of slots first_slot, first_slot + 1, ... live ON THE DEVICE, so that the matcher / Frame entry points
 * below can read them in place (on_device = 1) instead of being handed a copy: what operator() returned for a slot
 * (src/ORBextractor.cc:1043-1105) is kps_dev[(slot - first_slot) * cap + i], desc_dev[((slot - first_slot) * cap + i) * 32],
 * i < counts_dev[slot - first_slot], with *cap = orbx_max_keypoints().  Consecutive slots are a ready-made batch: e.g. the two
 * frames of ORBmatcher::SearchForInitialization extracted into slots 0 and 1 go to orbx_search_init_batch as
 * (kps_dev, desc_dev, counts_dev) and (kps_dev + cap, desc_dev + cap * 32, counts_dev + 1); orbx_undistort_grid_batch takes
 * kps_dev / counts_dev as they are.  The pointers stay valid until the handle is destroyed; their contents are those of
 * the last extraction into the slot (order your calls on one stream, or synchronise).  Any output may be NULL. */
orbx_status orbx_slot_results(const orbx_extractor* ext, int first_slot, const orbx_keypoint** kps_dev,
                              const uint8_t** desc_dev, const int** counts_dev, int* cap);

/* ORBextractor::mvImagePyramid[level] (public member, include/ORBextractor.h:85; read by
 * src/Frame.cc:473,563,575,580): device pointer to the level's inner ROI (w x h, rows `stride`
 * bytes apart; the 19-px reflect-101 border lies around it). */
orbx_status orbx_pyramid_level(const orbx_extractor* ext, int slot, int level,
                               const uint8_t** dev_ptr, int* w, int* h, int* stride);
/* Host copy of the bordered plane, tightly packed (w+38) x (h+38); synchronous. */
orbx_status orbx_download_level(orbx_extractor* ext, int slot, int level, uint8_t* dst);

/* Introspection for parity tests: the FAST candidates of one level in the order the reference
 * hands them to DistributeOctTree (src/ORBextractor.cc:820-825), packed x | y<<12 | response<<24
 * with x,y relative to the 16-px FAST border; and the post-quad-tree keypoints of one level in
 * list order (level coordinates).  Synchronous.  Returns the count in *n (may exceed cap). */
orbx_status orbx_debug_candidates(orbx_extractor* ext, int slot, int level, uint32_t* out, int cap, int* n);
orbx_status orbx_debug_level_keys(orbx_extractor* ext, int slot, int level, uint32_t* out, int cap, int* n);

/* ---- stereo ---------------------------------------------------------------------------- */

/* Frame::ComputeStereoMatches (src/Frame.cc:466-640) for n stereo pairs whose left images were
 * extracted into slots slotL0.. of extL and right images into slotR0.. of extR (extL may equal
 * extR).  bf and fx are Frame::mbf and Frame::fx; the search range is maxD = bf/(bf/fx)
 * (Frame::mb is read before it is assigned in the reference, src/Frame.cc:90 vs :114; it is
 * defined here as bf/fx).  Outputs per pair i, at [i*cap ..): mvuRight and mvDepth (-1 = no
 * match), n_matches[i] = surviving matches.  Asynchronous on `stream`.  Nothing is allocated:
 * the call works in scratch both handles own since create (SAD values in extL, the right keypoints
 * sorted by (octave, row) in extR), so calls that share a handle must not overlap on the GPU. */
orbx_status orbx_stereo_match_batch(orbx_extractor* extL, int slotL0, orbx_extractor* extR, int slotR0,
                                    int n, float bf, float fx, float* u_right, float* depth,
                                    int* n_matches, int cap, int dst_on_device, void* stream);

/* ---- ORBmatcher -------------------------------------------------------------------------- */

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663), host-side helper. */
int orbx_hamming256(const uint8_t* a, const uint8_t* b);

/* ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520) for n_pairs independent frame
 * pairs.  Pair p: frame 1 has n1[p] keypoints at kps1 + p*cap (desc1 likewise), frame 2 n2[p].
 * prev_xy (in/out) is vbPrevMatched as x,y pairs; matches12 receives vnMatches12 (-1 = none),
 * n_matches the return value.  grid_bounds[6] = frame 2's mnMinX, mnMaxX, mnMinY, mnMaxY,
 * mfGridElementWidthInv, mfGridElementHeightInv (src/Frame.cc:207-213; what orbx_undistort_grid_batch
 * returns as `bounds`; {0, cols, 0, rows, 64/cols, 48/rows} for an undistorted camera): a distorted
 * camera's undistorted keypoints may lie left of / above 0 and the 64x48 grid is laid over
 * [mnMinX, mnMaxX) x [mnMinY, mnMaxY) (PosInGrid :382-392, GetFeaturesInArea :327-380).
 * All pointers host or all device per on_device. */
orbx_status orbx_search_init_batch(const orbx_keypoint* kps1, const uint8_t* desc1, const int* n1,
                                   const orbx_keypoint* kps2, const uint8_t* desc2, const int* n2,
                                   int n_pairs, int cap, const float* grid_bounds, float* prev_xy,
                                   float nnratio, int window, int check_orientation,
                                   int* matches12, int* n_matches, int on_device, void* stream);

/* ORBmatcher::SearchByBoW(KeyFrame*,KeyFrame*,...) (src/ORBmatcher.cc:522-655) restricted to one
 * vocabulary bucket holding all features of each side in index order: brute-force best/second
 * best 256-bit Hamming, threshold (strict_lt: best < th as in the KF/KF form :598, else best <= th
 * as in the KF/Frame form :228), ratio test, greedy exclusion of matched B features, rotation
 * histogram.  Batched over n_pairs like orbx_search_init_batch; ang = keypoint angles.  th is 0..255
 * (the reference passes TH_LOW = 50); cap <= 6000. */
orbx_status orbx_knn_ratio_batch(const uint8_t* descA, const float* angA, const int* nA,
                                 const uint8_t* descB, const float* angB, const int* nB, int n_pairs,
                                 int cap, int th, float nnratio, int check_orientation, int strict_lt,
                                 int* matches12, int* n_matches, int on_device, void* stream);

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:45-129; the
 * local-map search of every tracked frame, called from src/Tracking.cc:1192) for n_frames independent
 * frames.  Frame f: n[f] keypoints at kps + f*cap (undistorted keypoints, descriptors, mvuRight),
 * grid_bounds[6] as in orbx_search_init_batch (the Frame's static bounds and inverse cell sizes),
 * scale_factors = mvScaleFactors.  Its m[f] map
 * points at index f*mcap carry what Frame::isInFrustum left in each MapPoint (src/Frame.cc:316-322):
 * mp_in_view (mbTrackInView && !isBad()), mp_proj_x/y/xr (mTrackProjX/Y/XR), mp_level
 * (mnTrackScaleLevel), mp_view_cos (mTrackViewCos), mp_desc (GetDescriptor()), mp_obs (Observations()).
 * frame_obs[f*cap + i] (in/out) is F.mvpMapPoints[i] on entry: -1 = none, else that point's
 * Observations(); assigned[f*cap + i] receives the index of the map point matched to keypoint i by this
 * call (-1 = none), n_matches[f] the return value.  nnratio/th as in the reference (0.8 and 1 or 3 at
 * src/Tracking.cc:1185-1192).  All pointers host or all device per on_device. */
orbx_status orbx_search_projection_batch(const orbx_keypoint* kps, const uint8_t* desc, const float* u_right,
                                         const int* n, int n_frames, int cap, const float* grid_bounds,
                                         const float* scale_factors, int nlevels, const uint8_t* mp_in_view,
                                         const float* mp_proj_x, const float* mp_proj_y, const float* mp_proj_xr,
                                         const int* mp_level, const float* mp_view_cos, const uint8_t* mp_desc,
                                         const int* mp_obs, const int* m, int mcap, float th, float nnratio,
                                         int* frame_obs, int* assigned, int* n_matches, int on_device,
                                         void* stream);

/* ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono) (src/ORBmatcher.cc:1328-1470;
 * the motion-model search, called from src/Tracking.cc:886,892) for n_frames independent frame pairs.
 * Current frame f as in orbx_search_projection_batch (kps/desc/u_right/n, grid_bounds, scale_factors); a
 * projection outside [mnMinX, mnMaxX] x [mnMinY, mnMaxY] is skipped (:1373-1376).
 * Last frame f: n_last[f] entries at index f*lcap: last_valid (mvpMapPoints[i] && !mvbOutlier[i]),
 * last_xw (the map point's world position, 3 floats), last_octave / last_angle (mvKeys[i].octave,
 * mvKeysUn[i].angle), mp_desc (GetDescriptor()), mp_obs (Observations()).  pose[f*12..] is
 * CurrentFrame.mTcw: Rcw row-major (9 floats) then tcw (3).  mode[f]: 0, 1 = bForward, 2 = bBackward
 * (the caller evaluates src/ORBmatcher.cc:1347-1348).  fx, fy, cx, cy, bf are the Frame's intrinsics.
 * frame_obs / assigned / n_matches as in orbx_search_projection_batch (assigned = index into the last
 * frame).  All pointers host or all device per on_device. */
orbx_status orbx_search_projection_frame_batch(const orbx_keypoint* kps, const uint8_t* desc, const float* u_right,
                                               const int* n, int n_frames, int cap, const float* grid_bounds,
                                               const float* scale_factors, int nlevels, const float* pose,
                                               const int* mode, float fx, float fy, float cx, float cy, float bf,
                                               const uint8_t* last_valid, const float* last_xw,
                                               const int* last_octave, const float* last_angle,
                                               const uint8_t* mp_desc, const int* mp_obs, const int* n_last,
                                               int lcap, float th, int check_orientation, int* frame_obs,
                                               int* assigned, int* n_matches, int on_device, void* stream);

/* ORBmatcher::SearchByBoW with the vocabulary buckets of DBoW2, both forms:
 *   SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)  src/ORBmatcher.cc:522-655: strict_lt = 1, validB = pMP2 && !isBad()
 *   SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) src/ORBmatcher.cc:159-288: strict_lt = 0, validB = all ones
 * for n_pairs independent pairs.  A side = the first keyframe: descA/angA (mvKeysUn[i].angle), validA[i] =
 * (pMP1 && !pMP1->isBad()), nodeA[i] = the FeatureVector node of feature i (the key under which
 * mFeatVec lists i; >= 0); B side likewise.  Features are visited in the reference's order (ascending node id,
 * then ascending index inside a node, Thirdparty/DBoW2/DBoW2/FeatureVector.cpp:28-45), candidates are the
 * still unmatched valid B features of the same node.  matches12[p*cap + i] = B index or -1 (for the frame
 * form vpMapPointMatches[matches12[i]] = vpMapPointsKF[i]); th is 0..255 (TH_LOW = 50), cap <= 4096. */
orbx_status orbx_search_by_bow_batch(const uint8_t* descA, const float* angA, const uint8_t* validA, const int* nodeA,
                                     const int* nA, const uint8_t* descB, const float* angB, const uint8_t* validB,
                                     const int* nodeB, const int* nB, int n_pairs, int cap, int th, float nnratio,
                                     int check_orientation, int strict_lt, int* matches12, int* n_matches,
                                     int on_device, void* stream);

/* ---- DBoW2 vocabulary transform ------------------------------------------ */

typedef struct orbx_vocabulary orbx_vocabulary;

/* A DBoW2 TemplatedVocabulary<FORB::TDescriptor, FORB> (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) copied to the
 * device as flat arrays.  n_nodes nodes, node 0 = root; node_desc = m_nodes[i].descriptor (32 bytes each);
 * the children of node i are children[child_first[i] .. child_first[i] + child_count[i]) in m_nodes[i].children
 * order (child_count 0 = leaf); node_word[i] / node_weight[i] = m_nodes[i].word_id / .weight (used for leaves).
 * L = m_L; weighting = m_weighting (0 TF_IDF, 1 TF, 2 IDF, 3 BINARY); norm = what
 * m_scoring_object->mustNormalize reports (0 no normalisation, 1 L1, 2 L2).  ORBvoc.txt: k 10, L 6, TF_IDF, L1. */
orbx_status orbx_vocabulary_create(const uint8_t* node_desc, const int* child_first, const int* child_count,
                                   const int* children, int n_children, const int* node_word,
                                   const double* node_weight, int n_nodes, int L, int weighting, int norm, int device,
                                   orbx_vocabulary** out);
void orbx_vocabulary_destroy(orbx_vocabulary* voc);

/* TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1127-1196) as Frame::ComputeBoW
 * and KeyFrame::ComputeBoW call it (src/Frame.cc:395-402, src/KeyFrame.cc:74-84; levelsup = 4) for n_frames frames:
 * frame f has n[f] descriptors at desc + f*cap*32.  word_of / node_of [f*cap + i] receive the word id of feature i
 * and its node at level L - levelsup (-1, -1 when the word is stopped, i.e. weight 0): mFeatVec is node_of read as
 * "node -> features in index order", the form orbx_search_by_bow_batch / orbx_search_triangulation_batch take.
 * mBowVec of frame f comes back as bow_count[f] (word, value) pairs at bow_word / bow_val + f*cap in ascending word
 * id, with the reference's double arithmetic (weights added per feature in feature order, normalised by the sum
 * taken in word order).  cap <= 8192.  All pointers host or all device per on_device. */
orbx_status orbx_bow_transform_batch(const orbx_vocabulary* voc, const uint8_t* desc, const int* n, int n_frames,
                                     int cap, int levelsup, int* word_of, int* node_of, int* bow_word,
                                     double* bow_val, int* bow_count, int on_device, void* stream);

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:657-823, with
 * CheckDistEpipolarLine :140-157; called from src/LocalMapping.cc:296) for n_pairs independent keyframe pairs.
 * Side 1 of pair p at index p*cap: kps1 = mvKeysUn, desc1, has_mp1[i] = (GetMapPoint(i) != NULL), u_right1 = mvuRight,
 * node1 = FeatureVector node per feature (as in orbx_search_by_bow_batch); side 2 likewise.  F12 + p*9 = the
 * fundamental matrix row-major; pose2 + p*12 = KF2's Rcw row-major then tcw; Cw + p*3 = KF1's camera centre;
 * cam2 + p*4 = fx, fy, cx, cy of KF2 (the epipole of :664-670 is computed from these);
 * scale_factors2 / level_sigma2_2 = KF2's mvScaleFactors / mvLevelSigma2 (nlevels entries).
 * matches12[p*cap + i] receives vMatches12[i] (the KF2 index or -1; vMatchedPairs is the list of (i, matches12[i])
 * with matches12[i] >= 0 in index order), n_matches[p] the return value.  cap <= 4096. */
orbx_status orbx_search_triangulation_batch(const orbx_keypoint* kps1, const uint8_t* desc1, const uint8_t* has_mp1,
                                            const float* u_right1, const int* node1, const int* n1,
                                            const orbx_keypoint* kps2, const uint8_t* desc2, const uint8_t* has_mp2,
                                            const float* u_right2, const int* node2, const int* n2, int n_pairs, int cap,
                                            const float* F12, const float* pose2, const float* Cw, const float* cam2,
                                            const float* scale_factors2, const float* level_sigma2_2, int nlevels,
                                            int only_stereo, int check_orientation, int* matches12, int* n_matches,
                                            int on_device, void* stream);

/* The search of ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:825-975; src/LocalMapping.cc:483-534) for n_kfs
 * independent (keyframe, candidate list) problems.  Keyframe p at index p*cap: kps_un = mvKeysUn, desc, u_right = mvuRight,
 * n[p] features; its feature grid as orbx_undistort_grid_batch returns it (grid_bounds[6], cell_start, cell_items) - or
 * cell_start = cell_items = NULL: the library then assigns the keypoints to cells itself with Frame::PosInGrid's rule and
 * grid_bounds (inside the kernel; on the host for capacities above ~3300 features, where device callers must pass the
 * grid).  The same holds for orbx_search_by_sim3_batch and orbx_search_projection_ordered_batch;
 * kf_bounds + p*4 = the KeyFrame's own mnMinX, mnMaxX, mnMinY, mnMaxY (ints in the reference, include/KeyFrame.h:185-188);
 * pose + p*12 = Rcw row-major then tcw; Ow + p*3 = GetCameraCenter(); fx..bf, scale_factors, inv_level_sigma2,
 * log_scale_factor = mfLogScaleFactor are the keyframe's.  Candidate k of problem p at index p*mcap + k: mp_valid =
 * (pMP && !pMP->isBad() && !pMP->IsInKeyFrame(pKF)), mp_xw = GetWorldPos(), mp_normal = GetNormal(), mp_min_dist /
 * mp_max_dist = mfMinDistance / mfMaxDistance (what Get{Min,Max}DistanceInvariance() return divided by 0.8 / 1.2),
 * mp_desc = GetDescriptor(); m[p] candidates.  best_idx[p*mcap + k] receives bestIdx when bestDist <= TH_LOW, else -1.
 * The map update of :944-966 (Replace / AddObservation + AddMapPoint, nFused++) stays with the caller, in candidate
 * order: the search of one candidate does not depend on the updates of the others.
 * sim3_variant = 1 selects the search of Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:977-1100, loop
 * closing): pose / Ow are then Rcw = sRcw/scw, tcw = Scw.col(3)/scw and Ow = -Rcw.t()*tcw as :986-991 compute them, mp_valid
 * = (!isBad() && !spAlreadyFound.count(pMP)), and there is no reprojection-error gate (u_right, inv_level_sigma2 unused). */
orbx_status orbx_fuse_search_batch(const orbx_keypoint* kps_un, const uint8_t* desc, const float* u_right, const int* n,
                                   int n_kfs, int cap, const float* grid_bounds, const int* cell_start,
                                   const int* cell_items, const float* kf_bounds, const float* pose, const float* Ow,
                                   float fx, float fy, float cx, float cy, float bf, const float* scale_factors,
                                   const float* inv_level_sigma2, int nlevels, float log_scale_factor,
                                   const uint8_t* mp_valid, const float* mp_xw, const float* mp_normal,
                                   const float* mp_min_dist, const float* mp_max_dist, const uint8_t* mp_desc,
                                   const int* m, int mcap, float th, int sim3_variant, int* best_idx, int on_device,
                                   void* stream);

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1102-1326; src/LoopClosing.cc:328)
 * for n_pairs independent keyframe pairs.  Keyframe 1 of pair p at index p*cap (kps1 = mvKeysUn, desc1, n1, its grid
 * cell_start1 / cell_items1, kf_bounds1, pose1 = Rcw then tcw), keyframe 2 likewise; grid_bounds, fx..cy (pKF1's, as the
 * reference uses them for both directions), scale_factors, log_scale_factor shared.  xf + p*24 = sR12 (9), t12 (3),
 * sR21 (9), t21 (3) exactly as :1119-1122 compute them.  The map points of keyframe 1 that take part (pMP && !isBad() &&
 * !vbAlreadyMatched1[i1]) are listed at index p*mcap: mp1_index = i1, mp1_xw, mp1_min_dist / mp1_max_dist (mfMinDistance /
 * mfMaxDistance), mp1_desc, m1[p] of them; keyframe 2 likewise with vbAlreadyMatched2.  matches12[p*cap + i1] receives the
 * feature index in keyframe 2 whose map point becomes vpMatches12[i1] (-1 = left as it was), n_found[p] the return value. */
orbx_status orbx_search_by_sim3_batch(const orbx_keypoint* kps1, const uint8_t* desc1, const int* n1, const int* cell_start1,
                                      const int* cell_items1, const float* kf_bounds1, const float* pose1,
                                      const orbx_keypoint* kps2, const uint8_t* desc2, const int* n2, const int* cell_start2,
                                      const int* cell_items2, const float* kf_bounds2, const float* pose2, int n_pairs,
                                      int cap, const float* grid_bounds, float fx, float fy, float cx, float cy,
                                      const float* scale_factors, int nlevels, float log_scale_factor, const float* xf,
                                      const int* mp1_index, const float* mp1_xw, const float* mp1_min_dist,
                                      const float* mp1_max_dist, const uint8_t* mp1_desc, const int* m1,
                                      const int* mp2_index, const float* mp2_xw, const float* mp2_min_dist,
                                      const float* mp2_max_dist, const uint8_t* mp2_desc, const int* m2, int mcap,
                                      float th, int* matches12, int* n_found, int on_device, void* stream);

/* The two SearchByProjection members that claim keypoints as they go, for n_problems independent (target, map point list)
 * problems:
 *   relocalisation = 0: SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:290-403, src/LoopClosing.cc:389):
 *     target = the keyframe, target_bounds = its mnMinX.. (ints), pose / Ow = Rcw, tcw, Ow as :298-303 compute them from Scw,
 *     occupied[i] = (vpMatched[i] != NULL), mp_valid = (!isBad() && !spAlreadyFound.count(pMP)), max_dist = TH_LOW;
 *   relocalisation = 1: SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (:1472-1599, src/Tracking.cc:1397,1412):
 *     target = the current frame, target_bounds = Frame::mnMinX.., pose = CurrentFrame.mTcw, Ow = -Rcw.t()*tcw, occupied[i] =
 *     (CurrentFrame.mvpMapPoints[i] != NULL), the list = pKF's map points in feature order with mp_valid = (pMP && !isBad() &&
 *     !sAlreadyFound.count(pMP)), mp_src_angle = pKF->mvKeysUn[i].angle, max_dist = ORBdist, rotation check per check_orientation.
 * Target p at index p*cap with its grid (orbx_undistort_grid_batch); map point k of problem p at index p*mcap + k: world position,
 * normal (relocalisation = 0 only), mfMinDistance / mfMaxDistance, descriptor.  Map points are taken in list order; assigned[p*cap + i]
 * receives the list position of the map point given to keypoint i by this call (-1 otherwise), n_matches[p] the return value. */
orbx_status orbx_search_projection_ordered_batch(const orbx_keypoint* kps_un, const uint8_t* desc, const int* n, int n_problems,
                                                 int cap, const float* grid_bounds, const int* cell_start, const int* cell_items,
                                                 const float* target_bounds, const float* pose, const float* Ow, float fx, float fy,
                                                 float cx, float cy, const float* scale_factors, int nlevels,
                                                 float log_scale_factor, const uint8_t* occupied, const uint8_t* mp_valid,
                                                 const float* mp_xw, const float* mp_normal, const float* mp_min_dist,
                                                 const float* mp_max_dist, const uint8_t* mp_desc, const float* mp_src_angle,
                                                 const int* m, int mcap, float th, int max_dist, int relocalisation,
                                                 int check_orientation, int* assigned, int* n_matches, int on_device,
                                                 void* stream);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-305) for n_points map points.  Map point m:
 * counts[m] observed descriptors (those of the keyframes that are not bad, in mObservations order) at
 * desc + m*kcap*32.  best_idx[m] receives the index of the descriptor with the least median Hamming distance
 * to all of them (the reference's sorted-row element [0.5*(N-1)], first index on ties; -1 when counts[m] == 0,
 * where the reference leaves mDescriptor alone); best_desc, if not NULL, receives that descriptor (32 bytes per
 * map point, untouched when counts[m] == 0).  kcap <= 1024. */
orbx_status orbx_distinctive_descriptors_batch(const uint8_t* desc, const int* counts, int n_points, int kcap,
                                               int* best_idx, uint8_t* best_desc, int on_device, void* stream);

/* ---- Frame construction after extraction --------------------------------- */

/* mK and mDistCoef of a Frame: fx, fy, cx, cy and OpenCV's k1, k2, p1, p2, k3 (src/Tracking.cc:57-87). */
typedef struct orbx_camera {
  float fx, fy, cx, cy, k1, k2, p1, p2, k3;
} orbx_camera;

#define ORBX_GRID_COLS 64 /* FRAME_GRID_COLS, include/Frame.h:38 */
#define ORBX_GRID_ROWS 48 /* FRAME_GRID_ROWS, include/Frame.h:37 */

/* Frame::UndistortKeyPoints (src/Frame.cc:404-434: cv::undistortPoints(mat, mat, mK, mDistCoef, Mat(), mK),
 * a plain copy when k1 == 0), Frame::ComputeImageBounds (src/Frame.cc:436-464, image size cols x rows) and
 * Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:230-245, 382-392) for n_frames frames of one camera.
 * Frame f: n[f] keypoints at kps + f*cap.  kps_un receives mvKeysUn in the same layout; bounds[6] receives
 * mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv; the grid comes back as
 * mGrid flattened: cell c = ix*48 + iy holds cell_items[f*cap + cell_start[f*3073 + c] ..
 * f*cap + cell_start[f*3073 + c + 1]) in the reference's push_back (index) order.  All pointers host or all
 * device per on_device. */
orbx_status orbx_undistort_grid_batch(const orbx_keypoint* kps, const int* n, int n_frames, int cap,
                                      const orbx_camera* camera, int cols, int rows, orbx_keypoint* kps_un,
                                      float* bounds, int* cell_start, int* cell_items, int on_device,
                                      void* stream);

/* Frame::GetFeaturesInArea(x, y, r, minLevel, maxLevel) (src/Frame.cc:327-380) for n_queries queries against
 * the frames prepared by orbx_undistort_grid_batch: query q asks frame q_frame[q].  out_idx[q*qcap ..] receives
 * the keypoint indices in the reference's order (cell column, cell row, index), out_cnt[q] their number; when
 * it exceeds qcap only the first qcap are written.  All pointers host or all device per on_device. */
orbx_status orbx_features_in_area_batch(const orbx_keypoint* kps_un, const int* n, int n_frames, int cap,
                                        const float* bounds, const int* cell_start, const int* cell_items,
                                        const int* q_frame, const float* q_x, const float* q_y, const float* q_r,
                                        const int* q_min_level, const int* q_max_level, int n_queries, int qcap,
                                        int* out_idx, int* out_cnt, int on_device, void* stream);

/* The brute-force search of orbx_knn_ratio_batch for ONE pair, for callers that must not
 * allocate.  It has no counterpart in the reference.  Some callers cannot afford or are not
 * allowed a CUDA allocation on this path: a tracking thread with a per-frame deadline
 * (cudaMalloc/cudaFree serialise on a device-wide lock and may stall behind another thread), a
 * process whose device memory is fully owned by a pool, a plug-in host that audits allocations.
 * orbx_knn_ratio_batch takes its scratch from cudaMallocAsync; this entry point calls no
 * cudaMalloc, cudaFree or any other allocating routine.  Its scratch is a fixed 20 KiB array that
 * is part of the library image: every GPU that loads liborbx.so holds it until the process exits.
 *   - all pointers are DEVICE pointers (inputs descA/angA/descB/angB, outputs matches12[nA] and
 *     n_matches[1]); the call is asynchronous on `stream`
 *   - at most ORBX_NOALLOC_MAX_FEATURES features per side, otherwise ORBX_ERR_CAPACITY; th is 0..255 as in
 *     orbx_knn_ratio_batch, otherwise ORBX_ERR_BAD_ARG; device indices 0..63
 *   - the first call on each device looks up the array's device address and raises the kernel's
 *     shared-memory limit; make that first call outside any stream capture
 *   - all calls on one GPU share the array (it keeps the last call's contents; the kernel
 *     re-clears what it needs zeroed): calls that overlap on the GPU race, so order them */
#define ORBX_NOALLOC_MAX_FEATURES 4096
orbx_status orbx_knn_ratio_noalloc(const uint8_t* descA, const float* angA, int nA, const uint8_t* descB,
                                   const float* angB, int nB, int th, float nnratio,
                                   int check_orientation, int strict_lt, int* matches12,
                                   int* n_matches, void* stream);

/* ---- misc -------------------------------------------------------------------------------- */

/* Per-stage device timing for roofline reports (no counterpart in the reference, which only
 * wall-clocks whole frames, Examples/Stereo/stereo_kitti.cc:80-97).  When enabled, extract_batch
 * and stereo_match_batch bracket each stage with CUDA events on the launching stream.
 * orbx_get_profile synchronises the device, adds up what was recorded since the last call and
 * resets: ms[s] = total milliseconds of stage s, calls[s] = number of recorded passes.
 * Stages: 0 pyramid (all level launches), 1 FAST cells, 2 quad-tree, 3 orientation + rBRIEF,
 * 4 stereo search+SAD, 5 stereo median cut, 6 Gaussian blur of the level planes. */
#define ORBX_N_STAGES 7
orbx_status orbx_set_profiling(orbx_extractor* ext, int enable);
orbx_status orbx_get_profile(orbx_extractor* ext, float* ms, int* calls);

/* cudaStreamSynchronize for callers that do not link the CUDA runtime themselves. */
orbx_status orbx_stream_synchronize(void* stream);
/* A non-blocking stream on `device` for such callers (no counterpart in the reference, which has no device).  The reference
 * runs its left and right extractor on two host threads (src/Frame.cc:78-81); two handles on two such streams run
 * concurrently on the GPU, whereas the legacy default stream (NULL) would serialise them. */
orbx_status orbx_stream_create(int device, void** stream);
void orbx_stream_destroy(void* stream);

/* Page-locked host memory for callers that do not link the CUDA runtime (no counterpart in the reference): images handed
 * to orbx_extract_batch from such a buffer are copied asynchronously at full PCIe rate (pageable memory is staged by the
 * driver and blocks the call).  write_combined = 1 asks for write-combined pages: faster for the GPU to read on some
 * hosts, very slow for the CPU to READ - only for upload buffers the CPU just fills. */
orbx_status orbx_host_alloc(size_t bytes, int write_combined, void** out);
void orbx_host_free(void* p);

/* Message of the last failing call on the calling thread ("" if none). */
const char* orbx_last_error_string(void);
/* Number of kernels launched by this library in the calling process so far (for bench.py). */
unsigned long long orbx_kernel_launch_count(void);
const char* orbx_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
