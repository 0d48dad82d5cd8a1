/* orbm.h — C ABI of the B200-native ORB descriptor matcher (part of liborbx.so).
 *
 * Drop-in boundary for the data-parallel part of ORB_SLAM::ORBmatcher (reference include/ORBmatcher.h:37-107,
 * src/ORBmatcher.cc) and the grid queries it calls (Frame::GetFeaturesInArea, src/Frame.cc:199-276;
 * KeyFrame::GetFeaturesInArea, src/KeyFrame.cc:612-652).  The reference passes Frame / KeyFrame / MapPoint objects;
 * here every routine takes flat arrays: the object graph (MapPoint pointers, isBad(), 3-D projection) stays with the
 * caller and is folded into per-keypoint `valid` flags and per-point (u, v, radius, level window) records.
 * There is NO CPU fallback: every entry point runs hand-written sm_100a kernels and returns ORBM_ERR_CUDA when no
 * usable device exists.
 *
 * Execution model: frames are uploaded once into a device-resident `orbm_frameset` (the 64x48 keypoint grid of
 * src/Frame.cc:108-122 is built on the device at upload); searches are BATCHED over lists of frame pairs, because a
 * single 1000-keypoint pair cannot fill one SM, let alone 148.  Candidate distances are computed in parallel; the
 * reference's loop-carried accept / steal / already-taken logic is replayed by an in-order commit pass, so results
 * equal the sequential reference bit for bit.
 *
 * Error convention: 0 on success, negative ORBM_ERR_* otherwise, text via orbm_last_error().  The reference
 * asserts (ROS_ASSERT, src/ORBmatcher.cc:240) instead.
 */
#ifndef ORBM_H_
#define ORBM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORBM_TH_HIGH = 100, ORBM_TH_LOW = 50, ORBM_HISTO_LENGTH = 30 }; /* src/ORBmatcher.cc:40-42 */
enum { ORBM_GRID_COLS = 64, ORBM_GRID_ROWS = 48 };                     /* include/Frame.h:35-36 */

enum {
  ORBM_OK = 0,
  ORBM_ERR_ARG = -1,
  ORBM_ERR_CUDA = -2,
  ORBM_ERR_UNSUPPORTED = -3,
  ORBM_ERR_OVERFLOW = -4,
  ORBM_ERR_COMM = -6 /* NCCL unavailable / failed (orbm_knn2_sharded only) */
};

/* What the matcher reads from a Frame / KeyFrame (all HOST pointers; copied at orbm_frameset_create). */
typedef struct orbm_frame {
  int32_t n;             /* number of keypoints, <= 65535 */
  const float* x;        /* mvKeysUn[i].pt.x */
  const float* y;        /* mvKeysUn[i].pt.y */
  const float* angle;    /* mvKeysUn[i].angle (degrees) */
  const int32_t* octave; /* mvKeysUn[i].octave */
  const uint8_t* desc;   /* mDescriptors: n x 32 bytes */
  const uint8_t* valid;  /* per routine, usually "mvpMapPoints[i] != NULL && !isBad()"; NULL = all 0 */
  float min_x, min_y;    /* Frame::mnMinX / mnMinY (src/Frame.cc:320-349) */
  float inv_w, inv_h;    /* mfGridElementWidthInv / mfGridElementHeightInv (src/Frame.cc:75-76) */
} orbm_frame;

/* DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>, Thirdparty/DBoW2/DBoW2/FeatureVector.h:21-22)
 * flattened: node ids strictly ascending; features of node k = items[start[k] .. start[k+1]). */
typedef struct orbm_featvec {
  int32_t nnodes;
  const int32_t* node_id;
  const int32_t* start; /* nnodes + 1 entries */
  const int32_t* items;
} orbm_featvec;

/* Projected map points for the SearchByProjection family (HOST pointers). */
typedef struct orbm_points {
  int32_t n;
  const float* u;            /* projection into the target image */
  const float* v;
  const float* radius;       /* search radius already scaled (e.g. r * mvScaleFactors[level], src/ORBmatcher.cc:72) */
  const int32_t* min_level;  /* level window handed to GetFeaturesInArea */
  const int32_t* max_level;
  const uint8_t* desc;       /* n x 32: MapPoint::GetDescriptor() / LastFrame.mDescriptors.row(i) */
  const float* angle;        /* keypoint angle of the point's source observation (rotation check); may be NULL */
  const uint8_t* valid;      /* point takes part (mbTrackInView, !isBad(), inside image ...); NULL = all 1 */
} orbm_points;

typedef struct orbm_handle orbm_handle;
typedef struct orbm_frameset orbm_frameset;

int orbm_create(int device, orbm_handle** out);
void orbm_destroy(orbm_handle* h);
const char* orbm_last_error(const orbm_handle* h); /* h == NULL: last failed orbm_create on this thread */
int64_t orbm_kernel_launches(const orbm_handle* h);
/* GPU milliseconds (CUDA events on the handle's stream, kernels only) of the last search / knn2 call. */
float orbm_last_kernel_ms(const orbm_handle* h);

/* ---- M1  ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1794-1810; same code: DBoW2 FORB::distance) --------
 * n independent pairs: dist[i] = Hamming(a + 32 i, b + 32 i).  HOST buffers. */
int orbm_descriptor_distance(orbm_handle* h, const uint8_t* a, const uint8_t* b, int n, int32_t* dist);

/* ---- frames on the device + M3 grid ------------------------------------------------------------------------- */
/* fvs may be NULL when no BoW search will be run on the set. */
int orbm_frameset_create(orbm_handle* h, const orbm_frame* frames, const orbm_featvec* fvs, int nframes,
                         orbm_frameset** out);
void orbm_frameset_destroy(orbm_frameset* s);

/* ---- N1: Frame glue on the device ---------------------------------------------------------------
 * What the Frame constructor does between the extractor call and the first search (src/Frame.cc:59-127):
 * UndistortKeyPoints (:288-318, cv::undistortPoints with K, D = (k1,k2,p1,p2,k3), P = K), ComputeImageBounds
 * (:320-349), grid element sizes (:75-76) and the 64x48 grid binning (:108-122, :266-276).  Values are the CV_32F
 * entries the reference stores (src/Tracking.cc:52-71). */
typedef struct orbm_camera {
  float fx, fy, cx, cy;
  float k1, k2, p1, p2, k3;
  int32_t width, height;   /* im.cols, im.rows */
} orbm_camera;
/* bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY}, inv = {mfGridElementWidthInv, mfGridElementHeightInv}.  Host only. */
int orbm_image_bounds(const orbm_camera* cam, int32_t bounds[4], float inv[2]);
/* Builds a frame set straight from the extractor's DEVICE outputs (orbx_extract_batch_device): d_kps
 * [nframes][kp_stride] 28-byte cv::KeyPoint records, d_desc [nframes][kp_stride][32], d_n [nframes] counts, all
 * written on `stream`.  Only the counts (4 bytes per frame) cross to the host.  valid flags start at 0, no feature
 * vectors. */
int orbm_frameset_from_extractor(orbm_handle* h, const void* d_kps, const uint8_t* d_desc, const int32_t* d_n,
                                 int nframes, int kp_stride, const orbm_camera* cam, void* stream,
                                 orbm_frameset** out);
/* Introspection: per-frame keypoint counts; the (undistorted) SoA columns of one frame (outputs may be NULL). */
int orbm_frameset_counts(orbm_frameset* s, int32_t* n_out);
int orbm_frameset_keypoints(orbm_frameset* s, int frame, float* x, float* y, float* angle, int32_t* octave,
                            uint8_t* desc);
/* Replaces the valid flags of one frame (MapPoint associations change between searches). */
int orbm_frameset_set_valid(orbm_frameset* s, int frame, const uint8_t* valid);
/* Sets the valid flag of EVERY keypoint of EVERY frame to value != 0 (asynchronous on the handle's stream). */
int orbm_frameset_fill_valid(orbm_frameset* s, int value);
/* The device-built grid as CSR in the reference's visiting order: cell = ix * 48 + iy (src/Frame.cc:232-258),
 * insertion order inside a cell.  cell_start: 3073 entries, items: n entries (count = cell_start[3072]). */
int orbm_frameset_grid(orbm_frameset* s, int frame, int32_t* cell_start, int32_t* items);
/* Frame::GetFeaturesInArea(x, y, r, minLevel, maxLevel) (src/Frame.cc:199-264) evaluated on the device;
 * minLevel = maxLevel = -1 gives KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:612-652). */
int orbm_features_in_area(orbm_frameset* s, int frame, float x, float y, float r, int min_level, int max_level,
                          int32_t* out, int cap, int* n_out);

/* ---- N2: BoW transform on the device ------------------------------------------------------------
 * Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:279-286, src/KeyFrame.cc) call
 * ORBVocabulary::transform(descriptors, BowVector, FeatureVector, levelsup = 4)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1194): every descriptor descends the k-ary vocabulary tree
 * (:1217-1259; FORB::distance = the 256-bit Hamming distance of M1, first strict minimum in child order), the words'
 * weights accumulate into the BowVector (BowVector.cpp:34-85, normalised as the scoring type asks) and the feature
 * index is filed under its ancestor `levelsup` levels above the leaves (FeatureVector.cpp:31-47).  Doubles are added
 * in the reference's order with individually rounded operations, so weights are BIT-exact.
 *
 * The vocabulary is handed over as flat arrays in the node numbering of loadFromTextFile (:1337-1424): node 0 is the
 * root, parent[i] < i, children of a node in ascending id = the reference's child order, word ids count the leaves in
 * id order.  scoring / weighting are the DBoW2 enums (BowVector.h:36-53).  Two defined behaviours where the
 * reference's is undefined: a blank line ends a text file (the reference turns the trailing newline into a node with
 * an uninitialised parent); a leaf above level L - levelsup is filed under itself (the reference leaves the NodeId
 * unset). */
typedef struct orbm_vocabulary orbm_vocabulary;
int orbm_vocabulary_create(orbm_handle* h, int k, int L, int scoring, int weighting, int nnodes,
                           const int32_t* parent, const uint8_t* desc /* nnodes x 32 */, const double* weight,
                           const uint8_t* is_leaf, orbm_vocabulary** out);
/* ORBVocabulary::loadFromTextFile (the ORBvoc.txt format). */
int orbm_vocabulary_load_text(orbm_handle* h, const char* path, orbm_vocabulary** out);
void orbm_vocabulary_destroy(orbm_vocabulary* v);
/* info = {k, L, scoring, weighting, number of words, number of nodes} */
int orbm_vocabulary_info(const orbm_vocabulary* v, int32_t info[6]);
/* transform() of ONE descriptor set given as a HOST buffer (n x 32).  Outputs in std::map order (ascending id):
 * bow_word / bow_weight [<= n], fv_node [<= n], fv_start [<= n + 1], fv_items [<= n]. */
int orbm_vocabulary_transform(orbm_vocabulary* v, const uint8_t* desc, int n, int levelsup, int32_t* bow_word,
                              double* bow_weight, int32_t* n_bow, int32_t* fv_node, int32_t* fv_start,
                              int32_t* fv_items, int32_t* n_nodes);
/* ComputeBoW for every frame of a device-resident set (descriptors never leave the GPU): replaces the set's feature
 * vectors (so orbm_search_by_bow can run on it) and stores the BowVectors. */
int orbm_frameset_compute_bow(orbm_frameset* s, const orbm_vocabulary* v, int levelsup);
/* Per-frame sizes after compute_bow / create-with-feature-vectors: words in the BowVector (0 when none was computed),
 * nodes and items of the FeatureVector.  Any output may be NULL. */
int orbm_frameset_bow_sizes(orbm_frameset* s, int32_t* n_words, int32_t* n_nodes, int32_t* n_items);
int orbm_frameset_bowvec(orbm_frameset* s, int frame, int32_t* word, double* weight);
int orbm_frameset_featvec(orbm_frameset* s, int frame, int32_t* node_id, int32_t* start, int32_t* items);
/* L1Scoring::score (ScoringObject.cpp:23-67) between BowVectors of frame pairs, summed in word order: bit-exact. */
int orbm_bow_score_l1(orbm_frameset* set1, orbm_frameset* set2, const int32_t* idx1, const int32_t* idx2, int npairs,
                      double* score);

/* ---- N3: Frame::isInFrustum batched (src/Frame.cc:136-197; Tracking::SearchReferencePointsInFrustum,
 * src/Tracking.cc:701-752) ------------------------------------------------------------------------------------
 * A view is what a Frame holds after UpdatePoseMatrices (src/Frame.cc:129-134): mRcw (row major), mtcw, mOw, the
 * intrinsics, the static image bounds (ints) and the scale pyramid.  Every (view, map point) pair is tested: positive
 * depth, projection inside the bounds, distance inside [min_dist, max_dist] (MapPoint::GetMin/MaxDistanceInvariance),
 * viewing cosine >= cos_limit (0.5 at the call site), predicted level = lower_bound(mvScaleFactors, dist/min_dist)
 * clamped to nlevels - 1.  Outputs are [nviews][n]: in_view = mbTrackInView, u/v = mTrackProjX/Y, level =
 * mnTrackScaleLevel, view_cos = mTrackViewCos (0 / -1 where in_view is 0).  Float results are BIT-exact: the cv::Mat
 * expressions are evaluated as OpenCV does (gemm small-matrix branch, double accumulation in norm / dot). */
typedef struct orbm_frustum_view {
  float Rcw[9], tcw[3], Ow[3];
  float fx, fy, cx, cy;
  int32_t min_x, max_x, min_y, max_y;
  int32_t nlevels;             /* 1..16 */
  float scale_factors[16];
} orbm_frustum_view;
int orbm_is_in_frustum(orbm_handle* h, const orbm_frustum_view* views, int nviews, const float* pos /* n x 3 */,
                       const float* normal /* n x 3 */, const float* min_dist, const float* max_dist, int n,
                       float cos_limit, uint8_t* in_view, float* u, float* v, int32_t* level, float* view_cos);

/* ---- N4: KeyFrameDatabase candidate scoring ------------------------------------------------------
 * KeyFrameDatabase::DetectRelocalisationCandidates (src/KeyFrameDatabase.cc:198-308) and ::DetectLoopCandidates
 * (:75-196) up to the point where they start walking the covisibility graph: for every query frame and every keyframe
 * of the database the number of shared vocabulary words (mnRelocWords / mnLoopWords, via the inverted file of
 * KeyFrameDatabase::add, :40-46), maxCommonWords over the candidates, minCommonWords = (int)(maxCommonWords * 0.8f)
 * (:231-238) and mpVoc->score(query, keyframe) - L1Scoring::score as the float the reference stores - for candidates
 * with more than minCommonWords shared words (:247-255).
 *   orbm_bowdb_create   builds the inverted file of a keyframe set that holds BowVectors (orbm_frameset_compute_bow);
 *                       the set must outlive the database
 *   orbm_bowdb_query    queries = frames qidx[0..nq) of a set with BowVectors.  exclude [nq][nkf] (may be NULL): != 0 =
 *                       the keyframe is no candidate for that query (DetectLoopCandidates drops the query's connected
 *                       keyframes, :95).  Outputs, HOST: common [nq][nkf] shared words (0 = none; counted for excluded
 *                       keyframes too), max_common / min_common [nq], score [nq][nkf] = the score where
 *                       common > min_common and not excluded, else -1.
 * Accumulating the scores over GetBestCovisibilityKeyFrames(10) and the 0.75 * best cut (:262-305) stay with the host. */
typedef struct orbm_bowdb orbm_bowdb;
int orbm_bowdb_create(orbm_handle* h, const orbm_frameset* keyframes, orbm_bowdb** out);
void orbm_bowdb_destroy(orbm_bowdb* d);
int orbm_bowdb_query(orbm_bowdb* d, const orbm_frameset* queries, const int32_t* qidx, int nq, const uint8_t* exclude,
                     int32_t* common, int32_t* max_common, int32_t* min_common, float* score);

/* ---- N4: MapPoint::ComputeDistinctiveDescriptors batched (src/MapPoint.cc:185-250) ---------------------------
 * Point p observes the descriptors desc[start[p] .. start[p+1]) (HOST buffers; the caller has already dropped bad
 * keyframes, :205-209).  best[p] = index inside the point's list of the descriptor with the least median Hamming
 * distance to the others (median = sorted row [(int)(0.5*(N-1))] with the self-distance 0 included, first strictly
 * smaller median wins, :231-243); -1 for a point without observations.  out_desc (npoints x 32, may be NULL) receives
 * that descriptor = the new mDescriptor (zeros where best = -1). */
int orbm_distinctive_descriptors(orbm_handle* h, const uint8_t* desc, const int32_t* start, int npoints,
                                 int32_t* best, uint8_t* out_desc);

/* ---- pair searches: pair p = (frame idx1[p] of set1, frame idx2[p] of set2) ----------------------------------
 * Per-pair outputs are rows of `stride` int32 (stride >= the largest frame of the written side); unused tail
 * entries are set to -1.  nmatches[p] = the routine's return value. */

/* M2  SearchForInitialization (src/ORBmatcher.cc:598-713): prev_xy [npairs][stride][2] = vbPrevMatched (in/out),
 * matches12 [npairs][stride] = vnMatches12. */
int orbm_search_for_initialization(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2,
                                   const int32_t* idx1, const int32_t* idx2, int npairs, float* prev_xy, int stride,
                                   int window, float ratio, int check_ori, int32_t* matches12, int32_t* nmatches);

/* M4  WindowSearch (src/ORBmatcher.cc:409-516): set1 valid = F1 keypoint has a good MapPoint.
 * matches21[p][i2] = i1 whose MapPoint is assigned to F2 keypoint i2. */
int orbm_window_search(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2, const int32_t* idx1,
                       const int32_t* idx2, int npairs, int window, float ratio, int check_ori, int min_level,
                       int max_level, int32_t* matches21, int stride, int32_t* nmatches);

/* M4  SearchByProjection(Frame&, Frame&, windowSize, ...) (src/ORBmatcher.cc:519-594): uv [npairs][stride][2] =
 * projection of F1's MapPoint i1 into F2; set1 valid = MapPoint present, good and not already found in F2;
 * set2 valid = F2 keypoint already has a MapPoint.  matches21 = NEW assignments only. */
int orbm_search_by_projection_frames(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2,
                                     const int32_t* idx1, const int32_t* idx2, int npairs, const float* uv,
                                     int stride, int window, float ratio, int32_t* matches21, int32_t* nmatches);

/* M6  projected-point candidate scoring, problem p = (points[p], frame idx[p] of set).
 *  variant ORBM_PROJ_LOCAL_MAP  src/ORBmatcher.cc:49-125: best + second with levels, <= TH_HIGH, ratio test only
 *                               when best and second share a level (`best > ratio * second` rejects)
 *  variant ORBM_PROJ_BEST       best only, accept iff best <= th_dist; rotation check when check_ori:
 *                               :1507-1620 (last frame, TH_HIGH), :1622-1746 (relocalisation, ORBdist),
 *                               :286-407 (Sim3 loop closing, TH_LOW, no orientation)
 *    For these two: set valid = keypoint already has a MapPoint (skipped as candidate), a keypoint accepted by an
 *    earlier point is skipped by the later ones; match_out[p][idx] = point index (row of the TARGET frame).
 *  variant ORBM_PROJ_BEST_INDEPENDENT  the best-only loops of Fuse (:1094-1113 and :1223-1242) and SearchBySim3
 *                               (:1373-1398, :1458-1483): every grid candidate inside the level window is scored,
 *                               keypoints that hold a MapPoint are NOT skipped, points do not see each other (two
 *                               points may pick the same keypoint: Fuse then replaces, :1116-1131), accept iff
 *                               best <= th_dist (TH_LOW in Fuse, TH_HIGH in SearchBySim3); no rotation check, ratio and
 *                               set valid unused.  match_out[p][k] = best keypoint of POINT k (row of the points,
 *                               stride >= points[p].n), nmatches[p] = accepted points (= nFused). */
enum { ORBM_PROJ_LOCAL_MAP = 0, ORBM_PROJ_BEST = 1, ORBM_PROJ_BEST_INDEPENDENT = 2 };
int orbm_search_by_projection(orbm_handle* h, const orbm_frameset* set, const int32_t* idx,
                              const orbm_points* points, int nproblems, int variant, int th_dist, float ratio,
                              int check_ori, int32_t* match_out, int stride, int32_t* nmatches);

/* M6  SearchBySim3 (src/ORBmatcher.cc:1267-1505), both directions + the agreement check (:1487-1502) for npairs keyframe
 * pairs (frames idx1[p], idx2[p] of one set).  points12[p]: one entry per keypoint of keyframe 1 (n == its keypoint
 * count) - its MapPoint projected into keyframe 2 with radius th * scale[level] and level window [pred-1, pred]
 * (:1334-1360), valid = MapPoint present, good, not already matched and inside keyframe 2's image / scale range;
 * points21[p] likewise for keyframe 2's points in keyframe 1.  match12[p][i1] = i2 where vnMatch1[i1] == i2 and
 * vnMatch2[i2] == i1, else -1 (stride >= keypoints of keyframe 1); nfound[p] = nFound. */
int orbm_search_by_sim3(orbm_handle* h, const orbm_frameset* set, const int32_t* idx1, const int32_t* idx2,
                        const orbm_points* points12, const orbm_points* points21, int npairs, int th_dist,
                        int32_t* match12, int stride, int32_t* nfound);

/* M5  SearchByBoW.  ORBM_BOW_KF_F: (KeyFrame*, Frame&) src/ORBmatcher.cc:155-284, set1 = keyframes (valid = good
 * MapPoint), match_out[p][idxF] = idxKF, accept best <= TH_LOW.  ORBM_BOW_KF_KF: (KeyFrame*, KeyFrame*) :715-850,
 * valid = good MapPoint on both sides, match_out[p][idx1] = idx2, accept best < TH_LOW. */
enum { ORBM_BOW_KF_F = 0, ORBM_BOW_KF_KF = 1 };
int orbm_search_by_bow(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2, const int32_t* idx1,
                       const int32_t* idx2, int npairs, int variant, float ratio, int check_ori, int32_t* match_out,
                       int stride, int32_t* nmatches);

/* M9  SearchForTriangulation (src/ORBmatcher.cc:852-1014) incl. CheckDistEpipolarLine (:136-153).
 * valid = keypoint already has a MapPoint (skipped, both sides).  F12 [npairs][9] row-major, sigma2 [nlevels] =
 * pKF2->GetSigma2(level).  match12[p][idx1] = idx2. */
int orbm_search_for_triangulation(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2,
                                  const int32_t* idx1, const int32_t* idx2, int npairs, const float* F12,
                                  const float* sigma2, int nlevels, int check_ori, int32_t* match12, int stride,
                                  int32_t* nmatches);

/* ---- M8  all-pairs 2-NN over a descriptor database (BASELINE config 5) ---------------------------------------
 * Semantics of the reference's shared inner loop (src/ORBmatcher.cc:201-222) without exclusions: strict `<`, the
 * lowest index wins ties, the second best is the multiset second minimum.  top2[2 q], top2[2 q + 1] =
 * (dist << 23) | (idx_base + index) keys of the best and second best (0xFFFFFFFF = none); idx_base + ndb <= 2^23.
 * Two implementations with bit-identical results, chosen by job size (ORBM_KNN_TC=0|1 at orbm_create forces one): jobs
 * of 4 M query x database pairs and more run on the tensor cores (knn2_tc_kernel: every bit a +-1 FP8 value, distance =
 * (256 - dot) / 2, sums exact), smaller ones on the XOR + POPC kernels. */
#define ORBM_KEY_NONE 0xFFFFFFFFu
#define ORBM_KEY_DIST(k) ((int)((k) >> 23))
#define ORBM_KEY_INDEX(k) ((int)((k) & 0x7FFFFFu))
int orbm_knn2(orbm_handle* h, const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int64_t idx_base,
              uint32_t* top2); /* HOST buffers */
int orbm_knn2_device(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb, int64_t idx_base,
                     uint32_t* d_top2, void* stream); /* DEVICE buffers, asynchronous */
/* orbm_knn2_device for callers that may not allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same result as orbm_knn2_device; the per-chunk partial keys and the error flag live in
 * a file-scope device array of ORBM_KNN2_STATIC_SCRATCH_BYTES (held by every GPU that loads the library until exit)
 * instead of the handle's grown-on-demand buffers, so the call issues no cudaMalloc / cudaFree / allocating routine.
 * Limit: 256 + nchunks * nq * 8 bytes must fit the array (nchunks = database chunks of the launch plan, about
 * 2 * SMs / ceil(nq / 1024)); on a 148-SM B200 that holds for every nq <= 200,000 whatever ndb is.  Beyond it the
 * call returns ORBM_ERR_UNSUPPORTED with a message and launches nothing.  All calls on one device share the array:
 * calls that overlap on the GPU (different streams, no ordering) race; calls on one stream are safe.  The first
 * call on a device looks the array's address up (cudaGetSymbolAddress) - make it once before capturing a graph. */
#define ORBM_KNN2_STATIC_SCRATCH_BYTES (4u << 20)
int orbm_knn2_device_static(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb,
                            int64_t idx_base, uint32_t* d_top2, void* stream);
/* Test hook, no device needed: the work split of the tensor-core kernel for nq queries x ndb descriptors on num_sms SMs.
 * out5 = {query-tile pairs, database tiles, clusters, tile steps per cluster, rows of the partial-result buffer}. */
int orbm_debug_knn2_tc_plan(int nq, int64_t ndb, int num_sms, int64_t* out5);
/* The same search for databases beyond the 2^23 indices of the 32-bit keys: d_top2 [nq][2] receives 64-bit keys
 * (dist << 40) | (idx_base + index), all ones = none; idx_base + ndb < 2^40.  The database is searched in segments of
 * 2^23 descriptors with the kernels above and merged on the device. */
#define ORBM_KEY64_NONE 0xFFFFFFFFFFFFFFFFull
#define ORBM_KEY64_DIST(k) ((int)((k) >> 40))
#define ORBM_KEY64_INDEX(k) ((int64_t)((k) & 0xFFFFFFFFFFull))
int orbm_knn2_device64(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb, int64_t idx_base,
                       uint64_t* d_top2, void* stream);
/* Waits for the handle's stream and reports what the asynchronous entry points above and below could not return: a
 * database tile or tensor-pipe signal that never arrived in knn2_kernel / knn2_tc_kernel (ORBM_ERR_CUDA; the affected queries hold ORBM_KEY_NONE keys) and an
 * asynchronous NCCL error on the communicator last used by orbm_knn2_sharded (ncclCommGetAsyncError, ORBM_ERR_COMM).
 * Work enqueued on a caller's stream must have been synchronised by the caller first.  The counterpart of
 * orbx_synchronize. */
int orbm_synchronize(orbm_handle* h);
/* Merge of per-shard results d_gathered [nshards][nq][2] (the all-gather receive buffer) into d_top2 [nq][2], plus
 * the accept rule of :224-226: d_match[q] = best index if best <= th_dist && (float)best < ratio * (float)second,
 * else -1 (d_match may be NULL). */
int orbm_merge_top2_device(orbm_handle* h, const uint32_t* d_gathered, int nshards, int nq, int th_dist, float ratio,
                           uint32_t* d_top2, int32_t* d_match, void* stream);
/* Database sharded over the ranks of an NCCL communicator (one process per GPU): local 2-NN over this rank's shard
 * (global indices = idx_base + local), ncclAllGather of the [nq][2] keys, merge.  comm = ncclComm_t.  d_gathered is
 * caller-provided scratch [nranks][nq][2].  NCCL is resolved at run time from the process (dlsym). */
/* Communicator helpers for callers without an NCCL dependency of their own: rank 0 creates the id, every rank
 * receives it out of band (e.g. one broadcast over the launcher's control channel) and joins. */
#define ORBM_COMM_ID_BYTES 128
int orbm_comm_unique_id(uint8_t id[ORBM_COMM_ID_BYTES]);
int orbm_comm_init_rank(orbm_handle* h, int nranks, const uint8_t id[ORBM_COMM_ID_BYTES], int rank, void** comm);
void orbm_comm_destroy(void* comm);
void orbm_comm_release(orbm_handle* h, void* comm); /* destroy + make orbm_synchronize(h) forget it */
int orbm_knn2_sharded(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db_shard, int64_t ndb_shard,
                      int64_t idx_base, int nranks, void* comm, uint32_t* d_gathered, int th_dist, float ratio,
                      uint32_t* d_top2, int32_t* d_match, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ORBM_H_ */
