// ORACLE / test infrastructure: C wrapper around the UNMODIFIED reference front-end — src/ORBmatcher.cc, src/Frame.cc,
// src/KeyFrame.cc, src/MapPoint.cc, src/Map.cc, src/KeyFrameDatabase.cc, src/ORBextractor.cc and the vendored DBoW2 are
// compiled from /root/reference where they lie (oracle/Makefile target `ref` -> oracle/_ref/libref_orbslam.so) against
// the type-only stand-ins in oracle/ref_shim_cv/ (OpenCV types, boost::mutex, ros macros, a Converter.h that drops
// Eigen/g2o).  The OpenCV ROUTINES on the path (resize, copyMakeBorder, GaussianBlur, FAST, fastAtan2,
// undistortPoints) are served by the installed cv2 through oracle/ref_extractor.py.
//
// Objects are built by the reference's own constructors: Frame(im, t, extractor, voc, K, distCoef) extracts,
// undistorts, computes the image bounds and bins the grid (src/Frame.cc:59-127); KeyFrame(F, map, db) copies it
// (src/KeyFrame.cc:30-52); MapPoint(pos, refKF, map).  "valid" flags of the flat-array world become real MapPoint
// objects hung on mvpMapPoints.  Results come back as indices.
//
// Pins oracle/match_oracle.cc (M2-M5), oracle/frame_oracle.cc (N1) and oracle/bow_oracle.cc (N2 inside Frame::ComputeBoW)
// — tests/test_ref_matcher.py — and generates tests/golden/ref_match.npz.
#include <cstdint>
#include <cstring>
#include <map>
#include <new>
#include <set>
#include <string>
#include <vector>

#include <opencv/cv.h>

#include "Frame.h"
#include "KeyFrame.h"
#include "KeyFrameDatabase.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBVocabulary.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "Converter.h"

using namespace ORB_SLAM;

// the one Converter function the compiled sources call (src/Converter.cc:27-36: one row per descriptor)
std::vector<cv::Mat> Converter::toDescriptorVector(const cv::Mat& Descriptors) {
  std::vector<cv::Mat> vDesc;
  vDesc.reserve(Descriptors.rows);
  for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
  return vDesc;
}

namespace {

struct World {                       // what owns the object graph of one test scene
  ORBVocabulary* voc = nullptr;
  Map map;
  KeyFrameDatabase* db = nullptr;
  std::vector<MapPoint*> points;
  std::vector<KeyFrame*> keyframes;
  KeyFrame* anchor = nullptr;        // MapPoint's constructor wants a reference keyframe
  float inv_w = 0.f, inv_h = 0.f;    // grid element sizes of the first frame (see ref_m_frame_create)
};

struct FrameBox {
  World* w;
  Frame* f;
  std::map<MapPoint*, int> index;    // MapPoint -> keypoint index in this frame
};

cv::Mat mat_from(const float* v, int r, int c) {
  cv::Mat m(r, c, CV_32F);
  for (int i = 0; i < r; i++) for (int j = 0; j < c; j++) m.at<float>(i, j) = v[i * c + j];
  return m;
}

}  // namespace

extern "C" {

void* ref_m_world_create(const char* voc_path) {
  World* w = new World();
  w->voc = new ORBVocabulary();
  if (voc_path && *voc_path && !w->voc->loadFromTextFile(voc_path)) { delete w->voc; delete w; return nullptr; }
  w->db = new KeyFrameDatabase(*w->voc);
  return w;
}

// Frame::Frame(im, timeStamp, extractor, voc, K, distCoef).  reset_static: the image bounds / grid element sizes are
// statics computed for the first frame only (mbInitialComputations, src/Frame.cc:74-87)
void* ref_m_frame_create(void* world, void* extractor, const uint8_t* img, int width, int height, const float* K9,
                         const float* D5, int reset_static) {
  World* w = static_cast<World*>(world);
  cv::Mat im(height, width, CV_8UC1);
  for (int y = 0; y < height; y++) memcpy(im.ptr(y), img + (size_t)y * width, width);
  cv::Mat K = mat_from(K9, 3, 3), D = mat_from(D5, 5, 1);
  if (reset_static) Frame::mbInitialComputations = true;
  FrameBox* b = new FrameBox();
  b->w = w;
  // The reference's constructor sets mfGridElementWidthInv / HeightInv (plain float members) only while
  // mbInitialComputations is true, i.e. for the FIRST frame (src/Frame.cc:72-87); later frames leave them
  // uninitialised and read whatever the storage held.  In the running system every Frame is a temporary built in the
  // same stack slot and then copied (src/Tracking.cc: mCurrentFrame = Frame(...)), so the first frame's values
  // survive.  Same here: later frames are constructed in storage pre-loaded with those two values.
  void* mem = ::operator new(sizeof(Frame));
  memset(mem, 0, sizeof(Frame));
  if (!Frame::mbInitialComputations) {
    Frame* shell = static_cast<Frame*>(mem);
    memcpy(reinterpret_cast<char*>(&shell->mfGridElementWidthInv), &w->inv_w, 4);
    memcpy(reinterpret_cast<char*>(&shell->mfGridElementHeightInv), &w->inv_h, 4);
  }
  b->f = new (mem) Frame(im, 0.0, static_cast<ORBextractor*>(extractor), w->voc, K, D);
  if (reset_static) { w->inv_w = b->f->mfGridElementWidthInv; w->inv_h = b->f->mfGridElementHeightInv; }
  b->f->mTcw = cv::Mat::eye(4, 4, CV_32F);
  return b;
}

int ref_m_frame_n(void* fb) { return static_cast<FrameBox*>(fb)->f->N; }

void ref_m_frame_bounds(void* fb, int32_t bounds[4], float inv[2]) {
  Frame* f = static_cast<FrameBox*>(fb)->f;
  bounds[0] = Frame::mnMinX; bounds[1] = Frame::mnMaxX; bounds[2] = Frame::mnMinY; bounds[3] = Frame::mnMaxY;
  inv[0] = f->mfGridElementWidthInv; inv[1] = f->mfGridElementHeightInv;
}

// mvKeys (28-byte records), undistorted positions, descriptors
void ref_m_frame_read(void* fb, void* kps, float* xu, float* yu, uint8_t* desc) {
  Frame* f = static_cast<FrameBox*>(fb)->f;
  static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
  if (f->N) memcpy(kps, f->mvKeys.data(), (size_t)f->N * 28);
  for (int i = 0; i < f->N; i++) {
    xu[i] = f->mvKeysUn[i].pt.x; yu[i] = f->mvKeysUn[i].pt.y;
    memcpy(desc + (size_t)i * 32, f->mDescriptors.ptr(i), 32);
  }
}

// the grid as CSR in visiting order (ix outer, iy inner), as orbm_frameset_grid returns it
void ref_m_frame_grid(void* fb, int32_t* cell_start, int32_t* items) {
  Frame* f = static_cast<FrameBox*>(fb)->f;
  int k = 0;
  for (int ix = 0; ix < FRAME_GRID_COLS; ix++)
    for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) {
      cell_start[ix * FRAME_GRID_ROWS + iy] = k;
      for (size_t v : f->mGrid[ix][iy]) items[k++] = (int32_t)v;
    }
  cell_start[FRAME_GRID_COLS * FRAME_GRID_ROWS] = k;
}

int ref_m_features_in_area(void* fb, float x, float y, float r, int minl, int maxl, int32_t* out) {
  std::vector<size_t> v = static_cast<FrameBox*>(fb)->f->GetFeaturesInArea(x, y, r, minl, maxl);
  for (size_t i = 0; i < v.size(); i++) out[i] = (int32_t)v[i];
  return (int)v.size();
}

// ORBmatcher(nnratio, checkOri).SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
int ref_m_search_for_initialization(void* fb1, void* fb2, float* prev_xy, int32_t* matches12, int window, float ratio,
                                    int check_ori) {
  Frame* f1 = static_cast<FrameBox*>(fb1)->f;
  Frame* f2 = static_cast<FrameBox*>(fb2)->f;
  std::vector<cv::Point2f> prev(f1->N);
  for (int i = 0; i < f1->N; i++) prev[i] = cv::Point2f(prev_xy[2 * i], prev_xy[2 * i + 1]);
  std::vector<int> m12;
  ORBmatcher matcher(ratio, check_ori != 0);
  const int n = matcher.SearchForInitialization(*f1, *f2, prev, m12, window);
  for (int i = 0; i < f1->N; i++) { matches12[i] = m12[i]; prev_xy[2 * i] = prev[i].x; prev_xy[2 * i + 1] = prev[i].y; }
  return n;
}

// valid[i] != 0 -> keypoint i carries a (good) MapPoint; valid[i] == 2 -> a MapPoint whose isBad() is true
void ref_m_frame_set_valid(void* fb, const uint8_t* valid) {
  FrameBox* b = static_cast<FrameBox*>(fb);
  World* w = b->w;
  if (!w->anchor) {
    w->anchor = new KeyFrame(*b->f, &w->map, w->db);
    w->keyframes.push_back(w->anchor);
  }
  b->index.clear();
  for (int i = 0; i < b->f->N; i++) {
    b->f->mvpMapPoints[i] = nullptr;
    if (!valid[i]) continue;
    MapPoint* p = new MapPoint(cv::Mat::zeros(3, 1, CV_32F), w->anchor, &w->map);
    if (valid[i] == 2) p->SetBadFlag();
    w->points.push_back(p);
    b->f->mvpMapPoints[i] = p;
    b->index[p] = i;
  }
}

// WindowSearch(F1, F2, windowSize, vpMapPointMatches2, minScaleLevel, maxScaleLevel): matches21[i2] = i1 or -1
int ref_m_window_search(void* fb1, void* fb2, int window, int minl, int maxl, float ratio, int check_ori,
                        int32_t* matches21) {
  FrameBox* b1 = static_cast<FrameBox*>(fb1);
  FrameBox* b2 = static_cast<FrameBox*>(fb2);
  std::vector<MapPoint*> out;
  ORBmatcher matcher(ratio, check_ori != 0);
  const int n = matcher.WindowSearch(*b1->f, *b2->f, window, out, minl, maxl);
  for (int i = 0; i < b2->f->N; i++) matches21[i] = out[i] ? b1->index.at(out[i]) : -1;
  return n;
}

// Frame::ComputeBoW (src/Frame.cc:279-286) and its results in std::map order
void ref_m_frame_compute_bow(void* fb) { static_cast<FrameBox*>(fb)->f->ComputeBoW(); }

void ref_m_frame_bow(void* fb, int32_t* word, double* weight, int32_t* n_words, int32_t* node, int32_t* start,
                     int32_t* items, int32_t* n_nodes) {
  Frame* f = static_cast<FrameBox*>(fb)->f;
  int i = 0;
  for (DBoW2::BowVector::const_iterator it = f->mBowVec.begin(); it != f->mBowVec.end(); ++it, ++i) {
    word[i] = (int32_t)it->first; weight[i] = it->second;
  }
  *n_words = i;
  int k = 0, pos = 0;
  for (DBoW2::FeatureVector::const_iterator it = f->mFeatVec.begin(); it != f->mFeatVec.end(); ++it, ++k) {
    node[k] = (int32_t)it->first; start[k] = pos;
    for (size_t j = 0; j < it->second.size(); j++) items[pos++] = (int32_t)it->second[j];
  }
  start[k] = pos;
  *n_nodes = k;
}

// KeyFrame(F, map, db): takes the frame's keypoints, descriptors, MapPoints, BoW vectors and grid
void* ref_m_keyframe_create(void* fb) {
  FrameBox* b = static_cast<FrameBox*>(fb);
  KeyFrame* kf = new KeyFrame(*b->f, &b->w->map, b->w->db);
  b->w->keyframes.push_back(kf);
  return kf;
}

// SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches): matchesF[iF] = keypoint index in the keyframe, or -1.
// kf_box = the frame the keyframe was made from (for the MapPoint -> index map)
int ref_m_search_by_bow_kf_f(void* kf, void* kf_box, void* fb, float ratio, int check_ori, int32_t* matchesF) {
  FrameBox* bk = static_cast<FrameBox*>(kf_box);
  FrameBox* bf = static_cast<FrameBox*>(fb);
  std::vector<MapPoint*> out;
  ORBmatcher matcher(ratio, check_ori != 0);
  const int n = matcher.SearchByBoW(static_cast<KeyFrame*>(kf), *bf->f, out);
  for (int i = 0; i < bf->f->N; i++) matchesF[i] = out[i] ? bk->index.at(out[i]) : -1;
  return n;
}

// SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12): matches12[i1] = keypoint index in keyframe 2, or -1
int ref_m_search_by_bow_kf_kf(void* kf1, void* kf2, void* kf2_box, int n1, float ratio, int check_ori,
                              int32_t* matches12) {
  FrameBox* b2 = static_cast<FrameBox*>(kf2_box);
  std::vector<MapPoint*> out;
  ORBmatcher matcher(ratio, check_ori != 0);
  const int n = matcher.SearchByBoW(static_cast<KeyFrame*>(kf1), static_cast<KeyFrame*>(kf2), out);
  for (int i = 0; i < n1; i++) matches12[i] = out[i] ? b2->index.at(out[i]) : -1;
  return n;
}

// SearchByProjection(Frame& F, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:49-125), the local-map tracking search.
// Point k: a MapPoint whose descriptor is keypoint desc_idx[k] of the keyframe made from `desc_fb` (one observation ->
// ComputeDistinctiveDescriptors picks it), with the tracking fields isInFrustum would have written.
// matchesF[i] = index of the point newly hung on keypoint i of F, else -1.
#ifndef REF_DROPIN   // not part of include/shim/ORBmatcher.h (served through the flat views of orbm.h)
int ref_m_search_by_projection_local(void* fb, void* desc_fb, int npts, const float* u, const float* v,
                                     const int32_t* level, const float* view_cos, const uint8_t* in_view,
                                     const uint8_t* bad, const int32_t* desc_idx, float th, float ratio,
                                     int32_t* matchesF) {
  FrameBox* b = static_cast<FrameBox*>(fb);
  FrameBox* bd = static_cast<FrameBox*>(desc_fb);
  World* w = b->w;
  KeyFrame* kf = new KeyFrame(*bd->f, &w->map, w->db);
  w->keyframes.push_back(kf);
  std::vector<MapPoint*> pts(npts);
  std::map<MapPoint*, int> which;
  for (int k = 0; k < npts; k++) {
    MapPoint* p = new MapPoint(cv::Mat::zeros(3, 1, CV_32F), kf, &w->map);
    w->points.push_back(p);
    p->AddObservation(kf, desc_idx[k]);
    p->ComputeDistinctiveDescriptors();
    p->mbTrackInView = in_view[k] != 0;
    p->mTrackProjX = u[k]; p->mTrackProjY = v[k];
    p->mnTrackScaleLevel = level[k];
    p->mTrackViewCos = view_cos[k];
    if (bad[k]) p->SetBadFlag();
    pts[k] = p;
    which[p] = k;
  }
  ORBmatcher matcher(ratio, true);
  const int n = matcher.SearchByProjection(*b->f, pts, th);
  for (int i = 0; i < b->f->N; i++) {
    MapPoint* p = b->f->mvpMapPoints[i];
    std::map<MapPoint*, int>::iterator it = p ? which.find(p) : which.end();
    matchesF[i] = it == which.end() ? -1 : it->second;
  }
  return n;
}
#endif

// SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th) (src/ORBmatcher.cc:1507-1622) with the current
// pose left at identity, so that Rcw*x3Dw+tcw returns the world point unchanged (adding zeros is exact) and
// u = fx*x/z + cx is the float expression of :1535-1538 only.  pos: last.N x 3 world positions for last's MapPoints.
// matchesCur[i2] = index in the last frame, else -1 (only matches made by this call).
#ifndef REF_DROPIN   // not part of include/shim/ORBmatcher.h (served through the flat views of orbm.h)
int ref_m_search_by_projection_last(void* cur_fb, void* last_fb, const float* pos, float th, float ratio, int check_ori,
                                    int32_t* matchesCur) {
  FrameBox* bc = static_cast<FrameBox*>(cur_fb);
  FrameBox* bl = static_cast<FrameBox*>(last_fb);
  for (int i = 0; i < bl->f->N; i++) {
    MapPoint* p = bl->f->mvpMapPoints[i];
    if (!p) continue;
    cv::Mat x(3, 1, CV_32F);
    for (int k = 0; k < 3; k++) x.at<float>(k) = pos[3 * i + k];
    p->SetWorldPos(x);
  }
  ORBmatcher matcher(ratio, check_ori != 0);
  const int n = matcher.SearchByProjection(*bc->f, *bl->f, th);
  for (int i = 0; i < bc->f->N; i++) {
    MapPoint* p = bc->f->mvpMapPoints[i];
    std::map<MapPoint*, int>::iterator it = p ? bl->index.find(p) : bl->index.end();
    matchesCur[i] = it == bl->index.end() ? -1 : it->second;
  }
  return n;
}
#endif

// Geometry state of the MapPoint on keypoint i of `fb` (protected members; this file is compiled with
// -fno-access-control): world position, scale-invariance distances (what UpdateNormalAndDepth writes,
// src/MapPoint.cc:265-312) and viewing normal.  pos / normal: n x 3; min_dist / max_dist: n.  Call BEFORE the frame is
// turned into a KeyFrame or used as a point source.
void ref_m_frame_set_point_geometry(void* fb, const float* pos, const float* min_dist, const float* max_dist,
                                    const float* normal) {
  FrameBox* b = static_cast<FrameBox*>(fb);
  for (int i = 0; i < b->f->N; i++) {
    MapPoint* p = b->f->mvpMapPoints[i];
    if (!p) continue;
    cv::Mat x(3, 1, CV_32F), nrm(3, 1, CV_32F);
    for (int k = 0; k < 3; k++) { x.at<float>(k) = pos[3 * i + k]; nrm.at<float>(k) = normal ? normal[3 * i + k] : 0.f; }
    p->SetWorldPos(x);
    p->mNormalVector = nrm;
    p->mfMinDistance = min_dist[i];
    p->mfMaxDistance = max_dist[i];
  }
}

// Gives the MapPoint on keypoint i of `fb` the descriptor of that keypoint (one observation in `kf`, the keyframe made
// from fb -> ComputeDistinctiveDescriptors picks it, src/MapPoint.cc:185-250).
void ref_m_frame_points_take_descriptors(void* fb, void* kf) {
  FrameBox* b = static_cast<FrameBox*>(fb);
  for (int i = 0; i < b->f->N; i++) {
    MapPoint* p = b->f->mvpMapPoints[i];
    if (!p) continue;
    p->AddObservation(static_cast<KeyFrame*>(kf), i);
    p->ComputeDistinctiveDescriptors();
  }
}

// SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, th, ORBdist)
// (src/ORBmatcher.cc:1622-1746, relocalisation) with the current pose at identity: Rcw*x3Dw+tcw and x3Dw-Ow return the
// world point unchanged, so the float projection (:1652-1653), the bounds tests, dist3D/minDistance -> lower_bound level
// prediction (:1660-1667), the radius, the area query, the best-only scan and the rotation check are the reference's.
// already[i] != 0 -> keyframe MapPoint i is in sAlreadyFound.  matchesCur[i2] = keyframe keypoint index, else -1.
#ifndef REF_DROPIN   // not part of include/shim/ORBmatcher.h (served through the flat views of orbm.h)
int ref_m_search_by_projection_reloc(void* cur_fb, void* kf, void* kf_box, const uint8_t* already, float th,
                                     int orb_dist, float ratio, int check_ori, int32_t* matchesCur) {
  FrameBox* bc = static_cast<FrameBox*>(cur_fb);
  FrameBox* bk = static_cast<FrameBox*>(kf_box);
  std::set<MapPoint*> found;
  for (int i = 0; i < bk->f->N; i++)
    if (already[i] && bk->f->mvpMapPoints[i]) found.insert(bk->f->mvpMapPoints[i]);
  bc->f->mTcw = cv::Mat::eye(4, 4, CV_32F);
  ORBmatcher matcher(ratio, check_ori != 0);
  const int n = matcher.SearchByProjection(*bc->f, static_cast<KeyFrame*>(kf), found, th, orb_dist);
  for (int i = 0; i < bc->f->N; i++) {
    MapPoint* p = bc->f->mvpMapPoints[i];
    std::map<MapPoint*, int>::iterator it = p ? bk->index.find(p) : bk->index.end();
    matchesCur[i] = it == bk->index.end() ? -1 : it->second;
  }
  return n;
}
#endif

// SearchByProjection(KeyFrame* pKF, cv::Mat Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:286-407, loop closing)
// with Scw = identity (scw = 1).  vpPoints = the MapPoints of src_fb's keypoints sel[0..npts) (in that order);
// vpMatched starts as the MapPoints already on the keyframe's own frame kf_box (set_valid) plus, for pre[k] >= 0, point
// sel[k] pre-assigned to keyframe keypoint pre[k] ("already found").  matched[idx] = position k in sel of the point
// NEWLY matched to keyframe keypoint idx, else -1.
#ifndef REF_DROPIN   // not part of include/shim/ORBmatcher.h (served through the flat views of orbm.h)
int ref_m_search_by_projection_sim3(void* kf, void* kf_box, void* src_fb, const int32_t* sel, const int32_t* pre,
                                    int npts, int th, float ratio, int32_t* matched) {
  FrameBox* bk = static_cast<FrameBox*>(kf_box);
  FrameBox* bs = static_cast<FrameBox*>(src_fb);
  std::vector<MapPoint*> pts(npts), vm(bk->f->N);
  std::map<MapPoint*, int> which;
  for (int i = 0; i < bk->f->N; i++) vm[i] = bk->f->mvpMapPoints[i];
  for (int k = 0; k < npts; k++) {
    pts[k] = bs->f->mvpMapPoints[sel[k]];
    which[pts[k]] = k;
    if (pre[k] >= 0) vm[pre[k]] = pts[k];
  }
  std::vector<MapPoint*> before = vm;
  ORBmatcher matcher(ratio, true);
  const int n = matcher.SearchByProjection(static_cast<KeyFrame*>(kf), cv::Mat::eye(4, 4, CV_32F), pts, vm, th);
  for (int i = 0; i < bk->f->N; i++) matched[i] = (vm[i] && vm[i] != before[i]) ? which.at(vm[i]) : -1;
  return n;
}
#endif

// ORBmatcher::Fuse(KeyFrame* pKF, vector<MapPoint*>& vpMapPoints, float th) (src/ORBmatcher.cc:1016-1134), pKF's pose at
// identity.  vpMapPoints = the MapPoints on src_fb's keypoints sel[k] (each observes its own keypoint in src_kf, the
// keyframe made from src_fb: points_take_descriptors).  What the call leaves behind, per point k:
//   added[k]    keypoint of pKF the point now observes (the AddObservation branch, :1124-1128), else -1
//   went_bad[k] 1 if the point was good before and is bad now (the Replace branch, :1119-1122: pMP->Replace(pMPinKF))
//   repl[k]     after a Replace the point's old observation src_kf[sel[k]] holds pMPinKF: its keypoint index in pKF's
//               own frame kf_box; -1 when that slot was erased (pMPinKF already observed src_kf through an earlier
//               Replace, src/MapPoint.cc:148-151) or nothing was replaced
// Returns nFused.
#ifndef REF_DROPIN   // not part of include/shim/ORBmatcher.h (served through the flat views of orbm.h)
int ref_m_fuse(void* kf, void* kf_box, void* src_fb, void* src_kf, const int32_t* sel, int npts, float th,
               int32_t* added, uint8_t* went_bad, int32_t* repl) {
  KeyFrame* pKF = static_cast<KeyFrame*>(kf);
  KeyFrame* sKF = static_cast<KeyFrame*>(src_kf);
  FrameBox* bk = static_cast<FrameBox*>(kf_box);
  FrameBox* bs = static_cast<FrameBox*>(src_fb);
  std::vector<MapPoint*> pts(npts);
  std::vector<char> was_bad(npts);
  for (int k = 0; k < npts; k++) { pts[k] = bs->f->mvpMapPoints[sel[k]]; was_bad[k] = pts[k]->isBad(); }
  ORBmatcher matcher(0.6f, true);
  const int n = matcher.Fuse(pKF, pts, th);
  for (int k = 0; k < npts; k++) {
    added[k] = pts[k]->GetIndexInKeyFrame(pKF);
    went_bad[k] = (!was_bad[k] && pts[k]->isBad()) ? 1 : 0;
    repl[k] = -1;
    if (went_bad[k]) {
      MapPoint* now = sKF->GetMapPoint(sel[k]);
      std::map<MapPoint*, int>::iterator it = now ? bk->index.find(now) : bk->index.end();
      if (it != bk->index.end()) repl[k] = it->second;
    }
  }
  return n;
}
#endif

// ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, float th) (src/ORBmatcher.cc:1136-1265)
// with Scw = identity.  added[k] as above (:1253-1257); target_went_bad[idx] = 1 when the MapPoint on pKF's keypoint idx
// was good before and is bad now (pMPinKF->Replace(pMP), :1248-1251).  Returns nFused.
#ifndef REF_DROPIN   // not part of include/shim/ORBmatcher.h (served through the flat views of orbm.h)
int ref_m_fuse_sim3(void* kf, void* kf_box, void* src_fb, const int32_t* sel, int npts, float th, int32_t* added,
                    uint8_t* target_went_bad) {
  KeyFrame* pKF = static_cast<KeyFrame*>(kf);
  FrameBox* bk = static_cast<FrameBox*>(kf_box);
  FrameBox* bs = static_cast<FrameBox*>(src_fb);
  std::vector<MapPoint*> pts(npts);
  for (int k = 0; k < npts; k++) pts[k] = bs->f->mvpMapPoints[sel[k]];
  const int N = bk->f->N;
  std::vector<MapPoint*> before(N);
  std::vector<char> was_bad(N, 0);
  for (int i = 0; i < N; i++) { before[i] = pKF->GetMapPoint(i); was_bad[i] = before[i] ? before[i]->isBad() : 0; }
  ORBmatcher matcher(0.6f, true);
  const int n = matcher.Fuse(pKF, cv::Mat::eye(4, 4, CV_32F), pts, th);
  for (int k = 0; k < npts; k++) added[k] = pts[k]->GetIndexInKeyFrame(pKF);
  for (int i = 0; i < N; i++) target_went_bad[i] = (before[i] && !was_bad[i] && before[i]->isBad()) ? 1 : 0;
  return n;
}
#endif

// ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12 = 1, R12 = I, t12 = 0, th) (src/ORBmatcher.cc:1267-1505), both
// keyframe poses at identity.  pre12[i1] >= 0: vpMatches12[i1] starts as keyframe 2's MapPoint on keypoint pre12[i1]
// (already matched, :1301-1311).  out12[i1] = keypoint of keyframe 2 whose MapPoint vpMatches12[i1] holds afterwards
// (pre-set entries included), else -1.  Returns nFound.
#ifndef REF_DROPIN   // not part of include/shim/ORBmatcher.h (served through the flat views of orbm.h)
int ref_m_search_by_sim3(void* kf1, void* box1, void* kf2, void* box2, const int32_t* pre12, float th, int32_t* out12) {
  FrameBox* b1 = static_cast<FrameBox*>(box1);
  FrameBox* b2 = static_cast<FrameBox*>(box2);
  const int N1 = b1->f->N;
  std::vector<MapPoint*> m12(N1, static_cast<MapPoint*>(nullptr));
  for (int i = 0; i < N1; i++)
    if (pre12[i] >= 0) m12[i] = b2->f->mvpMapPoints[pre12[i]];
  ORBmatcher matcher(0.75f, true);
  const float s12 = 1.0f;
  const int n = matcher.SearchBySim3(static_cast<KeyFrame*>(kf1), static_cast<KeyFrame*>(kf2), m12, s12,
                                     cv::Mat::eye(3, 3, CV_32F), cv::Mat::zeros(3, 1, CV_32F), th);
  for (int i = 0; i < N1; i++) {
    std::map<MapPoint*, int>::iterator it = m12[i] ? b2->index.find(m12[i]) : b2->index.end();
    out12[i] = it == b2->index.end() ? -1 : it->second;
  }
  return n;
}
#endif

// ORBmatcher::SearchByProjection(Frame& F1, Frame& F2, int windowSize, vpMapPointMatches2) (src/ORBmatcher.cc:519-594),
// F2's pose at identity (Rc2w*x3Dw+tc2w returns the world point; u2 = fx*x*invz+cx is the float expression of :540-550).
// pos: F1.N x 3 world positions of F1's MapPoints.  matches21[i2] = i1 whose MapPoint was NEWLY assigned to F2's
// keypoint i2, else -1.
int ref_m_search_by_projection_ff(void* fb1, void* fb2, const float* pos, int window, float ratio, int check_ori,
                                  int32_t* matches21) {
  FrameBox* b1 = static_cast<FrameBox*>(fb1);
  FrameBox* b2 = static_cast<FrameBox*>(fb2);
  for (int i = 0; i < b1->f->N; i++) {
    MapPoint* p = b1->f->mvpMapPoints[i];
    if (!p) continue;
    cv::Mat x(3, 1, CV_32F);
    for (int k = 0; k < 3; k++) x.at<float>(k) = pos[3 * i + k];
    p->SetWorldPos(x);
  }
  b2->f->mTcw = cv::Mat::eye(4, 4, CV_32F);
  std::vector<MapPoint*> out;
  ORBmatcher matcher(ratio, check_ori != 0);
  const int n = matcher.SearchByProjection(*b1->f, *b2->f, window, out);
  for (int i = 0; i < b2->f->N; i++) {
    std::map<MapPoint*, int>::iterator it = (out[i] && out[i] != b2->f->mvpMapPoints[i]) ? b1->index.find(out[i]) : b1->index.end();
    matches21[i] = it == b1->index.end() ? -1 : it->second;
  }
  return n;
}

// KeyFrameDatabase (src/KeyFrameDatabase.cc): add(pKF) puts the keyframe into the inverted file (:40-46).
void ref_m_kfdb_add(void* world, void* kf) { static_cast<World*>(world)->db->add(static_cast<KeyFrame*>(kf)); }

// DetectRelocalisationCandidates(F) (:198-308) for the frame in `fb` against everything add()ed.  For each of the nkf
// keyframes in kfs: touched[k] = 1 when the call counted it (mnRelocQuery == F.mnId), words[k] = mnRelocWords,
// score[k] = mRelocScore (meaningful only above the cut), cand[k] = 1 when the keyframe is among the returned candidates.
// Returns the number of candidates.  The frame's mnId must not be 0 (keyframes start with mnRelocQuery == 0).
int ref_m_detect_reloc(void* world, void* fb, void** kfs, int nkf, uint8_t* touched, int32_t* words, float* score,
                       uint8_t* cand) {
  World* w = static_cast<World*>(world);
  Frame* F = static_cast<FrameBox*>(fb)->f;
  std::vector<KeyFrame*> r = w->db->DetectRelocalisationCandidates(F);
  std::set<KeyFrame*> rs(r.begin(), r.end());
  for (int k = 0; k < nkf; k++) {
    KeyFrame* p = static_cast<KeyFrame*>(kfs[k]);
    touched[k] = p->mnRelocQuery == F->mnId ? 1 : 0;
    words[k] = p->mnRelocWords; score[k] = p->mRelocScore; cand[k] = rs.count(p) ? 1 : 0;
  }
  return (int)r.size();
}

// DetectLoopCandidates(pKF, minScore) (:75-196); connected[k] != 0: keyframe k is made a covisibility neighbour of the
// query first (KeyFrame::AddConnection), which the search then skips (:95).  Outputs as above from mnLoopQuery /
// mnLoopWords / mLoopScore.
int ref_m_detect_loop(void* world, void* query_kf, void** kfs, int nkf, const uint8_t* connected, float min_score,
                      uint8_t* touched, int32_t* words, float* score, uint8_t* cand) {
  World* w = static_cast<World*>(world);
  KeyFrame* q = static_cast<KeyFrame*>(query_kf);
  for (int k = 0; k < nkf; k++)
    if (connected[k]) q->AddConnection(static_cast<KeyFrame*>(kfs[k]), 20);
  std::vector<KeyFrame*> r = w->db->DetectLoopCandidates(q, min_score);
  std::set<KeyFrame*> rs(r.begin(), r.end());
  for (int k = 0; k < nkf; k++) {
    KeyFrame* p = static_cast<KeyFrame*>(kfs[k]);
    touched[k] = p->mnLoopQuery == q->mnId ? 1 : 0;
    words[k] = p->mnLoopWords; score[k] = p->mLoopScore; cand[k] = rs.count(p) ? 1 : 0;
  }
  return (int)r.size();
}

int ref_m_frame_id(void* fb) { return (int)static_cast<FrameBox*>(fb)->f->mnId; }
int ref_m_keyframe_id(void* kf) { return (int)static_cast<KeyFrame*>(kf)->mnId; }

// Frame::isInFrustum(pMP, viewingCosLimit) (src/Frame.cc:136-197) of frame `view_fb` (pose = identity through the
// reference's own UpdatePoseMatrices, :129-134) for the MapPoint on every keypoint of `src_fb` (geometry from
// ref_m_frame_set_point_geometry).  Outputs per source keypoint: the return value and the tracking fields it writes
// (left at the values below for keypoints without a MapPoint).
void ref_m_is_in_frustum(void* view_fb, void* src_fb, float cos_limit, uint8_t* ret, uint8_t* in_view, float* u, float* v,
                         int32_t* level, float* view_cos) {
  Frame* f = static_cast<FrameBox*>(view_fb)->f;
  FrameBox* bs = static_cast<FrameBox*>(src_fb);
  f->mTcw = cv::Mat::eye(4, 4, CV_32F);
  f->UpdatePoseMatrices();
  for (int i = 0; i < bs->f->N; i++) {
    MapPoint* p = bs->f->mvpMapPoints[i];
    ret[i] = 0; in_view[i] = 0; u[i] = v[i] = view_cos[i] = 0.f; level[i] = 0;
    if (!p) continue;
    ret[i] = f->isInFrustum(p, cos_limit) ? 1 : 0;
    in_view[i] = p->mbTrackInView ? 1 : 0;
    if (p->mbTrackInView) { u[i] = p->mTrackProjX; v[i] = p->mTrackProjY; level[i] = p->mnTrackScaleLevel; view_cos[i] = p->mTrackViewCos; }
  }
}

// SearchForTriangulation(pKF1, pKF2, F12, vMatchedKeys1, vMatchedKeys2, vMatchedPairs) (src/ORBmatcher.cc:852-1014):
// match12[i1] = i2 from vMatchedPairs, else -1
#ifndef REF_DROPIN   // not part of include/shim/ORBmatcher.h (served through the flat views of orbm.h)
int ref_m_search_for_triangulation(void* kf1, void* kf2, const float* F12, int n1, float ratio, int check_ori,
                                   int32_t* match12) {
  cv::Mat F = mat_from(F12, 3, 3);
  std::vector<cv::KeyPoint> k1, k2;
  std::vector<std::pair<size_t, size_t> > pairs;
  ORBmatcher matcher(ratio, check_ori != 0);
  const int n = matcher.SearchForTriangulation(static_cast<KeyFrame*>(kf1), static_cast<KeyFrame*>(kf2), F, k1, k2, pairs);
  for (int i = 0; i < n1; i++) match12[i] = -1;
  for (size_t k = 0; k < pairs.size(); k++) match12[pairs[k].first] = (int32_t)pairs[k].second;
  return n;
}
#endif

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:185-250) for a MapPoint observed at keypoint idx[k] of
// keyframe kfs[k].  The reference walks its std::map<KeyFrame*, size_t> in POINTER order; `order` reports that order
// (positions into kfs/idx) so the caller can hand the oracle the descriptors in the same sequence.  out32 = mDescriptor.
void ref_m_distinctive_desc(void* world, void** kfs, const int32_t* idx, int nobs, uint8_t* out32, int32_t* order) {
  World* w = static_cast<World*>(world);
  MapPoint* p = new MapPoint(cv::Mat::zeros(3, 1, CV_32F), static_cast<KeyFrame*>(kfs[0]), &w->map);
  w->points.push_back(p);
  std::map<KeyFrame*, size_t> obs;
  for (int k = 0; k < nobs; k++) { p->AddObservation(static_cast<KeyFrame*>(kfs[k]), idx[k]); obs[static_cast<KeyFrame*>(kfs[k])] = k; }
  int pos = 0;
  for (std::map<KeyFrame*, size_t>::iterator it = obs.begin(); it != obs.end(); ++it) order[pos++] = (int32_t)it->second;
  p->ComputeDistinctiveDescriptors();
  cv::Mat d = p->GetDescriptor();
  memcpy(out32, d.ptr(0), 32);
}

}  // extern "C"
