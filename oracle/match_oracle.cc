// ORACLE — TEST INFRASTRUCTURE ONLY (never linked into, imported by or executed from the product path).
//
// CPU restatement of the reference's descriptor-matching hot path, worxli/ORB_SLAM src/ORBmatcher.cc, on flat
// arrays: the reference's Frame / KeyFrame / MapPoint object graph is replaced by per-keypoint arrays plus
// `valid` flags standing for "has a MapPoint that is not bad" (SURVEY.md §8a M1-M9).  Each function cites the
// reference lines it follows; loops, comparison operators and float conversions are restated literally
// (SURVEY Appendix A items 12-15).  Built by oracle/Makefile with -ffp-contract=off.
//
// Parity pin: the reference ships no tests or golden vectors for this path, and ORBmatcher.cc cannot
// be compiled here (OpenCV C++ / ROS / boost absent).  Pins available and used in tests/:
// DescriptorDistance vs bit counting, hand-made semantic cases derived from the source (ties, steal, thresholds),
// and an independent pure-Python restatement (tests/pyref_matcher.py) compared on random inputs.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <utility>
#include <vector>

namespace {

constexpr int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;  // src/ORBmatcher.cc:40-42
constexpr int GRID_COLS = 64, GRID_ROWS = 48;                 // include/Frame.h:35-36

}  // namespace

extern "C" {

// Flat view of what the matcher reads from a Frame / KeyFrame.
struct mo_frame {
  int32_t n;
  const float* x;         // mvKeysUn[i].pt.x
  const float* y;
  const float* angle;     // mvKeysUn[i].angle
  const int32_t* octave;  // mvKeysUn[i].octave
  const uint8_t* desc;    // mDescriptors, n x 32
  const uint8_t* valid;   // per routine: mvpMapPoints[i] != NULL && !isBad()  (may be NULL = all zero)
  float min_x, min_y;     // mnMinX, mnMinY (ints in the reference, include/Frame.h:121)
  float inv_w, inv_h;     // mfGridElementWidthInv / HeightInv (src/Frame.cc:75-76)
};

// DBoW2::FeatureVector = std::map<NodeId, std::vector<unsigned>> (Thirdparty/DBoW2/DBoW2/FeatureVector.h:21-22)
// flattened: node ids strictly ascending, items of node k = items[start[k] .. start[k+1]).
struct mo_featvec {
  int32_t nnodes;
  const int32_t* node_id;
  const int32_t* start;
  const int32_t* items;
};

}  // extern "C"

namespace {

// ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1794-1810 (literal SWAR popcount on 8 int32 words)
int descriptor_distance(const uint8_t* a, const uint8_t* b) {
  int32_t pa[8], pb[8];
  memcpy(pa, a, 32);
  memcpy(pb, b, 32);
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    unsigned int v = pa[i] ^ pb[i];
    v = v - ((v >> 1) & 0x55555555);
    v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
    dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
  }
  return dist;
}

struct Grid {
  std::vector<int> cell[GRID_COLS][GRID_ROWS];
};

// Frame::PosInGrid + the binning loop, src/Frame.cc:108-122, 266-276
void build_grid(const mo_frame& f, Grid* g) {
  for (int i = 0; i < f.n; i++) {
    const int posX = (int)roundf((f.x[i] - f.min_x) * f.inv_w);
    const int posY = (int)roundf((f.y[i] - f.min_y) * f.inv_h);
    if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
    g->cell[posX][posY].push_back(i);
  }
}

// Frame::GetFeaturesInArea, src/Frame.cc:199-264.  KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:612-652) is the
// same walk without a level filter (minLevel = maxLevel = -1 here).
std::vector<int> features_in_area(const mo_frame& f, const Grid& g, float x, float y, float r, int minLevel,
                                  int maxLevel) {
  std::vector<int> out;
  int nMinCellX = (int)floorf((x - f.min_x - r) * f.inv_w);
  nMinCellX = std::max(0, nMinCellX);
  if (nMinCellX >= GRID_COLS) return out;
  int nMaxCellX = (int)ceilf((x - f.min_x + r) * f.inv_w);
  nMaxCellX = std::min(GRID_COLS - 1, nMaxCellX);
  if (nMaxCellX < 0) return out;
  int nMinCellY = (int)floorf((y - f.min_y - r) * f.inv_h);
  nMinCellY = std::max(0, nMinCellY);
  if (nMinCellY >= GRID_ROWS) return out;
  int nMaxCellY = (int)ceilf((y - f.min_y + r) * f.inv_h);
  nMaxCellY = std::min(GRID_ROWS - 1, nMaxCellY);
  if (nMaxCellY < 0) return out;
  bool bCheckLevels = true, bSameLevel = false;
  if (minLevel == -1 && maxLevel == -1) bCheckLevels = false;
  else if (minLevel == maxLevel) bSameLevel = true;
  for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
    for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
      const std::vector<int>& vCell = g.cell[ix][iy];
      for (size_t j = 0; j < vCell.size(); j++) {
        const int k = vCell[j];
        if (bCheckLevels && !bSameLevel) {
          if (f.octave[k] < minLevel || f.octave[k] > maxLevel) continue;
        } else if (bSameLevel) {
          if (f.octave[k] != minLevel) continue;
        }
        if (fabsf(f.x[k] - x) > r || fabsf(f.y[k] - y) > r) continue;
        out.push_back(k);
      }
    }
  return out;
}

// ORBmatcher::ComputeThreeMaxima, src/ORBmatcher.cc:1748-1789
void compute_three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; i++) {
    const int s = (int)histo[i].size();
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

// the rotation-bin expression shared by all search routines, e.g. src/ORBmatcher.cc:234-241
int rot_bin(float a1, float a2) {
  const float factor = 1.0f / HISTO_LENGTH;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)roundf(rot * factor);
  if (bin == HISTO_LENGTH) bin = 0;
  return bin;
}

bool valid_at(const mo_frame& f, int i) { return f.valid && f.valid[i]; }

}  // namespace

extern "C" {

int mo_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

// N4: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:185-250) for a batch of map points.
// Point p observes descriptors desc[start[p] .. start[p+1]); best[p] = index (within the point's list) of the
// descriptor with the least median distance to the rest: row i sorted ascending (the 0 self-distance included),
// median = row[(int)(0.5 * (N - 1))], first strictly smaller median wins.  best[p] = -1 for a point without
// observations (the reference returns before touching mDescriptor).
void mo_distinctive_descriptors(const uint8_t* desc, const int32_t* start, int npoints, int32_t* best) {
  for (int p = 0; p < npoints; p++) {
    const int n = start[p + 1] - start[p];
    const uint8_t* d = desc + (size_t)start[p] * 32;
    best[p] = -1;
    if (n <= 0) continue;
    std::vector<float> dist((size_t)n * n);   // `float Distances[N][N]` in the reference; values are exact integers
    for (int i = 0; i < n; i++) {
      dist[(size_t)i * n + i] = 0;
      for (int j = i + 1; j < n; j++) {
        const int dij = descriptor_distance(d + (size_t)i * 32, d + (size_t)j * 32);
        dist[(size_t)i * n + j] = (float)dij;
        dist[(size_t)j * n + i] = (float)dij;
      }
    }
    int best_median = INT_MAX, best_idx = 0;
    for (int i = 0; i < n; i++) {
      std::vector<int> v(dist.begin() + (size_t)i * n, dist.begin() + (size_t)(i + 1) * n);
      std::sort(v.begin(), v.end());
      const int median = v[(size_t)(0.5 * (n - 1))];
      if (median < best_median) { best_median = median; best_idx = i; }
    }
    best[p] = best_idx;
  }
}

// Grid introspection for tests: CSR in the reference's visiting order (ix outer, iy inner, insertion order).
void mo_grid_csr(const mo_frame* f, int32_t* cell_start /*3073*/, int32_t* items /*n*/) {
  Grid g;
  build_grid(*f, &g);
  int k = 0;
  for (int ix = 0; ix < GRID_COLS; ix++)
    for (int iy = 0; iy < GRID_ROWS; iy++) {
      cell_start[ix * GRID_ROWS + iy] = k;
      for (int v : g.cell[ix][iy]) items[k++] = v;
    }
  cell_start[GRID_COLS * GRID_ROWS] = k;
}

int mo_features_in_area(const mo_frame* f, float x, float y, float r, int minLevel, int maxLevel, int32_t* out) {
  Grid g;
  build_grid(*f, &g);
  std::vector<int> v = features_in_area(*f, g, x, y, r, minLevel, maxLevel);
  for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
  return (int)v.size();
}

void mo_three_maxima(const int32_t* sizes, int L, int32_t* out3) {
  std::vector<std::vector<int>> h(L);
  for (int i = 0; i < L; i++) h[i].resize(sizes[i]);
  int a = -1, b = -1, c = -1;
  compute_three_maxima(h.data(), L, a, b, c);
  out3[0] = a; out3[1] = b; out3[2] = c;
}

// ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:598-713.  prev_xy = vbPrevMatched (in/out, n1 x 2),
// matches12 = vnMatches12 (out, n1).
int mo_search_for_initialization(const mo_frame* F1, const mo_frame* F2, float* prev_xy, int windowSize, float ratio,
                                 int checkOri, int32_t* vnMatches12) {
  int nmatches = 0;
  for (int i = 0; i < F1->n; i++) vnMatches12[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  std::vector<int> vMatchedDistance(F2->n, INT_MAX), vnMatches21(F2->n, -1);
  Grid g2;
  build_grid(*F2, &g2);
  for (int i1 = 0; i1 < F1->n; i1++) {
    const int level1 = F1->octave[i1];
    if (level1 > 0) continue;
    std::vector<int> vIndices2 =
        features_in_area(*F2, g2, prev_xy[2 * i1], prev_xy[2 * i1 + 1], (float)windowSize, level1, level1);
    if (vIndices2.empty()) continue;
    const uint8_t* d1 = F1->desc + 32 * (size_t)i1;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      const int dist = descriptor_distance(d1, F2->desc + 32 * (size_t)i2);
      if (vMatchedDistance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) bestDist2 = dist;
    }
    if (bestDist <= TH_LOW) {
      if (bestDist < (float)bestDist2 * ratio) {
        if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
        vnMatches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (checkOri) rotHist[rot_bin(F1->angle[i1], F2->angle[bestIdx2])].push_back(i1);
      }
    }
  }
  if (checkOri) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i])
        if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
    }
  }
  for (int i1 = 0; i1 < F1->n; i1++)
    if (vnMatches12[i1] >= 0) {
      prev_xy[2 * i1] = F2->x[vnMatches12[i1]];
      prev_xy[2 * i1 + 1] = F2->y[vnMatches12[i1]];
    }
  return nmatches;
}

// ORBmatcher::WindowSearch, src/ORBmatcher.cc:409-516.  F1.valid = "F1.mvpMapPoints[i1] && !isBad()".
// matches21[i2] = i1 whose MapPoint lands in vpMapPointMatches2[i2] (-1 = NULL).
int mo_window_search(const mo_frame* F1, const mo_frame* F2, int windowSize, float ratio, int checkOri,
                     int minScaleLevel, int maxScaleLevel, int32_t* matches21) {
  int nmatches = 0;
  for (int i = 0; i < F2->n; i++) matches21[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  const bool bMinLevel = minScaleLevel > 0, bMaxLevel = maxScaleLevel < INT_MAX;
  Grid g2;
  build_grid(*F2, &g2);
  for (int i1 = 0; i1 < F1->n; i1++) {
    if (!valid_at(*F1, i1)) continue;
    const int level1 = F1->octave[i1];
    if (bMinLevel && level1 < minScaleLevel) continue;
    if (bMaxLevel && level1 > maxScaleLevel) continue;
    std::vector<int> vIndices2 = features_in_area(*F2, g2, F1->x[i1], F1->y[i1], (float)windowSize, level1, level1);
    if (vIndices2.empty()) continue;
    const uint8_t* d1 = F1->desc + 32 * (size_t)i1;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      if (matches21[i2] >= 0) continue;
      const int dist = descriptor_distance(d1, F2->desc + 32 * (size_t)i2);
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) bestDist2 = dist;
    }
    if (bestDist <= bestDist2 * ratio && bestDist <= TH_HIGH) {
      matches21[bestIdx2] = i1;
      nmatches++;
      rotHist[rot_bin(F1->angle[i1], F2->angle[bestIdx2])].push_back(bestIdx2);
    }
  }
  if (checkOri) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (int i2 : rotHist[i]) { matches21[i2] = -1; nmatches--; }
  }
  return nmatches;
}

// ORBmatcher::SearchByProjection(Frame&F1, Frame&F2, windowSize, vpMapPointMatches2), src/ORBmatcher.cc:519-594.
// (u2,v2)[i1] = projection of F1's MapPoint i1 into F2 (the 3-D math :540-550 stays host-side).
// F1.valid = MapPoint present, not bad, not already found in F2; F2.valid = F2.mvpMapPoints[i2] != NULL (taken).
// matches21[i2] = i1 for the NEW assignments, -1 elsewhere.
int mo_search_by_projection_ff(const mo_frame* F1, const mo_frame* F2, const float* u2, const float* v2,
                               int windowSize, float ratio, int32_t* matches21) {
  int nmatches = 0;
  for (int i = 0; i < F2->n; i++) matches21[i] = -1;
  Grid g2;
  build_grid(*F2, &g2);
  for (int i1 = 0; i1 < F1->n; i1++) {
    if (!valid_at(*F1, i1)) continue;
    const int level1 = F1->octave[i1];
    std::vector<int> vIndices2 = features_in_area(*F2, g2, u2[i1], v2[i1], (float)windowSize, level1, level1);
    if (vIndices2.empty()) continue;
    const uint8_t* d1 = F1->desc + 32 * (size_t)i1;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      if (valid_at(*F2, i2) || matches21[i2] >= 0) continue;
      const int dist = descriptor_distance(d1, F2->desc + 32 * (size_t)i2);
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) bestDist2 = dist;
    }
    if ((float)bestDist <= (float)bestDist2 * ratio && bestDist <= TH_HIGH) {
      matches21[bestIdx2] = i1;
      nmatches++;
    }
  }
  return nmatches;
}

// Projected-point candidate scoring shared by the SearchByProjection family.  Points come with
// their image projection (u,v), search radius and level window already computed by the caller:
//   variant 0  local map  src/ORBmatcher.cc:49-125   best+second with levels, ratio only if same level, <= TH_HIGH
//   variant 1  best only  :1507-1620 (last frame, th_dist = TH_HIGH), :1622-1746 (relocalisation, th_dist =
//              ORBdist), :286-407 (Sim3, th_dist = TH_LOW, no orientation), Fuse / SearchBySim3 inner loops
// F.valid = F.mvpMapPoints[idx] != NULL on entry (taken).  p_valid = point takes part (mbTrackInView, !isBad ...).
// match_out[idx] = index of the point assigned to keypoint idx (-1 = none new).
int mo_search_by_projection(const mo_frame* F, int npoints, const float* pu, const float* pv, const float* pr,
                            const int32_t* pminl, const int32_t* pmaxl, const uint8_t* pdesc, const float* pangle,
                            const uint8_t* p_valid, int variant, int th_dist, float ratio, int checkOri,
                            int32_t* match_out) {
  int nmatches = 0;
  for (int i = 0; i < F->n; i++) match_out[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  Grid g;
  build_grid(*F, &g);
  for (int ip = 0; ip < npoints; ip++) {
    if (p_valid && !p_valid[ip]) continue;
    std::vector<int> vIdx = features_in_area(*F, g, pu[ip], pv[ip], pr[ip], pminl[ip], pmaxl[ip]);
    if (vIdx.empty()) continue;
    const uint8_t* dMP = pdesc + 32 * (size_t)ip;
    int bestDist = INT_MAX, bestLevel = -1, bestDist2 = INT_MAX, bestLevel2 = -1, bestIdx = -1;
    for (int idx : vIdx) {
      if (valid_at(*F, idx) || match_out[idx] >= 0) continue;
      const int dist = descriptor_distance(dMP, F->desc + 32 * (size_t)idx);
      if (dist < bestDist) {
        bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F->octave[idx]; bestIdx = idx;
      } else if (dist < bestDist2) { bestLevel2 = F->octave[idx]; bestDist2 = dist; }
    }
    if (variant == 0) {
      if (bestDist <= TH_HIGH) {
        if (bestLevel == bestLevel2 && bestDist > ratio * bestDist2) continue;
        match_out[bestIdx] = ip;
        nmatches++;
      }
    } else {
      if (bestDist <= th_dist) {
        match_out[bestIdx] = ip;
        nmatches++;
        if (checkOri) rotHist[rot_bin(pangle[ip], F->angle[bestIdx])].push_back(bestIdx);
      }
    }
  }
  if (variant != 0 && checkOri) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (int idx : rotHist[i]) { match_out[idx] = -1; nmatches--; }
  }
  return nmatches;
}

}  // extern "C"

namespace {

// The merge-join over two FeatureVectors (src/ORBmatcher.cc:176-259): equal ids -> body, else lower_bound of the
// lagging side.  On sorted unique ids lower_bound is "advance to the first id >= the other", so the shared nodes are
// visited in ascending id order.  Returns (k1, k2) pairs of node positions.
std::vector<std::pair<int, int>> shared_nodes(const mo_featvec& a, const mo_featvec& b) {
  std::vector<std::pair<int, int>> out;
  int i = 0, j = 0;
  while (i < a.nnodes && j < b.nnodes) {
    if (a.node_id[i] == b.node_id[j]) { out.push_back({i, j}); i++; j++; }
    else if (a.node_id[i] < b.node_id[j]) i = (int)(std::lower_bound(a.node_id, a.node_id + a.nnodes, b.node_id[j]) - a.node_id);
    else j = (int)(std::lower_bound(b.node_id, b.node_id + b.nnodes, a.node_id[i]) - b.node_id);
  }
  return out;
}

}  // namespace

extern "C" {

// ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches), src/ORBmatcher.cc:155-284.
// KF.valid = vpMapPointsKF[i] && !isBad().  matchF[idxF] = idxKF whose MapPoint is assigned (-1 = NULL).
int mo_search_by_bow_kf_f(const mo_frame* KF, const mo_frame* F, const mo_featvec* fvKF, const mo_featvec* fvF,
                          float ratio, int checkOri, int32_t* matchF) {
  for (int i = 0; i < F->n; i++) matchF[i] = -1;
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];
  for (auto [kKF, kF] : shared_nodes(*fvKF, *fvF)) {
    for (int a = fvKF->start[kKF]; a < fvKF->start[kKF + 1]; a++) {
      const int realIdxKF = fvKF->items[a];
      if (!valid_at(*KF, realIdxKF)) continue;
      const uint8_t* dKF = KF->desc + 32 * (size_t)realIdxKF;
      int bestDist1 = INT_MAX, bestIdxF = -1, bestDist2 = INT_MAX;
      for (int b = fvF->start[kF]; b < fvF->start[kF + 1]; b++) {
        const int realIdxF = fvF->items[b];
        if (matchF[realIdxF] >= 0) continue;
        const int dist = descriptor_distance(dKF, F->desc + 32 * (size_t)realIdxF);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
        else if (dist < bestDist2) bestDist2 = dist;
      }
      if (bestDist1 <= TH_LOW) {
        if ((float)bestDist1 < ratio * (float)bestDist2) {
          matchF[bestIdxF] = realIdxKF;
          if (checkOri) rotHist[rot_bin(KF->angle[realIdxKF], F->angle[bestIdxF])].push_back(bestIdxF);
          nmatches++;
        }
      }
    }
  }
  if (checkOri) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx : rotHist[i]) { matchF[idx] = -1; nmatches--; }
    }
  }
  return nmatches;
}

// ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12), src/ORBmatcher.cc:715-850.
// valid = MapPoint present and not bad, on both sides.  match12[idx1] = idx2 (-1 = NULL).  Note `< TH_LOW` (:791).
int mo_search_by_bow_kf_kf(const mo_frame* K1, const mo_frame* K2, const mo_featvec* fv1, const mo_featvec* fv2,
                           float ratio, int checkOri, int32_t* match12) {
  for (int i = 0; i < K1->n; i++) match12[i] = -1;
  std::vector<char> vbMatched2(K2->n, 0);
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  for (auto [k1, k2] : shared_nodes(*fv1, *fv2)) {
    for (int a = fv1->start[k1]; a < fv1->start[k1 + 1]; a++) {
      const int idx1 = fv1->items[a];
      if (!valid_at(*K1, idx1)) continue;
      const uint8_t* d1 = K1->desc + 32 * (size_t)idx1;
      int bestDist1 = INT_MAX, bestIdx2 = -1, bestDist2 = INT_MAX;
      for (int b = fv2->start[k2]; b < fv2->start[k2 + 1]; b++) {
        const int idx2 = fv2->items[b];
        if (vbMatched2[idx2] || !valid_at(*K2, idx2)) continue;
        const int dist = descriptor_distance(d1, K2->desc + 32 * (size_t)idx2);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
        else if (dist < bestDist2) bestDist2 = dist;
      }
      if (bestDist1 < TH_LOW) {
        if ((float)bestDist1 < ratio * (float)bestDist2) {
          match12[idx1] = bestIdx2;
          vbMatched2[bestIdx2] = 1;
          if (checkOri) rotHist[rot_bin(K1->angle[idx1], K2->angle[bestIdx2])].push_back(idx1);
          nmatches++;
        }
      }
    }
  }
  if (checkOri) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx : rotHist[i]) { match12[idx] = -1; nmatches--; }
    }
  }
  return nmatches;
}

// ORBmatcher::SearchForTriangulation, src/ORBmatcher.cc:852-1014, with CheckDistEpipolarLine :136-153.
// valid = "already has a MapPoint" on both sides (such keypoints are SKIPPED here).  F12 row-major 3x3,
// sigma2[level] = pKF2->GetSigma2(level).  match12[idx1] = idx2 (vMatches12).
int mo_search_for_triangulation(const mo_frame* K1, const mo_frame* K2, const mo_featvec* fv1, const mo_featvec* fv2,
                                const float* F12, const float* sigma2, int checkOri, int32_t* match12) {
  for (int i = 0; i < K1->n; i++) match12[i] = -1;
  std::vector<char> vbMatched2(K2->n, 0);
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  for (auto [k1, k2] : shared_nodes(*fv1, *fv2)) {
    for (int a = fv1->start[k1]; a < fv1->start[k1 + 1]; a++) {
      const int idx1 = fv1->items[a];
      if (valid_at(*K1, idx1)) continue;
      const uint8_t* d1 = K1->desc + 32 * (size_t)idx1;
      std::vector<std::pair<int, size_t>> vDistIndex;
      for (int b = fv2->start[k2]; b < fv2->start[k2 + 1]; b++) {
        const int idx2 = fv2->items[b];
        if (vbMatched2[idx2] || valid_at(*K2, idx2)) continue;
        const int dist = descriptor_distance(d1, K2->desc + 32 * (size_t)idx2);
        if (dist > TH_LOW) continue;
        vDistIndex.push_back(std::make_pair(dist, (size_t)idx2));
      }
      if (vDistIndex.empty()) continue;
      std::sort(vDistIndex.begin(), vDistIndex.end());
      const int BestDist = vDistIndex.front().first;
      const int DistTh = (int)round(2 * BestDist);
      for (size_t id = 0; id < vDistIndex.size(); id++) {
        if (vDistIndex[id].first > DistTh) break;
        const int currentIdx2 = (int)vDistIndex[id].second;
        // CheckDistEpipolarLine(kp1, kp2, F12, pKF2)
        const float x1 = K1->x[idx1], y1 = K1->y[idx1], x2 = K2->x[currentIdx2], y2 = K2->y[currentIdx2];
        const float ea = x1 * F12[0] + y1 * F12[3] + F12[6];
        const float eb = x1 * F12[1] + y1 * F12[4] + F12[7];
        const float ec = x1 * F12[2] + y1 * F12[5] + F12[8];
        const float num = ea * x2 + eb * y2 + ec;
        const float den = ea * ea + eb * eb;
        bool ok = false;
        if (den != 0) {
          const float dsqr = num * num / den;
          ok = dsqr < 3.84 * sigma2[K2->octave[currentIdx2]];
        }
        if (ok) {
          vbMatched2[currentIdx2] = 1;
          match12[idx1] = currentIdx2;
          nmatches++;
          if (checkOri) rotHist[rot_bin(K1->angle[idx1], K2->angle[currentIdx2])].push_back(idx1);
          break;
        }
      }
    }
  }
  if (checkOri) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx : rotHist[i]) { match12[idx] = -1; nmatches--; }
    }
  }
  return nmatches;
}

// All-pairs 2-NN (BASELINE config 5): the reference's shared inner loop (src/ORBmatcher.cc:201-222)
// run over the whole database, no exclusions: strict `<`, so the lowest index wins ties and the second best is the
// multiset second minimum.  Results as (dist << 23) | index keys, index = idx_base + position (UINT32_MAX = none).
void mo_knn2(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int64_t idx_base, uint32_t* top2) {
  for (int i = 0; i < nq; i++) {
    int bestDist1 = INT_MAX, bestDist2 = INT_MAX;
    int64_t bestIdx = -1, secondIdx = -1;
    for (int64_t j = 0; j < ndb; j++) {
      const int dist = descriptor_distance(q + 32 * (size_t)i, db + 32 * (size_t)j);
      if (dist < bestDist1) { bestDist2 = bestDist1; secondIdx = bestIdx; bestDist1 = dist; bestIdx = j; }
      else if (dist < bestDist2) { bestDist2 = dist; secondIdx = j; }
    }
    top2[2 * i] = bestIdx < 0 ? 0xFFFFFFFFu : ((uint32_t)bestDist1 << 23) | (uint32_t)(idx_base + bestIdx);
    top2[2 * i + 1] = secondIdx < 0 ? 0xFFFFFFFFu : ((uint32_t)bestDist2 << 23) | (uint32_t)(idx_base + secondIdx);
  }
}

}  // extern "C"

// ---- multi-threaded drivers for the timed CPU baseline (bench.py cpu_baseline / --impl reference) ------------------
#include <thread>

extern "C" void mo_knn2_mt(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int64_t idx_base, uint32_t* top2,
                           int nthreads) {
  nthreads = std::max(1, std::min(nthreads, nq));
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++) {
    const int a = (int)((int64_t)nq * t / nthreads), b = (int)((int64_t)nq * (t + 1) / nthreads);
    th.emplace_back([=] { mo_knn2(q + 32 * (size_t)a, b - a, db, ndb, idx_base, top2 + 2 * (size_t)a); });
  }
  for (auto& x : th) x.join();
}

// pairs (F1[p], F2[p]); prev_xy / matches12 rows of `stride`
extern "C" void mo_search_for_initialization_many(const mo_frame* F1, const mo_frame* F2, int npairs, float* prev_xy,
                                                  int stride, int windowSize, float ratio, int checkOri,
                                                  int32_t* matches12, int32_t* nmatches, int nthreads) {
  nthreads = std::max(1, std::min(nthreads, npairs));
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++) {
    th.emplace_back([=] {
      for (int p = t; p < npairs; p += nthreads)
        nmatches[p] = mo_search_for_initialization(&F1[p], &F2[p], prev_xy + (size_t)p * stride * 2, windowSize, ratio,
                                                   checkOri, matches12 + (size_t)p * stride);
    });
  }
  for (auto& x : th) x.join();
}
