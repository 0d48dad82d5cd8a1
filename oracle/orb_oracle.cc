// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product; never linked, imported or called by
// orb_slam_b200/ (only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg use it).
//
// CPU restatement of the reference extractor, worxli/ORB_SLAM src/ORBextractor.cc (+ include/ORBextractor.h),
// following the reference line by line.  The reference cannot be compiled here (needs OpenCV C++ headers and ROS,
// SURVEY.md §8c), so the OpenCV primitives it calls are restated as integer / float models, each one cross-checked
// bit-for-bit against cv2 4.13 by tests/test_oracle_vs_cv2.py (resize, FAST+NMS, Gaussian blur, fastAtan2,
// copyMakeBorder) and KeyPointsFilter::retainBest calls the REAL libstdc++ std::nth_element / std::partition.
// Parity status: "pinned to cv2 4.13 primitives + source-derived KATs" (the reference itself ships no tests or
// golden vectors — SURVEY.md §4).
//
// Build: see oracle/Makefile (g++ -O2 -ffp-contract=off for the canonical checker; -O3 -march=native
// -ffp-contract=off for the timed CPU baseline).
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

// ---------------------------------------------------------------------------------------------------------------
// OpenCV scalar helpers
// ---------------------------------------------------------------------------------------------------------------
// cvRound: round-half-to-even (SSE2 cvtsd2si / lrint under the default rounding mode).
inline int cvRoundD(double v) { return (int)std::nearbyint(v); }
inline int cvRoundF(float v) { return (int)std::nearbyintf(v); }
inline int cvFloorD(double v) { int i = (int)v; return i - (i > v); }
inline int cvCeilD(double v) { int i = (int)v; return i + (i < v); }

struct KeyPoint {  // == cv::KeyPoint memory layout (28 bytes), include/ORBextractor.h:43-45
  float x, y, size, angle, response;
  int32_t octave, class_id;
};

// reflect-101 index map used by copyMakeBorder(BORDER_REFLECT_101): ... 2 1 | 0 1 2 ... n-1 | n-2 n-3 ...
inline int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) {
    if (i < 0) i = -i;
    else i = 2 * (n - 1) - i;
  }
  return i;
}

// ---------------------------------------------------------------------------------------------------------------
// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) for CV_8UC1 — 11-bit fixed-point model.
// ---------------------------------------------------------------------------------------------------------------
void resize_linear_8u(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh,
                      size_t dstride) {
  const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
  std::vector<int> xofs(dw), yofs(dh);
  std::vector<short> ialpha(dw * 2), ibeta(dh * 2);
  for (int dx = 0; dx < dw; dx++) {
    float fx = (float)((dx + 0.5) * scale_x - 0.5);
    int sx = cvFloorD(fx);
    fx -= sx;
    if (sx < 0) { fx = 0; sx = 0; }
    if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
    xofs[dx] = sx;
    ialpha[dx * 2] = (short)cvRoundF((1.f - fx) * 2048.f);
    ialpha[dx * 2 + 1] = (short)cvRoundF(fx * 2048.f);
  }
  for (int dy = 0; dy < dh; dy++) {
    float fy = (float)((dy + 0.5) * scale_y - 0.5);
    int sy = cvFloorD(fy);
    fy -= sy;
    yofs[dy] = sy;  // no weight clamp vertically: the row INDICES are clipped below, the weights stay
    ibeta[dy * 2] = (short)cvRoundF((1.f - fy) * 2048.f);
    ibeta[dy * 2 + 1] = (short)cvRoundF(fy * 2048.f);
  }
  std::vector<int> row0(dw), row1(dw);
  for (int dy = 0; dy < dh; dy++) {
    // rows sy and sy+1 are clipped into [0, sh-1] (resizeGeneric_Invoker: clip(sy0 + k, 0, ssize.height))
    int sy0 = std::min(std::max(yofs[dy], 0), sh - 1);
    int sy1 = std::min(std::max(yofs[dy] + 1, 0), sh - 1);
    const uint8_t* S0 = src + (size_t)sy0 * sstride;
    const uint8_t* S1 = src + (size_t)sy1 * sstride;
    for (int dx = 0; dx < dw; dx++) {
      int sx = xofs[dx];
      int a0 = ialpha[dx * 2], a1 = ialpha[dx * 2 + 1];
      int sx1 = sx + 1 < sw ? sx + 1 : sx;  // a1 == 0 whenever sx == sw-1
      row0[dx] = S0[sx] * a0 + S0[sx1] * a1;
      row1[dx] = S1[sx] * a0 + S1[sx1] * a1;
    }
    int b0 = ibeta[dy * 2], b1 = ibeta[dy * 2 + 1];
    uint8_t* D = dst + (size_t)dy * dstride;
    for (int dx = 0; dx < dw; dx++)
      D[dx] = (uint8_t)((((b0 * (row0[dx] >> 4)) >> 16) + ((b1 * (row1[dx] >> 4)) >> 16) + 2) >> 2);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// cv::FAST(img, kps, threshold, nonmaxSuppression=true)  (TYPE_9_16) — SURVEY §0.3 P4.
// Scores pixels >= 3 px inside the sub-image it is handed; 3x3 strict NMS inside that region; row-major output.
// ---------------------------------------------------------------------------------------------------------------
const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// threshold-free corner score: max over the 16 arcs of 9 contiguous ring pixels of min(d) resp. min(-d), minus 1
inline int fast_score_9_16(const uint8_t* p, size_t stride) {
  int d[25];
  int v = p[0];
  for (int k = 0; k < 16; k++) d[k] = v - p[(ptrdiff_t)kRingDy[k] * (ptrdiff_t)stride + kRingDx[k]];
  for (int k = 16; k < 25; k++) d[k] = d[k - 16];
  int best = INT_MIN;
  for (int k = 0; k < 16; k++) {
    int mn = d[k], mx = d[k];
    for (int j = 1; j < 9; j++) { mn = std::min(mn, d[k + j]); mx = std::max(mx, d[k + j]); }
    best = std::max(best, std::max(mn, -mx));
  }
  return best - 1;
}

struct FastPt { int x, y, score; };

void fast_9_16_nms(const uint8_t* img, int w, int h, size_t stride, int threshold, std::vector<FastPt>& out) {
  out.clear();
  if (w < 7 || h < 7) return;
  std::vector<int> sc((size_t)w * h, 0);
  const ptrdiff_t st = (ptrdiff_t)stride;
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      const uint8_t* p = img + (size_t)y * stride + x;
      // cv::FAST's own early-out (a 9-arc holds one pixel of every opposite pair), so the timed CPU baseline does
      // the work OpenCV does, not 16 full windows per pixel
      const int v = p[0], lo = v - threshold, hi = v + threshold;
      const int a = p[3 * st], b = p[-3 * st], c = p[3], e = p[-3];
      if ((a >= lo && a <= hi && b >= lo && b <= hi) || (c >= lo && c <= hi && e >= lo && e <= hi)) continue;
      int s = fast_score_9_16(p, stride);
      sc[(size_t)y * w + x] = s >= threshold ? s : 0;  // non-corners count as 0 in the NMS buffers
    }
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      int s = sc[(size_t)y * w + x];
      if (s < threshold || s == 0) continue;
      const int* c = &sc[(size_t)y * w + x];
      if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] && s > c[w] &&
          s > c[w + 1])
        out.push_back({x, y, s});
    }
}

// ---------------------------------------------------------------------------------------------------------------
// cv::fastAtan2(y, x) — float32 polynomial, every op individually rounded
// ---------------------------------------------------------------------------------------------------------------
float fast_atan2(float y, float x) {
  const float s = (float)(180.0 / 3.14159265358979323846);
  const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s, p5 = 0.1555786518463281f * s,
              p7 = -0.04432655554792128f * s;
  float ax = std::fabs(x), ay = std::fabs(y), a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + (float)DBL_EPSILON);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + (float)DBL_EPSILON);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

// ---------------------------------------------------------------------------------------------------------------
// cv::KeyPointsFilter::retainBest — real libstdc++ nth_element / partition on the response.
// retain_mode 0: OpenCV 2.4.x   nth_element(begin, begin+n,   end)   (the reference's era; default)
// retain_mode 1: OpenCV >= 3.4  nth_element(begin, begin+n-1, end)
// ---------------------------------------------------------------------------------------------------------------
struct RespGreater {
  bool operator()(const KeyPoint& a, const KeyPoint& b) const { return a.response > b.response; }
};
struct RespGE {
  float v;
  bool operator()(const KeyPoint& k) const { return k.response >= v; }
};
void retain_best(std::vector<KeyPoint>& kps, int n_points, int retain_mode) {
  if (n_points >= 0 && kps.size() > (size_t)n_points) {
    if (n_points == 0) { kps.clear(); return; }
    std::nth_element(kps.begin(), kps.begin() + (retain_mode == 0 ? n_points : n_points - 1), kps.end(),
                     RespGreater());
    float ambiguous = kps[n_points - 1].response;
    auto new_end = std::partition(kps.begin() + n_points, kps.end(), RespGE{ambiguous});
    kps.resize(new_end - kps.begin());
  }
}

// ---------------------------------------------------------------------------------------------------------------
// 7x7 sigma=2 Gaussian blur of an 8-bit ROI that lives inside a bordered parent (non-isolated ROI: the filter
// reads the parent's real border pixels, which here are the unblurred reflect-101 copies).
//   mode 0 (default) OpenCV 2.4.x sepFilter2D 8-bit fixed-point path: taps cvRound(k*256) = 18,34,49,55,49,34,18
//   mode 1           OpenCV >= 3.4.2 GaussianBlurFixedPoint:           taps                  18,34,48,56,48,34,18
// both: dst = (sum_j k_j sum_i k_i p + 32768) >> 16   
// ---------------------------------------------------------------------------------------------------------------
const int kTaps[2][7] = {{18, 34, 49, 55, 49, 34, 18}, {18, 34, 48, 56, 48, 34, 18}};

void blur7_roi_inplace(uint8_t* parent, size_t pstride, int ph, int rx, int ry, int w, int h, int mode) {
  const int* k = kTaps[mode];
  // row-filtered intermediate for ROI rows -3 .. h+2 (reads the parent border; border is 16 px >= 3)
  std::vector<int> hbuf((size_t)(h + 6) * w);
  (void)ph;
  for (int y = -3; y < h + 3; y++) {
    const uint8_t* row = parent + (size_t)(ry + y) * pstride + rx;
    int* H = &hbuf[(size_t)(y + 3) * w];
    for (int x = 0; x < w; x++) {
      int s = 0;
      for (int i = 0; i < 7; i++) s += k[i] * row[x + i - 3];
      H[x] = s;
    }
  }
  for (int y = 0; y < h; y++) {
    uint8_t* D = parent + (size_t)(ry + y) * pstride + rx;
    for (int x = 0; x < w; x++) {
      int s = 0;
      for (int j = 0; j < 7; j++) s += k[j] * hbuf[(size_t)(y + j) * w + x];
      int v = (s + 32768) >> 16;
      D[x] = (uint8_t)(v > 255 ? 255 : v);
    }
  }
}

// BRIEF sampling pattern (data table shared with the product; crc32-pinned by tests/test_kat.py)
const int8_t kOrbPattern31[1024] = {
#include "../orb_slam_b200/csrc/orb_pattern_31.inc"
};

// ---------------------------------------------------------------------------------------------------------------
// The extractor, restated (src/ORBextractor.cc:457-822)
// ---------------------------------------------------------------------------------------------------------------
const float HARRIS_K = 0.04f;   // :73
const int PATCH_SIZE = 31;       // :75
const int HALF_PATCH_SIZE = 15;  // :76
const int EDGE_THRESHOLD = 16;   // :77

struct Mat8 {  // bordered level: `temp` of ComputePyramid (:789) and the ROI inside it (:790)
  int w = 0, h = 0;  // ROI size
  size_t stride = 0;
  std::vector<uint8_t> temp;  // (w+32) x (h+32)
  uint8_t* roi() { return temp.data() + EDGE_THRESHOLD * stride + EDGE_THRESHOLD; }
  const uint8_t* roi() const { return temp.data() + EDGE_THRESHOLD * stride + EDGE_THRESHOLD; }
};

struct LevelDump {
  int cols = 0, rows = 0, cellW = 0, cellH = 0, nfeaturesCell = 0;
  std::vector<int> nTotal, nToRetain;              // [rows*cols]
  std::vector<std::vector<KeyPoint>> cellKeys;     // FAST output per cell before retainBest (cell-local coords)
  std::vector<KeyPoint> kept;                      // level keypoints after the level cap, level coords, with angle
  std::vector<uint8_t> unblurred;                  // ROI copy before blur (w*h)
};

struct Extractor {
  int nfeatures;
  double scaleFactor;  // include/ORBextractor.h:62 — double member initialised from a float argument
  int nlevels, scoreType, fastTh;
  int blur_mode, retain_mode;
  std::vector<int> mnFeaturesPerLevel, umax;
  std::vector<float> mvScaleFactor, mvInvScaleFactor;
  std::vector<Mat8> pyr;
  std::vector<LevelDump> dump;
  bool keep_dump = true;

  Extractor(int _nfeatures, float _scaleFactor, int _nlevels, int _scoreType, int _fastTh, int _blur, int _retain)
      : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), scoreType(_scoreType),
        fastTh(_fastTh), blur_mode(_blur), retain_mode(_retain) {
    // :462-471
    mvScaleFactor.resize(nlevels);
    mvScaleFactor[0] = 1;
    for (int i = 1; i < nlevels; i++) mvScaleFactor[i] = mvScaleFactor[i - 1] * scaleFactor;
    float invScaleFactor = 1.0f / scaleFactor;
    mvInvScaleFactor.resize(nlevels);
    mvInvScaleFactor[0] = 1;
    for (int i = 1; i < nlevels; i++) mvInvScaleFactor[i] = mvInvScaleFactor[i - 1] * invScaleFactor;
    pyr.resize(nlevels);
    dump.resize(nlevels);
    // :476-487
    mnFeaturesPerLevel.resize(nlevels);
    float factor = (float)(1.0 / scaleFactor);
    float nDesiredFeaturesPerScale = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sumFeatures = 0;
    for (int level = 0; level < nlevels - 1; level++) {
      mnFeaturesPerLevel[level] = cvRoundF(nDesiredFeaturesPerScale);
      sumFeatures += mnFeaturesPerLevel[level];
      nDesiredFeaturesPerScale *= factor;
    }
    mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sumFeatures, 0);
    // :495-510
    umax.resize(HALF_PATCH_SIZE + 1);
    int v, v0, vmax = cvFloorD(HALF_PATCH_SIZE * sqrt(2.f) / 2 + 1);
    int vmin = cvCeilD(HALF_PATCH_SIZE * sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) umax[v] = cvRoundD(sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
      while (umax[v0] == umax[v0 + 1]) ++v0;
      umax[v] = v0;
      ++v0;
    }
  }

  // :781-822 (image pyramid only; the mask pyramid is built by the reference but never read by detection)
  void ComputePyramid(const uint8_t* image, int cols, int rows, size_t stride) {
    for (int level = 0; level < nlevels; ++level) {
      float scale = mvInvScaleFactor[level];
      int sw = cvRoundF((float)cols * scale), sh = cvRoundF((float)rows * scale);
      Mat8& m = pyr[level];
      m.w = sw; m.h = sh; m.stride = sw + EDGE_THRESHOLD * 2;
      m.temp.assign(m.stride * (size_t)(sh + EDGE_THRESHOLD * 2), 0);
      if (level != 0) {
        resize_linear_8u(pyr[level - 1].roi(), pyr[level - 1].w, pyr[level - 1].h, pyr[level - 1].stride, m.roi(),
                         sw, sh, m.stride);
      } else {
        for (int y = 0; y < rows; y++) memcpy(m.roi() + (size_t)y * m.stride, image + (size_t)y * stride, cols);
      }
      // copyMakeBorder(..., BORDER_REFLECT_101) :806 / :814
      for (int y = -EDGE_THRESHOLD; y < sh + EDGE_THRESHOLD; y++) {
        int yy = reflect101(y, sh);
        uint8_t* D = m.roi() + (ptrdiff_t)y * (ptrdiff_t)m.stride;
        const uint8_t* S = m.roi() + (ptrdiff_t)yy * (ptrdiff_t)m.stride;
        for (int x = -EDGE_THRESHOLD; x < sw + EDGE_THRESHOLD; x++) {
          if (y >= 0 && y < sh && x >= 0 && x < sw) continue;
          D[x] = S[reflect101(x, sw)];
        }
      }
    }
  }

  // :79-120
  static void HarrisResponses(const uint8_t* img, size_t step_, std::vector<KeyPoint>& pts, int blockSize,
                              float harris_k) {
    const int step = (int)step_;
    int r = blockSize / 2;
    float scale = (1 << 2) * blockSize * 255.0f;
    scale = 1.0f / scale;
    float scale_sq_sq = scale * scale * scale * scale;
    for (size_t ptidx = 0; ptidx < pts.size(); ptidx++) {
      int x0 = cvRoundF(pts[ptidx].x - r);
      int y0 = cvRoundF(pts[ptidx].y - r);
      const uint8_t* ptr0 = img + (ptrdiff_t)y0 * step + x0;
      int a = 0, b = 0, c = 0;
      for (int i = 0; i < blockSize; i++)
        for (int j = 0; j < blockSize; j++) {
          const uint8_t* ptr = ptr0 + i * step + j;
          int Ix = (ptr[1] - ptr[-1]) * 2 + (ptr[-step + 1] - ptr[-step - 1]) + (ptr[step + 1] - ptr[step - 1]);
          int Iy = (ptr[step] - ptr[-step]) * 2 + (ptr[step - 1] - ptr[-step - 1]) + (ptr[step + 1] - ptr[-step + 1]);
          a += Ix * Ix;
          b += Iy * Iy;
          c += Ix * Iy;
        }
      pts[ptidx].response =
          ((float)a * b - (float)c * c - harris_k * ((float)a + b) * ((float)a + b)) * scale_sq_sq;
    }
  }

  // :124-151
  float IC_Angle(const Mat8& image, float ptx, float pty) const {
    int m_01 = 0, m_10 = 0;
    const int step = (int)image.stride;
    const uint8_t* center = image.roi() + (ptrdiff_t)cvRoundF(pty) * step + cvRoundF(ptx);
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
      int v_sum = 0;
      int d = umax[v];
      for (int u = -d; u <= d; ++u) {
        int val_plus = center[u + v * step], val_minus = center[u - v * step];
        v_sum += (val_plus - val_minus);
        m_10 += u * (val_plus + val_minus);
      }
      m_01 += v * v_sum;
    }
    return fast_atan2((float)m_01, (float)m_10);
  }

  // :155-194.  cosf/sinf are the box's glibc; products/sums are individually rounded (-ffp-contract=off).
  static void computeOrbDescriptor(const KeyPoint& kpt, const Mat8& img, uint8_t* desc) {
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float angle = (float)kpt.angle * factorPI;
    float a = (float)cosf(angle), b = (float)sinf(angle);
    const int step = (int)img.stride;
    const uint8_t* center = img.roi() + (ptrdiff_t)cvRoundF(kpt.y) * step + cvRoundF(kpt.x);
    const int8_t* pattern = kOrbPattern31;
    for (int i = 0; i < 32; ++i, pattern += 32) {
      int val = 0;
      for (int j = 0; j < 8; j++) {
        float x0 = pattern[4 * j], y0 = pattern[4 * j + 1], x1 = pattern[4 * j + 2], y1 = pattern[4 * j + 3];
        int t0 = center[cvRoundF(x0 * b + y0 * a) * step + cvRoundF(x0 * a - y0 * b)];
        int t1 = center[cvRoundF(x1 * b + y1 * a) * step + cvRoundF(x1 * a - y1 * b)];
        val |= (t0 < t1) << j;
      }
      desc[i] = (uint8_t)val;
    }
  }

  // :522-707.  Returns <0 when the reference itself would divide by zero / assert (degenerate grid).
  int ComputeKeyPoints(std::vector<std::vector<KeyPoint>>& allKeypoints) {
    allKeypoints.assign(nlevels, {});
    float imageRatio = (float)pyr[0].w / pyr[0].h;
    for (int level = 0; level < nlevels; ++level) {
      const int nDesiredFeatures = mnFeaturesPerLevel[level];
      const int levelCols = sqrtf((float)nDesiredFeatures / (5 * imageRatio));
      const int levelRows = imageRatio * levelCols;
      if (levelCols <= 0 || levelRows <= 0) return -1;
      const int minBorderX = EDGE_THRESHOLD;
      const int minBorderY = minBorderX;
      const int maxBorderX = pyr[level].w - EDGE_THRESHOLD;
      const int maxBorderY = pyr[level].h - EDGE_THRESHOLD;
      const int W = maxBorderX - minBorderX;
      const int H = maxBorderY - minBorderY;
      const int cellW = ceilf((float)W / levelCols);
      const int cellH = ceilf((float)H / levelRows);
      const int nCells = levelRows * levelCols;
      const int nfeaturesCell = ceilf((float)nDesiredFeatures / nCells);

      std::vector<std::vector<KeyPoint>> cellKeyPoints(nCells);
      std::vector<int> nToRetain(nCells, 0), nTotal(nCells, 0);
      std::vector<char> bNoMore(nCells, 0);
      std::vector<int> iniXCol(levelCols), iniYRow(levelRows);
      int nNoMore = 0, nToDistribute = 0;
      float hY = cellH + 6;
      std::vector<FastPt> fp;
      const Mat8& L = pyr[level];
      for (int i = 0; i < levelRows; i++) {
        const float iniY = minBorderY + i * cellH - 3;
        iniYRow[i] = iniY;
        if (i == levelRows - 1) {
          hY = maxBorderY + 3 - iniY;
          if (hY <= 0) continue;
        }
        float hX = cellW + 6;
        for (int j = 0; j < levelCols; j++) {
          float iniX;
          if (i == 0) { iniX = minBorderX + j * cellW - 3; iniXCol[j] = iniX; }
          else iniX = iniXCol[j];
          if (j == levelCols - 1) {
            hX = maxBorderX + 3 - iniX;
            if (hX <= 0) continue;
          }
          // cellImage = level.rowRange(iniY, iniY+hY).colRange(iniX, iniX+hX)  (:599) — cv::Mat asserts the range
          const int cx0 = (int)iniX, cy0 = (int)iniY, cw = (int)(iniX + hX) - cx0, ch = (int)(iniY + hY) - cy0;
          if (cx0 + cw > L.w || cy0 + ch > L.h) return -2;
          const uint8_t* cellImage = L.roi() + (ptrdiff_t)cy0 * (ptrdiff_t)L.stride + cx0;
          std::vector<KeyPoint>& ck = cellKeyPoints[i * levelCols + j];
          fast_9_16_nms(cellImage, cw, ch, L.stride, fastTh, fp);  // :607
          if (fp.size() <= 3) fast_9_16_nms(cellImage, cw, ch, L.stride, 7, fp);  // :609-614
          ck.resize(fp.size());
          for (size_t k = 0; k < fp.size(); k++)
            ck[k] = KeyPoint{(float)fp[k].x, (float)fp[k].y, 7.f, -1.f, (float)fp[k].score, 0, -1};
          if (scoreType == 0) HarrisResponses(cellImage, L.stride, ck, 7, HARRIS_K);  // :616-620
          const int nKeys = ck.size();
          nTotal[i * levelCols + j] = nKeys;
          if (nKeys > nfeaturesCell) {
            nToRetain[i * levelCols + j] = nfeaturesCell;
            bNoMore[i * levelCols + j] = false;
          } else {
            nToRetain[i * levelCols + j] = nKeys;
            nToDistribute += nfeaturesCell - nKeys;
            bNoMore[i * levelCols + j] = true;
            nNoMore++;
          }
        }
      }
      // :644-670
      while (nToDistribute > 0 && nNoMore < nCells) {
        int nNewFeaturesCell = nfeaturesCell + ceilf((float)nToDistribute / (nCells - nNoMore));
        nToDistribute = 0;
        for (int c = 0; c < nCells; c++) {
          if (!bNoMore[c]) {
            if (nTotal[c] > nNewFeaturesCell) {
              nToRetain[c] = nNewFeaturesCell;
              bNoMore[c] = false;
            } else {
              nToRetain[c] = nTotal[c];
              nToDistribute += nNewFeaturesCell - nTotal[c];
              bNoMore[c] = true;
              nNoMore++;
            }
          }
        }
      }
      if (keep_dump) {
        LevelDump& d = dump[level];
        d.cols = levelCols; d.rows = levelRows; d.cellW = cellW; d.cellH = cellH; d.nfeaturesCell = nfeaturesCell;
        d.nTotal = nTotal; d.nToRetain = nToRetain; d.cellKeys = cellKeyPoints;
      }
      std::vector<KeyPoint>& keypoints = allKeypoints[level];
      keypoints.reserve(nDesiredFeatures * 2);
      const int scaledPatchSize = PATCH_SIZE * mvScaleFactor[level];
      // :678-696
      for (int i = 0; i < levelRows; i++)
        for (int j = 0; j < levelCols; j++) {
          std::vector<KeyPoint>& keysCell = cellKeyPoints[i * levelCols + j];
          retain_best(keysCell, nToRetain[i * levelCols + j], retain_mode);
          if ((int)keysCell.size() > nToRetain[i * levelCols + j]) keysCell.resize(nToRetain[i * levelCols + j]);
          for (size_t k = 0, kend = keysCell.size(); k < kend; k++) {
            keysCell[k].x += iniXCol[j];
            keysCell[k].y += iniYRow[i];
            keysCell[k].octave = level;
            keysCell[k].size = scaledPatchSize;
            keypoints.push_back(keysCell[k]);
          }
        }
      // :697-701
      if ((int)keypoints.size() > nDesiredFeatures) {
        retain_best(keypoints, nDesiredFeatures, retain_mode);
        keypoints.resize(nDesiredFeatures);
      }
    }
    // :705-706
    for (int level = 0; level < nlevels; ++level)
      for (auto& kp : allKeypoints[level]) kp.angle = IC_Angle(pyr[level], kp.x, kp.y);
    return 0;
  }

  // :718-779
  int extract(const uint8_t* image, int cols, int rows, size_t stride, KeyPoint* out_kps, uint8_t* out_desc,
              int cap) {
    if (cols <= 0 || rows <= 0 || !image) return 0;  // :721-722 (empty image: nothing happens)
    ComputePyramid(image, cols, rows, stride);
    std::vector<std::vector<KeyPoint>> allKeypoints;
    int rc = ComputeKeyPoints(allKeypoints);
    if (rc < 0) return rc;
    int nkeypoints = 0;
    for (int level = 0; level < nlevels; ++level) nkeypoints += (int)allKeypoints[level].size();
    if (nkeypoints > cap) return -3;
    int offset = 0;
    for (int level = 0; level < nlevels; ++level) {
      std::vector<KeyPoint>& keypoints = allKeypoints[level];
      int nkeypointsLevel = (int)keypoints.size();
      if (keep_dump) {
        LevelDump& d = dump[level];
        d.kept = keypoints;
        d.unblurred.resize((size_t)pyr[level].w * pyr[level].h);
        for (int y = 0; y < pyr[level].h; y++)
          memcpy(&d.unblurred[(size_t)y * pyr[level].w], pyr[level].roi() + (size_t)y * pyr[level].stride,
                 pyr[level].w);
      }
      if (nkeypointsLevel == 0) continue;  // :755-756 — level not blurred
      Mat8& workingMat = pyr[level];
      blur7_roi_inplace(workingMat.temp.data(), workingMat.stride, workingMat.h + 2 * EDGE_THRESHOLD,
                        EDGE_THRESHOLD, EDGE_THRESHOLD, workingMat.w, workingMat.h, blur_mode);  // :760
      for (int i = 0; i < nkeypointsLevel; i++)
        computeOrbDescriptor(keypoints[i], workingMat, out_desc + (size_t)(offset + i) * 32);  // :764
      if (level != 0) {  // :769-775
        float scale = mvScaleFactor[level];
        for (auto& kp : keypoints) { kp.x *= scale; kp.y *= scale; }
      }
      memcpy(out_kps + offset, keypoints.data(), sizeof(KeyPoint) * nkeypointsLevel);  // :777
      offset += nkeypointsLevel;
    }
    return nkeypoints;
  }
};

}  // namespace

// =================================================================================================================
// C interface for ctypes (tests / bench cpu_baseline only)
// =================================================================================================================
extern "C" {

void* orb_oracle_create(int nfeatures, float scaleFactor, int nlevels, int scoreType, int fastTh, int blur_mode,
                        int retain_mode) {
  return new Extractor(nfeatures, scaleFactor, nlevels, scoreType, fastTh, blur_mode, retain_mode);
}
void orb_oracle_destroy(void* h) { delete (Extractor*)h; }
void orb_oracle_keep_dump(void* h, int on) { ((Extractor*)h)->keep_dump = on != 0; }

int orb_oracle_extract(void* h, const uint8_t* img, int w, int h_, long stride, void* kps, uint8_t* desc, int cap) {
  return ((Extractor*)h)->extract(img, w, h_, (size_t)stride, (KeyPoint*)kps, desc, cap);
}

// tables (KATs)
void orb_oracle_tables(void* h, int* features_per_level, float* scale, float* inv_scale, int* umax16) {
  Extractor* e = (Extractor*)h;
  for (int i = 0; i < e->nlevels; i++) {
    features_per_level[i] = e->mnFeaturesPerLevel[i];
    scale[i] = e->mvScaleFactor[i];
    inv_scale[i] = e->mvInvScaleFactor[i];
  }
  for (int i = 0; i < 16; i++) umax16[i] = e->umax[i];
}

// intermediates of the last extract() -------------------------------------------------------------------------
void orb_oracle_level_size(void* h, int level, int* w, int* h_) {
  Extractor* e = (Extractor*)h;
  *w = e->pyr[level].w; *h_ = e->pyr[level].h;
}
// which: 0 = unblurred ROI, 1 = ROI as left by operator() (blurred iff the level had keypoints)
void orb_oracle_level_image(void* h, int level, int which, uint8_t* out) {
  Extractor* e = (Extractor*)h;
  const Mat8& m = e->pyr[level];
  if (which == 0) { memcpy(out, e->dump[level].unblurred.data(), (size_t)m.w * m.h); return; }
  for (int y = 0; y < m.h; y++) memcpy(out + (size_t)y * m.w, m.roi() + (size_t)y * m.stride, m.w);
}
void orb_oracle_level_grid(void* h, int level, int* out6) {
  const LevelDump& d = ((Extractor*)h)->dump[level];
  out6[0] = d.cols; out6[1] = d.rows; out6[2] = d.cellW; out6[3] = d.cellH; out6[4] = d.nfeaturesCell;
  out6[5] = (int)d.kept.size();
}
void orb_oracle_level_cells(void* h, int level, int* nTotal, int* nToRetain) {
  const LevelDump& d = ((Extractor*)h)->dump[level];
  memcpy(nTotal, d.nTotal.data(), sizeof(int) * d.nTotal.size());
  memcpy(nToRetain, d.nToRetain.data(), sizeof(int) * d.nToRetain.size());
}
// FAST candidates of one cell (cell-local coordinates, before retainBest): xs, ys, responses
int orb_oracle_cell_keys(void* h, int level, int cell, int* xs, int* ys, float* resp, int cap) {
  const LevelDump& d = ((Extractor*)h)->dump[level];
  const auto& v = d.cellKeys[cell];
  int n = std::min((int)v.size(), cap);
  for (int i = 0; i < n; i++) { xs[i] = (int)v[i].x; ys[i] = (int)v[i].y; resp[i] = v[i].response; }
  return (int)v.size();
}
int orb_oracle_level_kept(void* h, int level, void* kps, int cap) {
  const LevelDump& d = ((Extractor*)h)->dump[level];
  int n = std::min((int)d.kept.size(), cap);
  memcpy(kps, d.kept.data(), sizeof(KeyPoint) * n);
  return (int)d.kept.size();
}

// primitives ---------------------------------------------------------------------------------------------------
void orb_oracle_resize(const uint8_t* src, int sw, int sh, long sstride, uint8_t* dst, int dw, int dh, long dstride) {
  resize_linear_8u(src, sw, sh, sstride, dst, dw, dh, dstride);
}
// FAST + NMS on a sub-image; returns count, fills up to cap
int orb_oracle_fast(const uint8_t* img, int w, int h, long stride, int threshold, int* xs, int* ys, int* sc,
                    int cap) {
  std::vector<FastPt> v;
  fast_9_16_nms(img, w, h, stride, threshold, v);
  int n = std::min((int)v.size(), cap);
  for (int i = 0; i < n; i++) { xs[i] = v[i].x; ys[i] = v[i].y; sc[i] = v[i].score; }
  return (int)v.size();
}
// threshold-free score map (0 where < 1); only pixels >= 3 px inside are scored
void orb_oracle_fast_score_map(const uint8_t* img, int w, int h, long stride, uint8_t* out) {
  memset(out, 0, (size_t)w * h);
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      int s = fast_score_9_16(img + (size_t)y * stride + x, stride);
      out[(size_t)y * w + x] = (uint8_t)std::max(s, 0);
    }
}
float orb_oracle_fast_atan2(float y, float x) { return fast_atan2(y, x); }
// blur of a w x h image with reflect-101 border (builds the bordered parent like ComputePyramid does)
void orb_oracle_blur7(const uint8_t* src, int w, int h, long stride, uint8_t* dst, int mode) {
  Mat8 m;
  m.w = w; m.h = h; m.stride = w + 32;
  m.temp.assign(m.stride * (size_t)(h + 32), 0);
  for (int y = -16; y < h + 16; y++)
    for (int x = -16; x < w + 16; x++)
      m.roi()[(ptrdiff_t)y * (ptrdiff_t)m.stride + x] = src[(size_t)reflect101(y, h) * stride + reflect101(x, w)];
  blur7_roi_inplace(m.temp.data(), m.stride, h + 32, 16, 16, w, h, mode);
  for (int y = 0; y < h; y++) memcpy(dst + (size_t)y * w, m.roi() + (size_t)y * m.stride, w);
}
// retainBest + truncate on (response, tag) records; returns the new size, reorders resp/tag in place
int orb_oracle_retain_best(float* resp, int* tag, int n, int n_points, int retain_mode) {
  std::vector<KeyPoint> v(n);
  for (int i = 0; i < n; i++) { v[i] = KeyPoint{0, 0, 0, 0, resp[i], tag[i], 0}; }
  retain_best(v, n_points, retain_mode);
  if ((int)v.size() > n_points) v.resize(n_points);
  for (size_t i = 0; i < v.size(); i++) { resp[i] = v[i].response; tag[i] = v[i].octave; }
  return (int)v.size();
}
void orb_oracle_sincos(float angle_rad, float* c, float* s) { *c = cosf(angle_rad); *s = sinf(angle_rad); }
const int8_t* orb_oracle_pattern() { return kOrbPattern31; }

}  // extern "C"

// Multi-threaded driver for the timed CPU baseline: independent frames over `nthreads` std::threads, one private
// Extractor per thread (the reference itself is single-threaded: one Tracking thread calls operator()).
#include <atomic>
#include <thread>
extern "C" int orb_oracle_extract_many(void* h, const uint8_t* imgs, int nframes, int w, int hgt, long stride,
                                       long frame_stride, void* kps, uint8_t* desc, int* n_out, int cap,
                                       int nthreads) {
  Extractor* proto = (Extractor*)h;
  std::atomic<int> next(0), bad(0);
  auto work = [&]() {
    Extractor e(proto->nfeatures, (float)proto->scaleFactor, proto->nlevels, proto->scoreType, proto->fastTh,
                proto->blur_mode, proto->retain_mode);
    e.keep_dump = false;
    for (;;) {
      int f = next.fetch_add(1);
      if (f >= nframes) break;
      int n = e.extract(imgs + (size_t)f * frame_stride, w, hgt, (size_t)stride, (KeyPoint*)kps + (size_t)f * cap,
                        desc + (size_t)f * cap * 32, cap);
      if (n < 0) { bad = 1; n = 0; }
      n_out[f] = n;
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < nthreads; t++) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
  return bad ? -1 : 0;
}
