// ORACLE / test infrastructure.  Stand-in for the OpenCV headers that the reference's src/ORBextractor.cc includes
// (<opencv/cv.h>, <opencv2/core/core.hpp>, <opencv2/highgui/highgui.hpp>), so that the reference file compiles where
// it lies WITHOUT the OpenCV C++ library, and runs with the REAL OpenCV doing the image primitives:
//
//   * data types (Mat with ROI views that share one buffer, Point_, Size, Rect, KeyPoint, AutoBuffer, InputArray /
//     OutputArray as Mat references) are written here with the semantics the reference relies on: `create` keeps an
//     existing buffer of the right shape (so resize / copyMakeBorder / `= Mat::zeros` write THROUGH a ROI view into
//     its parent, as in ComputePyramid :781-822 and computeDescriptors :709-716);
//   * the library routines cv::resize, cv::copyMakeBorder, cv::GaussianBlur, cv::FAST and cv::fastAtan2 are NOT
//     restated: they forward to function pointers which oracle/ref_extractor.py points at the installed cv2 (4.13);
//   * cvRound / cvFloor / cvCeil are OpenCV's inline definitions (round-half-even via lrint, as its SSE2 path);
//   * KeyPointsFilter::retainBest follows OpenCV's published source (features2d/keypoint.cpp): std::nth_element +
//     std::partition on KeyPoint; the element chosen for nth_element differs between OpenCV 2.4 (begin + n) and
//     >= 3.4 (begin + n - 1), selected by ref_cv::retain_mode.
// Used only to build oracle/_ref/libref_orbextractor.so (oracle/Makefile target `ref`).
#pragma once
#include <assert.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <limits.h>

#include <algorithm>
#include <cmath>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <string>
#include <vector>

typedef unsigned char uchar;
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_PI 3.1415926535897932384626433832795
#define CV_Assert(expr) assert(expr)

inline int cvRound(double value) { return (int)lrint(value); }
inline int cvFloor(double value) { int i = (int)value; return i - (i > value); }
inline int cvCeil(double value) { int i = (int)value; return i + (i < value); }

namespace ref_cv {   // the bridge to the real OpenCV (set from Python, oracle/ref_extractor.py)
typedef void (*resize_fn)(const uchar* src, int sw, int sh, long sstep, uchar* dst, int dw, int dh, long dstep, int interp);
typedef void (*border_fn)(const uchar* src, int w, int h, long sstep, uchar* dst, long dstep, int top, int bottom,
                          int left, int right, int border_type);
typedef void (*blur_fn)(uchar* img, int w, int h, long step, int ksize, double sigma, int border_type);
typedef int (*fast_fn)(const uchar* img, int w, int h, long step, int threshold, int nonmax, float* xyr, int cap);
typedef void (*undistort_fn)(float* xy, int n, const float* K9, const float* D, int nd);
typedef float (*atan2_fn)(float y, float x);
extern undistort_fn undistort;
extern resize_fn resize;
extern border_fn border;
extern blur_fn blur;
extern fast_fn fast;
extern atan2_fn atan2f32;
extern int retain_mode;   // 0: OpenCV 2.4 (nth = begin + n), 1: OpenCV >= 3.4 (nth = begin + n - 1)
}  // namespace ref_cv

namespace cv {

template <class T> struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T a, T b) : x(a), y(b) {}
};
template <class T> inline Point_<T>& operator*=(Point_<T>& a, float b) { a.x = (T)(a.x * b); a.y = (T)(a.y * b); return a; }
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
  int width, height;
  Size() : width(0), height(0) {}
  Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
  int x, y, width, height;
  Rect() : x(0), y(0), width(0), height(0) {}
  Rect(int a, int b, int w, int h) : x(a), y(b), width(w), height(h) {}
};

struct KeyPoint {
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
      : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};

template <class T> class AutoBuffer {
 public:
  explicit AutoBuffer(size_t n) : v_(n) {}
  operator T*() { return v_.data(); }
 private:
  std::vector<T> v_;
};

class Mat {
 public:
  int rows, cols;
  uchar* data;
  size_t step;
  struct ZerosExpr { int rows, cols, type; };

  Mat() : rows(0), cols(0), data(nullptr), step(0), type_(0) {}
  Mat(int r, int c, int t) : Mat() { create(r, c, t); }
  Mat(Size s, int t) : Mat() { create(s.height, s.width, t); }
  Mat(const Mat& m, const Rect& r) : Mat(m) { narrow(r); }
  void create(int r, int c, int t) {
    if (data && r == rows && c == cols && t == type_) return;   // cv::Mat::create: same shape and type -> keep
    rows = r; cols = c; type_ = t;
    step = (size_t)c * elemSize1();
    buf_ = std::make_shared<std::vector<uchar>>((size_t)r * step + 64);
    data = buf_->data();
  }
  void create(Size s, int t) { create(s.height, s.width, t); }
  void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = 0; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return type_; }
  size_t elemSize1() const { return type_ == CV_32F ? 4 : 1; }
  size_t step1() const { return step / elemSize1(); }
  Size size() const { return Size(cols, rows); }
  Mat getMat() const { return *this; }
  Mat clone() const {
    Mat m;
    if (!empty()) {
      m.create(rows, cols, type_);
      for (int r = 0; r < rows; r++) memcpy(m.data + (size_t)r * m.step, data + (size_t)r * step, (size_t)cols * elemSize1());
    }
    return m;
  }
  Mat operator()(const Rect& r) const { return Mat(*this, r); }
  Mat rowRange(int a, int b) const { return Mat(*this, Rect(0, a, cols, b - a)); }
  Mat colRange(int a, int b) const { return Mat(*this, Rect(a, 0, b - a, rows)); }
  uchar* ptr(int r = 0) { return data + (size_t)r * step; }
  const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
  template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
  template <class T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
  template <class T> T& at(int r, int c) { return reinterpret_cast<T*>(data + (size_t)r * step)[c]; }
  template <class T> const T& at(int r, int c) const { return reinterpret_cast<const T*>(data + (size_t)r * step)[c]; }
  // single-index access of a row or column vector (cv::Mat::at(int i0))
  template <class T> T& at(int i) {
    if (rows == 1) return reinterpret_cast<T*>(data)[i];
    if (cols == 1) return *reinterpret_cast<T*>(data + (size_t)i * step);
    return at<T>(i / cols, i % cols);
  }
  template <class T> const T& at(int i) const { return const_cast<Mat*>(this)->at<T>(i); }
  Mat row(int r) const { return Mat(*this, Rect(0, r, cols, 1)); }
  Mat col(int c) const { return Mat(*this, Rect(c, 0, 1, rows)); }
  Mat reshape(int) const { return *this; }   // channels are not modelled: an N x 2 float matrix stays N x 2
  void copyTo(Mat&& dst) const { Mat d = dst; copyTo(d); }   // into a temporary VIEW: create() keeps its buffer
  void copyTo(Mat& dst) const {
    if (empty()) { dst.release(); return; }
    dst.create(rows, cols, type_);
    for (int r = 0; r < rows; r++) memcpy(dst.data + (size_t)r * dst.step, data + (size_t)r * step, (size_t)cols * elemSize1());
  }
  Mat(const ZerosExpr& e) : Mat() { *this = e; }
  // ---- small float matrix algebra (poses, normals).  Plain loops: NOT a model of cv::gemm; nothing that is pinned
  // against this build goes through it (the searches read keypoints, descriptors, grids and flags only) ----------
  static Mat eye(int r, int c, int t) { Mat m = ZerosExpr{r, c, t}; for (int i = 0; i < std::min(r, c); i++) m.at<float>(i, i) = 1.f; return m; }
  Mat t() const { Mat m(cols, rows, type_); for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) m.at<float>(c, r) = at<float>(r, c); return m; }
  double dot(const Mat& o) const { double s = 0; for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) s += (double)at<float>(r, c) * o.at<float>(r, c); return s; }
  static ZerosExpr zeros(int r, int c, int t) { return ZerosExpr{r, c, t}; }
  // MatExpr assignment (MatOp_Initializer::assign): create() on the destination, then fill -> writes through a view
  Mat& operator=(const ZerosExpr& e) {
    create(e.rows, e.cols, e.type);
    for (int r = 0; r < rows; r++) memset(data + (size_t)r * step, 0, (size_t)cols * elemSize1());
    return *this;
  }

 private:
  void narrow(const Rect& r) {
    data += (size_t)r.y * step + (size_t)r.x * elemSize1();
    rows = r.height; cols = r.width;
  }
  int type_;
  std::shared_ptr<std::vector<uchar>> buf_;
};

inline Mat operator*(const Mat& a, const Mat& b) {
  Mat m(a.rows, b.cols, CV_32F);
  for (int r = 0; r < a.rows; r++) for (int c = 0; c < b.cols; c++) { float s = 0; for (int k = 0; k < a.cols; k++) s += a.at<float>(r, k) * b.at<float>(k, c); m.at<float>(r, c) = s; }
  return m;
}
template <class F> inline Mat mat_zip(const Mat& a, const Mat& b, F f) {
  Mat m(a.rows, a.cols, CV_32F);
  for (int r = 0; r < a.rows; r++) for (int c = 0; c < a.cols; c++) m.at<float>(r, c) = f(a.at<float>(r, c), b.at<float>(r, c));
  return m;
}
inline Mat operator+(const Mat& a, const Mat& b) { return mat_zip(a, b, [](float x, float y) { return x + y; }); }
inline Mat operator-(const Mat& a, const Mat& b) { return mat_zip(a, b, [](float x, float y) { return x - y; }); }
inline Mat operator-(const Mat& a) { return mat_zip(a, a, [](float x, float) { return -x; }); }
inline Mat operator*(double s, const Mat& a) { return mat_zip(a, a, [s](float x, float) { return (float)(x * s); }); }
inline Mat operator*(const Mat& a, double s) { return s * a; }
inline Mat operator/(const Mat& a, double s) { return mat_zip(a, a, [s](float x, float) { return (float)(x / s); }); }
inline double norm(const Mat& a) { return std::sqrt(a.dot(a)); }

// inert: the YAML members of DBoW2::TemplatedVocabulary are virtual and get instantiated, never called
class FileNode {
 public:
  size_t size() const { return 0; }
  FileNode operator[](int) const { return FileNode(); }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
  operator int() const { return 0; }
  operator float() const { return 0.f; }
  operator double() const { return 0.0; }
  operator std::string() const { return std::string(); }
};
class FileStorage {
 public:
  enum { READ = 0, WRITE = 1 };
  FileStorage(const std::string&, int) {}
  bool isOpened() const { return false; }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
};
template <class T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }

typedef const Mat& InputArray;
typedef Mat& OutputArray;

enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { BORDER_CONSTANT = 0, BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };

struct ORB { enum { HARRIS_SCORE = 0, FAST_SCORE = 1 }; };

inline float fastAtan2(float y, float x) { return ref_cv::atan2f32(y, x); }

inline void resize(const Mat& src, Mat& dst, Size dsize, double, double, int interpolation) {
  dst.create(dsize.height, dsize.width, src.type());
  ref_cv::resize(src.data, src.cols, src.rows, (long)src.step, dst.data, dst.cols, dst.rows, (long)dst.step, interpolation);
}

inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int borderType) {
  dst.create(src.rows + top + bottom, src.cols + left + right, src.type());
  ref_cv::border(src.data, src.cols, src.rows, (long)src.step, dst.data, (long)dst.step, top, bottom, left, right,
                 borderType);
}

inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sigmaX, double sigmaY, int borderType) {
  assert(src.data == dst.data && ksize.width == ksize.height && sigmaX == sigmaY);   // the one call the reference makes
  ref_cv::blur(dst.data, dst.cols, dst.rows, (long)dst.step, ksize.width, sigmaX, borderType);
}

inline void FAST(const Mat& image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
  const int cap = std::max(image.rows * image.cols, 1);
  std::vector<float> xyr((size_t)cap * 3);
  const int n = ref_cv::fast(image.data, image.cols, image.rows, (long)image.step, threshold, nonmaxSuppression ? 1 : 0,
                             xyr.data(), cap);
  keypoints.clear();
  for (int i = 0; i < n; i++) keypoints.push_back(KeyPoint(xyr[3 * i], xyr[3 * i + 1], 7.f, -1, xyr[3 * i + 2]));
}

// cv::undistortPoints(src, dst, K, D, R = Mat(), P): N x 2 float points, forwarded to the real OpenCV.  P must be K and
// R empty (the only call the reference makes, src/Frame.cc:306 and :333).
inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& D, const Mat& R, const Mat& P) {
  assert(src.data == dst.data && R.empty() && P.data == K.data && src.cols == 2);
  std::vector<float> xy((size_t)src.rows * 2), k(9), d;
  for (int i = 0; i < src.rows; i++) { xy[2 * i] = src.at<float>(i, 0); xy[2 * i + 1] = src.at<float>(i, 1); }
  for (int i = 0; i < 9; i++) k[i] = K.at<float>(i / 3, i % 3);
  for (int i = 0; i < D.rows * D.cols; i++) d.push_back(D.at<float>(i));
  ref_cv::undistort(xy.data(), src.rows, k.data(), d.data(), (int)d.size());
  for (int i = 0; i < src.rows; i++) { dst.at<float>(i, 0) = xy[2 * i]; dst.at<float>(i, 1) = xy[2 * i + 1]; }
}

struct KeypointResponseGreaterThanThreshold {
  explicit KeypointResponseGreaterThanThreshold(float v) : value(v) {}
  bool operator()(const KeyPoint& kpt) const { return kpt.response >= value; }
  float value;
};
struct KeypointResponseGreater {
  bool operator()(const KeyPoint& a, const KeyPoint& b) const { return a.response > b.response; }
};

struct KeyPointsFilter {
  // features2d/keypoint.cpp: take keypoints less than threshold, keep every keypoint whose response ties the last kept
  static void retainBest(std::vector<KeyPoint>& keypoints, int n_points) {
    if (n_points >= 0 && keypoints.size() > (size_t)n_points) {
      if (n_points == 0) { keypoints.clear(); return; }
      std::nth_element(keypoints.begin(), keypoints.begin() + (ref_cv::retain_mode == 0 ? n_points : n_points - 1),
                       keypoints.end(), KeypointResponseGreater());
      float ambiguous_response = keypoints[n_points - 1].response;
      std::vector<KeyPoint>::const_iterator new_end =
          std::partition(keypoints.begin() + n_points, keypoints.end(),
                         KeypointResponseGreaterThanThreshold(ambiguous_response));
      keypoints.resize(new_end - keypoints.begin());
    }
  }
};

}  // namespace cv
