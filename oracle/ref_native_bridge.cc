// ORACLE / test + baseline infrastructure: a NATIVE bridge for the compiled reference extractor.
// oracle/ref_orbextractor.cc compiles the UNMODIFIED /root/reference/src/ORBextractor.cc against type-only OpenCV
// stand-ins and forwards its five OpenCV library calls through function pointers.  ref_extractor.py points them at
// cv2 (the parity pin; every call crosses into Python, 65 ms per frame).  This file points them at the oracle's
// restatements of the same routines (orb_oracle.cc: resize_linear_8u, fast_9_16_nms, blur7, fast_atan2 — each pinned
// bit for bit against cv2 in tests/test_oracle_vs_cv2.py), so that the reference's own operator() — its pyramid
// loop, cell grid, two-threshold FAST retry, quota redistribution, retainBest, IC_Angle, rotated-BRIEF sampling and
// keypoint scaling, compiled with its own -O3 — runs natively end to end.  That is the timed CPU baseline
// (`cpu_baseline.kind = "reference"`, bench.py); tests/test_ref_extractor.py checks it returns exactly what the
// cv2-bridged build returns.
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "ORBextractor.h"

extern "C" {
void orb_oracle_resize(const uint8_t* src, int sw, int sh, long sstride, uint8_t* dst, int dw, int dh, long dstride);
int orb_oracle_fast(const uint8_t* img, int w, int h, long stride, int threshold, int* xs, int* ys, int* sc, int cap);
float orb_oracle_fast_atan2(float y, float x);
void orb_oracle_blur7_roi(uint8_t* roi, long step, int w, int h, int mode);
void ref_orb_set_bridge(ref_cv::resize_fn r, ref_cv::border_fn b, ref_cv::blur_fn g, ref_cv::fast_fn f,
                        ref_cv::atan2_fn a, int retain_mode);
int ref_orb_extract(void* h, const uint8_t* img, int width, int height, long stride, void* kps, uint8_t* desc, int cap);
}

namespace {

int g_blur_mode = 1;   // Gaussian taps: 0 = OpenCV 2.4, 1 = OpenCV >= 3.4 (what cv2 4.13 computes)

inline int refl101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
  return i;
}

void n_resize(const uchar* src, int sw, int sh, long sstep, uchar* dst, int dw, int dh, long dstep, int) {
  orb_oracle_resize(src, sw, sh, sstep, dst, dw, dh, dstep);   // INTER_LINEAR, the only mode the reference uses (:800)
}

// cv::copyMakeBorder(src, dst, t, b, l, r, BORDER_REFLECT_101 [+ BORDER_ISOLATED]); src may be a view into dst (:806)
void n_border(const uchar* src, int w, int h, long sstep, uchar* dst, long dstep, int top, int bottom, int left,
              int right, int) {
  thread_local std::vector<uchar> tmp;
  tmp.resize((size_t)w * h);
  for (int y = 0; y < h; y++) memcpy(&tmp[(size_t)y * w], src + (size_t)y * sstep, w);
  const int W = w + left + right, H = h + top + bottom;
  for (int y = 0; y < H; y++) {
    const uchar* S = &tmp[(size_t)refl101(y - top, h) * w];
    uchar* D = dst + (size_t)y * dstep;
    for (int x = 0; x < left; x++) D[x] = S[refl101(x - left, w)];
    memcpy(D + left, S, w);
    for (int x = left + w; x < W; x++) D[x] = S[refl101(x - left, w)];
  }
}

void n_blur(uchar* img, int w, int h, long step, int, double, int) { orb_oracle_blur7_roi(img, step, w, h, g_blur_mode); }

int n_fast(const uchar* img, int w, int h, long step, int th, int, float* xyr, int cap) {
  thread_local std::vector<int> xs, ys, sc;
  if ((int)xs.size() < cap) { xs.resize(cap); ys.resize(cap); sc.resize(cap); }
  int n = orb_oracle_fast(img, w, h, step, th, xs.data(), ys.data(), sc.data(), cap);
  if (n > cap) n = cap;
  for (int i = 0; i < n; i++) { xyr[3 * i] = (float)xs[i]; xyr[3 * i + 1] = (float)ys[i]; xyr[3 * i + 2] = (float)sc[i]; }
  return n;
}

}  // namespace

extern "C" {

void ref_orb_use_native_bridge(int retain_mode, int blur_mode) {
  g_blur_mode = blur_mode;
  ref_orb_set_bridge(n_resize, n_border, n_blur, n_fast, orb_oracle_fast_atan2, retain_mode);
}

// Independent frames over `nthreads` std::threads, one private ORB_SLAM::ORBextractor per thread (the reference itself
// is single-threaded: one Tracking thread calls operator(), src/Frame.cc:59).  kps [nframes][cap] 28-byte records,
// desc [nframes][cap][32], n_out [nframes].  Returns 0, or -1 when a frame produced more than cap keypoints.
int ref_orb_extract_many(int nfeatures, float scaleFactor, int nlevels, int scoreType, int fastTh, const uint8_t* imgs,
                         int nframes, int w, int h, long stride, long frame_stride, void* kps, uint8_t* desc,
                         int* n_out, int cap, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  std::vector<int> bad(nthreads, 0);
  auto work = [&](int t) {
    ORB_SLAM::ORBextractor ex(nfeatures, scaleFactor, nlevels, scoreType, fastTh);
    for (int f = t; f < nframes; f += nthreads) {
      const int n = ref_orb_extract(&ex, imgs + (size_t)f * frame_stride, w, h, stride,
                                    (char*)kps + (size_t)f * cap * 28, desc + (size_t)f * cap * 32, cap);
      if (n < 0) { bad[t] = 1; n_out[f] = 0; } else n_out[f] = n;
    }
  };
  if (nthreads == 1) work(0);
  else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(work, t);
    for (auto& x : th) x.join();
  }
  for (int b : bad) if (b) return -1;
  return 0;
}

}  // extern "C"
