// ORACLE / test infrastructure: C wrapper around the UNMODIFIED reference extractor — src/ORBextractor.cc is compiled
// from /root/reference where it lies (oracle/Makefile target `ref` -> oracle/_ref/libref_orbextractor.so), against the
// type-only OpenCV stand-in oracle/ref_shim_cv/, with cv::resize / copyMakeBorder / GaussianBlur / FAST / fastAtan2
// forwarded to the REAL OpenCV (cv2 4.13) through callbacks installed by oracle/ref_extractor.py.  Everything else
// that shapes the output — pyramid + borders, cell grid, two-threshold FAST retry, quota redistribution, retainBest
// call order, IC_Angle, the rotated BRIEF sampling with glibc cos/sin, keypoint scaling — is the reference's own code.
// Pins oracle/orb_oracle.cc end to end (tests/test_ref_extractor.py) and generates tests/golden/ref_extract_*.npz.
#include <cstdint>
#include <cstring>
#include <vector>

#include "ORBextractor.h"

namespace ref_cv {
resize_fn resize = nullptr;
border_fn border = nullptr;
blur_fn blur = nullptr;
fast_fn fast = nullptr;
atan2_fn atan2f32 = nullptr;
undistort_fn undistort = nullptr;
int retain_mode = 0;
}  // namespace ref_cv

extern "C" {

void ref_orb_set_bridge(ref_cv::resize_fn r, ref_cv::border_fn b, ref_cv::blur_fn g, ref_cv::fast_fn f,
                        ref_cv::atan2_fn a, int retain_mode) {
  ref_cv::resize = r; ref_cv::border = b; ref_cv::blur = g; ref_cv::fast = f; ref_cv::atan2f32 = a;
  ref_cv::retain_mode = retain_mode;
}

void ref_orb_set_undistort(ref_cv::undistort_fn u) { ref_cv::undistort = u; }

void* ref_orb_create(int nfeatures, float scaleFactor, int nlevels, int scoreType, int fastTh) {
  return new ORB_SLAM::ORBextractor(nfeatures, scaleFactor, nlevels, scoreType, fastTh);
}
void ref_orb_destroy(void* h) { delete static_cast<ORB_SLAM::ORBextractor*>(h); }

// kps: 28-byte cv::KeyPoint records {x, y, size, angle, response, octave, class_id}; returns the keypoint count
// (or -needed when cap is too small)
int ref_orb_extract(void* h, const uint8_t* img, int width, int height, long stride, void* kps, uint8_t* desc,
                    int cap) {
  cv::Mat image(height, width, CV_8UC1);
  for (int y = 0; y < height; y++) memcpy(image.ptr(y), img + (size_t)y * stride, width);
  std::vector<cv::KeyPoint> keypoints;
  cv::Mat descriptors;
  (*static_cast<ORB_SLAM::ORBextractor*>(h))(image, cv::Mat(), keypoints, descriptors);
  const int n = (int)keypoints.size();
  if (n > cap) return -n;
  static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
  if (n) memcpy(kps, keypoints.data(), (size_t)n * 28);
  for (int i = 0; i < n; i++) memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
  return n;
}

}  // extern "C"
