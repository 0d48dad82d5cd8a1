// ORACLE — TEST INFRASTRUCTURE ONLY.  Frame glue between extraction and matching:
//   Frame::UndistortKeyPoints   reference src/Frame.cc:288-318
//   Frame::ComputeImageBounds   reference src/Frame.cc:320-349  (+ grid element sizes, :75-76)
// cv::undistortPoints is an OpenCV routine (un-vendored dependency, CMakeLists.txt:19); its published algorithm
// (cvUndistortPoints: normalise, five fixed-point iterations of the Brown-Conrady inverse in double, re-project with
// P) is restated here for the call the reference makes: K and D are CV_32F, D = (k1, k2, p1, p2, k3), R empty, P = K.
// Pinned bit-for-bit against cv2 4.13 by tests/test_frame_oracle.py.  Built with -ffp-contract=off.
#include <cmath>
#include <cstdint>

extern "C" {

struct fo_camera {
  float fx, fy, cx, cy;
  float k1, k2, p1, p2, k3;
};

void fo_undistort_points(const fo_camera* c, const float* xy_in, int n, float* xy_out) {
  const double fx = c->fx, fy = c->fy, cx = c->cx, cy = c->cy;
  const double ifx = 1. / fx, ify = 1. / fy;
  const double k[5] = {c->k1, c->k2, c->p1, c->p2, c->k3};
  for (int i = 0; i < n; i++) {
    const double u = xy_in[2 * i], v = xy_in[2 * i + 1];
    double x = (u - cx) * ifx, y = (v - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
      const double r2 = x * x + y * y;
      const double icdist = 1. / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
      if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
      const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x);
      const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y;
      x = (x0 - deltaX) * icdist;
      y = (y0 - deltaY) * icdist;
    }
    xy_out[2 * i] = (float)(fx * x + cx);
    xy_out[2 * i + 1] = (float)(fy * y + cy);
  }
}

// out = {mnMinX, mnMaxX, mnMinY, mnMaxY}; inv = {mfGridElementWidthInv, mfGridElementHeightInv}
void fo_image_bounds(const fo_camera* c, int cols, int rows, int32_t* out, float* inv) {
  if (c->k1 != 0.0f) {
    const float corners[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
    float m[8];
    fo_undistort_points(c, corners, 4, m);
    out[0] = (int32_t)std::fmin(std::floor(m[0]), std::floor(m[4]));
    out[1] = (int32_t)std::fmax(std::ceil(m[2]), std::ceil(m[6]));
    out[2] = (int32_t)std::fmin(std::floor(m[1]), std::floor(m[3]));
    out[3] = (int32_t)std::fmax(std::ceil(m[5]), std::ceil(m[7]));
  } else {
    out[0] = 0; out[1] = cols; out[2] = 0; out[3] = rows;
  }
  inv[0] = 64.0f / (float)(out[1] - out[0]);
  inv[1] = 48.0f / (float)(out[3] - out[2]);
}

}  // extern "C"
