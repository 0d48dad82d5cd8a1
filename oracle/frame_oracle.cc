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


// ---- N3: Frame::isInFrustum (reference src/Frame.cc:136-197), the projection / culling step before
// SearchByProjection (src/Tracking.cc:701-752).  The cv::Mat arithmetic is restated as OpenCV evaluates it for these
// shapes (pinned against cv2 4.13 by tests/test_frame_oracle.py):
//   mRcw*P+mtcw  -> cv::gemm(R, P, 1, t, 1): 3x3 * 3x1 takes the small-matrix branch (matmul.cpp, len == 3):
//                   t = a0*b0 + a1*b1 + a2*b2 in FLOAT, then (float)(t*alpha + c*beta) with double alpha, beta;
//   cv::norm(PO) -> sum of squares accumulated in DOUBLE in element order, sqrt, narrowed to float by the assignment;
//   PO.dot(Pn)   -> products accumulated in DOUBLE in element order; the division by dist happens in double.
struct fo_view {
  float Rcw[9], tcw[3], Ow[3];
  float fx, fy, cx, cy;
  int32_t min_x, max_x, min_y, max_y;   // Frame::mnMinX ... are static ints (include/Frame.h:121)
  int32_t nlevels;
  float scale[16];                      // mvScaleFactors
};

void fo_is_in_frustum(const fo_view* f, const float* pos, const float* normal, const float* min_dist,
                      const float* max_dist, int n, float cos_limit, uint8_t* in_view, float* u_out, float* v_out,
                      int32_t* level, float* view_cos) {
  for (int i = 0; i < n; i++) {
    in_view[i] = 0; u_out[i] = 0.f; v_out[i] = 0.f; level[i] = -1; view_cos[i] = 0.f;
    const float* P = pos + 3 * i;
    float Pc[3];
    for (int r = 0; r < 3; r++) {
      const float t = f->Rcw[3 * r] * P[0] + f->Rcw[3 * r + 1] * P[1] + f->Rcw[3 * r + 2] * P[2];
      Pc[r] = (float)(t * 1.0 + f->tcw[r] * 1.0);
    }
    if (Pc[2] < 0.0) continue;
    const float invz = 1.0 / Pc[2];
    const float u = f->fx * Pc[0] * invz + f->cx;
    const float v = f->fy * Pc[1] * invz + f->cy;
    if (u < f->min_x || u > f->max_x) continue;
    if (v < f->min_y || v > f->max_y) continue;
    const float PO[3] = {P[0] - f->Ow[0], P[1] - f->Ow[1], P[2] - f->Ow[2]};
    double s = 0;
    for (int k = 0; k < 3; k++) s += (double)PO[k] * PO[k];
    const float dist = std::sqrt(s);
    if (dist < min_dist[i] || dist > max_dist[i]) continue;
    double dot = 0;
    for (int k = 0; k < 3; k++) dot += (double)PO[k] * normal[3 * i + k];
    const float vc = dot / dist;
    if (vc < cos_limit) continue;
    const float ratio = dist / min_dist[i];
    int lv = 0;
    while (lv < f->nlevels && f->scale[lv] < ratio) lv++;   // std::lower_bound on the ascending mvScaleFactors
    if (lv >= f->nlevels) lv = f->nlevels - 1;
    in_view[i] = 1; u_out[i] = u; v_out[i] = v; level[i] = lv; view_cos[i] = vc;
  }
}
}  // extern "C"
