// N1: Frame::UndistortKeyPoints / ComputeImageBounds, i.e. cv::undistortPoints(src, dst, K, D, Mat(), K)
// for the 5-coefficient model the reference builds (src/Tracking.cc:65-71: k1, k2, p1, p2, k3; K and D are CV_32F).
// Restated from the published OpenCV algorithm (cvUndistortPoints, calib3d/imgproc undistort): everything in double,
// FIVE fixed-point iterations (the default TermCriteria(MAX_ITER, 5, 0.01) of the overload the reference calls; the
// same loop count is hard-coded in OpenCV 2.4), identity R, P = K.  Every operation is individually rounded, in
// source order: the device path uses the _rn intrinsics so that nvcc cannot contract a*b+c into an FMA, the host
// path (image bounds, four points per camera) is plain C++ built without FMA contraction (x86-64 baseline).
// Terms that are exactly zero for this model (k4..k6, s1..s4, tilt) are omitted: they add +0.0 or multiply by 1.
#pragma once
#include <cuda_runtime.h>

namespace orbm {

struct Camera {     // values as the CV_32F matrices hold them
  float fx, fy, cx, cy;
  float k1, k2, p1, p2, k3;
};

#ifdef __CUDA_ARCH__
#define ORBM_DMUL(a, b) __dmul_rn((a), (b))
#define ORBM_DADD(a, b) __dadd_rn((a), (b))
#define ORBM_DSUB(a, b) __dsub_rn((a), (b))
#define ORBM_DDIV(a, b) __ddiv_rn((a), (b))
#else
#define ORBM_DMUL(a, b) ((a) * (b))
#define ORBM_DADD(a, b) ((a) + (b))
#define ORBM_DSUB(a, b) ((a) - (b))
#define ORBM_DDIV(a, b) ((a) / (b))
#endif

__host__ __device__ inline void undistort_point(const Camera& c, float xin, float yin, float* xo, float* yo) {
  const double fx = (double)c.fx, fy = (double)c.fy, cx = (double)c.cx, cy = (double)c.cy;
  const double ifx = ORBM_DDIV(1.0, fx), ify = ORBM_DDIV(1.0, fy);
  const double k1 = (double)c.k1, k2 = (double)c.k2, p1 = (double)c.p1, p2 = (double)c.p2, k3 = (double)c.k3;
  const double u = (double)xin, v = (double)yin;
  double x = ORBM_DMUL(ORBM_DSUB(u, cx), ifx), y = ORBM_DMUL(ORBM_DSUB(v, cy), ify);
  const double x0 = x, y0 = y;
  for (int j = 0; j < 5; j++) {
    const double r2 = ORBM_DADD(ORBM_DMUL(x, x), ORBM_DMUL(y, y));
    // icdist = (1 + 0) / (1 + ((k3*r2 + k2)*r2 + k1)*r2)
    const double den = ORBM_DADD(1.0, ORBM_DMUL(ORBM_DADD(ORBM_DMUL(ORBM_DADD(ORBM_DMUL(k3, r2), k2), r2), k1), r2));
    const double icdist = ORBM_DDIV(1.0, den);
    if (icdist < 0) {  // OpenCV >= 3.4.x gives up on a diverging point and returns the normalised input
      x = ORBM_DMUL(ORBM_DSUB(u, cx), ifx);
      y = ORBM_DMUL(ORBM_DSUB(v, cy), ify);
      break;
    }
    // deltaX = 2*p1*x*y + p2*(r2 + 2*x*x);  deltaY = p1*(r2 + 2*y*y) + 2*p2*x*y   (left-to-right products)
    const double dX = ORBM_DADD(ORBM_DMUL(ORBM_DMUL(ORBM_DMUL(2.0, p1), x), y),
                                ORBM_DMUL(p2, ORBM_DADD(r2, ORBM_DMUL(ORBM_DMUL(2.0, x), x))));
    const double dY = ORBM_DADD(ORBM_DMUL(p1, ORBM_DADD(r2, ORBM_DMUL(ORBM_DMUL(2.0, y), y))),
                                ORBM_DMUL(ORBM_DMUL(ORBM_DMUL(2.0, p2), x), y));
    x = ORBM_DMUL(ORBM_DSUB(x0, dX), icdist);
    y = ORBM_DMUL(ORBM_DSUB(y0, dY), icdist);
  }
  // P = K, R = I:  xx = fx*x + 0*y + cx, ww = 1 / (0*x + 0*y + 1) = 1
  *xo = (float)ORBM_DADD(ORBM_DMUL(fx, x), cx);
  *yo = (float)ORBM_DADD(ORBM_DMUL(fy, y), cy);
}

}  // namespace orbm
