// Device views + the literal float arithmetic of the reference's grid queries, shared by orbm_frames.cu and
// orbm_match.cu.
#pragma once
#include <climits>

#include "orbm_internal.cuh"

namespace orbm {

struct FrameSetView {
  const int32_t* frame_off;
  const float *x, *y, *angle;
  const int32_t* octave;
  const uint4* desc;
  const uint8_t* valid;
  const float4* gridp;
  int32_t* cell_start;
  int32_t* items;
  const int32_t *node_off, *fvitem_off, *node_id, *node_start, *fv_items;
};

FrameSetView view_of(const orbm_frameset* s);

// Frame::PosInGrid (src/Frame.cc:266-276): round() = half away from zero; -1 when outside the 64x48 grid
__device__ __forceinline__ int grid_cell_of(float x, float y, const float4& gp) {
  const int px = (int)roundf(__fmul_rn(__fsub_rn(x, gp.x), gp.z));
  const int py = (int)roundf(__fmul_rn(__fsub_rn(y, gp.y), gp.w));
  if (px < 0 || px >= kGridCols || py < 0 || py >= kGridRows) return -1;
  return px * kGridRows + py;
}

// Cell window of Frame::GetFeaturesInArea (src/Frame.cc:204-222); false = the early `return vIndices`
__device__ __forceinline__ bool area_cells(float x, float y, float r, const float4& gp, int* cx0, int* cx1, int* cy0,
                                           int* cy1) {
  int a = (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, gp.x), r), gp.z));
  a = max(0, a);
  if (a >= kGridCols) return false;
  int b = (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, gp.x), r), gp.z));
  b = min(kGridCols - 1, b);
  if (b < 0) return false;
  int c = (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, gp.y), r), gp.w));
  c = max(0, c);
  if (c >= kGridRows) return false;
  int d = (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, gp.y), r), gp.w));
  d = min(kGridRows - 1, d);
  if (d < 0) return false;
  *cx0 = a; *cx1 = b; *cy0 = c; *cy1 = d;
  return true;
}

// Level filter + box test of src/Frame.cc:224-255 (minLevel == maxLevel == -1: no level check, which is also
// KeyFrame::GetFeaturesInArea, src/KeyFrame.cc:645)
__device__ __forceinline__ bool area_accept(float kx, float ky, int oct, float x, float y, float r, int minl,
                                            int maxl) {
  if (!(minl == -1 && maxl == -1)) {
    if (minl == maxl) { if (oct != minl) return false; }
    else if (oct < minl || oct > maxl) return false;
  }
  return !(fabsf(__fsub_rn(kx, x)) > r || fabsf(__fsub_rn(ky, y)) > r);
}

}  // namespace orbm
