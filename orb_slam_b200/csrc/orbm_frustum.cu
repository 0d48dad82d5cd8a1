// N3: Frame::isInFrustum batched (reference src/Frame.cc:136-197; caller
// Tracking::SearchReferencePointsInFrustum src/Tracking.cc:701-752).  include/orbm.h.
// One thread per (view, map point).  The cv::Mat expressions are evaluated the way OpenCV evaluates them for these
// shapes, with individually rounded operations (no FMA contraction):
//   Pc = mRcw*P + mtcw   cv::gemm small-matrix branch: float dot of a row, then (float)(t*1.0 + c*1.0) in double
//   dist = cv::norm(PO)  double sum of squares in element order, sqrt, narrowed to float
//   viewCos              double dot in element order, divided by (double)dist, narrowed to float
#include "orbm_internal.cuh"

using namespace orbm;

namespace {

__global__ void __launch_bounds__(256) frustum_kernel(const orbm_frustum_view* __restrict__ views, int nviews,
                                                      const float* __restrict__ pos, const float* __restrict__ normal,
                                                      const float* __restrict__ min_dist,
                                                      const float* __restrict__ max_dist, int n, float cos_limit,
                                                      uint8_t* __restrict__ in_view, float* __restrict__ u_out,
                                                      float* __restrict__ v_out, int32_t* __restrict__ level,
                                                      float* __restrict__ view_cos) {
  __shared__ orbm_frustum_view s_view;
  const int w = blockIdx.y;
  for (int k = threadIdx.x; k < (int)(sizeof(orbm_frustum_view) / 4); k += blockDim.x)
    reinterpret_cast<uint32_t*>(&s_view)[k] = reinterpret_cast<const uint32_t*>(views + w)[k];
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const orbm_frustum_view& f = s_view;
  const int64_t o = (int64_t)w * n + i;
  uint8_t ok = 0;
  float u = 0.f, v = 0.f, vc = 0.f;
  int lv = -1;
  const float P0 = pos[3 * i], P1 = pos[3 * i + 1], P2 = pos[3 * i + 2];
  float Pc[3];
#pragma unroll
  for (int r = 0; r < 3; r++) {
    const float t = __fadd_rn(__fadd_rn(__fmul_rn(f.Rcw[3 * r], P0), __fmul_rn(f.Rcw[3 * r + 1], P1)),
                              __fmul_rn(f.Rcw[3 * r + 2], P2));
    Pc[r] = (float)__dadd_rn((double)t, (double)f.tcw[r]);   // t*1.0 and c*1.0 are exact
  }
  do {
    if (Pc[2] < 0.0f) break;
    const float invz = (float)__ddiv_rn(1.0, (double)Pc[2]);
    u = __fadd_rn(__fmul_rn(__fmul_rn(f.fx, Pc[0]), invz), f.cx);
    v = __fadd_rn(__fmul_rn(__fmul_rn(f.fy, Pc[1]), invz), f.cy);
    if (u < (float)f.min_x || u > (float)f.max_x) break;
    if (v < (float)f.min_y || v > (float)f.max_y) break;
    const float PO0 = __fsub_rn(P0, f.Ow[0]), PO1 = __fsub_rn(P1, f.Ow[1]), PO2 = __fsub_rn(P2, f.Ow[2]);
    double s = __dmul_rn((double)PO0, (double)PO0);           // 0 + x*x is exact
    s = __dadd_rn(s, __dmul_rn((double)PO1, (double)PO1));
    s = __dadd_rn(s, __dmul_rn((double)PO2, (double)PO2));
    const float dist = (float)__dsqrt_rn(s);
    const float mn = min_dist[i];
    if (dist < mn || dist > max_dist[i]) break;
    double dot = __dmul_rn((double)PO0, (double)normal[3 * i]);
    dot = __dadd_rn(dot, __dmul_rn((double)PO1, (double)normal[3 * i + 1]));
    dot = __dadd_rn(dot, __dmul_rn((double)PO2, (double)normal[3 * i + 2]));
    vc = (float)__ddiv_rn(dot, (double)dist);
    if (vc < cos_limit) break;
    const float ratio = __fdiv_rn(dist, mn);
    lv = 0;
    while (lv < f.nlevels && f.scale_factors[lv] < ratio) lv++;
    if (lv >= f.nlevels) lv = f.nlevels - 1;
    ok = 1;
  } while (false);
  if (!ok) { u = 0.f; v = 0.f; vc = 0.f; lv = -1; }
  in_view[o] = ok; u_out[o] = u; v_out[o] = v; level[o] = lv; view_cos[o] = vc;
}

}  // namespace

extern "C" int orbm_is_in_frustum(orbm_handle* h, const orbm_frustum_view* views, int nviews, const float* pos,
                                  const float* normal, const float* min_dist, const float* max_dist, int n,
                                  float cos_limit, uint8_t* in_view, float* u, float* v, int32_t* level,
                                  float* view_cos) {
  if (!h) return ORBM_ERR_ARG;
  if (!views || nviews < 1 || n < 0 || (n > 0 && (!pos || !normal || !min_dist || !max_dist)) || !in_view || !u || !v ||
      !level || !view_cos) {
    h->err = "is_in_frustum: null buffers / bad sizes";
    return ORBM_ERR_ARG;
  }
  for (int w = 0; w < nviews; w++)
    if (views[w].nlevels < 1 || views[w].nlevels > 16) { h->err = "is_in_frustum: nlevels must be 1..16"; return ORBM_ERR_ARG; }
  if (n == 0) return ORBM_OK;
  if (nviews > 65535) { h->err = "is_in_frustum: at most 65535 views per call"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  const size_t N = (size_t)n, T = (size_t)nviews * N;
  auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
  // inputs: views | pos | normal | min | max     outputs: u | v | cos | level | in_view
  const size_t i_views = 0, i_pos = up(sizeof(orbm_frustum_view) * nviews), i_nrm = i_pos + up(N * 12),
               i_mn = i_nrm + up(N * 12), i_mx = i_mn + up(N * 4), i_end = i_mx + up(N * 4);
  const size_t o_u = 0, o_v = up(T * 4), o_c = 2 * up(T * 4), o_l = 3 * up(T * 4), o_i = 4 * up(T * 4),
               o_end = o_i + up(T);
  ORBM_CK(h->stage_a.reserve(i_end));
  ORBM_CK(h->stage_b.reserve(o_end));
  char* a = h->stage_a.as<char>();
  char* b = h->stage_b.as<char>();
  ORBM_CK(cudaMemcpyAsync(a + i_views, views, sizeof(orbm_frustum_view) * nviews, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(a + i_pos, pos, N * 12, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(a + i_nrm, normal, N * 12, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(a + i_mn, min_dist, N * 4, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(a + i_mx, max_dist, N * 4, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaEventRecord(h->ev0, st));
  frustum_kernel<<<dim3((unsigned)((n + 255) / 256), (unsigned)nviews), 256, 0, st>>>(
      reinterpret_cast<const orbm_frustum_view*>(a + i_views), nviews, reinterpret_cast<const float*>(a + i_pos),
      reinterpret_cast<const float*>(a + i_nrm), reinterpret_cast<const float*>(a + i_mn),
      reinterpret_cast<const float*>(a + i_mx), n, cos_limit, reinterpret_cast<uint8_t*>(b + o_i),
      reinterpret_cast<float*>(b + o_u), reinterpret_cast<float*>(b + o_v), reinterpret_cast<int32_t*>(b + o_l),
      reinterpret_cast<float*>(b + o_c));
  h->launches++;
  ORBM_CK(cudaGetLastError());
  ORBM_CK(cudaEventRecord(h->ev1, st));
  ORBM_CK(cudaMemcpyAsync(u, b + o_u, T * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(v, b + o_v, T * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(view_cos, b + o_c, T * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(level, b + o_l, T * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(in_view, b + o_i, T, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaStreamSynchronize(st));
  ORBM_CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
  return ORBM_OK;
}
