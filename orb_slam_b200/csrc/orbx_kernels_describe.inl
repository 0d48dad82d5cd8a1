// Included by orbx_kernels.cu.  E7 IC_Angle, E9 rotated BRIEF, E10 output assembly: one warp per keypoint slot.
namespace orbx {

__device__ __align__(16) int8_t d_pattern[1024] = {
#include "orb_pattern_31.inc"
};

// cv::fastAtan2(y, x): float32 polynomial, every operation individually rounded (reference :150)
__device__ __forceinline__ float fast_atan2_dev(float y, float x) {
  const float s = (float)(180.0 / 3.14159265358979323846);
  const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s, p5 = 0.1555786518463281f * s,
              p7 = -0.04432655554792128f * s;  // folded at compile time in float, like the host's statics
  const float eps = (float)DBL_EPSILON;
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3),
                                                      c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

// glibc >= 2.28 sinf/cosf (ARM optimized-routines sincosf: double polynomial, rounded once to float), the functions
// `(float)cos(angle)`, `(float)sin(angle)` resolve to at reference :159-160.  Verified over EVERY float in
// [0, 2*pi*1.0001] against the host libm (tools/check_sincosf.c; tests/test_gpu_extract.py re-checks on the device).
__device__ __forceinline__ void sincosf_glibc(float y, float* sn, float* cs) {
  const double hpi_inv = 0x1.45F306DC9C883p+23, hpi = 0x1.921FB54442D18p0;
  const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
               C4 = 0x1.99343027bf8c3p-16;
  const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
  double x = (double)y;
  const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
  int n = 0;
  double sgn = 1.0, neg = 1.0;
  if (top < ((__float_as_uint(0x1.921FB6p-1f) >> 20) & 0x7ff)) {
    if (top < ((__float_as_uint(0x1p-12f) >> 20) & 0x7ff)) { *sn = y; *cs = 1.0f; return; }
  } else {
    const double r = x * hpi_inv;
    n = ((int32_t)r + 0x800000) >> 24;
    x = fma(-(double)n, hpi, x);
    sgn = ((n + 1) & 2) ? -1.0 : 1.0;  // sign table {1,-1,-1,1}[n & 3]
    neg = (n & 2) ? -1.0 : 1.0;        // second coefficient table = first with the cosine coefficients negated
  }
  const double xs = x * sgn, x2 = x * x;
  // sine polynomial
  double x3 = xs * x2, s1 = fma(x2, S3, S2), x7 = x3 * x2, s = fma(x3, S1, xs);
  const float fs = (float)fma(x7, s1, s);
  // cosine polynomial
  double x4 = x2 * x2, c2 = fma(x2, neg * C4, neg * C3), c1 = fma(x2, neg * C1, neg * C0), x6 = x4 * x2,
         c = fma(x4, neg * C2, c1);
  const float fc = (float)fma(x6, c2, c);
  if ((n & 1) == 0) { *sn = fs; *cs = fc; }
  else { *sn = fc; *cs = fs; }
}

// umax of the 31-px disc (reference :495-510); build_plan() verifies that the table it computes with the reference's
// code equals this one.
__device__ constexpr int kUmax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

constexpr int kDescWarps = 8, kPatchR = 18, kPatchRows = 2 * kPatchR + 1, kPatchWords = 11, kPatchPitch = 4 * kPatchWords;

// One warp per keypoint slot.
//  E7  lane = COLUMN u of the 31-px disc: 31 row loads, each one coalesced 31-byte segment (the former lane = row
//      layout cost 31 scattered sectors per load instruction and made the kernel L1-wavefront bound).
//  E9  the 37x37 window of the BLURRED level that the rotated pattern can reach (|offset| <= 18) is staged into
//      shared memory by aligned 32-bit loads (2 rows per instruction) and the 512 samples are read from there;
//      the pattern lives in registers (two 16-byte loads per lane) instead of 32-way divergent __constant__ reads.
//      Windows touching the image border take the per-sample path: in-image samples read the blurred level, the
//      <= 2 px outside read the UNBLURRED reflect-101 border, which the reference's in-place ROI blur leaves
//      untouched (:759-760).
__global__ void __launch_bounds__(32 * kDescWarps) describe_kernel(ChunkBuffers b, orbx_keypoint* out_kps,
                                                                  uint8_t* out_desc, int32_t* out_n,
                                                                  int nslots_total) {
  __shared__ __align__(16) uint8_t s_patch[kDescWarps][kPatchRows * kPatchPitch];
  const DevicePlan* P = b.plan;
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (gw >= nslots_total) return;
  const int frame = gw / P->nfeatures, slot = gw - frame * P->nfeatures;
  // level of the slot and output offset = sum of the lower levels' counts: one parallel round of loads (lane = level)
  // instead of a chain of dependent ones
  const int32_t* cnt = b.kp_count + frame * kMaxLevels;
  const int my_cnt = lane < P->nlevels ? cnt[lane] : 0;
  const int my_base = lane < P->nlevels ? P->lv[lane].kp_base : 0x7fffffff;
  const int level = __popc(__ballot_sync(0xffffffffu, lane >= 1 && slot >= my_base));
  const int out_off = __reduce_add_sync(0xffffffffu, lane < level ? my_cnt : 0);
  const int level_cnt = __shfl_sync(0xffffffffu, my_cnt, level);
  const LevelPlan& L = P->lv[level];
  const int idx = slot - L.kp_base;
  {
    const int tot = __reduce_add_sync(0xffffffffu, my_cnt);
    if (slot == 0 && lane == 0) out_n[frame] = tot;
  }
  if (idx >= level_cnt) return;
  const Cand kp = b.kp[(int64_t)frame * P->nfeatures + slot];
  const int cx = kp.xy & 0xffff, cy = kp.xy >> 16;
  int pitch;
  const uint8_t* img = level_ptr(b, P, level, frame, &pitch);
  const int lw = L.w, lh = L.h, lpitch = L.pitch;
  const uint8_t* blur = b.blur + (int64_t)frame * P->blur_frame_bytes + L.off;

  // pattern of this lane's descriptor byte: 16 points x (x, y) int8
  const uint4 pa = reinterpret_cast<const uint4*>(d_pattern)[2 * lane], pb = reinterpret_cast<const uint4*>(d_pattern)[2 * lane + 1];
  const uint32_t pat[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w};

  // ---- E9 staging first (independent of the angle): the loads overlap the orientation arithmetic
  const bool fast_path = cx >= kPatchR + 3 && cx + kPatchR + 8 <= lpitch && cx + kPatchR < lw && cy >= kPatchR &&
                         cy + kPatchR < lh;
  int mis = 0;
  if (fast_path) {
    const uint8_t* w0 = blur + (int64_t)(cy - kPatchR) * lpitch + (cx - kPatchR);
    mis = (int)((uintptr_t)w0 & 3);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(w0 - mis);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_patch[warp]);
    const int half = lane >> 4, j = lane & 15, wpitch = lpitch >> 2;
    if (j < kPatchWords) {
#pragma unroll
      for (int r0 = 0; r0 < kPatchRows; r0 += 2) {
        const int r = r0 + half;
        if (r < kPatchRows) dst[r * kPatchWords + j] = src[r * wpitch + j];
      }
    }
  }

  // ---- E7 IC_Angle (:124-151): m10 = sum u*I, m01 = sum v*I over the 749-px disc, UNBLURRED level
  int m10 = 0, m01 = 0;
  if (lane < 31) {
    const int u = lane - kHalfPatch, au = u < 0 ? -u : u;
    const uint8_t* c = img + (int64_t)cy * pitch + cx + u;
    int colsum = 0;
#pragma unroll
    for (int v = -kHalfPatch; v <= kHalfPatch; v++) {
      if (au <= kUmax[v < 0 ? -v : v]) {
        const int I = c[v * pitch];
        colsum += I;
        m01 += v * I;
      }
    }
    m10 = u * colsum;
  }
  m10 = __reduce_add_sync(0xffffffffu, m10);
  m01 = __reduce_add_sync(0xffffffffu, m01);
  const float angle = fast_atan2_dev((float)m01, (float)m10);

  // ---- E9 computeOrbDescriptor (:155-194): lane = descriptor byte, 8 comparisons of rotated pattern samples
  const float factorPI = (float)(3.14159265358979323846 / 180.f);
  float sa, ca;
  sincosf_glibc(__fmul_rn(angle, factorPI), &sa, &ca);
  __syncwarp();
  const uint8_t* sp = s_patch[warp] + kPatchR * kPatchPitch + kPatchR + mis;
  int val = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    int t[2];
#pragma unroll
    for (int e = 0; e < 2; e++) {
      const uint32_t w = pat[j] >> (16 * e);
      // exact small-integer <-> float conversions through the 1.5 * 2^23 trick (FADD/IADD instead of the XU pipe's
      // I2F / F2I); float -> int rounds half to even exactly like cvRound
      const float px = __fsub_rn(__int_as_float(0x4B400000 + (int)(int8_t)(w & 0xff)), 12582912.f);
      const float py = __fsub_rn(__int_as_float(0x4B400000 + (int)(int8_t)((w >> 8) & 0xff)), 12582912.f);
      const int ry = __float_as_int(__fadd_rn(__fadd_rn(__fmul_rn(px, sa), __fmul_rn(py, ca)), 12582912.f)) - 0x4B400000;
      const int rx = __float_as_int(__fadd_rn(__fsub_rn(__fmul_rn(px, ca), __fmul_rn(py, sa)), 12582912.f)) - 0x4B400000;
      if (fast_path) {
        t[e] = sp[ry * kPatchPitch + rx];
      } else {
        const int x = cx + rx, y = cy + ry;
        if (x >= 0 && x < lw && y >= 0 && y < lh) t[e] = blur[(int64_t)y * lpitch + x];
        else t[e] = img[(int64_t)reflect101(y, lh) * pitch + reflect101(x, lw)];
      }
    }
    val |= (t[0] < t[1]) << j;
  }
  const int o = out_off + idx;
  out_desc[((int64_t)frame * P->nfeatures + o) * 32 + lane] = (uint8_t)val;

  // ---- E10 keypoint record (:687-694, :769-775)
  if (lane < 7) {
    float fx = (float)cx, fy = (float)cy;
    if (level != 0) { fx = __fmul_rn(fx, L.scale); fy = __fmul_rn(fy, L.scale); }
    uint32_t w;
    switch (lane) {
      case 0: w = __float_as_uint(fx); break;
      case 1: w = __float_as_uint(fy); break;
      case 2: w = __float_as_uint(L.size); break;
      case 3: w = __float_as_uint(angle); break;
      case 4: w = __float_as_uint(kp.resp); break;
      case 5: w = (uint32_t)level; break;
      default: w = 0xffffffffu; break;
    }
    reinterpret_cast<uint32_t*>(out_kps + (int64_t)frame * P->nfeatures + o)[lane] = w;
  }
}

void launch_describe(const ChunkBuffers& b, const DevicePlan& hp, int nframes, orbx_keypoint* d_kps, uint8_t* d_desc,
                     int32_t* d_n_out, cudaStream_t s) {
  const int nslots = hp.nfeatures * nframes;
  describe_kernel<<<(nslots + kDescWarps - 1) / kDescWarps, 32 * kDescWarps, 0, s>>>(b, d_kps, d_desc, d_n_out, nslots);
}

}  // namespace orbx
