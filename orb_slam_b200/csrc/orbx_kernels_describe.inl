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

// Persistent warps, one keypoint slot per warp per iteration, software-pipelined three deep so that no global-memory
// latency sits on a warp's critical path:
//   stage A (slot i+3)  plain loads of the slot's record (Cand) and of its frame's per-level counts
//   stage B (slot i+2)  decode (level, output index, centre), then cp.async of BOTH windows the slot needs into the
//                       warp's double-buffered shared memory: the 37x37 window of the BLURRED level the rotated
//                       pattern can reach (|offset| <= 18) and the 31x31 window of the UNBLURRED level for IC_Angle
//   stage C (slot i)    everything else, from shared memory only
//  E7  lane = COLUMN u of the 31-px disc, 31 byte reads per lane.
//  E9  lane = descriptor byte; the pattern sits in shared memory as float pairs (transposed, conflict-free), the 16
//      rotated samples are read from the staged window.  Windows touching the image border take the per-sample path:
//      in-image samples read the blurred level, the <= 2 px outside read the UNBLURRED reflect-101 border, which the
//      reference's in-place ROI blur leaves untouched (:759-760).
constexpr int kRawRows = 31, kRawWords = 9, kRawPitch = 4 * kRawWords;
constexpr int kBlurBufBytes = (kPatchRows * kPatchPitch + 15) & ~15, kRawBufBytes = (kRawRows * kRawPitch + 15) & ~15;
constexpr int kDescBufBytes = kBlurBufBytes + kRawBufBytes;
constexpr float kMagic = 12582912.f;          // 1.5 * 2^23: float(int) and cvRound(float) through FADD only
constexpr int kMagicBits = 0x4B400000;

struct DescRaw {   // stage A
  Cand kp;
  int my_cnt, frame, slot;   // slot < 0: past the end
};
struct DescRec {   // stage B -> C
  int frame, level, cx, cy, o, flags, mis;   // flags: 1 valid, 2 blurred window staged, 4 raw window staged
  float resp;
};

__device__ __forceinline__ void cp_async4(uint32_t dst_smem, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst_smem), "l"(__cvta_generic_to_global(src)) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// u8 x s8 dot product (IDP.4A)
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(32 * kDescWarps, 4) describe_kernel(ChunkBuffers b, orbx_keypoint* out_kps,
                                                                     uint8_t* out_desc, int32_t* out_n,
                                                                     int nslots_total) {
  __shared__ __align__(16) uint8_t s_buf[kDescWarps][2][kDescBufBytes];
  __shared__ float2 s_pat[16 * 32];            // [j][lane] = pattern point 16 * lane + j
  __shared__ uint32_t s_icm[8 * 32];           // [w][lane]: disc row v = lane - 15, byte k = 0xff iff column 4w+k-15 is inside
  const DevicePlan* P = b.plan;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  asm volatile("griddepcontrol.launch_dependents;");
  for (int t = threadIdx.x; t < 512; t += 32 * kDescWarps) {   // constant table: legal before the dependency wait
    const int pl = t & 31, pj = t >> 5, pt = 16 * pl + pj;
    s_pat[t] = make_float2((float)d_pattern[2 * pt], (float)d_pattern[2 * pt + 1]);
  }
  {
    const int t = threadIdx.x, w = t >> 5, l = t & 31;         // 256 threads = 8 words x 32 lanes
    uint32_t wm = 0;
    if (l < 31) {
      const int um = kUmax[l < kHalfPatch ? kHalfPatch - l : l - kHalfPatch];
      for (int k = 0; k < 4; k++) {
        const int u = 4 * w + k - kHalfPatch;
        if ((u < 0 ? -u : u) <= um) wm |= 0xffu << (8 * k);
      }
    }
    s_icm[t] = wm;
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");
  __syncthreads();
  const int nfeatures = P->nfeatures, nlevels = P->nlevels;
  const int my_base = lane < nlevels ? P->lv[lane].kp_base : 0x7fffffff;
  const int stride = gridDim.x * kDescWarps;
  const int stride_f = stride / nfeatures, stride_s = stride - stride_f * nfeatures;
  // stage A cursor: global slot gA = frameA * nfeatures + slotA, advanced by `stride` per iteration
  int gA = blockIdx.x * kDescWarps + warp;
  int frameA = gA / nfeatures, slotA = gA - frameA * nfeatures;

  const int bl_half = lane / kPatchWords, bl_j4 = 4 * (lane - bl_half * kPatchWords);   // blurred-window copy role
  const int rw_third = lane / kRawWords, rw_j4 = 4 * (lane - rw_third * kRawWords);     // raw-window copy role

  auto load_raw = [&]() {
    DescRaw r;
    r.kp.resp = 0.f; r.kp.xy = 0; r.my_cnt = 0; r.frame = frameA; r.slot = -1;
    if (gA < nslots_total) {
      r.slot = slotA;
      r.kp = b.kp[gA];                                        // slot gA of the [frame][nfeatures] array
      if (lane < nlevels) r.my_cnt = b.kp_count[frameA * kMaxLevels + lane];
    }
    gA += stride; frameA += stride_f; slotA += stride_s;
    if (slotA >= nfeatures) { slotA -= nfeatures; frameA++; }
    return r;
  };
  // stage B: decode the slot and start the asynchronous copies of its two windows into `buf`
  auto decode_issue = [&](const DescRaw& raw, uint8_t* buf) {
    DescRec r;
    r.flags = 0; r.frame = 0; r.level = 0; r.cx = 0; r.cy = 0; r.o = 0; r.mis = 0; r.resp = raw.kp.resp;
    if (raw.slot >= 0) {
      const int frame = raw.frame, slot = raw.slot;
      // level of the slot and output offset = sum of the lower levels' counts (lane = level)
      const int level = __popc(__ballot_sync(0xffffffffu, lane >= 1 && slot >= my_base));
      const int out_off = __reduce_add_sync(0xffffffffu, lane < level ? raw.my_cnt : 0);
      const int level_cnt = __shfl_sync(0xffffffffu, raw.my_cnt, level);
      const int idx = slot - __shfl_sync(0xffffffffu, my_base, level);
      if (slot == 0) {
        const int tot = __reduce_add_sync(0xffffffffu, raw.my_cnt);
        if (lane == 0) out_n[frame] = tot;
      }
      if (idx < level_cnt) {
        const LevelPlan& L = P->lv[level];
        const int cx = raw.kp.xy & 0xffff, cy = raw.kp.xy >> 16;
        const int lw = L.w, lh = L.h, lpitch = L.pitch;
        r.frame = frame; r.level = level; r.cx = cx; r.cy = cy; r.o = out_off + idx; r.flags = 1;
        const uint32_t sdst = (uint32_t)__cvta_generic_to_shared(buf);
        if (cx >= kPatchR + 3 && cx + kPatchR + 8 <= lpitch && cx + kPatchR < lw && cy >= kPatchR && cy + kPatchR < lh) {
          const uint8_t* w0 = b.blur + (int64_t)frame * P->blur_frame_bytes + L.off + (int64_t)(cy - kPatchR) * lpitch +
                              (cx - kPatchR);
          const int mis = (int)((uintptr_t)w0 & 3);
          const uint8_t* src = w0 - mis;
          r.mis = mis; r.flags |= 2;
          // two window rows per round: lanes 0-10 one row, lanes 11-21 the next
          if (bl_half < 2) {
            const uint8_t* p = src + bl_half * lpitch + bl_j4;
            const uint32_t d = sdst + bl_half * kPatchPitch + bl_j4;
#pragma unroll
            for (int r0 = 0; r0 < kPatchRows; r0 += 2) {
              if (r0 + bl_half < kPatchRows) cp_async4(d + r0 * kPatchPitch, p);
              p += 2 * lpitch;
            }
          }
        }
        int pitch;
        const uint8_t* img = level_ptr(b, P, level, frame, &pitch);
        // (a cell grid whose ceil() overshoots the border puts corners closer than 15 px to the right / bottom edge:
        // their disc reaches the reference's reflect-101 frame (:806-814) and takes the per-pixel path in stage C)
        if ((pitch & 3) == 0 && cx >= kHalfPatch && cx + kHalfPatch < lw && cy >= kHalfPatch && cy + kHalfPatch < lh) {
          // the window's misalignment is the same on every row
          const uint8_t* w0 = img + (int64_t)(cy - kHalfPatch) * pitch + (cx - kHalfPatch);
          const int mis = (int)((uintptr_t)w0 & 3);
          const uint8_t* src = w0 - mis;
          r.mis |= mis << 8; r.flags |= 4;
          // three window rows per round: lanes 0-8, 9-17, 18-26
          if (rw_third < 3) {
            const uint8_t* p = src + rw_third * pitch + rw_j4;
            const uint32_t d = sdst + kBlurBufBytes + rw_third * kRawPitch + rw_j4;
#pragma unroll
            for (int r0 = 0; r0 < kRawRows; r0 += 3) {
              if (r0 + rw_third < kRawRows) cp_async4(d + r0 * kRawPitch, p);
              p += 3 * pitch;
            }
          }
        }
      }
    }
    cp_async_commit();            // one group per slot, empty or not, so that wait_group counts slots
    return r;
  };
  // stage C: orientation, descriptor, keypoint record
  auto compute = [&](const DescRec& r, const uint8_t* buf) {
    if (!(r.flags & 1)) return;
    const int frame = r.frame, level = r.level, cx = r.cx, cy = r.cy;
    const LevelPlan& L = P->lv[level];
    // ---- E7 IC_Angle (:124-151): m10 = sum u*I, m01 = sum v*I over the 749-px disc, UNBLURRED level
    int m10 = 0, m01 = 0;
    if (r.flags & 4) {
      // lane = ROW v of the staged window: its 9 words are shifted to start at column -15 and masked to the disc, then
      // every word is two dot products (ones -> row sum, column numbers -> sum of u * I), 4 pixels per IDP
      const uint32_t* row = reinterpret_cast<const uint32_t*>(buf + kBlurBufBytes) + min(lane, kRawRows - 1) * kRawWords;
      uint32_t d[kRawWords];
#pragma unroll
      for (int w = 0; w < kRawWords; w++) d[w] = row[w];
      const int sh = ((r.mis >> 8) & 3) * 8;
      uint32_t rowsum = 0;
#pragma unroll
      for (int w = 0; w < 8; w++) {
        const uint32_t x = __funnelshift_r(d[w], d[w + 1], sh) & s_icm[w * 32 + lane];
        const int u0 = 4 * w - kHalfPatch;                     // columns u0 .. u0 + 3 as signed bytes
        const uint32_t cols = (uint32_t)(u0 & 0xff) | ((uint32_t)((u0 + 1) & 0xff) << 8) |
                              ((uint32_t)((u0 + 2) & 0xff) << 16) | ((uint32_t)((u0 + 3) & 0xff) << 24);
        rowsum = __dp4a(x, 0x01010101u, rowsum);
        m10 = dp4a_us(x, cols, m10);
      }
      m01 = (lane - kHalfPatch) * (int)rowsum;
    } else if (lane < 31) {
      const int u = lane - kHalfPatch, au = u < 0 ? -u : u;
      int colsum = 0;
      int pitch;
      const uint8_t* img = level_ptr(b, P, level, frame, &pitch);
      if (cx >= kHalfPatch && cx + kHalfPatch < L.w && cy >= kHalfPatch && cy + kHalfPatch < L.h) {   // warp-uniform
        const uint8_t* c = img + (int64_t)cy * pitch + cx + u;
#pragma unroll
        for (int v = -kHalfPatch; v <= kHalfPatch; v++) {
          if (au <= kUmax[v < 0 ? -v : v]) {
            const int I = c[v * pitch];
            colsum += I;
            m01 += v * I;
          }
        }
      } else {                                                  // the disc reaches the reflect-101 frame of the level
        const uint8_t* c = img + reflect101(cx + u, L.w);
#pragma unroll 1
        for (int v = -kHalfPatch; v <= kHalfPatch; v++) {
          if (au <= kUmax[v < 0 ? -v : v]) {
            const int I = c[(int64_t)reflect101(cy + v, L.h) * pitch];
            colsum += I;
            m01 += v * I;
          }
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
    int val = 0;
    if (r.flags & 2) {
      // sample (rx, ry) sits at window byte (ry + 18) * 44 + rx + 18 + mis >= 0; rx, ry come out of the magic-number
      // rounding as kMagicBits + value, so the constant part of the index is folded into the offset (mod 2^32)
      const uint32_t base_off = (uint32_t)(kPatchR * kPatchPitch + kPatchR + (r.mis & 3)) -
                                (uint32_t)(kPatchPitch + 1) * (uint32_t)kMagicBits;
#pragma unroll
      for (int j = 0; j < 8; j++) {
        int t[2];
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const float2 p = s_pat[(2 * j + e) * 32 + lane];
          // float -> int rounds half to even exactly like cvRound; every product and sum individually rounded (:165)
          const int ryb = __float_as_int(__fadd_rn(__fadd_rn(__fmul_rn(p.x, sa), __fmul_rn(p.y, ca)), kMagic));
          const int rxb = __float_as_int(__fadd_rn(__fsub_rn(__fmul_rn(p.x, ca), __fmul_rn(p.y, sa)), kMagic));
          t[e] = buf[(uint32_t)ryb * (uint32_t)kPatchPitch + (uint32_t)rxb + base_off];
        }
        val |= (t[0] < t[1]) << j;
      }
    } else {
      int pitch;
      const uint8_t* img = level_ptr(b, P, level, frame, &pitch);
      const int lw = L.w, lh = L.h, lpitch = L.pitch;
      const uint8_t* blur = b.blur + (int64_t)frame * P->blur_frame_bytes + L.off;
#pragma unroll 1
      for (int j = 0; j < 8; j++) {
        int t[2];
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const float2 p = s_pat[(2 * j + e) * 32 + lane];
          const int ry = __float_as_int(__fadd_rn(__fadd_rn(__fmul_rn(p.x, sa), __fmul_rn(p.y, ca)), kMagic)) - kMagicBits;
          const int rx = __float_as_int(__fadd_rn(__fsub_rn(__fmul_rn(p.x, ca), __fmul_rn(p.y, sa)), kMagic)) - kMagicBits;
          const int x = cx + rx, y = cy + ry;
          if (x >= 0 && x < lw && y >= 0 && y < lh) t[e] = blur[(int64_t)y * lpitch + x];
          else t[e] = img[(int64_t)reflect101(y, lh) * pitch + reflect101(x, lw)];
        }
        val |= (t[0] < t[1]) << j;
      }
    }
    out_desc[((int64_t)frame * nfeatures + r.o) * 32 + lane] = (uint8_t)val;

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
        case 4: w = __float_as_uint(r.resp); break;
        case 5: w = (uint32_t)level; break;
        default: w = 0xffffffffu; break;
      }
      reinterpret_cast<uint32_t*>(out_kps + (int64_t)frame * nfeatures + r.o)[lane] = w;
    }
  };

  // ---- pipeline ---------------------------------------------------------------------------------------------------------
  uint8_t* buf0 = s_buf[warp][0];
  uint8_t* buf1 = s_buf[warp][1];
  DescRaw raw = load_raw();
  DescRec rec0 = decode_issue(raw, buf0);
  raw = load_raw();
  DescRec rec1 = decode_issue(raw, buf1);
  raw = load_raw();
  for (int g = blockIdx.x * kDescWarps + warp; g < nslots_total; g += stride) {
    cp_async_wait<1>();           // every group but the newest has landed: this slot's windows are complete
    __syncwarp();
    compute(rec0, buf0);
    __syncwarp();                 // all lanes are done reading buf0 before it is refilled
    rec0 = rec1;
    rec1 = decode_issue(raw, buf0);
    raw = load_raw();
    uint8_t* tmp = buf0; buf0 = buf1; buf1 = tmp;
  }
  cp_async_wait<0>();
}

void launch_describe(const ChunkBuffers& b, const DevicePlan& hp, int nframes, orbx_keypoint* d_kps, uint8_t* d_desc,
                     int32_t* d_n_out, cudaStream_t s) {
  const int nslots = hp.nfeatures * nframes;
  static int max_ctas = 0;
  if (!max_ctas) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, describe_kernel, 32 * kDescWarps, 0);
    if (const char* e = getenv("ORBX_DESC_CTAS_PER_SM")) per_sm = std::max(1, std::min(per_sm, atoi(e)));   // co-residency experiments
    max_ctas = sms * (per_sm > 0 ? per_sm : 1);
  }
  const int ctas = (nslots + kDescWarps - 1) / kDescWarps;
  launch_k(describe_kernel, dim3(ctas < max_ctas ? ctas : max_ctas), dim3(32 * kDescWarps), 0, s, b, d_kps, d_desc, d_n_out, nslots);
}

}  // namespace orbx
