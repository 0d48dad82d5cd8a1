// Included by orbx_kernels.cu.  E3/E4 per-cell FAST detection and E5/E6 quota + retainBest selection.
namespace orbx {

// ------------------------------------------------------------------------------------------------------------------
// E3  FAST-9/16 score (threshold-free):  S = max over the 16 arcs of 9 contiguous ring pixels of
//     min(d) and min(-d), minus 1, d_k = I(p) - I(ring_k); p is a corner at threshold t  <=>  S >= t
//     (cv::FAST / cornerScore<16>).
// TWO horizontally adjacent pixels are scored at once: every value is a zero-extended u16 in one half of a 32-bit
// register and the min/max ladder runs on the DPX 3-input packed instructions (VIMNMX3.U16x2):
//     m3[k] = min3(r[k], r[k+1], r[k+2]);  m9[k] = min3(m3[k], m3[k+3], m3[k+6])  = min of the arc k..k+8
// 40 packed instructions per side per pixel PAIR instead of ~75 scalar ones per pixel.
// NOTE (CUDA 12.9 ptxas, sm_100a): max(min9(d), -max9(d)) with d = I(p) - ring is MISCOMPILED — ptxas fuses the
// neg.s32 into a 3-input VIMNMX3 and drops the negation (tools/ptxas_vimnmx3_neg_repro.cu).  The score is therefore
// computed on the RAW ring values with no negation anywhere:
//     S = max( max_k min9(ring) - v ,  v - min_k max9(ring) ) - 1.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void fast_score_x2(const uint32_t (&r)[16], uint32_t v2, int tmin, int* s0, int* s1) {
  uint32_t mn3[16], mx3[16];
#pragma unroll
  for (int k = 0; k < 16; k++) {
    mn3[k] = __vimin3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
    mx3[k] = __vimax3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
  }
  uint32_t mn9[16], mx9[16];
#pragma unroll
  for (int k = 0; k < 16; k++) {
    mn9[k] = __vimin3_u16x2(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]);
    mx9[k] = __vimax3_u16x2(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]);
  }
  uint32_t b = __vimax3_u16x2(mn9[0], mn9[1], mn9[2]), d = __vimin3_u16x2(mx9[0], mx9[1], mx9[2]);
#pragma unroll
  for (int k = 3; k < 15; k += 2) {
    b = __vimax3_u16x2(b, mn9[k], mn9[k + 1]);
    d = __vimin3_u16x2(d, mx9[k], mx9[k + 1]);
  }
  b = __vmaxu2(b, mn9[15]);
  d = __vminu2(d, mx9[15]);
  const int v0 = v2 & 0xffff, v1 = v2 >> 16;
  const int a0 = max((int)(b & 0xffff) - v0, v0 - (int)(d & 0xffff)) - 1;
  const int a1 = max((int)(b >> 16) - v1, v1 - (int)(d >> 16)) - 1;
  *s0 = a0 >= tmin ? a0 : 0;
  *s1 = a1 >= tmin ? a1 : 0;
}

// E4  HarrisResponses(cellImage, kps, 7, 0.04f)   (reference :79-120); p = level image at the keypoint
__device__ float harris_response(const uint8_t* p, int step) {
  int a = 0, b = 0, c = 0;
  for (int i = -3; i <= 3; i++)
    for (int j = -3; j <= 3; j++) {
      const uint8_t* q = p + i * step + j;
      const int Ix = (q[1] - q[-1]) * 2 + (q[-step + 1] - q[-step - 1]) + (q[step + 1] - q[step - 1]);
      const int Iy = (q[step] - q[-step]) * 2 + (q[step - 1] - q[-step - 1]) + (q[step + 1] - q[-step + 1]);
      a += Ix * Ix; b += Iy * Iy; c += Ix * Iy;
    }
  float scale = __fmul_rn((float)(1 << 2) * 7, 255.0f);
  scale = __fdiv_rn(1.0f, scale);
  const float s4 = __fmul_rn(__fmul_rn(__fmul_rn(scale, scale), scale), scale);
  const float fa = (float)a, fb = (float)b, fc = (float)c;
  const float apb = __fadd_rn(fa, fb);
  const float r = __fsub_rn(__fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc)), __fmul_rn(__fmul_rn(0.04f, apb), apb));
  return __fmul_rn(r, s4);
}

// ------------------------------------------------------------------------------------------------------------------
// One CTA per (cell, frame): FAST score of the cell's scored region, 3x3 strict NMS that never looks outside the
// region (the reference runs cv::FAST per cell sub-image, :599-607), the "<= 3 corners -> redo with threshold 7"
// rule (:609-614) on the same score map, then an ORDER-PRESERVING compaction (row-major, the input order of
// retainBest's nth_element) into the cell's candidate list.
//
// Shared memory: the cell sub-image as zero-extended u16 (sub-image column c at halfword c + 5, so scored column 0
// is 16-byte aligned), and the u8 score map with a zero ring.  A thread scores GROUPS of 4 adjacent pixels: per ring
// row it loads 12 halfwords (3 x LDS.64); ring values of the two pixel pairs are whole registers (even dx) or one
// PRMT of two neighbouring registers (odd dx).
// ------------------------------------------------------------------------------------------------------------------
constexpr int kFastThreads = 256;

__device__ __forceinline__ uint32_t hw_mid(uint32_t lo, uint32_t hi) { return __byte_perm(lo, hi, 0x5432); }

__global__ void __launch_bounds__(kFastThreads) fast_cells_kernel(ChunkBuffers b) {
  extern __shared__ __align__(16) uint8_t smem[];
  const DevicePlan* P = b.plan;
  const int cell = blockIdx.x, frame = blockIdx.y;
  const CellPlan c = b.cells[cell];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int32_t* n_total = b.n_total + (int64_t)frame * P->total_cells + cell;
  if (c.rw <= 0) {
    if (tid == 0) *n_total = 0;
    return;
  }
  __shared__ int s_cnt[2];
  __shared__ int s_warp[kFastThreads / 32];
  int pitch;
  const uint8_t* img = level_ptr(b, P, c.level, frame, &pitch);
  const int rw = c.rw, rh = c.rh, cw = rw + 6, ch = rh + 6;   // cw x ch = the sub-image handed to cv::FAST
  const int sp16 = fast_sp16(cw), scp = fast_scp(rw);
  uint16_t* s_img = reinterpret_cast<uint16_t*>(smem);          // [ch][sp16]
  uint8_t* s_sc = smem + fast_img_bytes(cw, ch);                // [rh + 2][scp], scored column x at byte 4 + x
  const int G = (rw + 3) >> 2;                                  // 4-pixel groups per scored row

  // ---- stage the sub-image: aligned 32-bit global loads, four u16 stores -------------------------------------------
  {
    const int wpr = (cw + 3 + 3) >> 2;  // aligned words covering one sub-image row (any misalignment 0..3)
    for (int i = tid; i < ch * wpr; i += kFastThreads) {
      const int r = i / wpr, j = i - r * wpr;
      const uint8_t* row = img + (int64_t)(c.y0 - 3 + r) * pitch + (c.x0 - 3);
      const int mis = (int)((uintptr_t)row & 3);
      if (4 * j >= mis + cw) continue;                          // this row needs one word less
      const uint32_t w = *reinterpret_cast<const uint32_t*>(row - mis + 4 * j);
      uint16_t* dst = s_img + r * sp16 + 5 + 4 * j - mis;       // >= 2; bytes outside the sub-image land in padding
      dst[0] = w & 0xff; dst[1] = (w >> 8) & 0xff; dst[2] = (w >> 16) & 0xff; dst[3] = w >> 24;
    }
    for (int i = tid; i < ((rh + 2) * scp) >> 2; i += kFastThreads) reinterpret_cast<uint32_t*>(s_sc)[i] = 0;
    if (tid < 2) s_cnt[tid] = 0;
  }
  __syncthreads();

  // ---- scores -----------------------------------------------------------------------------------------------------
  const int tA = P->fast_th, tB = kRetryTh, tmin = max(min(tA, tB), 1);
  const int ngroups = G * rh;
  for (int i = tid; i < ngroups; i += kFastThreads) {
    const int y = i / G, g = i - y * G;
    // W[dy + 3][j] = halfwords 4g + 4 + 2j, 4g + 5 + 2j of sub-image row y + 3 + dy
    uint32_t W[7][6];
#pragma unroll
    for (int dy = 0; dy < 7; dy++) {
      const uint2* src = reinterpret_cast<const uint2*>(s_img + (y + dy) * sp16 + 4 * g + 4);
      const uint2 a0 = src[0], a1 = src[1], a2 = src[2];
      W[dy][0] = a0.x; W[dy][1] = a0.y; W[dy][2] = a1.x; W[dy][3] = a1.y; W[dy][4] = a2.x; W[dy][5] = a2.y;
    }
    int sc[4];
#pragma unroll
    for (int h = 0; h < 2; h++) {  // pixel pair h: pixels 2h, 2h + 1 of the group
      // value at ring offset (dx, dy) for the pair = halfwords (4g + 8 + 2h + dx, +1) of row dy
#define RV(dx, dy) (((dx) & 1) ? hw_mid(W[(dy) + 3][(3 + (dx)) / 2 + h], W[(dy) + 3][(3 + (dx)) / 2 + h + 1]) \
                               : W[(dy) + 3][2 + (dx) / 2 + h])
      uint32_t r[16];
      r[0] = RV(0, 3);   r[1] = RV(1, 3);   r[2] = RV(2, 2);    r[3] = RV(3, 1);
      r[4] = RV(3, 0);   r[5] = RV(3, -1);  r[6] = RV(2, -2);   r[7] = RV(1, -3);
      r[8] = RV(0, -3);  r[9] = RV(-1, -3); r[10] = RV(-2, -2); r[11] = RV(-3, -1);
      r[12] = RV(-3, 0); r[13] = RV(-3, 1); r[14] = RV(-2, 2);  r[15] = RV(-1, 3);
      const uint32_t v2 = RV(0, 0);
#undef RV
      fast_score_x2(r, v2, tmin, &sc[2 * h], &sc[2 * h + 1]);
    }
    const int x0 = 4 * g;
    uint32_t packed = 0;
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (x0 + k < rw) packed |= (uint32_t)sc[k] << (8 * k);
    if (packed) *reinterpret_cast<uint32_t*>(s_sc + (y + 1) * scp + 4 + x0) = packed;
  }
  __syncthreads();

  // ---- NMS: survivors' scores into s_ks (aliases s_img), counts at both thresholds ------------------------------------
  uint8_t* s_ks = smem;  // [rh][4 G]
  int nA = 0, nB = 0;
  for (int i = tid; i < ngroups; i += kFastThreads) {
    const int y = i / G, g = i - y * G;
    const uint8_t* q0 = s_sc + (y + 1) * scp + 4 + 4 * g;
    const uint32_t cw4 = *reinterpret_cast<const uint32_t*>(q0);
    uint32_t keep4 = 0;
    if (cw4) {
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int s = (cw4 >> (8 * k)) & 255;
        if (s) {
          const uint8_t* q = q0 + k;
          const int m = max(max(max(q[-1], q[1]), max(q[-scp - 1], q[-scp])),
                            max(max(q[-scp + 1], q[scp - 1]), max(q[scp], q[scp + 1])));
          if (s > m) { keep4 |= (uint32_t)s << (8 * k); nA += s >= tA; nB += s >= tB; }
        }
      }
    }
    reinterpret_cast<uint32_t*>(s_ks)[i] = keep4;
  }
  // (s_ks aliases s_img, which is dead since the barrier after the score phase; the loop above reads only s_sc)
  nA = __reduce_add_sync(0xffffffffu, nA);
  nB = __reduce_add_sync(0xffffffffu, nB);
  if (lane == 0) { atomicAdd(&s_cnt[0], nA); atomicAdd(&s_cnt[1], nB); }
  __syncthreads();
  const int t = s_cnt[0] > 3 ? tA : tB;          // :609-614
  const int total = s_cnt[0] > 3 ? s_cnt[0] : s_cnt[1];
  if (tid == 0) *n_total = total;
  if (total > c.cand_cap) {                       // cannot happen (NMS survivors are non-adjacent); never truncate
    if (tid == 0) atomicExch(b.err_flag, 1);
    return;
  }
  if (total == 0) return;

  // ---- ordered compaction: rounds of 256 groups (1024 row-major pixels) ---------------------------------------------
  Cand* out = b.cand + (int64_t)frame * P->cand_per_frame + c.cand_off;
  int base = 0;
  for (int i0 = 0; i0 < ngroups; i0 += kFastThreads) {
    const int i = i0 + tid;
    uint32_t k4 = i < ngroups ? reinterpret_cast<const uint32_t*>(s_ks)[i] : 0;
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) cnt += (int)((k4 >> (8 * k)) & 255) >= t && ((k4 >> (8 * k)) & 255) != 0;
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int off = base + incl - cnt, round_total = 0;
#pragma unroll
    for (int wq = 0; wq < kFastThreads / 32; wq++) {
      const int v = s_warp[wq];
      if (wq < warp) off += v;
      round_total += v;
    }
    if (cnt) {
      const int y = i / G, g = i - y * G;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int s = (k4 >> (8 * k)) & 255;
        if (s >= t && s != 0) {
          const int lx = c.x0 + 4 * g + k, ly = c.y0 + y;
          Cand r;
          r.resp = P->score_type == ORBX_HARRIS_SCORE ? harris_response(img + (int64_t)ly * pitch + lx, pitch)
                                                      : (float)s;
          r.xy = ((uint32_t)ly << 16) | (uint32_t)lx;
          out[off++] = r;
        }
      }
    }
    base += round_total;
    __syncthreads();
  }
}

void launch_fast_cells(const ChunkBuffers& b, const DevicePlan& hp, int smem_bytes, int nframes, cudaStream_t s) {
  if (smem_bytes > 48 * 1024)
    cudaFuncSetAttribute(fast_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  dim3 grid(hp.total_cells, nframes);
  fast_cells_kernel<<<grid, kFastThreads, smem_bytes, s>>>(b);
}

}  // namespace orbx
#include "orbx_kernels_select.inl"
