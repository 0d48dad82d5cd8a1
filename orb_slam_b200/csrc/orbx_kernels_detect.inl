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
// rule (:609-614) on the same score map, then an ORDER-PRESERVING emission (row-major, the input order of
// retainBest's nth_element) into the cell's candidate list.
//
// Shared memory: the cell sub-image as zero-extended u16 (sub-image column c at halfword c + 5, so scored column 0
// is 16-byte aligned), the u8 score map with a zero ring, two survivor bitmaps (threshold fastTh / 7).
// Work is split DENSE-CHEAP / SPARSE-HEAVY with warp-local queues so that the heavy code always runs on full warps:
//   1  every 4-pixel group: opposite-pair reject  min(max(|r0-v|,|r8-v|), max(|r4-v|,|r12-v|)) > tmin  (an arc of 9
//      contains a pixel of every opposite pair) on packed pixel PAIRS; surviving pairs are queued
//   2  queued pairs, 32 at a time: the full VIMNMX3.U16x2 ladder -> score bytes
//   3  every score word: pixels with S >= tmin are queued; queued pixels, 32 at a time: strict 3x3 NMS -> bitmaps
//   4  popcount prefix over the bitmap words -> survivors in row-major order
// Pixels failing the reject keep score 0 instead of their true S < tmin: harmless, a neighbour below tmin can never
// suppress a survivor (S_c >= t >= tmin > S_n), which is also why one threshold-free map serves both thresholds.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kFastThreads = 256, kFastWarps = kFastThreads / 32, kFastQueue = 160;

__device__ __forceinline__ uint32_t hw_mid(uint32_t lo, uint32_t hi) { return __byte_perm(lo, hi, 0x5432); }
__device__ __forceinline__ uint32_t absdiff_u16x2(uint32_t a, uint32_t v) { return __vmaxu2(a, v) - __vminu2(a, v); }

// Full score of the pixel pair at scored position (x even, y): loads the 16 ring values of both pixels as packed
// halfwords straight from the staged sub-image (row pitch sp32 words).
__device__ __forceinline__ uint32_t fast_pair_score(const uint32_t* s_img32, int sp32, int x, int y) {
  const uint32_t* c0 = s_img32 + (y + 3) * sp32 + ((x + 8) >> 1);  // word holding the pair's centre pixels
#define L(dy, dw) c0[(dy) * sp32 + (dw)]
  uint32_t r[16];
  {
    const uint32_t a = L(3, -1), bq = L(3, 0), cq = L(3, 1);
    r[0] = bq; r[1] = hw_mid(bq, cq); r[15] = hw_mid(a, bq);
  }
  r[2] = L(2, 1); r[14] = L(2, -1);
  {
    const uint32_t a = L(1, -2), bq = L(1, -1), cq = L(1, 1), dq = L(1, 2);
    r[13] = hw_mid(a, bq); r[3] = hw_mid(cq, dq);
  }
  uint32_t v2;
  {
    const uint32_t a = L(0, -2), bq = L(0, -1), cq = L(0, 1), dq = L(0, 2);
    v2 = L(0, 0);
    r[12] = hw_mid(a, bq); r[4] = hw_mid(cq, dq);
  }
  {
    const uint32_t a = L(-1, -2), bq = L(-1, -1), cq = L(-1, 1), dq = L(-1, 2);
    r[11] = hw_mid(a, bq); r[5] = hw_mid(cq, dq);
  }
  r[6] = L(-2, 1); r[10] = L(-2, -1);
  {
    const uint32_t a = L(-3, -1), bq = L(-3, 0), cq = L(-3, 1);
    r[8] = bq; r[7] = hw_mid(bq, cq); r[9] = hw_mid(a, bq);
  }
#undef L
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
  uint32_t bb = __vimax3_u16x2(mn9[0], mn9[1], mn9[2]), dd = __vimin3_u16x2(mx9[0], mx9[1], mx9[2]);
#pragma unroll
  for (int k = 3; k < 15; k += 2) {
    bb = __vimax3_u16x2(bb, mn9[k], mn9[k + 1]);
    dd = __vimin3_u16x2(dd, mx9[k], mx9[k + 1]);
  }
  bb = __vmaxu2(bb, mn9[15]);
  dd = __vminu2(dd, mx9[15]);
  // S = max(bright - v, v - dark) - 1 per halfword (signed 16-bit lanes, |values| <= 255), clamped at 0
  const uint32_t m = __vmaxs2(__vsub2(bb, v2), __vsub2(v2, dd));
  return __vmaxs2(__vsub2(m, 0x00010001u), 0u);
}

__global__ void __launch_bounds__(kFastThreads, 4) fast_cells_kernel(ChunkBuffers b) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ uint32_t s_queue[kFastWarps][kFastQueue];
  __shared__ int s_cnt[2];
  __shared__ int s_warp[kFastWarps];
  const DevicePlan* P = b.plan;
  const int cell = blockIdx.x, frame = blockIdx.y;
  const CellPlan c = b.cells[cell];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned lt_mask = (1u << lane) - 1;
  int32_t* n_total = b.n_total + (int64_t)frame * P->total_cells + cell;
  if (c.rw <= 0) {
    if (tid == 0) *n_total = 0;
    return;
  }
  int pitch;
  const uint8_t* img = level_ptr(b, P, c.level, frame, &pitch);
  const int rw = c.rw, rh = c.rh, cw = rw + 6, ch = rh + 6;   // cw x ch = the sub-image handed to cv::FAST
  const int sp16 = fast_sp16(cw), sp32 = sp16 >> 1, scp = fast_scp(rw);
  const int npx = rw * rh, nbw = (npx + 31) >> 5;
  uint16_t* s_img = reinterpret_cast<uint16_t*>(smem);          // [ch][sp16]
  uint8_t* s_sc = smem + fast_img_bytes(cw, ch);                // [rh + 2][scp], scored column x at byte 4 + x
  uint32_t* s_bmA = reinterpret_cast<uint32_t*>(s_sc + fast_sc_bytes(rw, rh));  // survivors with S >= fastTh
  uint32_t* s_bmB = s_bmA + nbw;                                                // survivors with S >= 7
  uint32_t* s_bmC = s_bmB + nbw;                                                // pixels with S >= tmin (NMS inputs)
  const int G = (rw + 3) >> 2;                                  // 4-pixel groups per scored row
  const int ngroups = G * rh;

  // ---- 0  stage the sub-image: aligned 32-bit global loads, four u16 stores ------------------------------------------
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
    const int nz = (fast_sc_bytes(rw, rh) >> 2) + 3 * nbw;       // score map + the three bitmaps are contiguous
    for (int i = tid; i < nz; i += kFastThreads) reinterpret_cast<uint32_t*>(s_sc)[i] = 0;
    if (tid < 2) s_cnt[tid] = 0;
  }
  __syncthreads();

  const int tA = P->fast_th, tB = kRetryTh, tmin = max(min(tA, tB), 1);
  uint32_t* q = s_queue[warp];
  const uint32_t* s_img32 = reinterpret_cast<const uint32_t*>(s_img);
  const int gstep_y = kFastThreads / G, gstep_g = kFastThreads - gstep_y * G;
  // floor(i / d) == umulhi(i, floor(2^32 / d) + 1) for i * d < 2^32 (all indices here are < 2^17)
  const uint32_t rw_magic = 0xffffffffu / (uint32_t)rw + 1u, g_magic = 0xffffffffu / (uint32_t)G + 1u;

  // ---- 1 + 2  scores ---------------------------------------------------------------------------------------------------
  {
    auto heavy = [&](uint32_t e) {
      const int x = e & 0xffff, y = e >> 16;
      const uint32_t s2 = fast_pair_score(s_img32, sp32, x, y);
      uint32_t two = __byte_perm(s2, 0, 0x4420);               // (S(x), S(x+1)) as two bytes
      if (x + 1 >= rw) two &= 0xff;
      if (two) {
        *reinterpret_cast<uint16_t*>(s_sc + (y + 1) * scp + 4 + x) = (uint16_t)two;
        const int idx = y * rw + x;                             // row-major pixel index = bitmap bit
        if ((int)(two & 0xff) >= tmin) atomicOr(&s_bmC[idx >> 5], 1u << (idx & 31));
        if ((int)(two >> 8) >= tmin) atomicOr(&s_bmC[(idx + 1) >> 5], 1u << ((idx + 1) & 31));
      }
    };
    const uint32_t bias = (0x8000u - (uint32_t)(tmin + 1)) * 0x00010001u;  // bit 15 of (q + bias) <=> q > tmin
    int qlen = 0;
    int y = __umulhi((uint32_t)(warp * 32 + lane), g_magic), g = (warp * 32 + lane) - y * G;
    for (int i0 = warp * 32; i0 < ngroups; i0 += kFastThreads) {
      uint32_t flags = 0;
      if (i0 + lane < ngroups) {
        const uint2* mid = reinterpret_cast<const uint2*>(s_img + (y + 3) * sp16 + 4 * g + 4);
        const uint2 a0 = mid[0], a1 = mid[1], a2 = mid[2];     // halfwords 4g+4 .. 4g+15 of the centre row
        const uint2 up = *reinterpret_cast<const uint2*>(s_img + y * sp16 + 4 * g + 8);        // dy = -3, dx = 0
        const uint2 dn = *reinterpret_cast<const uint2*>(s_img + (y + 6) * sp16 + 4 * g + 8);  // dy = +3
        // pair 0: centre a1.x, left = hw(4g+5,4g+6), right = hw(4g+11,4g+12); pair 1: one word to the right
        const uint32_t qa = __vminu2(__vmaxu2(absdiff_u16x2(up.x, a1.x), absdiff_u16x2(dn.x, a1.x)),
                                     __vmaxu2(absdiff_u16x2(hw_mid(a0.x, a0.y), a1.x),
                                              absdiff_u16x2(hw_mid(a1.y, a2.x), a1.x)));
        const uint32_t qb = __vminu2(__vmaxu2(absdiff_u16x2(up.y, a1.y), absdiff_u16x2(dn.y, a1.y)),
                                     __vmaxu2(absdiff_u16x2(hw_mid(a0.y, a1.x), a1.y),
                                              absdiff_u16x2(hw_mid(a2.x, a2.y), a1.y)));
        if ((qa + bias) & 0x80008000u) flags |= 1;
        if (((qb + bias) & 0x80008000u) && 4 * g + 2 < rw) flags |= 2;
      }
      const unsigned b0 = __ballot_sync(0xffffffffu, flags & 1), b1 = __ballot_sync(0xffffffffu, flags & 2);
      if (b0 | b1) {
        // queue order is irrelevant: all "pair 0" entries first, then all "pair 1" entries
        if (flags & 1) q[qlen + __popc(b0 & lt_mask)] = ((uint32_t)y << 16) | (uint32_t)(4 * g);
        if (flags & 2) q[qlen + __popc(b0) + __popc(b1 & lt_mask)] = ((uint32_t)y << 16) | (uint32_t)(4 * g + 2);
        qlen += __popc(b0) + __popc(b1);
        __syncwarp();
        while (qlen >= 32) {
          qlen -= 32;
          heavy(q[qlen + lane]);
        }
        __syncwarp();
      }
      y += gstep_y; g += gstep_g;
      if (g >= G) { g -= G; y++; }
    }
    if (lane < qlen) heavy(q[lane]);
  }
  __syncthreads();

  // ---- 3  NMS of the pixels with S >= tmin -----------------------------------------------------------------------------
  int nA = 0, nB = 0;
  {
    auto nms = [&](uint32_t e) {
      const int x = e & 0xffff, y = e >> 16;
      const uint8_t* p = s_sc + (y + 1) * scp + 4 + x;
      const int s = p[0];
      const int m = max(max(max(p[-1], p[1]), max(p[-scp - 1], p[-scp])),
                        max(max(p[-scp + 1], p[scp - 1]), max(p[scp], p[scp + 1])));
      if (s > m) {
        const int idx = y * rw + x;
        if (s >= tA) { atomicOr(&s_bmA[idx >> 5], 1u << (idx & 31)); nA++; }
        if (s >= tB) { atomicOr(&s_bmB[idx >> 5], 1u << (idx & 31)); nB++; }
      }
    };
    // candidates come from the bitmap written by phase 2: 32 words (1024 pixels) per warp round
    int qlen = 0;
    for (int i0 = warp * 32; i0 < nbw; i0 += kFastThreads) {
      const int i = i0 + lane;
      uint32_t bits = i < nbw ? s_bmC[i] : 0u;
      const int cnt = __popc(bits);
      if (!__any_sync(0xffffffffu, cnt)) continue;
      int incl = cnt;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
      }
      const int round_total = __shfl_sync(0xffffffffu, incl, 31);
      if (qlen + round_total > kFastQueue) {
        // dense textures (e.g. noise): more candidates than the queue holds; each lane walks its own word
        while (bits) {
          const int idx = 32 * i + __ffs(bits) - 1;
          bits &= bits - 1;
          const int yy = __umulhi((uint32_t)idx, rw_magic);
          nms(((uint32_t)yy << 16) | (uint32_t)(idx - yy * rw));
        }
        continue;
      }
      int off = qlen + incl - cnt;
      while (bits) {
        const int idx = 32 * i + __ffs(bits) - 1;
        bits &= bits - 1;
        const int yy = __umulhi((uint32_t)idx, rw_magic);
        q[off++] = ((uint32_t)yy << 16) | (uint32_t)(idx - yy * rw);
      }
      qlen += round_total;
      __syncwarp();
      while (qlen >= 32) {
        qlen -= 32;
        nms(q[qlen + lane]);
      }
      __syncwarp();
    }
    if (lane < qlen) nms(q[lane]);
  }
  nA = __reduce_add_sync(0xffffffffu, nA);
  nB = __reduce_add_sync(0xffffffffu, nB);
  if (lane == 0) { atomicAdd(&s_cnt[0], nA); atomicAdd(&s_cnt[1], nB); }
  __syncthreads();
  const bool useA = s_cnt[0] > 3;                 // :609-614
  const int total = useA ? s_cnt[0] : s_cnt[1];
  const uint32_t* bm = useA ? s_bmA : s_bmB;
  if (tid == 0) *n_total = total;
  if (total > c.cand_cap) {                       // cannot happen (NMS survivors are non-adjacent); never truncate
    if (tid == 0) atomicExch(b.err_flag, 1);
    return;
  }
  if (total == 0) return;

  // ---- 4  ordered emission: popcount prefix over the bitmap words (bit index = y * rw + x, row-major) --------------------
  Cand* out = b.cand + (int64_t)frame * P->cand_per_frame + c.cand_off;
  int base = 0;
  for (int i0 = 0; i0 < nbw; i0 += kFastThreads) {
    const int i = i0 + tid;
    uint32_t bits = i < nbw ? bm[i] : 0;
    const int cnt = __popc(bits);
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
    for (int wq = 0; wq < kFastWarps; wq++) {
      const int v = s_warp[wq];
      if (wq < warp) off += v;
      round_total += v;
    }
    while (bits) {
      const int bit = __ffs(bits) - 1;
      bits &= bits - 1;
      const int idx = 32 * i + bit, y = __umulhi((uint32_t)idx, rw_magic), x = idx - y * rw;
      const int lx = c.x0 + x, ly = c.y0 + y;
      Cand r;
      r.resp = P->score_type == ORBX_HARRIS_SCORE ? harris_response(img + (int64_t)ly * pitch + lx, pitch)
                                                  : (float)s_sc[(y + 1) * scp + 4 + x];
      r.xy = ((uint32_t)ly << 16) | (uint32_t)lx;
      out[off++] = r;
    }
    base += round_total;
    __syncthreads();
  }
}

void launch_fast_cells(const ChunkBuffers& b, const DevicePlan& hp, int smem_bytes, int nframes, cudaStream_t s) {
  // per-device attributes; a few microseconds per chunk of up to max_batch frames
  cudaFuncSetAttribute(fast_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(fast_cells_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
  dim3 grid(hp.total_cells, nframes);
  fast_cells_kernel<<<grid, kFastThreads, smem_bytes, s>>>(b);
}

}  // namespace orbx
#include "orbx_kernels_select.inl"
