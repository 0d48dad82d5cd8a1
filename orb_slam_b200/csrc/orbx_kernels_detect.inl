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

// The same for a corner closer than 4 px to the level's edge (a cell grid whose ceil() overshoots the border, see
// describe_kernel): the 9x9 support reaches the reference's reflect-101 frame (:806-814).
__device__ __noinline__ float harris_response_edge(const uint8_t* img, int step, int x, int y, int w, int h) {
  uint8_t v[9][9];
  for (int i = 0; i < 9; i++)
    for (int j = 0; j < 9; j++) v[i][j] = img[(int64_t)reflect101(y + i - 4, h) * step + reflect101(x + j - 4, w)];
  int a = 0, b = 0, c = 0;
  for (int i = 1; i <= 7; i++)
    for (int j = 1; j <= 7; j++) {
      const int Ix = (v[i][j + 1] - v[i][j - 1]) * 2 + (v[i - 1][j + 1] - v[i - 1][j - 1]) + (v[i + 1][j + 1] - v[i + 1][j - 1]);
      const int Iy = (v[i + 1][j] - v[i - 1][j]) * 2 + (v[i + 1][j - 1] - v[i - 1][j - 1]) + (v[i + 1][j + 1] - v[i - 1][j + 1]);
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
// Work is split DENSE-CHEAP / SPARSE-HEAVY; the sparse stages run from per-warp queues so that the heavy code always
// executes on full warps:
//   0  stage the sub-image twice: zero-extended u16 (ladder operand) and u8 >> 1 (reject operand)
//   1  byte-SIMD reject, 8 pixels per thread: an arc of 9 ring pixels contains one pixel of every opposite pair, so a
//      corner at threshold t needs (|up-v| > t or |down-v| > t) and (|left-v| > t or |right-v| > t).  On the halved
//      image |a'-v'| >= ceil(t/2) is implied by |a-v| > t, differences fit 7 bits, and adding (0x80 - ceil(t/2)) to a
//      VABSDIFF4 result raises bit 7 of each byte exactly when the halved test passes: 2 instructions per 4 pixels
//      per ring direction.  The 8 pass bits go to a bitmap byte.  (Conservative: never drops a corner.)
//   2  flagged pixel PAIRS, 32 at a time: the full VIMNMX3.U16x2 ladder -> score bytes, bitmap C = (S >= tmin)
//   3  bitmap C pixels, 32 at a time: strict 3x3 NMS -> bitmaps A (S >= fastTh) and B (S >= 7)
//   4  popcount prefix over the bitmap words -> survivors in row-major order
// Pixels failing the reject keep score 0 instead of their true S < tmin: harmless, a neighbour below tmin can never
// suppress a survivor (S_c >= t >= tmin > S_n), which is also why one threshold-free map serves both thresholds.
// ------------------------------------------------------------------------------------------------------------------
#ifndef ORBX_FAST_THREADS
#define ORBX_FAST_THREADS 256
#endif
constexpr int kFastThreads = ORBX_FAST_THREADS, kFastWarps = kFastThreads / 32, kFastQueue = 160, kFastSub = 4;

__device__ __forceinline__ uint32_t hw_mid(uint32_t lo, uint32_t hi) { return __byte_perm(lo, hi, 0x5432); }

// Appends base + b for every set bit b of every lane's `bits` to the warp's queue (lane order, then bit order) and
// runs fn on each full group of 32 entries.  A round that would overflow the queue is split into sub-rounds of
// kFastSub lanes (<= 32 * kFastSub = 128 new entries on top of <= 31 left over).  Must be called by all 32 lanes.
template <class F>
__device__ __forceinline__ void queue_bits(uint32_t* q, int& qlen, uint32_t bits, uint32_t base, int lane, F&& fn) {
  const int cnt = __popc(bits);
  if (!__any_sync(0xffffffffu, cnt)) return;
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  const int excl = incl - cnt, total = __shfl_sync(0xffffffffu, incl, 31);
  if (qlen + total <= kFastQueue) {
    int off = qlen + excl;
    while (bits) {                                            // entry order is irrelevant: take the TOP bit, one FLO
      const int bit = 31 - __clz(bits);                       // (ffs = BREV + FLO, both quarter-rate)
      q[off++] = base + (uint32_t)bit;
      bits ^= 1u << bit;
    }
    qlen += total;
    __syncwarp();
    while (qlen >= 32) {
      qlen -= 32;
      fn(q[qlen + lane]);
    }
    __syncwarp();
    return;
  }
#pragma unroll 1
  for (int k = 0; k < 32; k += kFastSub) {
    const int lo = __shfl_sync(0xffffffffu, excl, k), hi = __shfl_sync(0xffffffffu, incl, k + kFastSub - 1);
    if (hi == lo) continue;
    if (lane >= k && lane < k + kFastSub) {
      int off = qlen + excl - lo;
      while (bits) {
        const int bit = 31 - __clz(bits);
        q[off++] = base + (uint32_t)bit;
        bits ^= 1u << bit;
      }
    }
    qlen += hi - lo;
    __syncwarp();
    while (qlen >= 32) {
      qlen -= 32;
      fn(q[qlen + lane]);
    }
    __syncwarp();
  }
}

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

// Per-level constants of the launch, passed by value: they sit in the constant bank, so the only global round trip
// before the image loads is the CellPlan record.
struct FastLevels {
  int fast_th, score_type, total_cells;
  int64_t cand_per_frame, frame_bytes;
  int pitch[kMaxLevels];
  int64_t off[kMaxLevels];
};

// kHarris: the Harris-score build carries the response code (and its rare reflect-101 variant, a real call); the
// FAST-score build, the default and the benchmarked one, is free of it
template <bool kHarris>
__global__ void __launch_bounds__(kFastThreads, 1024 / kFastThreads) fast_cells_kernel(ChunkBuffers b, FastLevels fl) {
  ORBX_PDL_ENTRY();
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ uint32_t s_queue[kFastWarps][kFastQueue];
  __shared__ int s_cnt[2];
  __shared__ int s_warp[kFastWarps];
  const int cell = blockIdx.x, frame = blockIdx.y;
  const CellPlan c = b.cells[cell];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int32_t* n_total = b.n_total + (int64_t)frame * fl.total_cells + cell;
  if (c.skipped & 2) return;                       // a cell beyond shared memory: fast_cells_big_kernel's
  if (c.rw <= 0) {
    if (tid == 0) *n_total = 0;
    return;
  }
  const int pitch = c.level == 0 ? b.lv0_pitch : fl.pitch[c.level];
  const uint8_t* img = c.level == 0 ? b.lv0 + (int64_t)frame * b.lv0_frame_stride
                                    : b.pyr + (int64_t)frame * fl.frame_bytes + fl.off[c.level];
  const int tA = fl.fast_th;
  const int rw = c.rw, rh = c.rh, cw = rw + 6, ch = rh + 6;   // cw x ch = the sub-image handed to cv::FAST
  const int sp16 = fast_sp16(cw), sp32 = sp16 >> 1, bp = fast_bp(rw), scp = fast_scp(rw);
  const int wpr = (rw + 31) >> 5, nbw = wpr * rh;               // bitmap words per scored row / per bitmap
  uint16_t* s_img = reinterpret_cast<uint16_t*>(smem);          // [ch][sp16]
  uint8_t* s_h8 = smem + fast_img_bytes(cw, ch);                // [ch][bp]
  uint8_t* s_sc = s_h8 + fast_h8_bytes(cw, ch);                 // [rh + 2][scp], scored column x at byte 4 + x
  uint32_t* s_bmA = reinterpret_cast<uint32_t*>(s_sc + fast_sc_bytes(rw, rh));  // survivors with S >= fastTh
  uint32_t* s_bmB = s_bmA + nbw;                                                // survivors with S >= 7
  uint32_t* s_bmC = s_bmB + nbw;                                                // pixels with S >= tmin (NMS inputs)
  uint32_t* s_bmF = s_bmC + nbw;                                                // pixels passing the reject

  // ---- 0  stage the sub-image: 8 pixels per thread from three aligned 32-bit global loads -----------------------------
  // group u of row r = sub-image columns 8u-5 .. 8u+2 (the first group starts left of the sub-image, inside the level:
  // cells start at column >= 13) -> halfwords / bytes 8u .. 8u+7 of the staged rows.  A thread keeps its column group
  // and walks down the rows, `rpr` rows per round of the CTA, so the per-round address math is three additions.
  {
    const int ngrp = fast_ngrp(cw);                              // <= kFastThreads (build_plan)
    const int rpr = magic_div(kFastThreads, c.m_grp);            // rows per round
    const int r0 = magic_div(tid, c.m_grp), u = tid - r0 * ngrp;
    if (r0 < rpr) {
      const uint8_t* img0 = img + (int64_t)(c.y0 - 3) * pitch + (c.x0 - 3) - 5;
      const int mis0 = (int)(reinterpret_cast<uintptr_t>(img0) & 3);
      const uint8_t* base0 = img0 - mis0;                        // 4-byte aligned
      int off = r0 * pitch + 8 * u + mis0;                       // byte offset of the group from base0
      uint8_t* d16 = reinterpret_cast<uint8_t*>(s_img) + r0 * (2 * sp16) + 16 * u;
      uint8_t* d8 = s_h8 + r0 * bp + 8 * u;
      const int s_off = rpr * pitch, s16 = rpr * 2 * sp16, s8 = rpr * bp;
      // three rows per trip, all nine loads issued before the first use
      for (int r = r0; r < ch; r += 3 * rpr) {
        uint32_t w[3][3];
        int sh[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          if (r + k * rpr < ch) {
            const uint32_t* a = reinterpret_cast<const uint32_t*>(base0 + ((off + k * s_off) & ~3));
            w[k][0] = a[0]; w[k][1] = a[1]; w[k][2] = a[2];
            sh[k] = ((off + k * s_off) & 3) * 8;
          }
        }
#pragma unroll
        for (int k = 0; k < 3; k++) {
          if (r + k * rpr < ch) {
            const uint32_t b0 = __funnelshift_r(w[k][0], w[k][1], sh[k]), b1 = __funnelshift_r(w[k][1], w[k][2], sh[k]);
            *reinterpret_cast<uint4*>(d16 + k * s16) =
                make_uint4(__byte_perm(b0, 0, 0x4140), __byte_perm(b0, 0, 0x4342), __byte_perm(b1, 0, 0x4140),
                           __byte_perm(b1, 0, 0x4342));
            *reinterpret_cast<uint2*>(d8 + k * s8) = make_uint2((b0 >> 1) & 0x7f7f7f7fu, (b1 >> 1) & 0x7f7f7f7fu);
          }
        }
        off += 3 * s_off; d16 += 3 * s16; d8 += 3 * s8;
      }
    }
    // score map + bitmaps A, B, C are contiguous; the rounding to 16 bytes may run into bitmap F (written in phase 1)
    const int nz = ((fast_sc_bytes(rw, rh) >> 2) + 3 * nbw + 3) >> 2;
    for (int i = tid; i < nz; i += kFastThreads) reinterpret_cast<uint4*>(s_sc)[i] = make_uint4(0, 0, 0, 0);
    if (tid < 2) s_cnt[tid] = 0;
  }
  __syncthreads();

  const int tB = kRetryTh;
  uint32_t* q = s_queue[warp];
  const uint32_t* s_img32 = reinterpret_cast<const uint32_t*>(s_img);
  // Like the reference (:599-614) the cell is first searched at fastTh only; the sparse stages then see the few
  // pixels with S >= fastTh instead of every pixel with S >= 7.  Only a cell with <= 3 survivors is searched again
  // at 7 (everything below is CTA-uniform).  fastTh <= 7: one pass at fastTh serves both thresholds.
  int tmin = max(tA > tB ? tA : min(tA, tB), 1);
  bool useA;
  for (;;) {
  // ---- 1  reject: 8 pixels per thread on the halved u8 image -----------------------------------------------------------
  {
    const int G8 = (rw + 7) >> 3, n8 = G8 * rh;
    const uint32_t bias4 = (0x80u - (uint32_t)((tmin + 1) >> 1)) * 0x01010101u;
    const int sy = magic_div(kFastThreads, c.m_g8), sg = kFastThreads - sy * G8;
    int y = magic_div(tid, c.m_g8), g = tid - y * G8;
    uint8_t* s_flag = reinterpret_cast<uint8_t*>(s_bmF);
    for (int i = tid; i < n8; i += kFastThreads) {
      const uint8_t* ctr = s_h8 + (y + 3) * bp + 8 * g + 8;     // centre bytes of scored pixels 8g .. 8g+7
      const uint2 C = *reinterpret_cast<const uint2*>(ctr);
      const uint2 U = *reinterpret_cast<const uint2*>(ctr - 3 * bp), D = *reinterpret_cast<const uint2*>(ctr + 3 * bp);
      const uint32_t W0 = *reinterpret_cast<const uint32_t*>(ctr - 4), W3 = *reinterpret_cast<const uint32_t*>(ctr + 8);
      const uint32_t Lx = __byte_perm(W0, C.x, 0x4321), Ly = __byte_perm(C.x, C.y, 0x4321);   // columns x - 3
      const uint32_t Rx = __byte_perm(C.x, C.y, 0x6543), Ry = __byte_perm(C.y, W3, 0x6543);   // columns x + 3
      const uint32_t mx = ((__vabsdiffu4(U.x, C.x) + bias4) | (__vabsdiffu4(D.x, C.x) + bias4)) &
                          ((__vabsdiffu4(Lx, C.x) + bias4) | (__vabsdiffu4(Rx, C.x) + bias4)) & 0x80808080u;
      const uint32_t my = ((__vabsdiffu4(U.y, C.y) + bias4) | (__vabsdiffu4(D.y, C.y) + bias4)) &
                          ((__vabsdiffu4(Ly, C.y) + bias4) | (__vabsdiffu4(Ry, C.y) + bias4)) & 0x80808080u;
      // bits 7/15/23/31 -> one nibble: the products of 0x00204081 land them on bits 28..31 without collisions
      s_flag[y * (4 * wpr) + g] = (uint8_t)(((mx * 0x00204081u) >> 28) | (((my * 0x00204081u) >> 24) & 0xf0u));
      y += sy; g += sg;
      if (g >= G8) { g -= G8; y++; }
    }
  }
  __syncthreads();

  // ---- 2  full score of the flagged pixel pairs -----------------------------------------------------------------------
  {
    auto heavy = [&](uint32_t e) {
      const int x = e & 0xffff, y = e >> 16;
      const uint32_t s2 = fast_pair_score(s_img32, sp32, x, y);
      uint32_t two = __byte_perm(s2, 0, 0x4420);               // (S(x), S(x+1)) as two bytes
      if (x + 1 >= rw) two &= 0xff;
      if (two) {
        *reinterpret_cast<uint16_t*>(s_sc + (y + 1) * scp + 4 + x) = (uint16_t)two;
        const uint32_t m = ((int)(two & 0xff) >= tmin ? 1u : 0u) | ((int)(two >> 8) >= tmin ? 2u : 0u);
        if (m) atomicOr(&s_bmC[y * wpr + (x >> 5)], m << (x & 31));   // x is even: both bits in one word
      }
    };
    int qlen = 0;
    // word i = (round * 32 + lane) * 8 + warp: every warp samples the whole cell, so the sparse work is balanced
    for (int i0 = 0; i0 < nbw; i0 += kFastThreads) {
      const int i = i0 + lane * kFastWarps + warp;
      uint32_t bits = 0, base = 0;
      if (i < nbw) {
        const int y = magic_div(i, c.m_wpr), wx = i - y * wpr;
        uint32_t f = s_bmF[i];
        const int nv = rw - 32 * wx;                            // columns >= rw were computed from padding
        if (nv < 32) f &= (1u << nv) - 1u;
        bits = (f | (f >> 1)) & 0x55555555u;                    // pair (x, x+1), x even, when either pixel passed
        base = ((uint32_t)y << 16) | (uint32_t)(32 * wx);
      }
      queue_bits(q, qlen, bits, base, lane, heavy);
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
        const int w = y * wpr + (x >> 5);
        if (s >= tA) { atomicOr(&s_bmA[w], 1u << (x & 31)); nA++; }
        if (s >= tB) { atomicOr(&s_bmB[w], 1u << (x & 31)); nB++; }
      }
    };
    int qlen = 0;
    for (int i0 = 0; i0 < nbw; i0 += kFastThreads) {
      const int i = i0 + lane * kFastWarps + warp;
      uint32_t bits = 0, base = 0;
      if (i < nbw) {
        const int y = magic_div(i, c.m_wpr);
        bits = s_bmC[i];
        base = ((uint32_t)y << 16) | (uint32_t)(32 * (i - y * wpr));
      }
      queue_bits(q, qlen, bits, base, lane, nms);
    }
    if (lane < qlen) nms(q[lane]);
  }
  nA = __reduce_add_sync(0xffffffffu, nA);
  nB = __reduce_add_sync(0xffffffffu, nB);
  if (lane == 0) { atomicAdd(&s_cnt[0], nA); atomicAdd(&s_cnt[1], nB); }
  __syncthreads();
  useA = s_cnt[0] > 3;                            // :609-614
  if (useA || tmin <= tB) break;
  // retry at 7: forget this pass's scores and bitmaps
  __syncthreads();                                // everyone has read s_cnt
  {
    const int nz = ((fast_sc_bytes(rw, rh) >> 2) + 3 * nbw + 3) >> 2;
    for (int i = tid; i < nz; i += kFastThreads) reinterpret_cast<uint4*>(s_sc)[i] = make_uint4(0, 0, 0, 0);
    if (tid < 2) s_cnt[tid] = 0;
  }
  tmin = tB;
  __syncthreads();
  }
  const int total = useA ? s_cnt[0] : s_cnt[1];
  const uint32_t* bm = useA ? s_bmA : s_bmB;
  if (tid == 0) *n_total = total;
  if (total > c.cand_cap) {                       // cannot happen (NMS survivors are non-adjacent); never truncate
    if (tid == 0) atomicExch(b.err_flag, 1);
    return;
  }
  if (total == 0) return;

  // ---- 4  ordered emission: popcount prefix over the bitmap words (row-major) ------------------------------------------
  Cand* out = b.cand + (int64_t)frame * fl.cand_per_frame + c.cand_off;
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
    if (bits) {
      const int y = magic_div(i, c.m_wpr), xb = 32 * (i - y * wpr);
      const int ly = c.y0 + y;
      while (bits) {
        const int x = xb + __ffs(bits) - 1;
        bits &= bits - 1;
        const int lx = c.x0 + x;
        Cand r;
        if constexpr (kHarris) {
          const LevelPlan& LP = b.plan->lv[c.level];
          r.resp = (lx >= 4 && ly >= 4 && lx + 4 < LP.w && ly + 4 < LP.h)
                       ? harris_response(img + (int64_t)ly * pitch + lx, pitch)
                       : harris_response_edge(img, pitch, lx, ly, LP.w, LP.h);
        } else {
          r.resp = (float)s_sc[(y + 1) * scp + 4 + x];
        }
        r.xy = ((uint32_t)ly << 16) | (uint32_t)lx;
        out[off++] = r;
      }
    }
    base += round_total;
    __syncthreads();
  }
}

void launch_fast_cells(const ChunkBuffers& b, const DevicePlan& hp, int smem_bytes, int nframes, cudaStream_t s) {
  // per-device attributes; a few microseconds per chunk of up to max_batch frames
  auto* kern = hp.score_type == ORBX_HARRIS_SCORE ? fast_cells_kernel<true> : fast_cells_kernel<false>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
  FastLevels fl;
  fl.fast_th = hp.fast_th; fl.score_type = hp.score_type; fl.total_cells = hp.total_cells;
  fl.cand_per_frame = hp.cand_per_frame; fl.frame_bytes = hp.pyr_frame_bytes;
  for (int l = 0; l < kMaxLevels; l++) { fl.pitch[l] = hp.lv[l].pitch; fl.off[l] = hp.lv[l].off; }
  dim3 grid(hp.total_cells, nframes);
  launch_k(kern, grid, dim3(kFastThreads), (size_t)smem_bytes, s, b, fl);
}

// ------------------------------------------------------------------------------------------------------------------
// Cells whose sub-image does not fit one CTA's shared memory (few features on a large image: e.g. 200 features at
// 975 x 686 give 2 x 2 cells of 472 x 327 px).  Same per-cell semantics, worked in GLOBAL memory: the score of EVERY
// pixel of the scored region straight from the level image (scalar ring walk, no reject), score map and survivor bitmaps
// in a per-frame scratch block, NMS, retry at 7, ordered emission by one warp.  Rare and not tuned: a 150 k-pixel cell
// takes a few hundred microseconds.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int fast_score_scalar(const uint8_t* p, int pitch) {
  const int v = p[0];
  const int r[16] = {p[3 * pitch], p[3 * pitch + 1], p[2 * pitch + 2], p[pitch + 3], p[3], p[-pitch + 3], p[-2 * pitch + 2],
                     p[-3 * pitch + 1], p[-3 * pitch], p[-3 * pitch - 1], p[-2 * pitch - 2], p[-pitch - 3], p[-3],
                     p[pitch - 3], p[2 * pitch - 2], p[3 * pitch - 1]};
  int bright = 0, dark = 255;                                 // max over arcs of min9(ring), min over arcs of max9(ring)
#pragma unroll
  for (int k = 0; k < 16; k++) {
    int mn = 255, mx = 0;
#pragma unroll
    for (int j = 0; j < 9; j++) { mn = min(mn, r[(k + j) & 15]); mx = max(mx, r[(k + j) & 15]); }
    bright = max(bright, mn); dark = min(dark, mx);
  }
  return max(max(bright - v, v - dark) - 1, 0);               // as fast_pair_score
}

template <bool kHarris>
__global__ void __launch_bounds__(1024) fast_cells_big_kernel(ChunkBuffers b, FastLevels fl, const int32_t* big_cells,
                                                             uint8_t* scratch, int64_t scratch_per_frame) {
  ORBX_PDL_ENTRY();
  __shared__ int s_cnt[2];
  const int cell = big_cells[blockIdx.x], frame = blockIdx.y;
  const CellPlan c = b.cells[cell];
  const int tid = threadIdx.x, lane = tid & 31, nthr = blockDim.x;
  const int level = c.level, rw = c.rw, rh = c.rh;
  const int pitch = level == 0 ? b.lv0_pitch : fl.pitch[level];
  const uint8_t* img = level == 0 ? b.lv0 + (int64_t)frame * b.lv0_frame_stride
                                  : b.pyr + (int64_t)frame * fl.frame_bytes + fl.off[level];
  const int scp = fast_scp(rw), wpr = (rw + 31) >> 5, nbw = wpr * rh;
  uint8_t* s_sc = scratch + (int64_t)frame * scratch_per_frame + c.big_off;      // [rh + 2][scp], pixel x at byte 4 + x
  uint32_t* bmA = reinterpret_cast<uint32_t*>(s_sc + fast_sc_bytes(rw, rh));
  uint32_t* bmB = bmA + nbw;
  const int tA = fl.fast_th, tB = kRetryTh;
  int tmin = max(tA > tB ? tA : min(tA, tB), 1);
  const int nz = (fast_sc_bytes(rw, rh) + 8 * nbw) >> 2;
  // the threshold-free score map is computed once; a retry at 7 only repeats the suppression
  for (int i = tid; i < nz; i += nthr) reinterpret_cast<uint32_t*>(s_sc)[i] = 0;
  if (tid < 2) s_cnt[tid] = 0;
  __syncthreads();
  for (int i = tid; i < rw * rh; i += nthr) {
    const int y = i / rw, x = i - y * rw;
    const int sc = fast_score_scalar(img + (int64_t)(c.y0 + y) * pitch + c.x0 + x, pitch);
    if (sc) s_sc[(y + 1) * scp + 4 + x] = (uint8_t)sc;
  }
  __syncthreads();
  bool useA;
  for (;;) {
    int nA = 0, nB = 0;
    for (int i = tid; i < rw * rh; i += nthr) {
      const int y = i / rw, x = i - y * rw;
      const uint8_t* p = s_sc + (y + 1) * scp + 4 + x;
      const int sc = p[0];
      if (sc < tmin) continue;
      const int m = max(max(max(p[-1], p[1]), max(p[-scp - 1], p[-scp])),
                        max(max(p[-scp + 1], p[scp - 1]), max(p[scp], p[scp + 1])));
      if (sc > m) {
        const int w = y * wpr + (x >> 5);
        if (sc >= tA) { atomicOr(&bmA[w], 1u << (x & 31)); nA++; }
        if (sc >= tB) { atomicOr(&bmB[w], 1u << (x & 31)); nB++; }
      }
    }
    nA = __reduce_add_sync(0xffffffffu, nA);
    nB = __reduce_add_sync(0xffffffffu, nB);
    if (lane == 0 && (nA | nB)) { atomicAdd(&s_cnt[0], nA); atomicAdd(&s_cnt[1], nB); }
    __syncthreads();
    useA = s_cnt[0] > 3;                          // :609-614
    if (useA || tmin <= tB) break;
    __syncthreads();
    for (int i = tid; i < 2 * nbw; i += nthr) bmA[i] = 0;
    if (tid < 2) s_cnt[tid] = 0;
    tmin = tB;
    __syncthreads();
  }
  const int total = useA ? s_cnt[0] : s_cnt[1];
  int32_t* n_total = b.n_total + (int64_t)frame * fl.total_cells + cell;
  if (tid == 0) *n_total = total;
  if (total > c.cand_cap) {
    if (tid == 0) atomicExch(b.err_flag, 1);
    return;
  }
  if (total == 0 || tid >= 32) return;
  const uint32_t* bm = useA ? bmA : bmB;
  const int per = (nbw + 31) >> 5, w0 = min(lane * per, nbw), w1 = min(w0 + per, nbw);
  int cnt = 0;
  for (int i = w0; i < w1; i++) cnt += __popc(bm[i]);
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  if (cnt == 0) return;
  Cand* out = b.cand + (int64_t)frame * fl.cand_per_frame + c.cand_off + (incl - cnt);
  for (int i = w0; i < w1; i++) {
    uint32_t bits = bm[i];
    if (!bits) continue;
    const int y = i / wpr, xb = 32 * (i - y * wpr);
    const int ly = c.y0 + y;
    while (bits) {
      const int x = xb + __ffs(bits) - 1;
      bits &= bits - 1;
      const int lx = c.x0 + x;
      Cand r;
      if constexpr (kHarris) {
        const LevelPlan& LP = b.plan->lv[level];
        r.resp = (lx >= 4 && ly >= 4 && lx + 4 < LP.w && ly + 4 < LP.h)
                     ? harris_response(img + (int64_t)ly * pitch + lx, pitch)
                     : harris_response_edge(img, pitch, lx, ly, LP.w, LP.h);
      } else {
        r.resp = (float)s_sc[(y + 1) * scp + 4 + x];
      }
      r.xy = ((uint32_t)ly << 16) | (uint32_t)lx;
      *out++ = r;
    }
  }
}

void launch_fast_cells_big(const ChunkBuffers& b, const DevicePlan& hp, const int32_t* d_big_cells, int nbig, uint8_t* d_scratch,
                           int64_t scratch_per_frame, int nframes, cudaStream_t s) {
  auto* kern = hp.score_type == ORBX_HARRIS_SCORE ? fast_cells_big_kernel<true> : fast_cells_big_kernel<false>;
  FastLevels fl;
  fl.fast_th = hp.fast_th; fl.score_type = hp.score_type; fl.total_cells = hp.total_cells;
  fl.cand_per_frame = hp.cand_per_frame; fl.frame_bytes = hp.pyr_frame_bytes;
  for (int l = 0; l < kMaxLevels; l++) { fl.pitch[l] = hp.lv[l].pitch; fl.off[l] = hp.lv[l].off; }
  launch_k(kern, dim3(nbig, nframes), dim3(1024), 0, s, b, fl, d_big_cells, d_scratch, scratch_per_frame);
}

#include "orbx_kernels_detect_tma.inl"

}  // namespace orbx
#include "orbx_kernels_select.inl"
