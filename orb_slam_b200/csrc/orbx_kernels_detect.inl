// Included by orbx_kernels.cu.  E3/E4 per-cell FAST detection and E5/E6 quota + retainBest selection.
namespace orbx {

// ------------------------------------------------------------------------------------------------------------------
// E3  FAST-9/16 score (threshold-free):  S = max over the 16 arcs of 9 contiguous ring pixels of
//     min(d) and min(-d), minus 1, d_k = I(p) - I(ring_k); p is a corner at threshold t  <=>  S >= t
//     (cv::FAST / cornerScore<16>).  Scores below `tmin` are reported as 0.
// ------------------------------------------------------------------------------------------------------------------
// NOTE (CUDA 12.9 ptxas, sm_100a): the textbook form max(min9(d), -max9(d)) with d = I(p) - ring is MISCOMPILED —
// ptxas fuses the neg.s32 into a 3-input VIMNMX3 and drops the negation (found with tools/dbg_fast2.cu, the PTX is
// correct, the SASS is not).  The score is therefore computed on the RAW ring values with no negation or abs
// anywhere:  S = max( max_k min9(ring) - v ,  v - min_k max9(ring) ) - 1.
__device__ __forceinline__ int fast_score(const uint8_t* p, int pitch, int tmin) {
  const int v = p[0];
  const int lo = v - tmin, hi = v + tmin;
  int r[16];
  r[0] = p[3 * pitch];   r[8] = p[-3 * pitch];
  r[4] = p[3];           r[12] = p[-3];
  // opposite-pair quick reject: an arc of 9 contains one pixel of every opposite pair, so a pair with both
  // pixels inside [v - tmin, v + tmin] rules out every score >= tmin
  if ((max(r[0], r[8]) <= hi && min(r[0], r[8]) >= lo) || (max(r[4], r[12]) <= hi && min(r[4], r[12]) >= lo))
    return 0;
  r[1] = p[3 * pitch + 1];   r[9] = p[-3 * pitch - 1];
  r[2] = p[2 * pitch + 2];   r[10] = p[-2 * pitch - 2];
  r[3] = p[pitch + 3];       r[11] = p[-pitch - 3];
  r[5] = p[-pitch + 3];      r[13] = p[pitch - 3];
  r[6] = p[-2 * pitch + 2];  r[14] = p[2 * pitch - 2];
  r[7] = p[-3 * pitch + 1];  r[15] = p[3 * pitch - 1];
  // min/max over every window of 9 consecutive ring pixels, via windows of 2, 4, 8 (+1)
  int mn2[16], mx2[16];
#pragma unroll
  for (int k = 0; k < 16; k++) { mn2[k] = min(r[k], r[(k + 1) & 15]); mx2[k] = max(r[k], r[(k + 1) & 15]); }
  int mn4[16], mx4[16];
#pragma unroll
  for (int k = 0; k < 16; k++) { mn4[k] = min(mn2[k], mn2[(k + 2) & 15]); mx4[k] = max(mx2[k], mx2[(k + 2) & 15]); }
  int brightest_min = 0, darkest_max = 255;
#pragma unroll
  for (int k = 0; k < 16; k++) {
    brightest_min = max(brightest_min, min(min(mn4[k], mn4[(k + 4) & 15]), r[(k + 8) & 15]));
    darkest_max = min(darkest_max, max(max(mx4[k], mx4[(k + 4) & 15]), r[(k + 8) & 15]));
  }
  const int s = max(brightest_min - v, v - darkest_max) - 1;
  return s >= tmin ? s : 0;
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
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) fast_cells_kernel(ChunkBuffers b, int smem_half) {
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
  __shared__ int s_warp[8];
  __shared__ int s_base;
  int pitch;
  const uint8_t* img = level_ptr(b, P, c.level, frame, &pitch);
  const int cw = c.rw + 6, ch = c.rh + 6;       // the cell sub-image the reference hands to cv::FAST
  const int sp = (cw + 3) & ~3;                 // shared pitch of the sub-image
  uint8_t* s_img = smem;                         // [ch][sp]   (later reused for the NMS survivors' scores)
  uint8_t* s_sc = smem + smem_half;              // [(rh+2)][(rw+2)] score map with a zero ring
  const int scp = c.rw + 2;
  const uint8_t* gsub = img + (int64_t)(c.y0 - 3) * pitch + (c.x0 - 3);
  for (int i = tid; i < ch * cw; i += 256) {
    const int r = i / cw, q = i - r * cw;
    s_img[r * sp + q] = gsub[(int64_t)r * pitch + q];
  }
  for (int i = tid; i < (c.rh + 2) * scp; i += 256) s_sc[i] = 0;
  if (tid < 2) s_cnt[tid] = 0;
  __syncthreads();
  const int tA = P->fast_th, tB = kRetryTh, tmin = max(min(tA, tB), 1);
  const int npx = c.rw * c.rh;
  for (int i = tid; i < npx; i += 256) {
    const int y = i / c.rw, x = i - y * c.rw;
    const int s = fast_score(s_img + (y + 3) * sp + (x + 3), sp, tmin);
    if (s) s_sc[(y + 1) * scp + (x + 1)] = (uint8_t)s;
  }
  __syncthreads();
  // NMS -> survivors' scores into s_ks (aliases s_img, no longer needed), counts at both thresholds
  uint8_t* s_ks = s_img;
  int nA = 0, nB = 0;
  for (int i = tid; i < npx; i += 256) {
    const int y = i / c.rw, x = i - y * c.rw;
    const uint8_t* q = s_sc + (y + 1) * scp + (x + 1);
    const int s = q[0];
    int keep = 0;
    if (s) {
      const int m = max(max(max(q[-1], q[1]), max(q[-scp - 1], q[-scp])),
                        max(max(q[-scp + 1], q[scp - 1]), max(q[scp], q[scp + 1])));
      keep = s > m;
    }
    s_ks[i] = keep ? (uint8_t)s : 0;
    nA += keep && s >= tA;
    nB += keep && s >= tB;
  }
  nA = __reduce_add_sync(0xffffffffu, nA);
  nB = __reduce_add_sync(0xffffffffu, nB);
  if (lane == 0) { atomicAdd(&s_cnt[0], nA); atomicAdd(&s_cnt[1], nB); }
  __syncthreads();
  const int t = s_cnt[0] > 3 ? tA : tB;          // :609-614
  const int total = s_cnt[0] > 3 ? s_cnt[0] : s_cnt[1];
  if (tid == 0) { *n_total = total; s_base = 0; }
  if (total > c.cand_cap) {                       // cannot happen (NMS survivors are non-adjacent); never truncate
    if (tid == 0) atomicExch(b.err_flag, 1);
    return;
  }
  Cand* out = b.cand + (int64_t)frame * P->cand_per_frame + c.cand_off;
  const int lvpitch = pitch;
  for (int base = 0; base < npx; base += 256) {   // ordered compaction, 256 row-major pixels per round
    __syncthreads();
    const int i = base + tid;
    const int s = i < npx ? s_ks[i] : 0;
    const bool on = s >= t && s > 0;
    const unsigned bal = __ballot_sync(0xffffffffu, on);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int off = s_base;
    for (int wq = 0; wq < warp; wq++) off += s_warp[wq];
    if (on) {
      const int y = i / c.rw, x = i - y * c.rw;
      const int lx = c.x0 + x, ly = c.y0 + y;
      Cand r;
      r.resp = P->score_type == ORBX_HARRIS_SCORE ? harris_response(img + (int64_t)ly * lvpitch + lx, lvpitch)
                                                  : (float)s;
      r.xy = ((uint32_t)ly << 16) | (uint32_t)lx;
      out[off + __popc(bal & ((1u << lane) - 1))] = r;
    }
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int wq = 0; wq < 8; wq++) tot += s_warp[wq];
      s_base += tot;
    }
  }
}

int fast_cells_smem_bytes(int max_cell_px) {
  // max_cell_px already counts the padded shared pitch; the ringed score map is smaller; halves aligned to 16
  const int half = (max_cell_px + 15) & ~15;
  return 2 * half;
}

void launch_fast_cells(const ChunkBuffers& b, const DevicePlan& hp, int max_cell_px, int nframes, cudaStream_t s) {
  const int bytes = fast_cells_smem_bytes(max_cell_px);
  if (bytes > 48 * 1024)
    cudaFuncSetAttribute(fast_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  dim3 grid(hp.total_cells, nframes);
  fast_cells_kernel<<<grid, 256, bytes, s>>>(b, bytes / 2);
}

}  // namespace orbx
#include "orbx_kernels_select.inl"
