// sm_100a kernels of the ORB extractor.  Every kernel cites the reference lines whose arithmetic it reproduces
// (worxli/ORB_SLAM src/ORBextractor.cc) and the OpenCV primitive model it implements.
// All image arithmetic is exact integer work on bytes; the only float code is fastAtan2 and the BRIEF rotation,
// written with explicitly rounded intrinsics (no FMA contraction) so results equal the CPU's bit for bit.
#include <cfloat>

#include "orbx_kernels.cuh"

namespace orbx {

// ------------------------------------------------------------------------------------------------------------------
// helpers
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int reflect101(int i, int n) {  // copyMakeBorder(BORDER_REFLECT_101), |overhang| < n
  i = i < 0 ? -i : i;
  return i >= n ? 2 * (n - 1) - i : i;
}

__device__ __forceinline__ const uint8_t* level_ptr(const ChunkBuffers& b, const DevicePlan* P, int level, int frame,
                                                    int* pitch) {
  if (level == 0) {
    *pitch = b.lv0_pitch;
    return b.lv0 + (int64_t)frame * b.lv0_frame_stride;
  }
  *pitch = P->lv[level].pitch;
  return b.pyr + (int64_t)frame * P->pyr_frame_bytes + P->lv[level].off;
}

// ------------------------------------------------------------------------------------------------------------------
// E2  image pyramid: level l = cv::resize(level l-1, INTER_LINEAR)           (reference :781-822, resize at :800)
// 11-bit fixed point: H = S[sx]*a0 + S[sx+1]*a1;  dst = (((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2
// One thread produces 4 horizontally adjacent destination pixels and stores them as one 32-bit word.
// The 16-px reflect border of the reference (copyMakeBorder :806/:814) is never materialised: consumers reflect
// indices instead.
// ------------------------------------------------------------------------------------------------------------------
// Fast path (aligned source rows, scale < ~2.3): the thread's 4 destination pixels read source columns
// sx0 .. sx3 + 1, at most 8 bytes, so each of the two source rows is fetched as three aligned 32-bit words, shifted
// once to start at sx0, and every pixel's (S0[sx], S0[sx+1], S1[sx], S1[sx+1]) quadruple is three PRMTs away; the two
// horizontal interpolations are then ONE DP2A each (u8 pair . s16 weight pair).
__global__ void __launch_bounds__(256) resize_level_kernel(ChunkBuffers b, int level) {
  const DevicePlan* P = b.plan;
  const LevelPlan& L = P->lv[level];
  const int frame = blockIdx.z;
  const int dx0 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int dy = blockIdx.y * 8 + threadIdx.y;
  if (dx0 >= L.w || dy >= L.h) return;
  int spitch;
  const uint8_t* src = level_ptr(b, P, level - 1, frame, &spitch);
  uint8_t* dst = b.pyr + (int64_t)frame * P->pyr_frame_bytes + L.off;
  const ResizeY ry = b.ry[L.tab_y + dy];
  const uint8_t* S0 = src + (int64_t)ry.sy0 * spitch;
  const uint8_t* S1 = src + (int64_t)ry.sy1 * spitch;
  const int sw = P->lv[level - 1].w;
  uint2 rx[4];
#pragma unroll
  for (int k = 0; k < 4; k++) rx[k] = reinterpret_cast<const uint2*>(b.rx)[L.tab_x + min(dx0 + k, L.w - 1)];
  const int sx0 = rx[0].x & 0xffff, sx3 = rx[3].x & 0xffff;
  const bool fast = ((spitch & 3) == 0) && (((uintptr_t)src & 3) == 0) && (sx3 + 1 - sx0 <= 7);
  uint32_t packed = 0;
  if (fast) {
    const int base4 = sx0 & ~3, sh = (sx0 & 3) * 8;
    const uint32_t* A = reinterpret_cast<const uint32_t*>(S0 + base4);
    const uint32_t* B = reinterpret_cast<const uint32_t*>(S1 + base4);
    const bool in1 = base4 + 4 < sw, in2 = base4 + 8 < sw;   // never start a load beyond the row
    const uint32_t a0 = A[0], a1 = in1 ? A[1] : 0u, a2 = in2 ? A[2] : 0u;
    const uint32_t c0 = B[0], c1 = in1 ? B[1] : 0u, c2 = in2 ? B[2] : 0u;
    const uint32_t wa0 = __funnelshift_r(a0, a1, sh), wa1 = __funnelshift_r(a1, a2, sh);  // bytes sx0.. of row 0
    const uint32_t wb0 = __funnelshift_r(c0, c1, sh), wb1 = __funnelshift_r(c1, c2, sh);  // ... of row 1
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int o = (int)(rx[k].x & 0xffff) - sx0;
      const uint32_t sel = (uint32_t)o | ((uint32_t)(o + 1) << 4);
      const uint32_t q = __byte_perm(__byte_perm(wa0, wa1, sel), __byte_perm(wb0, wb1, sel), 0x5410);
      const uint32_t h0 = __dp2a_lo(rx[k].y, q, 0u), h1 = __dp2a_hi(rx[k].y, q, 0u);
      const int v = (((ry.b0 * (int)(h0 >> 4)) >> 16) + ((ry.b1 * (int)(h1 >> 4)) >> 16) + 2) >> 2;
      packed |= (uint32_t)(v & 255) << (8 * k);
    }
  } else {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int sx = rx[k].x & 0xffff, sx1 = min(sx + 1, sw - 1);
      const int ca0 = (int)(short)(rx[k].y & 0xffff), ca1 = (int)(short)(rx[k].y >> 16);
      const int h0 = S0[sx] * ca0 + S0[sx1] * ca1;
      const int h1 = S1[sx] * ca0 + S1[sx1] * ca1;
      const int v = (((ry.b0 * (h0 >> 4)) >> 16) + ((ry.b1 * (h1 >> 4)) >> 16) + 2) >> 2;
      packed |= (uint32_t)(v & 255) << (8 * k);
    }
  }
  *reinterpret_cast<uint32_t*>(dst + (int64_t)dy * L.pitch + dx0) = packed;  // pitch % 16 == 0, dx0 % 4 == 0
}

void launch_resize_level(const ChunkBuffers& b, const DevicePlan& hp, int level, int nframes, cudaStream_t s) {
  const LevelPlan& L = hp.lv[level];
  dim3 block(32, 8), grid((L.w + 127) / 128, (L.h + 7) / 8, nframes);
  resize_level_kernel<<<grid, block, 0, s>>>(b, level);
}

// ------------------------------------------------------------------------------------------------------------------
// E8  7x7 Gaussian blur, sigma 2, BORDER_REFLECT_101                                            (reference :760)
// Separable integer taps, dst = (sum_j k_j sum_i k_i p + 32768) >> 16  (SURVEY §0.3 P3).  The reference blurs in
// place and keeps the unblurred border; here the blurred image goes to its own slab and the unblurred level stays
// intact (the descriptor kernel reads the unblurred reflect for the <= 2 px of patch overhang).
// One CTA = 64x64 output tile of one level of one frame; all levels in one launch.
//
// Both passes run on PACKED u16x2 lanes, two pixels per IMAD:
//   horizontal  pixels (x, x+2) share a register (even / odd bytes of a word are one AND away from that form);
//               h = k0 (S[-3] + S[3]) + k1 (S[-2] + S[2]) + k2 (S[-1] + S[1]) + k3 S[0] <= 255 * 257 = 65535 fits a lane
//   vertical    sum_j k_j h_j needs 25 bits, so h is split into its low and high BYTE planes at the end of the
//               horizontal pass: sum_j k_j hL_j and sum_j k_j hH_j each fit a lane again, and
//               v = 256 * VH + VL is rebuilt per pixel only for the final (v + 32768) >> 16.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kBlurT = 64, kBlurRows = kBlurT + 6, kBlurWords = 18;  // staged row = image columns x0-4 .. x0+67

__device__ __forceinline__ uint32_t packed_mid(uint32_t lo, uint32_t hi) { return __byte_perm(lo, hi, 0x5432); }

__global__ void __launch_bounds__(256) blur7_kernel(ChunkBuffers b, BlurTiles tiles) {
  const DevicePlan* P = b.plan;
  int level = 0;
  while (level + 1 < P->nlevels && (int)blockIdx.x >= tiles.base[level + 1]) level++;
  const LevelPlan& L = P->lv[level];
  const int t = blockIdx.x - tiles.base[level];
  const int tx = t % tiles.tiles_x[level], ty = t / tiles.tiles_x[level];
  const int frame = blockIdx.y;
  const int x0 = tx * kBlurT, y0 = ty * kBlurT;
  const int lw = L.w, lh = L.h;
  int spitch;
  const uint8_t* src = level_ptr(b, P, level, frame, &spitch);
  uint8_t* dst = b.blur + (int64_t)frame * P->blur_frame_bytes + L.off;

  __shared__ uint32_t s_in[kBlurRows][kBlurWords];
  __shared__ uint4 s_h[kBlurRows][kBlurT / 4];  // per 4 output columns: {L(o0,o2), L(o1,o3), H(o0,o2), H(o1,o3)}
  const int tid = threadIdx.x;
  const bool words_ok = ((spitch & 3) == 0) && (((uintptr_t)src & 3) == 0);
  for (int i = tid; i < kBlurRows * kBlurWords; i += 256) {
    const int r = i / kBlurWords, j = i - r * kBlurWords;
    const int yy = reflect101(min(y0 + r - 3, lh + 2), lh);
    const int xs = x0 - 4 + 4 * j;
    const uint8_t* row = src + (int64_t)yy * spitch;
    uint32_t w;
    if (words_ok && xs >= 0 && xs + 3 < lw) {
      w = *reinterpret_cast<const uint32_t*>(row + xs);
    } else {
      w = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) w |= (uint32_t)row[reflect101(min(max(xs + k, -3), lw + 2), lw)] << (8 * k);
    }
    s_in[r][j] = w;
  }
  __syncthreads();
  const uint32_t k0 = 18, k1 = 34, k2 = P->blur_mode == 0 ? 49 : 48, k3 = P->blur_mode == 0 ? 55 : 56;
  const uint32_t M = 0x00FF00FFu;
  for (int i = tid; i < kBlurRows * (kBlurT / 4); i += 256) {
    const int r = i >> 4, k = i & 15;
    const uint32_t w0 = s_in[r][k], w1 = s_in[r][k + 1], w2 = s_in[r][k + 2];
    // S_i = (B_i, B_{i+2}) with B_i = staged byte 4k + i; output o_j is centred on B_{4+j}
    const uint32_t E0 = w0 & M, E1 = w1 & M, E2 = w2 & M, O0 = (w0 >> 8) & M, O1 = (w1 >> 8) & M, O2 = (w2 >> 8) & M;
    const uint32_t S1 = O0, S2 = packed_mid(E0, E1), S3 = packed_mid(O0, O1), S4 = E1, S5 = O1,
                   S6 = packed_mid(E1, E2), S7 = packed_mid(O1, O2), S8 = E2;
    const uint32_t hA = k0 * (S1 + S7) + k1 * (S2 + S6) + k2 * (S3 + S5) + k3 * S4;  // (o0, o2)
    const uint32_t hB = k0 * (S2 + S8) + k1 * (S3 + S7) + k2 * (S4 + S6) + k3 * S5;  // (o1, o3)
    s_h[r][k] = make_uint4(hA & M, hB & M, (hA >> 8) & M, (hB >> 8) & M);
  }
  __syncthreads();
  {
    const int k = tid & 15, strip = tid >> 4;  // 4 columns x 4 rows per thread
    const int x = x0 + 4 * k;
    if (x >= lw || y0 + 4 * strip >= lh) return;
    uint4 a[10];
#pragma unroll
    for (int q = 0; q < 10; q++) a[q] = s_h[4 * strip + q][k];
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int y = y0 + 4 * strip + q;
      if (y >= lh) break;
#define BLURV(c) (k0 * (a[q].c + a[q + 6].c) + k1 * (a[q + 1].c + a[q + 5].c) + k2 * (a[q + 2].c + a[q + 4].c) + k3 * a[q + 3].c)
      const uint32_t AL = BLURV(x), BL = BLURV(y), AH = BLURV(z), BH = BLURV(w);
#undef BLURV
      const uint32_t r0 = min((((AH & 0xffff) << 8) + (AL & 0xffff) + 32768u) >> 16, 255u);
      const uint32_t r2 = min((((AH >> 16) << 8) + (AL >> 16) + 32768u) >> 16, 255u);
      const uint32_t r1 = min((((BH & 0xffff) << 8) + (BL & 0xffff) + 32768u) >> 16, 255u);
      const uint32_t r3 = min((((BH >> 16) << 8) + (BL >> 16) + 32768u) >> 16, 255u);
      *reinterpret_cast<uint32_t*>(dst + (int64_t)y * L.pitch + x) = r0 | (r1 << 8) | (r2 << 16) | (r3 << 24);
    }
  }
}

void launch_blur(const ChunkBuffers& b, const DevicePlan& hp, const BlurTiles& t, int nframes, cudaStream_t s) {
  dim3 grid(t.base[hp.nlevels], nframes);
  blur7_kernel<<<grid, 256, 0, s>>>(b, t);
}

// ------------------------------------------------------------------------------------------------------------------
// Synthetic frames (bench / test input; CPU twin: oracle/synth.cc).  Integer-only so both sides agree byte for byte.
// ------------------------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t seed, uint64_t i) {
  uint64_t z = seed * 0xD1342543DE82EF95ull + (i + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// pass 1: paint the rectangles (last one covering a pixel wins)
__global__ void __launch_bounds__(256) synth_paint_kernel(uint8_t* tmp, int w, int h, uint64_t seed0) {
  extern __shared__ int s_rect[];  // x0,y0,x1,y1,val per rectangle
  const int frame = blockIdx.y;
  const uint64_t seed = seed0 + frame;
  const int nrect = w * h / 768;
  for (int r = threadIdx.x; r < nrect; r += blockDim.x) {
    const uint64_t u = mix64(seed, (uint64_t)r);
    const int x0 = (int)(((u & 0xFFFF) * (uint64_t)w) >> 16);
    const int y0 = (int)((((u >> 16) & 0xFFFF) * (uint64_t)h) >> 16);
    const int ww = 8 + (int)((((u >> 32) & 0xFFFF) * 88ull) >> 16);
    const int hh = 8 + (int)((((u >> 48) & 0xFFFF) * 88ull) >> 16);
    s_rect[5 * r + 0] = x0; s_rect[5 * r + 1] = y0;
    s_rect[5 * r + 2] = min(x0 + ww, w); s_rect[5 * r + 3] = min(y0 + hh, h);
    s_rect[5 * r + 4] = (int)(mix64(seed, 0x10000ull + r) & 255);
  }
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < w * h; i += gridDim.x * blockDim.x) {
    const int x = i % w, y = i / w;
    int v = 128;
    for (int r = nrect - 1; r >= 0; r--)
      if (x >= s_rect[5 * r] && x < s_rect[5 * r + 2] && y >= s_rect[5 * r + 1] && y < s_rect[5 * r + 3]) {
        v = s_rect[5 * r + 4];
        break;
      }
    tmp[(int64_t)frame * w * h + i] = (uint8_t)v;
  }
}

// pass 2: 5x5 binomial blur (reflect-101) + uniform noise in [-2,2]
__global__ void __launch_bounds__(256) synth_finish_kernel(const uint8_t* tmp, uint8_t* out, int w, int h,
                                                           size_t stride, size_t frame_stride, uint64_t seed0) {
  const int frame = blockIdx.y;
  const uint64_t seed = seed0 + frame;
  const uint8_t* src = tmp + (int64_t)frame * w * h;
  const int k[5] = {1, 4, 6, 4, 1};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < w * h; i += gridDim.x * blockDim.x) {
    const int x = i % w, y = i / w;
    int s = 0;
#pragma unroll
    for (int j = 0; j < 5; j++) {
      const uint8_t* row = src + (int64_t)reflect101(y + j - 2, h) * w;
      int hs = 0;
#pragma unroll
      for (int q = 0; q < 5; q++) hs += k[q] * row[reflect101(x + q - 2, w)];
      s += k[j] * hs;
    }
    int v = (s + 128) >> 8;
    v += (int)(mix64(seed, 0x1000000ull + (uint64_t)y * w + x) % 5) - 2;
    out[(int64_t)frame * frame_stride + (int64_t)y * stride + x] = (uint8_t)min(max(v, 0), 255);
  }
}

void launch_synth(uint8_t* d_out, uint8_t* d_tmp, int nframes, int w, int h, size_t stride, size_t frame_stride,
                  uint64_t seed0, cudaStream_t s) {
  const int nrect = w * h / 768;
  dim3 grid(148, nframes);
  synth_paint_kernel<<<grid, 256, nrect * 5 * sizeof(int), s>>>(d_tmp, w, h, seed0);
  synth_finish_kernel<<<grid, 256, 0, s>>>(d_tmp, d_out, w, h, stride, frame_stride, seed0);
}

}  // namespace orbx

#include "orbx_kernels_detect.inl"
#include "orbx_kernels_describe.inl"
