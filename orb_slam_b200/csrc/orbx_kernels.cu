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
  uint32_t packed = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int dx = dx0 + k;
    if (dx < L.w) {
      const ResizeX rx = b.rx[L.tab_x + dx];
      const int sx = rx.sx, sx1 = min(sx + 1, sw - 1);
      const int h0 = S0[sx] * rx.a0 + S0[sx1] * rx.a1;
      const int h1 = S1[sx] * rx.a0 + S1[sx1] * rx.a1;
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
// One CTA = 64x32 output tile of one level of one frame; all levels in one launch.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kBlurTW = 64, kBlurTH = 32;

__global__ void __launch_bounds__(256) blur7_kernel(ChunkBuffers b, BlurTiles tiles) {
  const DevicePlan* P = b.plan;
  int level = 0;
  while (level + 1 < P->nlevels && (int)blockIdx.x >= tiles.base[level + 1]) level++;
  const LevelPlan& L = P->lv[level];
  const int t = blockIdx.x - tiles.base[level];
  const int tx = t % tiles.tiles_x[level], ty = t / tiles.tiles_x[level];
  const int frame = blockIdx.y;
  const int x0 = tx * kBlurTW, y0 = ty * kBlurTH;
  int spitch;
  const uint8_t* src = level_ptr(b, P, level, frame, &spitch);
  uint8_t* dst = b.blur + (int64_t)frame * P->blur_frame_bytes + L.off;

  __shared__ uint8_t s_in[kBlurTH + 6][kBlurTW + 8];
  __shared__ uint16_t s_h[kBlurTH + 6][kBlurTW];
  const int tid = threadIdx.x;
  for (int i = tid; i < (kBlurTH + 6) * (kBlurTW + 6); i += 256) {
    const int r = i / (kBlurTW + 6), c = i % (kBlurTW + 6);
    const int yy = reflect101(min(y0 + r - 3, L.h + 2), L.h), xx = reflect101(min(x0 + c - 3, L.w + 2), L.w);
    s_in[r][c] = src[(int64_t)yy * spitch + xx];
  }
  __syncthreads();
  const int k0 = 18, k1 = 34, k2 = P->blur_mode == 0 ? 49 : 48, k3 = P->blur_mode == 0 ? 55 : 56;
  for (int i = tid; i < (kBlurTH + 6) * kBlurTW; i += 256) {
    const int r = i / kBlurTW, c = i % kBlurTW;
    const uint8_t* p = &s_in[r][c];
    const int v = k0 * (p[0] + p[6]) + k1 * (p[1] + p[5]) + k2 * (p[2] + p[4]) + k3 * p[3];
    s_h[r][c] = (uint16_t)v;  // <= 255*257 = 65535
  }
  __syncthreads();
  // each thread: 4 adjacent columns of 2 rows -> 32-bit stores
  for (int i = tid; i < kBlurTH * (kBlurTW / 4); i += 256) {
    const int r = i / (kBlurTW / 4), c4 = (i % (kBlurTW / 4)) * 4;
    const int y = y0 + r;
    if (y >= L.h || x0 + c4 >= L.w) continue;
    uint32_t packed = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int c = c4 + k;
      const int v = k0 * (s_h[r][c] + s_h[r + 6][c]) + k1 * (s_h[r + 1][c] + s_h[r + 5][c]) +
                    k2 * (s_h[r + 2][c] + s_h[r + 4][c]) + k3 * s_h[r + 3][c];
      packed |= (uint32_t)min((v + 32768) >> 16, 255) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(dst + (int64_t)y * L.pitch + x0 + c4) = packed;  // padding columns are scratch
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
