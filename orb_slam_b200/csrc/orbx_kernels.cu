// sm_100a kernels of the ORB extractor.  Every kernel cites the reference lines whose arithmetic it reproduces
// (worxli/ORB_SLAM src/ORBextractor.cc) and the OpenCV primitive model it implements.
// All image arithmetic is exact integer work on bytes; the only float code is fastAtan2 and the BRIEF rotation,
// written with explicitly rounded intrinsics (no FMA contraction) so results equal the CPU's bit for bit.
#include <cfloat>

#include "orbx_kernels.cuh"

#include <algorithm>
#include <cstdlib>
#include <utility>

namespace orbx {

static thread_local bool g_launch_pdl = false;
void set_launch_pdl(bool on) { g_launch_pdl = on; }

// Launch with (optionally) the programmatic-stream-serialization attribute; see set_launch_pdl.
template <class... KArgs, class... Args>
static void launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = g_launch_pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// First statement of every chained kernel: let the successor start launching, then wait until the predecessor grid
// has completed and its writes are visible.  Both are no-ops for a launch without the attribute.
#define ORBX_PDL_ENTRY()                                         \
  do {                                                           \
    asm volatile("griddepcontrol.launch_dependents;");           \
    asm volatile("griddepcontrol.wait;" ::: "memory");           \
  } while (0)

// ------------------------------------------------------------------------------------------------------------------
// helpers
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int reflect101(int i, int n) {  // copyMakeBorder(BORDER_REFLECT_101), |overhang| < n
  i = i < 0 ? -i : i;
  return i >= n ? 2 * (n - 1) - i : i;
}

// floor(i / d) with m = fast_magic(d) (orbx_internal.cuh); m == 0 encodes d == 1
__device__ __forceinline__ int magic_div(int i, uint32_t m) { return m ? (int)__umulhi((uint32_t)i, m) : i; }

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
// One thread = 4 destination columns x kResizeRows destination rows; one warp = 128 columns of those rows.
// The column set-up (table loads, byte selectors) is paid once per thread.  The source rows a warp needs form one
// contiguous range [sy0(first row), sy1(last row)]: all of them are fetched up front (one memory round trip), each
// gets ONE horizontal pass, and a destination row is emitted when the walk reaches its lower source row - consecutive
// destination rows share a source row (sy1(dy) == sy0(dy + 1) for scale < 2), so ~1.3 horizontal passes per
// destination row instead of 2.
// Fast path (aligned source rows, scale < ~2.3): the thread's 4 destination pixels read source columns
// sx0 .. sx3 + 1, at most 8 bytes, so a source row is fetched as three aligned 32-bit words, shifted once to start
// at sx0; every pixel's (S[sx], S[sx+1]) pair is one PRMT away and its horizontal interpolation is ONE DP2A
// (u8 pair . s16 weight pair).
constexpr int kResizeRows = 6, kResizeSrcMax = 10;

__global__ void __launch_bounds__(256) resize_level_kernel(ChunkBuffers b, int level) {
  ORBX_PDL_ENTRY();
  const DevicePlan* P = b.plan;
  const LevelPlan& L = P->lv[level];
  const int frame = blockIdx.z, lane = threadIdx.x;
  const int lw = L.w, lh = L.h, dpitch = L.pitch;
  const int dx_raw = (blockIdx.x * 32 + lane) * 4;
  const int dy0 = (blockIdx.y * 8 + threadIdx.y) * kResizeRows;
  if (dy0 >= lh) return;                                      // whole warp
  const bool col_ok = dx_raw < lw;                            // lanes past the row keep computing (shuffles), never store
  const int dx0 = min(dx_raw, (lw - 1) & ~3);
  int spitch;
  const uint8_t* src = level_ptr(b, P, level - 1, frame, &spitch);
  uint8_t* drow = b.pyr + (int64_t)frame * P->pyr_frame_bytes + L.off + (int64_t)dy0 * dpitch + dx0;
  const int nrows = min(kResizeRows, lh - dy0);
  const uint2 myry = reinterpret_cast<const uint2*>(b.ry)[L.tab_y + dy0 + min(lane, nrows - 1)];  // lane r: row dy0 + r
  const int sw = P->lv[level - 1].w;
  uint2 rx[4];
#pragma unroll
  for (int k = 0; k < 4; k++) rx[k] = reinterpret_cast<const uint2*>(b.rx)[L.tab_x + min(dx0 + k, lw - 1)];
  const int sx0 = rx[0].x & 0xffff, sx3 = rx[3].x & 0xffff;
  const int s_first = (int)(short)(__shfl_sync(0xffffffffu, myry.x, 0) & 0xffff);
  const int s_last = (int)(short)(__shfl_sync(0xffffffffu, myry.x, nrows - 1) >> 16);
  const int ns = s_last - s_first + 1;
  const bool fast = __all_sync(0xffffffffu, ((spitch & 3) == 0) && (((uintptr_t)src & 3) == 0) && (sx3 + 1 - sx0 <= 7)) &&
                    ns >= 1 && ns <= kResizeSrcMax;
  if (fast) {
    const int base4 = sx0 & ~3, sh = (sx0 & 3) * 8;
    const bool in1 = base4 + 4 < sw, in2 = base4 + 8 < sw;   // never start a load beyond the row
    uint32_t raw[kResizeSrcMax][3];
    const uint8_t* scol = src + (int64_t)s_first * spitch + base4;
#pragma unroll
    for (int j = 0; j < kResizeSrcMax; j++) {
      if (j < ns) {
        const uint32_t* A = reinterpret_cast<const uint32_t*>(scol + (int64_t)j * spitch);
        raw[j][0] = A[0]; raw[j][1] = in1 ? A[1] : 0u; raw[j][2] = in2 ? A[2] : 0u;
      }
    }
    uint32_t sel[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t o = (rx[k].x & 0xffff) - (uint32_t)sx0;
      sel[k] = o | ((o + 1) << 4);                            // bytes (S[sx], S[sx+1], -, -): dp2a_lo reads two
    }
    // H[k] = (S[sx_k] * a0 + S[sx_k + 1] * a1) >> 4 of one source row
    auto hrow = [&](const uint32_t (&a)[3], uint32_t (&H)[4]) {
      const uint32_t w0 = __funnelshift_r(a[0], a[1], sh), w1 = __funnelshift_r(a[1], a[2], sh);
#pragma unroll
      for (int k = 0; k < 4; k++) H[k] = __dp2a_lo(rx[k].y, __byte_perm(w0, w1, sel[k]), 0u) >> 4;
    };
    // emits every destination row whose LOWER source row is s_first + j (upper row = Hp unless the row is clamped)
    int r = 0;
    auto emit = [&](int j, const uint32_t (&Hp)[4], const uint32_t (&Hq)[4]) {
      while (r < nrows) {
        const uint32_t yx = __shfl_sync(0xffffffffu, myry.x, r), yy = __shfl_sync(0xffffffffu, myry.y, r);
        const int sy0 = (int)(short)(yx & 0xffff), sy1 = (int)(short)(yx >> 16);
        if (sy1 != s_first + j) break;
        const int b0 = (int)(short)(yy & 0xffff), b1 = (int)(short)(yy >> 16);
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int h0 = (int)(sy0 == sy1 ? Hq[k] : Hp[k]);
          const int v = (((b0 * h0) >> 16) + ((b1 * (int)Hq[k]) >> 16) + 2) >> 2;
          packed |= (uint32_t)(v & 255) << (8 * k);
        }
        if (col_ok) *reinterpret_cast<uint32_t*>(drow) = packed;   // pitch % 16 == 0, dx0 % 4 == 0
        drow += dpitch;
        r++;
      }
    };
    uint32_t Hp[4], Hq[4];
    hrow(raw[0], Hp);
    emit(0, Hp, Hp);
#pragma unroll
    for (int j = 1; j < kResizeSrcMax; j++) {
      if (j >= ns) break;
      hrow(raw[j], Hq);
      emit(j, Hp, Hq);
#pragma unroll
      for (int k = 0; k < 4; k++) Hp[k] = Hq[k];
    }
  } else if (col_ok) {
    for (int r = 0; r < nrows; r++) {
      const ResizeY ry = b.ry[L.tab_y + dy0 + r];
      const uint8_t* S0 = src + (int64_t)ry.sy0 * spitch;
      const uint8_t* S1 = src + (int64_t)ry.sy1 * spitch;
      uint32_t packed = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int sx = rx[k].x & 0xffff, sx1 = min(sx + 1, sw - 1);
        const int ca0 = (int)(short)(rx[k].y & 0xffff), ca1 = (int)(short)(rx[k].y >> 16);
        const int h0 = S0[sx] * ca0 + S0[sx1] * ca1;
        const int h1 = S1[sx] * ca0 + S1[sx1] * ca1;
        const int v = (((ry.b0 * (h0 >> 4)) >> 16) + ((ry.b1 * (h1 >> 4)) >> 16) + 2) >> 2;
        packed |= (uint32_t)(v & 255) << (8 * k);
      }
      *reinterpret_cast<uint32_t*>(drow) = packed;
      drow += dpitch;
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// E2, TMA-staged form.  One CTA = a 128 x 64 tile of destination level l of one frame; its source rectangle (level
// l - 1) arrives as ONE 3-D tensor copy (cp.async.bulk.tensor -> UTMALDG, out-of-image bytes zero-filled, start column
// rounded down to the 16-byte boundary the engine wants).  A warp owns 128 columns x 16 rows; a thread 4 columns, walking
// down its rows with the horizontal interpolations of the two current source rows in registers (a source row is
// interpolated once: 1.2 horizontal passes per destination row at scale 1.2).  Compared with resize_level_kernel
// (4 x 6 pixels per thread, global loads): the column set-up is paid once per 64 pixels instead of 24, the source
// needs no global addressing, alignment guards or prefetch registers, and the row table comes from shared memory
// instead of shuffles.  Same arithmetic, same tables, byte-identical levels.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kResizeWarpRows = kResizeTileH / 4;

__global__ void __launch_bounds__(128) resize_tma_kernel(ChunkBuffers b, const __grid_constant__ ResizeTma tm, int level) {
  ORBX_PDL_ENTRY();
  extern __shared__ __align__(128) uint8_t smem_rs[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ int4 s_ry[kResizeTileH];                         // per destination row: staged rows of sy0, sy1; b0, b1
  const DevicePlan* P = b.plan;
  const LevelPlan& L = P->lv[level];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, frame = blockIdx.z;
  const int lw = L.w, lh = L.h, dpitch = L.pitch;
  const int dxb = blockIdx.x * kResizeTileW, dyb = blockIdx.y * kResizeTileH;
  const int rp = tm.box_w[level];
  const uint2* rxt = reinterpret_cast<const uint2*>(b.rx) + L.tab_x;
  const uint2* ryt = reinterpret_cast<const uint2*>(b.ry) + L.tab_y;
  const int sx_org = (int)(rxt[dxb].x & 0xffff) & ~15;
  const int sy_org = (int)(short)(ryt[dyb].x & 0xffff);
  const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
  const uint32_t simg = (uint32_t)__cvta_generic_to_shared(smem_rs);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(rp * tm.box_h[level])) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(simg), "l"(reinterpret_cast<uint64_t>(&tm.map[level])), "r"(sx_org), "r"(sy_org), "r"(frame), "r"(bar)
        : "memory");
  }
  if (tid < kResizeTileH) {
    const uint2 e = ryt[min(dyb + tid, lh - 1)];
    s_ry[tid] = make_int4((int)(short)(e.x & 0xffff) - sy_org, (int)(short)(e.x >> 16) - sy_org, (int)(short)(e.y & 0xffff),
                          (int)(short)(e.y >> 16));
  }
  // column set-up (overlaps the copy)
  const int dx_raw = dxb + lane * 4;
  const bool col_ok = dx_raw < lw;                            // lanes past the row keep computing, never store
  const int dx0 = min(dx_raw, (lw - 1) & ~3);
  uint2 rx[4];
#pragma unroll
  for (int k = 0; k < 4; k++) rx[k] = rxt[min(dx0 + k, lw - 1)];
  const int o0 = (int)(rx[0].x & 0xffff) - sx_org;            // byte offset of S[sx0] in a staged row
  const uint32_t scol = simg + (uint32_t)(o0 & ~3);           // shared address of the thread's first word in staged row 0
  const int sh = (o0 & 3) * 8;
  uint32_t sel[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint32_t o = (rx[k].x & 0xffff) - (rx[0].x & 0xffff);
    sel[k] = o | ((o + 1) << 4);                              // bytes (S[sx], S[sx+1], -, -): dp2a_lo reads two
  }
  __syncthreads();                                            // barrier initialised, row table staged
  {
    uint32_t ok = 0;
#pragma unroll 1
    for (int spin = 0; spin < (1 << 22) && !ok; ++spin)
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(ok) : "r"(bar) : "memory");
    if (!ok) {
      if (tid == 0) atomicExch(b.err_flag, 3);
      return;
    }
  }
  const int r0 = warp * kResizeWarpRows;
  const int nrows = min(kResizeWarpRows, lh - dyb - r0);
  if (nrows <= 0) return;
  // H[k] = (S[sx_k] * a0 + S[sx_k + 1] * a1) >> 4 of staged source row j (register selectors: plain PRMT, no masking)
  auto hrow = [&](int j, uint32_t (&H)[4]) {
    uint32_t a0, a1, a2;
    asm volatile("ld.shared.u32 %0, [%3];\n\tld.shared.u32 %1, [%3+4];\n\tld.shared.u32 %2, [%3+8];"
                 : "=r"(a0), "=r"(a1), "=r"(a2) : "r"(scol + (uint32_t)(j * rp)));
    const uint32_t w0 = __funnelshift_r(a0, a1, sh), w1 = __funnelshift_r(a1, a2, sh);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      uint32_t pr;
      asm("prmt.b32 %0, %1, %2, %3;" : "=r"(pr) : "r"(w0), "r"(w1), "r"(sel[k]));
      H[k] = __dp2a_lo(rx[k].y, pr, 0u) >> 4;
    }
  };
  uint8_t* drow = b.pyr + (int64_t)frame * P->pyr_frame_bytes + L.off + (int64_t)(dyb + r0) * dpitch + dx0;
  // dst = (((b0 * H(sy0)) >> 16) + ((b1 * H(sy1)) >> 16) + 2) >> 2
  auto vrow = [&](const int4& ry, const uint32_t (&H0)[4], const uint32_t (&H1)[4]) {
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = (uint32_t)((((ry.z * (int)H0[k]) >> 16) + ((ry.w * (int)H1[k]) >> 16) + 2) >> 2);
    const uint32_t packed = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
    if (col_ok) *reinterpret_cast<uint32_t*>(drow) = packed;  // pitch % 16 == 0, dx0 % 4 == 0
    drow += dpitch;
  };
  // Two register sets X, Y hold the interpolations of staged rows ix, iy.  A destination row needs rows sy0 and sy1
  // (sy1 = sy0 + 1, or sy0 at the clamped bottom edge): whichever set already holds sy0 stays, the other one takes sy1 -
  // no register moves, every source row is interpolated once (sy0 advances by one or two per destination row).
  uint32_t X[4], Y[4];
  int ix = -1, iy = -1;
#pragma unroll 1
  for (int r = 0; r < nrows; r++) {
    const int4 ry = s_ry[r0 + r];
    if (ix != ry.x && iy != ry.x) { hrow(ry.x, X); ix = ry.x; }
    if (ix == ry.x) {
      if (iy != ry.y) { hrow(ry.y, Y); iy = ry.y; }
      vrow(ry, X, Y);
    } else {
      if (ix != ry.y) { hrow(ry.y, X); ix = ry.y; }
      vrow(ry, Y, X);
    }
  }
}

void launch_resize_level_tma(const ChunkBuffers& b, const DevicePlan& hp, const ResizeTma& tm, int level, int nframes,
                             cudaStream_t s) {
  const LevelPlan& L = hp.lv[level];
  dim3 grid((L.w + kResizeTileW - 1) / kResizeTileW, (L.h + kResizeTileH - 1) / kResizeTileH, nframes);
  launch_k(resize_tma_kernel, grid, dim3(128), (size_t)(tm.box_w[level] * tm.box_h[level]), s, b, tm, level);
}

void launch_resize_level(const ChunkBuffers& b, const DevicePlan& hp, int level, int nframes, cudaStream_t s) {
  const LevelPlan& L = hp.lv[level];
  dim3 block(32, 8), grid((L.w + 127) / 128, (L.h + 8 * kResizeRows - 1) / (8 * kResizeRows), nframes);
  launch_k(resize_level_kernel, grid, block, 0, s, b, level);
}

// ------------------------------------------------------------------------------------------------------------------
// E8  7x7 Gaussian blur, sigma 2, BORDER_REFLECT_101                                            (reference :760)
// Separable integer taps, dst = min((sum_j k_j sum_i k_i p + 32768) >> 16, 255)  (SURVEY §0.3 P3).  The reference blurs
// in place and keeps the unblurred border; here the blurred image goes to its own slab and the unblurred level stays
// intact (the descriptor kernel reads the unblurred reflect for the <= 2 px of patch overhang).
// One CTA = 64 x 58 output tile (64 staged rows) of one level of one frame; all levels in one launch.
//
// Both passes are integer dot products on the IDP pipe (measured 2 warp-inst/clk/SM, tools/int_ops_microbench.cu),
// which runs beside the ALU pipe that does the byte shuffles:
//   horizontal  h(x) = dp4a(bytes x-4..x-1, (0,k0,k1,k2)) + dp4a(bytes x..x+3, (k3,k4,k5,k6)); the 12 byte windows of 8
//               consecutive outputs are 3 aligned words and 9 funnel shifts.  h <= 255 * 257 fits 16 bits; a thread
//               does the same 8 columns of two consecutive rows and stores (h[2r], h[2r+1]) as one u16x2 word
//   vertical    v(y) = four dp2a over those row-pair words; odd rows use the SAME words with the taps shifted by one
//               ((0,k0),(k1,k2),(k3,k4),(k5,k6)), so no re-pairing is needed.  v needs 25 bits: a plain 32-bit sum.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kBlurTW = 64, kBlurTH = 58, kBlurRows = 64, kBlurInW = 20;  // staged row = image columns x0-8 .. x0+71

__global__ void __launch_bounds__(256) blur7_kernel(ChunkBuffers b, BlurTiles tiles) {
  ORBX_PDL_ENTRY();
  int level = 0;
  while (level + 1 < tiles.nlevels && (int)blockIdx.x >= tiles.base[level + 1]) level++;
  const int t = blockIdx.x - tiles.base[level];
  const int ty = magic_div(t, tiles.magic_x[level]), tx = t - ty * tiles.tiles_x[level];
  const int frame = blockIdx.y;
  const int x0 = tx * kBlurTW, y0 = ty * kBlurTH;
  const int lw = tiles.w[level], lh = tiles.h[level], dpitch = tiles.pitch[level];
  const int spitch = level == 0 ? b.lv0_pitch : dpitch;
  const uint8_t* src = level == 0 ? b.lv0 + (int64_t)frame * b.lv0_frame_stride
                                  : b.pyr + (int64_t)frame * tiles.frame_bytes + tiles.off[level];
  uint8_t* dst = b.blur + (int64_t)frame * tiles.frame_bytes + tiles.off[level];
  const uint32_t k0 = 18, k1 = 34, k2 = tiles.blur_mode == 0 ? 49 : 48, k3 = tiles.blur_mode == 0 ? 55 : 56;

  __shared__ __align__(16) uint32_t s_in[kBlurRows][kBlurInW];
  __shared__ __align__(16) uint32_t s_v[kBlurRows / 2][kBlurTW];  // (h[2r][x], h[2r+1][x])
  const int tid = threadIdx.x;
  // ---- stage 64 rows x 80 bytes, 8 bytes per load ---------------------------------------------------------------------
  // Rows reflect through the address.  Columns: whole 8-byte groups are loaded wherever they lie inside the row pitch
  // (bytes at columns >= lw are padding), and the <= 3 reflected columns on either side of the image are patched in
  // shared memory afterwards - no per-byte path in the common case.
  const bool vec_ok = ((spitch & 7) == 0) && ((reinterpret_cast<uintptr_t>(src) & 7) == 0);
  {
    uint2 v[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const int i = tid + 256 * k;
      if (i >= kBlurRows * (kBlurInW / 2)) break;
      const int r = i / (kBlurInW / 2), j = i - r * (kBlurInW / 2);
      const int yy = reflect101(min(y0 + r - 3, lh + 2), lh);
      const int xs = x0 - 8 + 8 * j;
      const uint8_t* row = src + (int64_t)yy * spitch;
      if (vec_ok) {
        v[k] = (xs >= 0 && xs + 8 <= spitch) ? *reinterpret_cast<const uint2*>(row + xs) : make_uint2(0u, 0u);
      } else {                                               // caller's level-0 buffer with an odd base or stride
        uint32_t lo = 0, hi = 0;
        for (int q = 0; q < 4; q++) {
          lo |= (uint32_t)row[reflect101(min(max(xs + q, -3), lw + 2), lw)] << (8 * q);
          hi |= (uint32_t)row[reflect101(min(max(xs + 4 + q, -3), lw + 2), lw)] << (8 * q);
        }
        v[k] = make_uint2(lo, hi);
      }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const int i = tid + 256 * k;
      if (i >= kBlurRows * (kBlurInW / 2)) break;
      reinterpret_cast<uint2*>(&s_in[0][0])[i] = v[k];
    }
  }
  __syncthreads();
  if (vec_ok && (x0 == 0 || x0 + kBlurTW + 3 > lw)) {         // CTA-uniform: this tile touches an image side
    if (tid < kBlurRows) {
      uint8_t* rowb = reinterpret_cast<uint8_t*>(&s_in[tid][0]);  // byte i = column x0 - 8 + i
      if (x0 == 0) { rowb[5] = rowb[11]; rowb[6] = rowb[10]; rowb[7] = rowb[9]; }
      for (int k = 0; k < 3; k++) {                           // column lw + k <- column lw - 2 - k
        const int i = lw + k - x0 + 8;
        if (i >= 8 && i < 4 * kBlurInW) rowb[i] = rowb[lw - 2 - k - x0 + 8];
      }
    }
    __syncthreads();
  }
  // ---- horizontal: one thread = 8 columns x 2 rows --------------------------------------------------------------------
  {
    const int tc = tid & 7, rp = tid >> 3;
    const uint32_t cA = (k0 << 8) | (k1 << 16) | (k2 << 24), cB = k3 | (k2 << 8) | (k1 << 16) | (k0 << 24);
    uint32_t h[2][8];
#pragma unroll
    for (int rr = 0; rr < 2; rr++) {
      const uint32_t* in = &s_in[2 * rp + rr][2 * tc];       // byte 8 tc of the staged row = image column x0 + 8 tc - 8
      const uint32_t w1 = in[1], w4 = in[4];
      const uint2 w23 = *reinterpret_cast<const uint2*>(in + 2);
      uint32_t win[12];                                       // win[s] = staged bytes 8 tc + 4 + s .. + 3
      win[0] = w1; win[4] = w23.x; win[8] = w23.y;
#pragma unroll
      for (int q = 1; q < 4; q++) {
        win[q] = __funnelshift_r(w1, w23.x, 8 * q);
        win[4 + q] = __funnelshift_r(w23.x, w23.y, 8 * q);
        win[8 + q] = __funnelshift_r(w23.y, w4, 8 * q);
      }
#pragma unroll
      for (int i = 0; i < 8; i++) h[rr][i] = __dp4a(win[i], cA, __dp4a(win[i + 4], cB, 0u));
    }
    uint4* out = reinterpret_cast<uint4*>(&s_v[rp][8 * tc]);
#define PAIR(i) __byte_perm(h[0][i], h[1][i], 0x5410)
    out[0] = make_uint4(PAIR(0), PAIR(1), PAIR(2), PAIR(3));
    out[1] = make_uint4(PAIR(4), PAIR(5), PAIR(6), PAIR(7));
#undef PAIR
  }
  __syncthreads();
  // ---- vertical: one thread = 4 columns x 4 rows ----------------------------------------------------------------------
  {
    const int cg = tid & 15, strip = tid >> 4;
    const int x = x0 + 4 * cg, ys = y0 + 4 * strip;
    if (x >= lw || ys >= lh || 4 * strip >= kBlurTH) return;
    // tap pairs for dp2a_lo (low half) / dp2a_hi (high half): even output rows start on a row pair, odd rows do not
    const uint32_t cE0 = (k0 | (k1 << 8)) | ((k2 | (k3 << 8)) << 16);   // (k0,k1) (k2,k3)
    const uint32_t cE1 = (k2 | (k1 << 8)) | (k0 << 16);                 // (k4,k5) (k6,0)
    const uint32_t cO0 = (k0 << 8) | ((k1 | (k2 << 8)) << 16);          // (0,k0) (k1,k2)
    const uint32_t cO1 = (k3 | (k2 << 8)) | ((k1 | (k0 << 8)) << 16);   // (k3,k4) (k5,k6)
    uint4 a[5];
#pragma unroll
    for (int m = 0; m < 5; m++) a[m] = *reinterpret_cast<const uint4*>(&s_v[min(2 * strip + m, kBlurRows / 2 - 1)][4 * cg]);
    const int nrows = min(min(4, kBlurTH - 4 * strip), lh - ys);
    uint8_t* orow = dst + (int64_t)ys * dpitch + x;
#pragma unroll
    for (int d = 0; d < 4; d++) {
      if (d >= nrows) break;
      const int m0 = d >> 1;
      uint32_t acc[4];
#define BLURV(c, i)                                                                                        \
      acc[i] = (d & 1) ? __dp2a_hi(a[m0 + 3].c, cO1, __dp2a_lo(a[m0 + 2].c, cO1, __dp2a_hi(a[m0 + 1].c, cO0,     \
                             __dp2a_lo(a[m0].c, cO0, 32768u))))                                                 \
                       : __dp2a_hi(a[m0 + 3].c, cE1, __dp2a_lo(a[m0 + 2].c, cE1, __dp2a_hi(a[m0 + 1].c, cE0,     \
                             __dp2a_lo(a[m0].c, cE0, 32768u))));
      BLURV(x, 0) BLURV(y, 1) BLURV(z, 2) BLURV(w, 3)
#undef BLURV
      // bits 16..24 of each sum, clamped to 255 on u16x2 lanes, then packed to four bytes
      const uint32_t p01 = __vminu2(__byte_perm(acc[0], acc[1], 0x7632), 0x00ff00ffu);
      const uint32_t p23 = __vminu2(__byte_perm(acc[2], acc[3], 0x7632), 0x00ff00ffu);
      *reinterpret_cast<uint32_t*>(orow) = __byte_perm(p01, p23, 0x6420);
      orow += dpitch;
    }
  }
}

void launch_blur(const ChunkBuffers& b, const DevicePlan& hp, const BlurTiles& t, int nframes, cudaStream_t s) {
  dim3 grid(t.base[hp.nlevels], nframes);
  launch_k(blur7_kernel, grid, dim3(256), 0, s, b, t);
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
