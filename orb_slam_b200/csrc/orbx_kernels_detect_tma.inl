// Included by orbx_kernels_detect.inl (inside namespace orbx).  E3/E4, the TMA-staged form of fast_cells_kernel.
//
// Same per-cell semantics and the same arithmetic as fast_cells_kernel (byte-SIMD reject on the halved image, packed
// VIMNMX3.U16x2 ladder, strict 3x3 NMS inside the scored region, retry at threshold 7, row-major emission); what changed
// is everything AROUND the arithmetic, which was half of the instruction stream (profiles/r1j_fast_phases.txt):
//   staging   the cell's sub-image arrives as ONE 3-D TMA tensor copy (cp.async.bulk.tensor -> UTMALDG) of raw bytes,
//             issued by one thread: box = the level's largest cell, coordinates (x0 - 8 - lead, y0 - 3, frame) - the copy
//             must start on a 16-byte boundary of the image row, so the kernel works in shifted coordinates
//             xs = x + lead - out-of-image bytes zero-filled by the engine.  No per-thread address math, no misalignment shifts, no global loads on
//             the issue stream.  The u16 copy of the image is gone: the ladder expands ring bytes with PRMT as it
//             loads them, and the reject tests the raw bytes directly: ONE copy of the cell in shared memory.
//   queues    flagged pixel pairs / pixels above the threshold are appended to CTA-wide lists with one shared-memory
//             atomicAdd per thread that has any (sparse), then consumed thread = entry by full warps; the per-warp
//             bit queues with their prefix scans over every bitmap word are gone.  A list that overflows (noise
//             images) switches that phase to a dense walk over the cell - slower, same result.
//   score map lives in the bytes of the image, which is dead once the (few) pixels above the threshold are known: they
//             are kept as a list with their scores, scattered into the cleared map, and suppressed against it.  One
//             14 KB array per cell instead of image + score map: nine CTAs per SM instead of six (+13 % on its own).
//   emission  one warp: each lane owns a contiguous run of bitmap words, one shuffle scan gives its output offset.
// ------------------------------------------------------------------------------------------------------------------
// 128 threads: with one 35 KB copy of the cell per CTA six CTAs share an SM, and the sparse phases (few entries per
// cell) leave fewer idle warps at the barriers than 256-thread CTAs did: 14.6 instead of 15.4 ms per 10k frames
#ifndef ORBX_FAST_TMA_THREADS
#define ORBX_FAST_TMA_THREADS 128
#endif
constexpr int kFastTmaThreads = ORBX_FAST_TMA_THREADS;
constexpr int kFastQ1 = 1024, kFastQ2 = 512;   // list capacities (entries): flagged pairs / pixels with S >= tmin

__device__ __forceinline__ uint32_t ft_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// (I(q), I(q + 1)) as zero-extended halfwords; q 2-byte aligned / not
__device__ __forceinline__ uint32_t ft_pair_even(const uint8_t* q) {
  return __byte_perm((uint32_t)*reinterpret_cast<const uint16_t*>(q), 0u, 0x4140);
}
__device__ __forceinline__ uint32_t ft_pair_odd(const uint8_t* q) { return __byte_perm((uint32_t)q[0], (uint32_t)q[1], 0x1410); }

// fast_pair_score on the RAW byte image: pair at shifted scored position (x even, y); row pitch rp bytes, position x at
// byte 16 + x of sub-image row y + 3.  Ring order as in fast_pair_score: k -> (dx, dy) =
// (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3).
__device__ __forceinline__ uint32_t fast_pair_score_u8(const uint8_t* raw, int rp, int x, int y) {
  const uint8_t* c = raw + (y + 3) * rp + 16 + x;             // 2-byte aligned (x even, rp multiple of 16)
  uint32_t r[16];
  r[0] = ft_pair_even(c + 3 * rp);      r[1] = ft_pair_odd(c + 3 * rp + 1);   r[15] = ft_pair_odd(c + 3 * rp - 1);
  r[2] = ft_pair_even(c + 2 * rp + 2);  r[14] = ft_pair_even(c + 2 * rp - 2);
  r[3] = ft_pair_odd(c + rp + 3);       r[13] = ft_pair_odd(c + rp - 3);
  r[4] = ft_pair_odd(c + 3);            r[12] = ft_pair_odd(c - 3);
  const uint32_t v2 = ft_pair_even(c);
  r[5] = ft_pair_odd(c - rp + 3);       r[11] = ft_pair_odd(c - rp - 3);
  r[6] = ft_pair_even(c - 2 * rp + 2);  r[10] = ft_pair_even(c - 2 * rp - 2);
  r[8] = ft_pair_even(c - 3 * rp);      r[7] = ft_pair_odd(c - 3 * rp + 1);   r[9] = ft_pair_odd(c - 3 * rp - 1);
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
  const uint32_t m = __vmaxs2(__vsub2(bb, v2), __vsub2(v2, dd));   // see fast_pair_score (no negation anywhere)
  return __vmaxs2(__vsub2(m, 0x00010001u), 0u);
}

template <bool kHarris>
__global__ void __launch_bounds__(kFastTmaThreads, 1024 / kFastTmaThreads)
fast_cells_tma_kernel(ChunkBuffers b, FastLevels fl, const __grid_constant__ FastTma tm) {
  ORBX_PDL_ENTRY();
  extern __shared__ __align__(128) uint8_t smem_tma[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ int s_cnt[4];                                       // survivors >= fastTh, >= 7; list lengths 1, 2
  const int cell = blockIdx.x, frame = blockIdx.y;
  const CellPlan c = b.cells[cell];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int32_t* n_total = b.n_total + (int64_t)frame * fl.total_cells + cell;
  if (c.rw <= 0) {
    if (tid == 0) *n_total = 0;
    return;
  }
  const int level = c.level;
  const int rp = tm.box_w[level];                                // raw / halved image pitch = the TMA box width
  const int tA = fl.fast_th;
  // shifted coordinates: xs = x + lead (the tensor copy starts on a 16-byte boundary of the image row); the scored region
  // is columns [lead, rw) of rows of rw = lead + c.rw columns; position xs sits at byte 16 + xs of its row
  const int lead = c.lead, rw = c.rwt, rh = c.rh;
  const int scp = c.scp, wpr = c.wpr, nbw = c.nbw;
  uint8_t* s_raw = smem_tma;                                     // [box_h][rp] raw bytes
  uint8_t* s_sc = s_raw;                                         // [rh + 2][scp] score map, position xs at byte 4 + xs:
                                                                 // takes the image's place once the scores are known
  uint32_t* s_bmA = reinterpret_cast<uint32_t*>(s_raw + c.off_bm);   // survivors with S >= fastTh
  uint32_t* s_bmB = s_bmA + nbw;                                     // survivors with S >= 7
  uint32_t* s_q1 = reinterpret_cast<uint32_t*>(s_raw + c.off_q);     // flagged pairs  (y << 16 | xs)
  uint32_t* s_q2 = s_q1 + kFastQ1;                                   // pixels with S >= tmin
  uint8_t* s_flag = reinterpret_cast<uint8_t*>(s_q2);   // [rh][g16] one byte of pair flags per 16-pixel group (rounded up to
                                                        // a word); dead before list 2 is written (same bytes)
  const uint32_t bar = ft_smem_u32(&s_bar);

  // ---- 0  one thread asks the TMA engine for the sub-image; everybody else clears the score map and bitmaps ------------
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)c.off_bm) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(ft_smem_u32(s_raw)), "l"(reinterpret_cast<uint64_t>(&tm.map[level])), "r"((int)c.x0 - 16 - lead),
          "r"((int)c.y0 - 3), "r"(frame), "r"(bar)
        : "memory");
  }
  const int nz = c.nz16;                                         // uint4 stores that clear the score map
  const int nzb0 = (8 * nbw + 15) >> 4;                          // ... and bitmaps A + B
  for (int i = tid; i < nzb0; i += kFastTmaThreads) reinterpret_cast<uint4*>(s_bmA)[i] = make_uint4(0, 0, 0, 0);
  if (tid < 4) s_cnt[tid] = 0;
  __syncthreads();                                               // barrier initialised before anybody polls it
  {
    uint32_t ok = 0;
#pragma unroll 1
    for (int spin = 0; spin < (1 << 22) && !ok; ++spin)
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(ok) : "r"(bar) : "memory");
    if (!ok) {                                                   // a lost completion ends as a reported error
      if (tid == 0) { atomicExch(b.err_flag, 3); *n_total = 0; }
      return;
    }
  }
  const int tB = kRetryTh;
  // searched at fastTh first, again at 7 only when <= 3 corners survive (:599-614); fastTh <= 7: one pass serves both
  int tmin = max(tA > tB ? tA : min(tA, tB), 1);
  bool useA, dense = false;
  for (;;) {
    // ---- 1a  reject, 16 pixels per thread on the raw bytes, branch-free: one flag word per group ------------------------
    // An arc of 9 ring pixels holds one pixel of every opposite pair, so a corner at threshold t needs
    // (|up-v| > t or |down-v| > t) and (|left-v| > t or |right-v| > t): VABSDIFF4 + one add per direction and 4 pixels.
    // The four result words (bit 7 of byte k = pixel 4j + k of word j) are folded into one: bit 8k + j = pixel 4j + k.
    {
      const int G16 = c.g16, ng = c.ngrp;
      if (tid < 3 && ng + tid < ((ng + 3) & ~3)) s_flag[ng + tid] = 0;      // the tail of the last flag word
      const uint32_t bias4 = (0x7fu - (uint32_t)tmin) * 0x01010101u;
      const int sy = magic_div(kFastTmaThreads, c.m_g16), sg = kFastTmaThreads - sy * G16;
      int y = magic_div(tid, c.m_g16), g = tid - y * G16;
      for (int i = tid; i < ng; i += kFastTmaThreads) {
        const uint8_t* ctr = s_raw + (y + 3) * rp + 16 * g + 16;  // centre bytes of positions 16g .. 16g+15
        const uint4 C = *reinterpret_cast<const uint4*>(ctr);
        const uint4 U = *reinterpret_cast<const uint4*>(ctr - 3 * rp), D = *reinterpret_cast<const uint4*>(ctr + 3 * rp);
        const uint32_t W0 = *reinterpret_cast<const uint32_t*>(ctr - 4), W5 = *reinterpret_cast<const uint32_t*>(ctr + 16);
        // d = |a - v| per byte; d > t  <=>  bit 7 of (d + 0x7f - t), as long as that sum stays below 256, i.e. d < 128;
        // a byte with d >= 128 (> t: fastTh <= 120 here) is caught by its own bit 7, and the carry it may send into the
        // next byte can only ADD a flag there (the reject is conservative by design)
#define ORBX_GT(a, cc, out)                                            \
  { const uint32_t d_ = __vabsdiffu4(a, cc); out = (d_ + bias4) | d_; }
#define ORBX_REJ(u, d, l, r, cc, out)                                  \
  { uint32_t a_, b_, c_, e_; ORBX_GT(u, cc, a_) ORBX_GT(d, cc, b_) ORBX_GT(l, cc, c_) ORBX_GT(r, cc, e_)   \
    out = (a_ | b_) & (c_ | e_) & 0x80808080u; }
        uint32_t m0, m1, m2, m3;
        ORBX_REJ(U.x, D.x, __byte_perm(W0, C.x, 0x4321), __byte_perm(C.x, C.y, 0x6543), C.x, m0)
        ORBX_REJ(U.y, D.y, __byte_perm(C.x, C.y, 0x4321), __byte_perm(C.y, C.z, 0x6543), C.y, m1)
        ORBX_REJ(U.z, D.z, __byte_perm(C.y, C.z, 0x4321), __byte_perm(C.z, C.w, 0x6543), C.z, m2)
        ORBX_REJ(U.w, D.w, __byte_perm(C.z, C.w, 0x4321), __byte_perm(C.w, W5, 0x6543), C.w, m3)
#undef ORBX_REJ
#undef ORBX_GT
        const uint32_t t = (m0 >> 7) | (m1 >> 6) | (m2 >> 5) | (m3 >> 4);
        const uint32_t pr = (t | (t >> 8)) & 0x000f000fu;       // bit j: pair at 4j (pixels 4j, 4j+1); bit 16+j: pair at 4j+2
        s_flag[i] = (uint8_t)(pr | (pr >> 12));                 // bit j / bit 4+j
        y += sy; g += sg;
        if (g >= G16) { g -= G16; y++; }
      }
    }
    __syncthreads();
    // ---- 1b  flag bytes -> list 1 of pixel PAIRS (xs even) with either pixel flagged; a thread takes four groups ---------
    // (pairs outside [lead, rw) - padding / the neighbour cell - are dropped by the scorer: cheaper than masking here)
    {
      const int G16 = c.g16, nw = (c.ngrp + 3) >> 2;
      const uint32_t* fw = reinterpret_cast<const uint32_t*>(s_flag);
      for (int i = tid; i < nw; i += kFastTmaThreads) {
        uint32_t pr = fw[i];
        if (pr) {
          int off = atomicAdd(&s_cnt[2], __popc(pr));
          while (pr) {
            const int bit = 31 - __clz(pr);
            pr ^= 1u << bit;
            const int gi = 4 * i + (bit >> 3), y = magic_div(gi, c.m_g16), g = gi - y * G16, bb = bit & 7;
            if (off < kFastQ1) s_q1[off] = ((uint32_t)y << 16) | (uint32_t)(16 * g + 4 * (bb & 3) + 2 * (bb >> 2));
            off++;
          }
        }
      }
    }
    __syncthreads();

    // ---- 2  full score of the flagged pairs; pixels with S >= tmin go to list 2 as (y << 16 | xs << 8 | S) -------------------
    // Only those pixels matter to the NMS, as centres AND as neighbours: a neighbour below tmin cannot suppress anybody.
    {
      auto heavy = [&](int x, int y) {
        if (x + 1 < lead || x >= rw) return;                    // a pair of the padding / the neighbour cell
        const uint32_t s2 = fast_pair_score_u8(s_raw, rp, x, y);
        const int s0 = x >= lead ? (int)(s2 & 0xffff) : 0, s1 = x + 1 < rw ? (int)(s2 >> 16) : 0;
        const bool m0 = s0 >= tmin, m1 = s1 >= tmin;
        if (m0 | m1) {
          int off = atomicAdd(&s_cnt[3], (int)m0 + (int)m1);
          const uint32_t e = ((uint32_t)y << 16) | ((uint32_t)x << 8);
          if (m0) { if (off < kFastQ2) s_q2[off] = e | (uint32_t)s0; off++; }
          if (m1) { if (off < kFastQ2) s_q2[off] = (e + 0x100u) | (uint32_t)s1; }
        }
      };
      const int n1 = s_cnt[2];
      if (n1 <= kFastQ1) {
        for (int i = tid; i < n1; i += kFastTmaThreads) { const uint32_t e = s_q1[i]; heavy((int)(e & 0xffff), (int)(e >> 16)); }
      } else {                                                   // list 1 overflowed: every pair of the cell
        const int hp = (rw + 1) >> 1, np = hp * rh;
        for (int i = tid; i < np; i += kFastTmaThreads) { const int y = i / hp; heavy(2 * (i - y * hp), y); }
      }
    }
    __syncthreads();

    // ---- 3  strict 3x3 NMS of the pixels with S >= tmin ---------------------------------------------------------------------
    int nA = 0, nB = 0;
    const int n2 = s_cnt[3];
    dense = n2 > kFastQ2;
    if (!dense) {
      // The raw image is dead now: the score map takes its place (cleared, then list 2 scattered into it).  Not keeping a
      // second 11 KB array alive beside the image is what lets nine CTAs share an SM instead of six.
      for (int i = tid; i < nz; i += kFastTmaThreads) reinterpret_cast<uint4*>(s_sc)[i] = make_uint4(0, 0, 0, 0);
      __syncthreads();
      for (int i = tid; i < n2; i += kFastTmaThreads) {
        const uint32_t e = s_q2[i];
        s_sc[((e >> 16) + 1) * scp + 4 + ((e >> 8) & 0xff)] = (uint8_t)e;
      }
      __syncthreads();
      for (int i = tid; i < n2; i += kFastTmaThreads) {
        const uint32_t e = s_q2[i];
        const int x = (e >> 8) & 0xff, y = e >> 16, sc = e & 0xff;
        const uint8_t* p = s_sc + (y + 1) * scp + 4 + x;
        const int m = max(max(max(p[-1], p[1]), max(p[-scp - 1], p[-scp])),
                          max(max(p[-scp + 1], p[scp - 1]), max(p[scp], p[scp + 1])));
        if (sc > m) {
          const int w = y * wpr + (x >> 5);
          if (sc >= tA) { atomicOr(&s_bmA[w], 1u << (x & 31)); nA++; }
          if (sc >= tB) { atomicOr(&s_bmB[w], 1u << (x & 31)); nB++; }
        }
      }
    } else {
      // List 2 overflowed (noise-like cells): dense walk in bands of rows whose score map fits the bytes of the two lists;
      // the raw image stays intact (every score is recomputed from it, the band's halo rows included).  Slow, rare, exact.
      uint8_t* s_band = reinterpret_cast<uint8_t*>(s_q1);
      const int brows = (kFastQ1 + kFastQ2) * 4 / scp - 2;       // rows per band (the map has one halo row on either side)
      const int hp = (rw + 1) >> 1;
      for (int y0 = 0; y0 < rh; y0 += brows) {
        const int y1 = min(y0 + brows, rh);                     // band rows [y0, y1), map rows y0 - 1 .. y1
        const int nzb = ((y1 - y0 + 2) * scp + 15) >> 4;
        for (int i = tid; i < nzb; i += kFastTmaThreads) reinterpret_cast<uint4*>(s_band)[i] = make_uint4(0, 0, 0, 0);
        __syncthreads();
        const int ya = max(y0 - 1, 0), yb = min(y1 + 1, rh);
        for (int i = tid; i < hp * (yb - ya); i += kFastTmaThreads) {
          const int yy = i / hp, x = 2 * (i - yy * hp), y = ya + yy;
          const uint32_t s2 = fast_pair_score_u8(s_raw, rp, x, y);
          uint32_t two = __byte_perm(s2, 0, 0x4420);
          if (x + 1 >= rw) two &= 0xff;
          if (x < lead) two &= x + 1 < lead ? 0u : 0xff00u;
          *reinterpret_cast<uint16_t*>(s_band + (y - y0 + 1) * scp + 4 + x) = (uint16_t)two;
        }
        __syncthreads();
        for (int i = tid; i < rw * (y1 - y0); i += kFastTmaThreads) {
          const int yy = i / rw, x = i - yy * rw, y = y0 + yy;
          const uint8_t* p = s_band + (yy + 1) * scp + 4 + x;
          const int sc = p[0];
          if (sc < tmin) continue;
          const int m = max(max(max(p[-1], p[1]), max(p[-scp - 1], p[-scp])),
                            max(max(p[-scp + 1], p[scp - 1]), max(p[scp], p[scp + 1])));
          if (sc > m) {
            const int w = y * wpr + (x >> 5);
            if (sc >= tA) { atomicOr(&s_bmA[w], 1u << (x & 31)); nA++; }
            if (sc >= tB) { atomicOr(&s_bmB[w], 1u << (x & 31)); nB++; }
          }
        }
        __syncthreads();
      }
    }
    nA = __reduce_add_sync(0xffffffffu, nA);
    nB = __reduce_add_sync(0xffffffffu, nB);
    if (lane == 0 && (nA | nB)) { atomicAdd(&s_cnt[0], nA); atomicAdd(&s_cnt[1], nB); }
    __syncthreads();
    useA = s_cnt[0] > 3;                          // :609-614
    if (useA || tmin <= tB) break;
    // retry at 7: forget this pass's bitmaps and lists; the score map may have overwritten the image: fetch it again
    __syncthreads();                              // everyone has read s_cnt and is done with the score map
    for (int i = tid; i < nzb0; i += kFastTmaThreads) reinterpret_cast<uint4*>(s_bmA)[i] = make_uint4(0, 0, 0, 0);
    if (tid < 4) s_cnt[tid] = 0;
    tmin = tB;
    if (!dense) {
      if (tid == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // generic writes (score map) before the async copy
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)c.off_bm) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
            ::"r"(ft_smem_u32(s_raw)), "l"(reinterpret_cast<uint64_t>(&tm.map[level])), "r"((int)c.x0 - 16 - lead),
              "r"((int)c.y0 - 3), "r"(frame), "r"(bar)
            : "memory");
      }
      uint32_t ok = 0;
#pragma unroll 1
      for (int spin = 0; spin < (1 << 22) && !ok; ++spin)
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 1;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(bar) : "memory");
      if (!ok) {
        if (tid == 0) { atomicExch(b.err_flag, 3); *n_total = 0; }
        return;
      }
    }
    __syncthreads();
  }
  const int total = useA ? s_cnt[0] : s_cnt[1];
  if (tid == 0) *n_total = total;
  if (total > c.cand_cap) {                       // cannot happen (NMS survivors are non-adjacent); never truncate
    if (tid == 0) atomicExch(b.err_flag, 1);
    return;
  }
  if (total == 0 || warp != 0) return;

  // ---- 4  ordered emission by ONE warp: lane = a contiguous run of bitmap words (row-major), one scan for the offsets ----
  const uint32_t* bm = useA ? s_bmA : s_bmB;
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
  [[maybe_unused]] const int pitch = level == 0 ? b.lv0_pitch : fl.pitch[level];
  [[maybe_unused]] const uint8_t* img = level == 0 ? b.lv0 + (int64_t)frame * b.lv0_frame_stride
                                                   : b.pyr + (int64_t)frame * fl.frame_bytes + fl.off[level];
  for (int i = w0; i < w1; i++) {
    uint32_t bits = bm[i];
    if (!bits) continue;
    const int y = magic_div(i, c.m_wprt), xb = 32 * (i - y * wpr) - lead;
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
        if (!dense) r.resp = (float)s_sc[(y + 1) * scp + 4 + lead + x];
        else {                                  // the dense walk keeps no score map: recompute from the intact image
          const int xs = lead + x;
          const uint32_t s2 = fast_pair_score_u8(s_raw, rp, xs & ~1, y);
          r.resp = (float)((xs & 1) ? (s2 >> 16) : (s2 & 0xffff));
        }
      }
      r.xy = ((uint32_t)ly << 16) | (uint32_t)lx;
      *out++ = r;
    }
  }
}

void launch_fast_cells_tma(const ChunkBuffers& b, const DevicePlan& hp, const FastTma& tm, int nframes, cudaStream_t s) {
  auto* kern = hp.score_type == ORBX_HARRIS_SCORE ? fast_cells_tma_kernel<true> : fast_cells_tma_kernel<false>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
  FastLevels fl;
  fl.fast_th = hp.fast_th; fl.score_type = hp.score_type; fl.total_cells = hp.total_cells;
  fl.cand_per_frame = hp.cand_per_frame; fl.frame_bytes = hp.pyr_frame_bytes;
  for (int l = 0; l < kMaxLevels; l++) { fl.pitch[l] = hp.lv[l].pitch; fl.off[l] = hp.lv[l].off; }
  dim3 grid(hp.total_cells, nframes);
  static const int pad = [] { const char* e = getenv("ORBX_FAST_SMEM_PAD"); return e ? atoi(e) : 0; }();   // occupancy experiments
  launch_k(kern, grid, dim3(kFastTmaThreads), (size_t)(tm.smem_bytes + pad), s, b, fl, tm);
}
