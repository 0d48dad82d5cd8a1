// Included by orbm_knn.cu after orbm_knn_kernels.inl (mbarrier helpers, top2_insert, kNone).
// ------------------------------------------------------------------------------------------------------------------
// M8 on the tensor cores: all-pairs 256-bit Hamming 2-NN as a +-1 GEMM (tcgen05 / TMEM, CTA pairs).
//
// Every descriptor bit becomes one FP8 (E4M3) value, +1.0 (0x38) for a clear bit and -1.0 (0xB8) for a set bit, so
// that for two descriptors  dot = (#equal bits) - (#different bits) = 256 - 2 * hamming.  Products are +-1 and the sums
// are integers of magnitude <= 256, exact in the FP32 accumulator: the result is bit-for-bit the popcount distance
// (src/ORBmatcher.cc:1794-1810).  tcgen05.mma.cta_group::2 (kind::f8f6f4, M = 256 queries over the two CTAs of a
// cluster, N = 256 database descriptors, K = 32 bytes per instruction, 8 instructions per tile) does 8192 pair-bits
// per clock and SM, against 16 lanes of POPC - the bound of knn2_kernel.
//
// A cluster of two CTAs (one TPC) holds two tiles of 128 queries, each expanded once into its CTA's shared memory
// (operand A, 32 KB), and streams database tiles of 256 descriptors; each CTA expands HALF of every tile (operand B,
// 32 KB per stage, 4 stages), the pair's MMA reads both halves.  Why pairs: with one CTA per tile the tensor core
// re-reads A and B from shared memory at 96 B/clk and leaves the expanders' stores the rest of the 128 B/clk
// (measured: 6.0 instead of 4.9 ms without any epilogue); in a pair each SM reads 64 B/clk and stores half as much.
// Warp roles per CTA (17 warps):
//   warps 0-7    epilogue: warp w owns TMEM lanes 32 (w & 3) .. + 31 (= queries) and column half w >> 2 of the
//                128 x 256 FP32 accumulator; 64 columns per round (two tcgen05.ld.32x32b.x32), folded with 3-input
//                integer max on the float bit patterns; only when some lane sees a dot ABOVE its query's current
//                second-best (a later index never wins a tie - the strict `<` of the sequential scan, :212-221) the
//                flagged 16-column quarters are re-read from tensor memory and inserted as packed keys
//   warp 8       allocates the 512 TMEM columns (two accumulator buffers).  CTA 0: one lane issues the pair's MMAs
//                and commits them (multicast) to the "stage free" and "accumulator full" barriers of both CTAs.
//                CTA 1: forwards "my half of the stage is written" to CTA 0
//   warps 9-16   expanders: two threads per database descriptor (16 packed bytes each, fetched two tiles ahead)
//                write 2 x 128 operand bytes in the canonical K-major SWIZZLE_128B layout (two 128-byte K blocks of
//                [rows][128 B], 16-byte chunk index XOR (row & 7)): one shift + one LOP3 per output word
// While the tensor core works on tile i+1 the epilogue drains accumulator i and the expanders fill a free stage.
// The bit -> K-position mapping is the expanders' own (word i, bit 8b+t -> chunk 2i + t/4, word t%4, byte b); it is
// the same for A and B, and a dot product does not depend on the order of its terms.
// Output: [piece][nq] (best, second) key pairs in the format of knn2_kernel, merged by merge_top2_kernel.
// ------------------------------------------------------------------------------------------------------------------
namespace {

constexpr int kTcQ = 128, kTcN = 256;                       // queries per CTA, database descriptors per tile
constexpr int kTcEpiWarps = 8, kTcExpWarps = 8, kTcThreads = 32 * (kTcEpiWarps + 1 + kTcExpWarps);
constexpr int kTcABytes = kTcQ * 256;

// one lane of a converged warp, as a predicate ptxas keeps uniform: `if (lane == 0)` made it wrap every UTCQMMA in an
// ELECT / BRA.U.ANY loop over the active lanes, ~100 clocks of dependent predicate logic per MMA on the issuing lane
__device__ __forceinline__ bool tc_elect_one() {
  uint32_t p;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
  return p != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// K-major SWIZZLE_128B shared-memory matrix descriptor: start address >> 4, LBO (unused for swizzled K-major) = 1,
// SBO = 1024 B between 8-row groups, descriptor version 1 (sm_100), layout type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// NW packed 32-bit words (descriptor words wbase .. wbase + NW - 1) -> 32 * NW operand bytes of row `row` of a
// [2][nrows][128 B] swizzled tile at shared address `tile`.
template <int NW>
__device__ __forceinline__ void tc_expand_words(uint32_t tile, int nrows, int row, int wbase, const uint32_t (&w)[NW]) {
  const uint32_t rowbase = tile + (uint32_t)row * 128u, rx = ((uint32_t)row & 7u) << 4;
  uint32_t c80 = 0x80808080u, c38 = 0x38383838u;
  asm volatile("" : "+r"(c80), "+r"(c38));            // keep them in registers
#pragma unroll
  for (int i = 0; i < NW; i++) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
      uint32_t o[4];
#pragma unroll
      for (int t = 0; t < 4; t++)   // (x & 0x80808080) | 0x38383838 as ONE lop3 (both constants in registers; ptxas emits two otherwise)
        asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(o[t]) : "r"(w[i] << (7 - (4 * h + t))), "r"(c80), "r"(c38));
      const int chunk = 2 * (wbase + i) + h;               // wbase is a multiple of 4: the K block is uniform per call
      const uint32_t a = rowbase + (uint32_t)(chunk >> 3) * (uint32_t)nrows * 128u + ((((uint32_t)chunk & 7u) << 4) ^ rx);
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3])
                   : "memory");
    }
  }
}
template <int NW>
__device__ __forceinline__ void tc_zero_words(uint32_t tile, int nrows, int row, int wbase) {
#pragma unroll 1
  for (int c = 2 * wbase; c < 2 * (wbase + NW); c++)
    asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(tile + (uint32_t)(c >> 3) * (uint32_t)nrows * 128u + (uint32_t)row * 128u + (uint32_t)(c & 7) * 16u), "r"(0u) : "memory");
}

__device__ __forceinline__ void tc_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr) : "memory");
}
// Largest of 16 accumulator columns of one query, as a SIGNED INTEGER max of the float bit patterns (VIMNMX3; the
// float 3-input max issues several times slower).  Non-negative floats order like their bit patterns and every
// negative float is a negative integer, so the result is the true maximum whenever that maximum is >= 0; the caller
// takes the exact path unconditionally while its threshold is negative.
__device__ __forceinline__ int tc_max16(const float* v) {
  const int* x = reinterpret_cast<const int*>(v);
  int m = max(max(x[0], x[1]), x[2]);
#pragma unroll
  for (int c = 3; c + 1 < 16; c += 2) m = max(max(m, x[c]), x[c + 1]);
  return max(m, x[15]);
}
// The rare path (about 1 % of the groups): a column of this warp's 64 beats some lane's second-best.  The quarters
// (16 columns) flagged in `qmask` are read again from tensor memory eight columns at a time and EVERY column is
// inserted as a packed key - exact whatever the order or the threshold.  A short loop on purpose: walking the 64
// registers of the fast path needs the walk unrolled (registers cannot be indexed), 40 KB of code that was slower.
__device__ __noinline__ uint2 tc_epi_walk(uint32_t taddr, int valid, uint32_t kbase, uint32_t qmask, uint32_t best,
                                          uint32_t second) {
#pragma unroll 1
  for (int c0 = 0; c0 < 64 && c0 < valid; c0 += 8) {
    if (!((qmask >> (c0 >> 4)) & 1u)) continue;
    float v[8];
    tc_ld8(taddr + (uint32_t)c0, v);
    tc_wait_ld();
#pragma unroll
    for (int j = 0; j < 8; j++) {
      // key = dist << 23 | index, dist = (256 - dot) / 2 with 256 - dot even
      const uint32_t key = (uint32_t)(256 - (int)v[j]) * (1u << 22) + (kbase + (uint32_t)(c0 + j));
      top2_insert(best, second, c0 + j < valid ? key : kNone);
    }
  }
  return make_uint2(best, second);
}

// cluster helpers (kCtas == 2: the CTA pair of one TPC works as one 256-row MMA, tcgen05 cta_group::2)
__device__ __forceinline__ uint32_t tc_cta_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void tc_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// remote arrives on barriers of CTA 0 (the MMA-issuing CTA)
__device__ __forceinline__ uint32_t tc_leader_addr(uint32_t bar) {   // the barrier at this offset in CTA 0, as a cluster address
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(ra) : "r"(bar));
  return ra;
}
__device__ __forceinline__ void tc_arrive_remote(uint32_t ra) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(ra) : "memory");
}
// without release semantics: for the accumulator-free signal, whose only prior accesses are tcgen05.ld reads already
// ordered by tcgen05.wait::ld + tcgen05.fence::before_thread_sync (a release at cluster scope costs hundreds of clocks)
__device__ __forceinline__ void tc_arrive_remote_relaxed(uint32_t ra) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(ra) : "memory");
}
// Bounded waits: a lost signal must end as a reported error, never as a hung GPU.  The bound is TIME (about ten
// seconds of SM clock - generous, so that a time-sliced or debugged GPU does not trip it), not a try count: one try_wait may itself sleep for a system-defined while, and 2^20 of those
// was a twenty-minute "bounded" wait.  A suspend-time hint on try_wait was measured and dropped: the waiting warps wake
// later (8.21 -> 8.45 ms on the 8,000 x 5 M job).
constexpr long long kTcWaitClocks = 20000000000ll;
__device__ __forceinline__ bool tc_wait(uint32_t bar, uint32_t parity) {
  long long t0 = 0;
#pragma unroll 1
  for (;;) {
#pragma unroll 1
    for (int spin = 0; spin < 64; ++spin) {                  // nothing but the try in the polling loop
      uint32_t ok;
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
      if (ok) return true;
    }
    const long long now = clock64();
    if (t0 == 0) t0 = now;
    else if (now - t0 > kTcWaitClocks) return false;
  }
}

constexpr int kTcStages = 4;                                 // shared-memory stages of operand B
constexpr int kTcBRows = kTcN / 2;                           // database descriptors each CTA of the pair expands per tile
constexpr int kTcBBytes = kTcBRows * 256;
constexpr int kTcSmemBytes = kTcABytes + kTcStages * kTcBBytes + 1024;   // + slack for the 1024-byte alignment
// instruction descriptor: D = F32 (bits 4-5 = 1), A = B = E4M3 (0), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
// (M = 256: the two CTAs' 128 rows each)
constexpr uint32_t kTcIdesc = (1u << 4) | ((uint32_t)(kTcN >> 3) << 17) | ((uint32_t)((2 * kTcQ) >> 4) << 24);

__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(kTcIdesc), "r"(accumulate) : "memory");
}
// arrives (once the MMAs issued so far have completed) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}

// Launched with cluster dimension (2, 1, 1); grid.x = 2 * clusters.  The work - (query-tile pairs) x (database tiles),
// linearised pair-major - is cut into `per` consecutive tile steps per cluster, so every SM pair gets the same amount
// whatever the shapes; a cluster's range may cross from one query-tile pair into the next (a "segment" each).  Segment
// results go to part[piece][nq] with piece = cluster - (first cluster working on that pair); the host presets the
// buffer to 'none' keys because the last pieces of a pair need not exist.
__global__ void __launch_bounds__(kTcThreads, 1) knn2_tc_kernel(const uint4* __restrict__ q, int nq,
                                                                const uint4* __restrict__ db, int64_t ndb, int ntiles,
                                                                int64_t per, int64_t total, uint32_t idx_base,
                                                                uint32_t* __restrict__ part, int* __restrict__ err) {
  extern __shared__ __align__(1024) uint8_t smem_tc[];
  __shared__ __align__(8) uint64_t s_full[kTcStages], s_peer_full[kTcStages], s_free[kTcStages], s_acc_full[2], s_acc_free[2];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint32_t s_xch[2 * kTcQ];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = tc_cta_rank();
  const int64_t cluster = blockIdx.x >> 1;
  const int64_t lin0 = cluster * per, lin1 = min(total, lin0 + per);
  const uint32_t sbase = (smem_u32(smem_tc) + 1023u) & ~1023u;
  const uint32_t sA = sbase, sB = sbase + kTcABytes;

  if (tid == 0) {
    for (int s = 0; s < kTcStages; s++) {
      mbar_init(smem_u32(&s_full[s]), kTcExpWarps);
      mbar_init(smem_u32(&s_peer_full[s]), 1);
      mbar_init(smem_u32(&s_free[s]), 1);
    }
    for (int s = 0; s < 2; s++) {
      mbar_init(smem_u32(&s_acc_full[s]), 1);
      mbar_init(smem_u32(&s_acc_free[s]), kTcEpiWarps * 2);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kTcEpiWarps) {   // the MMA warp owns the tensor memory: 512 columns = two 128 x 256 FP32 accumulators
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s_tmem)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  tc_cluster_sync();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  // barrier addresses once: taking them inside the tile loops cost a dozen instructions per use (S2UR of the cluster
  // window, LEA, MAPA) on warps whose per-tile instruction count is the latency of the accumulator hand-back
  const uint32_t b_full = smem_u32(&s_full[0]), b_peer = smem_u32(&s_peer_full[0]), b_free = smem_u32(&s_free[0]);
  const uint32_t b_acc_full = smem_u32(&s_acc_full[0]), b_acc_free = smem_u32(&s_acc_free[0]);
  const uint32_t r_acc_free = tc_leader_addr(b_acc_free), r_peer = tc_leader_addr(b_peer);
  bool ok = true;
  int it0 = 0;                                               // tile steps done so far: the barriers' phases run on

  for (int64_t lin = lin0; lin < lin1;) {
    const int pair = (int)(lin / ntiles), t0 = (int)(lin - (int64_t)pair * ntiles);
    const int nseg = (int)min((int64_t)(ntiles - t0), lin1 - lin);
    const int qtile = 2 * pair + (int)rank;
    const int piece = (int)(cluster - ((int64_t)pair * ntiles) / per);
    const int seg_rows = (int)(ndb - (int64_t)t0 * kTcN);   // database rows from the segment's first tile on (< 2^31)

    if (warp > kTcEpiWarps) {  // operand A: the CTA's 128 queries of this segment (rows past nq: +0.0 bytes)
      const int e = tid - 32 * (kTcEpiWarps + 1);
      const int row = e & (kTcQ - 1), wbase = 4 * (e >> 7);  // 256 threads: two per query, four words each
      const int qi = qtile * kTcQ + row;
      if (qi < nq) {
        const uint4 v = q[2 * qi + (e >> 7)];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        tc_expand_words<4>(sA, kTcQ, row, wbase, w);
      } else tc_zero_words<4>(sA, kTcQ, row, wbase);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    tc_cluster_sync();

    if (warp < kTcEpiWarps) {
      // ================================ epilogue: top-2 per query (= TMEM lane) ======================================
      // warp w: TMEM lanes 32 (w & 3) .. + 31 (the hardware ties a warp to the lane quarter warp-id % 4), column half
      // w >> 2.  thr_i: bit pattern of the dot product of the query's current second-best while that is >= 0, else
      // INT_MIN (= every group takes the exact path); a later index never wins a tie, so only dots ABOVE it matter.
      uint32_t best = kNone, second = kNone;
      int thr_i = INT_MIN;
      const int quarter = warp & 3, half = warp >> 2;
      constexpr int kHalfCols = kTcN / 2;
      const uint32_t lane_addr = tmem + ((uint32_t)(32 * quarter) << 16) + (uint32_t)(half * kHalfCols);
      for (int i = 0; i < nseg && ok; ++i) {
        const int it = it0 + i, a = it & 1, t = t0 + i;
        ok = tc_wait(b_acc_full + 8u * (uint32_t)a, (uint32_t)(it >> 1) & 1u);
        if (!ok) break;
        tc_fence_after();
        const int cnt = min(kTcN, seg_rows - i * kTcN) - half * kHalfCols;   // valid columns of this half
        const uint32_t kb = idx_base + (uint32_t)t * kTcN + (uint32_t)(half * kHalfCols);
#pragma unroll 1
        for (int g = 0; g < kHalfCols / 64; g++) {
          float v0[32], v1[32];
          const uint32_t ta = lane_addr + (uint32_t)(a * kTcN + g * 64);
          tc_ld32(ta, v0);
          tc_ld32(ta + 32u, v1);
          tc_wait_ld();
          const int valid = cnt - g * 64;
          if (valid <= 0) continue;                          // warp-uniform
          const int m0 = tc_max16(v0), m1 = tc_max16(v0 + 16), m2 = tc_max16(v1), m3 = tc_max16(v1 + 16);
          if (__any_sync(0xffffffffu, max(max(m0, m1), max(m2, m3)) > thr_i) || valid < 64) {
            // quarters to walk: a warp-uniform decision (the votes must not sit under a per-lane condition: thr_i
            // differs between lanes; a lane still at INT_MIN flags every quarter by itself)
            uint32_t qmask = 0xFu;
            if (valid >= 64)
              qmask = (__any_sync(0xffffffffu, m0 > thr_i) ? 1u : 0u) | (__any_sync(0xffffffffu, m1 > thr_i) ? 2u : 0u) |
                      (__any_sync(0xffffffffu, m2 > thr_i) ? 4u : 0u) | (__any_sync(0xffffffffu, m3 > thr_i) ? 8u : 0u);
            const uint2 r = tc_epi_walk(ta, valid, kb + (uint32_t)(g * 64), qmask, best, second);
            best = r.x; second = r.y;
            if (second != kNone) {
              const int d2 = 256 - 2 * (int)(second >> 23);
              thr_i = d2 >= 0 ? __float_as_int((float)d2) : INT_MIN;
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) tc_arrive_remote_relaxed(r_acc_free + 8u * (uint32_t)a);
      }
      if (!ok) { atomicExch(err, 3); best = kNone; second = kNone; }
      // the two column halves of a query meet in shared memory
      uint2* xch = reinterpret_cast<uint2*>(s_xch);
      if (half == 1) xch[32 * quarter + lane] = make_uint2(best, second);
      asm volatile("bar.sync 1, %0;" ::"n"(32 * kTcEpiWarps) : "memory");
      if (half == 0) {
        const uint2 o = xch[32 * quarter + lane];
        top2_insert(best, second, o.x);
        top2_insert(best, second, o.y);
        const int qi = qtile * kTcQ + 32 * quarter + lane;
        if (qi < nq) reinterpret_cast<uint2*>(part)[(int64_t)piece * nq + qi] = make_uint2(best, second);
      }
    } else if (warp == kTcEpiWarps) {
      if (rank == 0) {
        // ================================ MMA issue: CTA 0 of the pair ===============================================
        // (this one lane's instruction stream is the pace of the tensor pipe: keep it short)
        const uint64_t adesc0 = tc_smem_desc(sA), bdesc0 = tc_smem_desc(sB);
        for (int i = 0; i < nseg && ok; ++i) {
          const int it = it0 + i, s = it % kTcStages, a = it & 1;
          const uint32_t ph = (uint32_t)(it / kTcStages) & 1u, pa = (uint32_t)(it >> 1) & 1u;
          // CTA-scope acquires on purpose: what the remote arrives order is shared memory read by the tensor core (async
          // proxy, fenced by its writers), nothing this thread loads; a cluster-scope acquire costs a CCTL.IVALL - the
          // whole L1 invalidated - on every successful wait (12 % of this warp's time in the ncu samples)
          ok = tc_wait(b_full + 8u * (uint32_t)s, ph) && tc_wait(b_peer + 8u * (uint32_t)s, ph) &&
               tc_wait(b_acc_free + 8u * (uint32_t)a, pa ^ 1u);
          if (!ok) break;
          tc_fence_after();
          if (tc_elect_one()) {
            // descriptors differ in their start-address field only (16-byte units, low word): one add per operand
            const uint32_t blo = (uint32_t)bdesc0 + (uint32_t)s * (kTcBBytes >> 4), d = tmem + (uint32_t)(a * kTcN);
#pragma unroll
            for (int k = 0; k < 8; k++) {
              const uint32_t ka = ((uint32_t)(k >> 2) * (kTcQ * 128u) + (uint32_t)(k & 3) * 32u) >> 4;
              const uint32_t kb = ((uint32_t)(k >> 2) * (kTcBRows * 128u) + (uint32_t)(k & 3) * 32u) >> 4;
              tc_mma(d, (adesc0 & 0xFFFFFFFF00000000ull) | ((uint32_t)adesc0 + ka),
                     (bdesc0 & 0xFFFFFFFF00000000ull) | (blo + kb), k > 0);
            }
            tc_commit(b_free + 8u * (uint32_t)s);       // the stage may be refilled once these MMAs have read it
            tc_commit(b_acc_full + 8u * (uint32_t)a);   // and the accumulator is complete
          }
          __syncwarp();
        }
      } else {
        // CTA 1: its expander warps signal its OWN stage barrier (a release at CTA scope); this otherwise idle warp
        // forwards each completed stage to CTA 0 with one cluster-scope release.  (Every expander warp arriving
        // remotely cost it 700 clocks per tile and made the expanders the bottleneck.)
        for (int i = 0; i < nseg && ok; ++i) {
          const int it = it0 + i, s = it % kTcStages;
          ok = tc_wait(b_full + 8u * (uint32_t)s, (uint32_t)(it / kTcStages) & 1u);
          if (!ok) break;
          if (lane == 0) tc_arrive_remote(r_peer + 8u * (uint32_t)s);
          __syncwarp();
        }
      }
      if (!ok && lane == 0) atomicExch(err, 3);
    } else {
      // ================================ expanders: packed database tile -> operand B ==================================
      const int e = tid - 32 * (kTcEpiWarps + 1);
      const int row = e & (kTcBRows - 1), wbase = 4 * (e >> 7);   // row of this CTA's half of the tile, word group
      const int trow = (int)rank * kTcBRows + row;                // row of the whole 256-descriptor tile
      // packed descriptors are fetched two tile steps ahead into registers (a deeper register pipeline and L2 prefetch
      // instructions sixteen steps ahead were measured: no change, the expanders do not wait for memory)
      const uint4* dbp = db + 2 * ((int64_t)t0 * kTcN + trow) + (e >> 7);   // this thread's 16 bytes of tile step 0
      auto fetch = [&](int i) {                                    // tile step i of the segment
        uint4 v = make_uint4(0, 0, 0, 0);
        if (i < nseg && trow < seg_rows - i * kTcN) v = __ldg(dbp + (size_t)i * (2 * kTcN));
        return v;
      };
      uint4 w0 = fetch(0), w1 = fetch(1);
      for (int i = 0; i < nseg && ok; ++i) {
        const int it = it0 + i, s = it % kTcStages;
        const uint4 w2 = fetch(i + 2);
        if (it >= kTcStages) ok = tc_wait(b_free + 8u * (uint32_t)s, (uint32_t)(it / kTcStages - 1) & 1u);
        if (!ok) break;
        const uint32_t b0 = sB + (uint32_t)s * kTcBBytes;
        if (trow < seg_rows - i * kTcN) {
          const uint32_t w[4] = {w0.x, w0.y, w0.z, w0.w};
          tc_expand_words<4>(b0, kTcBRows, row, wbase, w);
        } else tc_zero_words<4>(b0, kTcBRows, row, wbase);        // past the end of the database: +0.0, ignored by the epilogue
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(b_full + 8u * (uint32_t)s);
        w0 = w1; w1 = w2;
      }
      if (!ok && lane == 0) atomicExch(err, 3);
    }
    // every MMA of the segment has completed (the epilogue saw its last accumulator): operand A may be replaced
    tc_fence_before();
    tc_cluster_sync();
    tc_fence_after();
    lin += nseg;
    it0 += nseg;
  }
  if (warp == kTcEpiWarps) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

}  // namespace
