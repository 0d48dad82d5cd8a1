// Included by orbm_knn.cu.
// ------------------------------------------------------------------------------------------------------------------
// M8  all-pairs 256-bit Hamming 2-NN.
//
// Bound: integer issue (XOR + POPC + adds per 32-bit word), not HBM: each CTA keeps 256*QPT query descriptors in
// REGISTERS and streams its chunk of the database through shared memory in 4 KB tiles, so a database byte is read
// from HBM/L2 once per 256*QPT queries.  Tiles arrive by TMA 1-D bulk copies (cp.async.bulk -> UBLKCP) issued by one
// thread into a 4-stage ring guarded by full/empty mbarriers; every thread of a warp reads the same database
// descriptor (shared-memory broadcast, conflict-free).  Top-2 per query is kept as packed (dist << 23) | index keys:
// key order == (distance, index) order == the reference's sequential scan with strict `<`
// (src/ORBmatcher.cc:212-221), so ties resolve to the lowest index exactly as on the CPU.
// grid = (database chunks, query tiles); per-chunk partial results are merged by merge_top2_kernel.
// ------------------------------------------------------------------------------------------------------------------
namespace {

constexpr int kKnnThreads = 256, kKnnTile = 128, kKnnStages = 4;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a lost completion must end as a reported error, never as a hung GPU.
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
#pragma unroll 1
  for (int spin = 0; spin < (1 << 20); ++spin) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return true;
  }
  return false;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

__device__ __forceinline__ void top2_insert(uint32_t& best, uint32_t& second, uint32_t key) {
  second = min(second, max(best, key));
  best = min(best, key);
}

// Hamming distance of one 256-bit pair, three instruction mixes (all exact):
//   DV 0  8 XOR + 8 POPC                          POPC-pipe bound: POPC issues 16 lanes/clk/SM on sm_100
//                                                 (tools/popc_microbench.cu), XOR/IADD3 64 lanes/clk/SM
//   DV 1  two 3:2 carry-save adders  -> 6 POPC    sum = a^b^c, carry = maj(a,b,c): one LOP3 each
//   DV 2  three carry-save adders    -> 5 POPC    balances the POPC pipe (5/16 clk) against the ALU pipe
// dist = popc(ones) + 2 * popc(twos).
// lop3 through inline PTX: ptxas otherwise re-derives its own (longer) LOP3 decomposition of the adder tree
__device__ __forceinline__ uint32_t lop3_xor3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint32_t lop3_maj(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}

template <int DV>
__device__ __forceinline__ uint32_t hamming_dv(const uint4& qa, const uint4& qb, const uint4& d0, const uint4& d1) {
  const uint32_t x0 = qa.x ^ d0.x, x1 = qa.y ^ d0.y, x2 = qa.z ^ d0.z, x3 = qa.w ^ d0.w;
  const uint32_t x4 = qb.x ^ d1.x, x5 = qb.y ^ d1.y, x6 = qb.z ^ d1.z, x7 = qb.w ^ d1.w;
  if (DV == 0) {
    return __popc(x0) + __popc(x1) + __popc(x2) + __popc(x3) + __popc(x4) + __popc(x5) + __popc(x6) + __popc(x7);
  } else {
    const uint32_t sa = lop3_xor3(x0, x1, x2), ca = lop3_maj(x0, x1, x2);
    const uint32_t sb = lop3_xor3(x3, x4, x5), cb = lop3_maj(x3, x4, x5);
    if (DV == 1) {
      return __popc(sa) + __popc(sb) + __popc(x6) + __popc(x7) + 2 * (__popc(ca) + __popc(cb));
    } else {
      const uint32_t sc = lop3_xor3(sa, sb, x6), cc = lop3_maj(sa, sb, x6);
      return __popc(sc) + __popc(x7) + 2 * (__popc(ca) + __popc(cb) + __popc(cc));
    }
  }
}

template <int QPT, int DV>
__global__ void __launch_bounds__(kKnnThreads) knn2_kernel(const uint4* __restrict__ q, int nq,
                                                           const uint4* __restrict__ db, int64_t ndb, int chunk_len,
                                                           uint32_t idx_base, uint32_t* __restrict__ part,
                                                           int* __restrict__ err) {
  __shared__ __align__(128) uint4 s_tile[kKnnStages][kKnnTile * 2];
  __shared__ __align__(8) uint64_t s_full[kKnnStages], s_empty[kKnnStages];
  const int tid = threadIdx.x, lane = tid & 31;
  const int chunk = blockIdx.x, qtile = blockIdx.y;
  const int64_t c0 = (int64_t)chunk * chunk_len;
  const int clen = (int)min((int64_t)chunk_len, ndb - c0);
  const int ntiles = (clen + kKnnTile - 1) / kKnnTile;

  uint4 qa[QPT], qb[QPT];
  uint32_t best[QPT], second[QPT];
#pragma unroll
  for (int k = 0; k < QPT; k++) {
    const int qi = (qtile * QPT + k) * kKnnThreads + tid;
    if (qi < nq) { qa[k] = q[2 * qi]; qb[k] = q[2 * qi + 1]; }
    else { qa[k] = make_uint4(0, 0, 0, 0); qb[k] = qa[k]; }
    best[k] = kNone; second[k] = kNone;
  }
  if (tid == 0) {
    for (int s = 0; s < kKnnStages; s++) {
      mbar_init(smem_u32(&s_full[s]), 1);
      mbar_init(smem_u32(&s_empty[s]), kKnnThreads / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int tile, int stage) {
    const int cnt = min(kKnnTile, clen - tile * kKnnTile);
    const uint32_t bytes = (uint32_t)cnt * 32u;
    const uint32_t bar = smem_u32(&s_full[stage]);
    mbar_expect_tx(bar, bytes);
    bulk_g2s(smem_u32(&s_tile[stage][0]), db + 2 * (c0 + (int64_t)tile * kKnnTile), bytes, bar);
  };
  if (tid == 0)
    for (int s = 0; s < kKnnStages && s < ntiles; s++) issue(s, s);

  bool ok = true;
  for (int it = 0; it < ntiles; ++it) {
    const int s = it % kKnnStages;
    ok = mbar_wait(smem_u32(&s_full[s]), (it / kKnnStages) & 1);
    if (!ok) break;
    const int cnt = min(kKnnTile, clen - it * kKnnTile);
    const uint4* t = s_tile[s];
    const uint32_t kb = idx_base + (uint32_t)(c0 + (int64_t)it * kKnnTile);
#pragma unroll 2
    for (int j = 0; j < cnt; j++) {
      const uint4 d0 = t[2 * j], d1 = t[2 * j + 1];
#pragma unroll
      for (int k = 0; k < QPT; k++) {
        const uint32_t dist = hamming_dv<DV>(qa[k], qb[k], d0, d1);
        top2_insert(best[k], second[k], (dist << 23) + (kb + j));
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&s_empty[s]));
    if (tid == 0 && it >= 1 && it - 1 + kKnnStages < ntiles) {  // refill the stage every warp left one tile ago
      const int ps = (it - 1) % kKnnStages;
      ok = mbar_wait(smem_u32(&s_empty[ps]), ((it - 1) / kKnnStages) & 1);
      if (!ok) break;
      issue(it - 1 + kKnnStages, ps);
    }
  }
  if (!ok) {                       // a lost tile: report it and leave 'none' keys, never stale ones, for the merge
    atomicExch(err, 3);
#pragma unroll
    for (int k = 0; k < QPT; k++) { best[k] = kNone; second[k] = kNone; }
  }
#pragma unroll
  for (int k = 0; k < QPT; k++) {
    const int qi = (qtile * QPT + k) * kKnnThreads + tid;
    if (qi < nq) {
      uint2* o = reinterpret_cast<uint2*>(part) + ((int64_t)chunk * nq + qi);
      *o = make_uint2(best[k], second[k]);
    }
  }
}

// Merge [nshards][nq] (best, second) key pairs into one pair per query, plus the accept rule of
// src/ORBmatcher.cc:224-226: best <= th && (float)best < ratio * (float)second (second = INT_MAX when absent).
__global__ void __launch_bounds__(256) merge_top2_kernel(const uint2* __restrict__ in, int nshards, int nq, int th,
                                                         float ratio, uint2* __restrict__ out,
                                                         int32_t* __restrict__ match) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  uint32_t best = kNone, second = kNone;
  for (int s = 0; s < nshards; s++) {
    const uint2 v = in[(int64_t)s * nq + qi];
    top2_insert(best, second, v.x);
    top2_insert(best, second, v.y);
  }
  if (out) out[qi] = make_uint2(best, second);
  if (match) {
    int m = -1;
    if (best != kNone) {
      const int bd = (int)(best >> 23);
      const float sd = second == kNone ? (float)INT_MAX : (float)(int)(second >> 23);
      if (bd <= th && (float)bd < __fmul_rn(ratio, sd)) m = (int)(best & 0x7FFFFFu);
    }
    match[qi] = m;
  }
}

}  // namespace
