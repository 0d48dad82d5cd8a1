// Included by orbm_match.cu: kernels of the search routines (M2, M4, M5, M6, M7, M9).
//
// Every routine of the reference is "for each query in order: ordered candidate list -> Hamming distances ->
// best / second -> thresholds, with a loop-carried already-taken / steal dependence" (SURVEY H3).  Here:
//   prep    (parallel)  query records + an upper bound of each candidate list
//   offsets (warp/problem) scan -> candidate block of the problem inside one shared arena
//   fill    (warp/query) candidates in the REFERENCE'S VISITING ORDER with their distances, (dist << 22) | target
//   commit  (warp/problem) in-order replay of the accept / steal / taken logic; lanes share one query's list and
//           reduce packed (dist, position) keys, whose order equals the sequential scan with strict `<`
//   emit    output rows in the layout of include/orbm.h
namespace {

// Programmatic dependent launch (see orbx_kernels.cu): first statement of every kernel of the search chain.
#define ORBM_PDL_ENTRY()                                         \
  do {                                                           \
    asm volatile("griddepcontrol.launch_dependents;");           \
    asm volatile("griddepcontrol.wait;" ::: "memory");           \
  } while (0)

// K_PROJ_INDEP: the best-only loops of Fuse (src/ORBmatcher.cc:1094-1113, 1223-1242) and SearchBySim3 (:1373-1398,
// :1458-1483): no target is taken on entry, nothing an earlier point accepted is skipped, every point reports its own
// best keypoint (two points may pick the same one) - no loop-carried state at all.
enum Kind { K_INIT = 0, K_WINDOW, K_PROJ_FF, K_PROJ_LOCAL, K_PROJ_BEST, K_BOW_KF_F, K_BOW_KF_KF, K_TRIANG, K_PROJ_INDEP };

struct Problem {
  int nq, nt;          // queries (may be rewritten by bow_prep), target keypoints
  int qstart, tstart;  // offsets into the per-query / per-target workspaces
  int qsrc;            // index of source element 0 in the query source arrays (set1 keypoints or points)
  int nsrc;            // number of source elements (rows of the per-source output)
  int tbase;           // set2.frame_off[target frame]
  int tframe;          // target frame in set2
  int f1;              // source frame in set1 (-1 for point problems)
  int cand_base;       // written by offsets_kernel
  int cand_total;
  int nmatches;        // written by commit_kernel
};

struct __align__(16) QRec {
  float x, y, r;  // window query
  int lv;         // (minLevel & 0xffff) | (maxLevel << 16)
  int cells;      // cx0 | cx1 << 8 | cy0 << 16 | cy1 << 24, -1 = empty window
  int src;        // source index: keypoint of F1 / point / feature-vector item
  int active;
  int cnt;        // upper bound after prep, exact after fill
  int off;        // offset inside the problem's candidate block
  int match;      // matched target (-1 none)
  int bin;        // rotation-histogram bin of the accepted match (-1 none)
  int list2;      // BoW: absolute start of the shared node's target list in set2.fv_items
};

struct QuerySrc {  // where query descriptors / angles / positions come from
  const uint4* desc;
  const float *angle, *x, *y;
  const int32_t* octave;
  const uint8_t* valid;
};

struct SearchArgs {
  int kind, nproblems;
  Problem* prob;
  QRec* q;
  int2* tstate;      // per target: (owner source index: -1 free, -2 taken on entry ; matched distance)
  uint32_t* cand;
  int64_t cand_cap;
  unsigned long long* cursor;  // [0] arena cursor, [1] overflow flag
  QuerySrc src;
  FrameSetView s1, s2;
  const float* qxy;  // K_INIT: prev_xy rows, K_PROJ_FF: uv rows  [problem][stride][2]
  const float* pu; const float* pv; const float* pr; const int32_t* pminl; const int32_t* pmaxl;  // point problems
  int stride;
  int window, min_level, max_level, check_ori, th_dist;
  float ratio;
  const float* F12;     // K_TRIANG [problem][9]
  const float* sigma2;  // K_TRIANG [nlevels]
  int nlevels;
  int32_t* out;         // [problem][stride]
  float* out_xy;        // K_INIT: prev_xy (same buffer as qxy)
  int32_t* nmatches;    // [problem]
};

__device__ __forceinline__ int pack_lv(int a, int b) { return (a & 0xffff) | (b << 16); }
__device__ __forceinline__ int lv_min(int lv) { return (int)(short)(lv & 0xffff); }
__device__ __forceinline__ int lv_max(int lv) { return lv >> 16; }

__device__ __forceinline__ int find_problem(const Problem* prob, int np, int idx, bool by_target) {
  int lo = 0, hi = np - 1;  // last problem whose start <= idx
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    const int st = by_target ? prob[mid].tstart : prob[mid].qstart;
    if (st <= idx) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// per-target state on entry
__global__ void __launch_bounds__(256) target_init_kernel(SearchArgs a, int ttotal) {
  ORBM_PDL_ENTRY();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ttotal) return;
  const Problem& P = a.prob[find_problem(a.prob, a.nproblems, i, true)];
  const int t = i - P.tstart;
  const bool v = a.s2.valid[P.tbase + t] != 0;
  bool taken = false;
  switch (a.kind) {
    case K_PROJ_FF: case K_PROJ_LOCAL: case K_PROJ_BEST: case K_TRIANG: taken = v; break;  // already has a MapPoint
    case K_BOW_KF_KF: taken = !v; break;                                                  // :763-767
    default: break;
  }
  a.tstate[i] = make_int2(taken ? -2 : -1, INT_MAX);
}

// grid family: query record + candidate upper bound (sum of the CSR column ranges of the cell window)
__global__ void __launch_bounds__(256) grid_prep_kernel(SearchArgs a, int qtotal) {
  ORBM_PDL_ENTRY();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= qtotal) return;
  const int p = find_problem(a.prob, a.nproblems, i, false);
  const Problem& P = a.prob[p];
  const int q = i - P.qstart, g = P.qsrc + q;
  QRec r;
  r.src = q; r.match = -1; r.bin = -1; r.list2 = 0; r.off = 0;
  switch (a.kind) {
    case K_INIT: {  // :611-618
      const int level1 = a.src.octave[g];
      r.active = !(level1 > 0);
      r.x = a.qxy[((int64_t)p * a.stride + q) * 2]; r.y = a.qxy[((int64_t)p * a.stride + q) * 2 + 1];
      r.r = (float)a.window; r.lv = pack_lv(level1, level1);
      break;
    }
    case K_WINDOW: {  // :425-443
      const int level1 = a.src.octave[g];
      r.active = a.src.valid[g] != 0;
      if (a.min_level > 0 && level1 < a.min_level) r.active = 0;
      if (a.max_level < INT_MAX && level1 > a.max_level) r.active = 0;
      r.x = a.src.x[g]; r.y = a.src.y[g]; r.r = (float)a.window; r.lv = pack_lv(level1, level1);
      break;
    }
    case K_PROJ_FF: {  // :532-553
      const int level1 = a.src.octave[g];
      r.active = a.src.valid[g] != 0;
      r.x = a.qxy[((int64_t)p * a.stride + q) * 2]; r.y = a.qxy[((int64_t)p * a.stride + q) * 2 + 1];
      r.r = (float)a.window; r.lv = pack_lv(level1, level1);
      break;
    }
    default: {  // K_PROJ_LOCAL / K_PROJ_BEST: points
      r.active = a.src.valid ? a.src.valid[g] != 0 : 1;
      r.x = a.pu[g]; r.y = a.pv[g]; r.r = a.pr[g]; r.lv = pack_lv(a.pminl[g], a.pmaxl[g]);
      break;
    }
  }
  int ub = 0;
  r.cells = -1;
  if (r.active) {
    int cx0, cx1, cy0, cy1;
    if (area_cells(r.x, r.y, r.r, a.s2.gridp[P.tframe], &cx0, &cx1, &cy0, &cy1)) {
      r.cells = cx0 | (cx1 << 8) | (cy0 << 16) | (cy1 << 24);
      const int32_t* cs = a.s2.cell_start + (int64_t)P.tframe * (kGridCells + 1);
      for (int ix = cx0; ix <= cx1; ix++) ub += cs[ix * kGridRows + cy1 + 1] - cs[ix * kGridRows + cy0];
    }
  }
  r.cnt = ub;
  a.q[i] = r;
}

// one warp per problem: exclusive scan of the per-query bounds, block allocation in the shared candidate arena
__global__ void __launch_bounds__(128) offsets_kernel(SearchArgs a) {
  ORBM_PDL_ENTRY();
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= a.nproblems) return;
  Problem& P = a.prob[p];
  QRec* q = a.q + P.qstart;
  int run = 0;
  for (int i0 = 0; i0 < P.nq; i0 += 128) {      // four consecutive queries per lane: independent loads, one round trip
    int n[4], sum = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int i = i0 + 4 * lane + k;
      n[k] = i < P.nq ? q[i].cnt : 0;
      sum += n[k];
    }
    int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    int o = run + incl - sum;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int i = i0 + 4 * lane + k;
      if (i < P.nq) q[i].off = o;
      o += n[k];
    }
    run += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) {
    const unsigned long long base = atomicAdd(a.cursor, (unsigned long long)run);
    P.cand_total = run;
    if (base + run > (unsigned long long)a.cand_cap) { a.cursor[1] = 1; P.cand_base = -1; }
    else P.cand_base = (int)base;
  }
}

// grid family: one warp per query.  Every column of the cell window is ONE contiguous CSR range; the ranges are
// concatenated (prefix sums in shared memory) and walked 32 items at a time with all lanes busy - a 21 x 21 cell
// window holds ~6 items per column, so a column-at-a-time walk would leave 80 % of the lanes idle.  Order = column by
// column, CSR order inside a column = Frame::GetFeaturesInArea's order; survivors keep it (ballot prefix) and get
// their Hamming distance attached.
__global__ void __launch_bounds__(256) grid_fill_kernel(SearchArgs a, int qtotal) {
  ORBM_PDL_ENTRY();
  __shared__ int s_pre[8][kGridCols + 1];   // exclusive prefix of the column lengths, [ncol] = total
  __shared__ int s_s0[8][kGridCols];        // CSR start of each column's range
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (i >= qtotal) return;
  QRec& r = a.q[i];
  if (r.cnt == 0) return;  // inactive query or empty window: the bound is already 0
  const Problem& P = a.prob[find_problem(a.prob, a.nproblems, i, false)];
  if (P.cand_base < 0) return;
  const int g = P.qsrc + r.src;
  const uint4 d0 = a.src.desc[2 * (int64_t)g], d1 = a.src.desc[2 * (int64_t)g + 1];
  const int cx0 = r.cells & 255, cx1 = (r.cells >> 8) & 255, cy0 = (r.cells >> 16) & 255, cy1 = (r.cells >> 24) & 255;
  const int minl = lv_min(r.lv), maxl = lv_max(r.lv);
  const float x = r.x, y = r.y, rad = r.r;
  const int32_t* cs = a.s2.cell_start + (int64_t)P.tframe * (kGridCells + 1);
  const int32_t* items = a.s2.items + P.tbase;
  uint32_t* out = a.cand + P.cand_base + r.off;
  const int ncol = cx1 - cx0 + 1;           // <= kGridCols
  int* pre = s_pre[warp];
  int* st0 = s_s0[warp];
  int total = 0;
  for (int c0 = 0; c0 < ncol; c0 += 32) {
    const int c = c0 + lane;
    int s0 = 0, len = 0;
    if (c < ncol) {
      s0 = cs[(cx0 + c) * kGridRows + cy0];
      len = cs[(cx0 + c) * kGridRows + cy1 + 1] - s0;
    }
    int incl = len;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (c < ncol) { st0[c] = s0; pre[c] = total + incl - len; }
    total += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) pre[ncol] = total;
  __syncwarp();
  int n = 0;
  for (int k0 = 0; k0 < total; k0 += 32) {
    const int k = k0 + lane;
    bool on = false;
    int t = 0;
    if (k < total) {
      // the column holding item k: the largest c with pre[c] <= k (empty columns repeat a prefix; the last one wins)
      int c = 0;
#pragma unroll
      for (int step = 32; step >= 1; step >>= 1)
        if (c + step < ncol && pre[c + step] <= k) c += step;
      t = items[st0[c] + (k - pre[c])];
      const int gt = P.tbase + t;
      on = area_accept(a.s2.x[gt], a.s2.y[gt], a.s2.octave[gt], x, y, rad, minl, maxl);
    }
    const unsigned bal = __ballot_sync(0xffffffffu, on);
    if (on) {
      const int gt = P.tbase + t;
      const int dist = hamming256(d0, d1, a.s2.desc[2 * (int64_t)gt], a.s2.desc[2 * (int64_t)gt + 1]);
      out[n + __popc(bal & ((1u << lane) - 1))] = ((uint32_t)dist << kCandShift) | (uint32_t)t;
    }
    n += __popc(bal);
  }
  __syncwarp();
  if (lane == 0) r.cnt = n;
}

// BoW family: one warp per problem.  Shared vocabulary nodes in ascending id order (the merge-join of
// src/ORBmatcher.cc:176-259 visits exactly these); queries = the source features of every shared node, in order.
__global__ void __launch_bounds__(128) bow_prep_kernel(SearchArgs a) {
  ORBM_PDL_ENTRY();
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= a.nproblems) return;
  Problem& P = a.prob[p];
  const int f1 = P.f1, f2 = P.tframe;
  const int nn1 = a.s1.node_off[f1 + 1] - a.s1.node_off[f1], nn2 = a.s2.node_off[f2 + 1] - a.s2.node_off[f2];
  const int32_t* id1 = a.s1.node_id + a.s1.node_off[f1];
  const int32_t* id2 = a.s2.node_id + a.s2.node_off[f2];
  const int32_t* st1 = a.s1.node_start + a.s1.node_off[f1] + f1;
  const int32_t* st2 = a.s2.node_start + a.s2.node_off[f2] + f2;
  const int32_t* it1 = a.s1.fv_items + a.s1.fvitem_off[f1];
  const int list2_base = a.s2.fvitem_off[f2];
  QRec* q = a.q + P.qstart;
  int run = 0;
  for (int k0 = 0; k0 < nn1; k0 += 32) {
    const int k1 = k0 + lane;
    int k2 = -1, len = 0;
    if (k1 < nn1) {
      const int id = id1[k1];
      int lo = 0, hi = nn2;  // lower_bound
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (id2[mid] < id) lo = mid + 1; else hi = mid; }
      if (lo < nn2 && id2[lo] == id) { k2 = lo; len = st1[k1 + 1] - st1[k1]; }
    }
    int incl = len;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    const int off = run + incl - len;
    for (int j = 0; j < len; j++) {
      const int idx1 = it1[st1[k1] + j];
      const bool v = a.src.valid[P.qsrc + idx1] != 0;
      QRec r;
      r.x = r.y = r.r = 0.f; r.lv = 0; r.cells = -1; r.off = 0;
      r.src = idx1; r.match = -1; r.bin = -1;
      r.active = a.kind == K_TRIANG ? !v : v;  // :191-195, :750-754 / :891-895
      r.list2 = list2_base + st2[k2];
      r.cnt = r.active ? st2[k2 + 1] - st2[k2] : 0;
      q[off + j] = r;
    }
    run += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) P.nq = run;
}

__global__ void __launch_bounds__(256) bow_fill_kernel(SearchArgs a, int qtotal) {
  ORBM_PDL_ENTRY();
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (i >= qtotal) return;
  const Problem& P = a.prob[find_problem(a.prob, a.nproblems, i, false)];
  if (i - P.qstart >= P.nq || P.cand_base < 0) return;
  const QRec& r = a.q[i];
  if (r.cnt == 0) return;
  const int g = P.qsrc + r.src;
  const uint4 d0 = a.src.desc[2 * (int64_t)g], d1 = a.src.desc[2 * (int64_t)g + 1];
  uint32_t* out = a.cand + P.cand_base + r.off;
  for (int j = lane; j < r.cnt; j += 32) {
    const int t = a.s2.fv_items[r.list2 + j];
    const int gt = P.tbase + t;
    const int dist = hamming256(d0, d1, a.s2.desc[2 * (int64_t)gt], a.s2.desc[2 * (int64_t)gt + 1]);
    out[j] = ((uint32_t)dist << kCandShift) | (uint32_t)t;
  }
}

// rotation bin (e.g. src/ORBmatcher.cc:234-241): factor = 1/HISTO_LENGTH, round() half away from zero
__device__ __forceinline__ int rot_bin(float a1, float a2) {
  const float factor = 1.0f / ORBM_HISTO_LENGTH;
  float rot = __fsub_rn(a1, a2);
  if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
  int bin = (int)roundf(__fmul_rn(rot, factor));
  if (bin == ORBM_HISTO_LENGTH) bin = 0;
  return bin;
}

__device__ __forceinline__ void warp_top2(uint32_t& b, uint32_t& s) {
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    const uint32_t ob = __shfl_xor_sync(0xffffffffu, b, d), os = __shfl_xor_sync(0xffffffffu, s, d);
    s = min(min(s, os), max(b, ob));
    b = min(b, ob);
  }
}

// ORBmatcher::CheckDistEpipolarLine (src/ORBmatcher.cc:136-153), individually rounded float ops in source order
__device__ __forceinline__ bool epipolar_ok(float x1, float y1, float x2, float y2, const float* F, float sigma2) {
  const float ea = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[0]), __fmul_rn(y1, F[3])), F[6]);
  const float eb = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[1]), __fmul_rn(y1, F[4])), F[7]);
  const float ec = __fadd_rn(__fadd_rn(__fmul_rn(x1, F[2]), __fmul_rn(y1, F[5])), F[8]);
  const float num = __fadd_rn(__fadd_rn(__fmul_rn(ea, x2), __fmul_rn(eb, y2)), ec);
  const float den = __fadd_rn(__fmul_rn(ea, ea), __fmul_rn(eb, eb));
  if (den == 0) return false;
  const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
  return (double)dsqr < __dmul_rn(3.84, (double)sigma2);
}


// What follows the query loop, by ONE warp: rotation histogram over the accept events, ComputeThreeMaxima + removal of
// the other bins, copy-out of a shared-memory target state (copy_to != nullptr) and the match count.
__device__ void commit_tail(const SearchArgs& a, Problem& P, int p, QRec* qr, int2* ts, int2* copy_to, int* s_hist_row,
                            bool use_hist, int lane) {
  // rotation histogram over the accept events (rotHist[bin].push_back at :665-676 and its siblings), in parallel
  int hist = 0;                             // lane b (< 30) = rotHist[b].size()
  if (use_hist) {
    s_hist_row[lane] = 0;
    __syncwarp();
    for (int q = lane; q < P.nq; q += 32) {
      const int t = qr[q].bin;
      if (t >= 0) {
        const int bin = rot_bin(a.src.angle[P.qsrc + qr[q].src], a.s2.angle[P.tbase + t]);
        qr[q].bin = bin;
        atomicAdd(&s_hist_row[bin], 1);
      }
    }
    __syncwarp();
    hist = s_hist_row[lane];
  }
  // rotation consistency (ComputeThreeMaxima, src/ORBmatcher.cc:1748-1789) + removal of the other bins
  if (a.check_ori && use_hist) {
    int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
    for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
      const int sz = __shfl_sync(0xffffffffu, hist, i);
      if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
      else if (sz > max3) { max3 = sz; ind3 = i; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
    for (int q = lane; q < P.nq; q += 32) {
      const int bin = qr[q].bin;
      if (bin >= 0 && bin != ind1 && bin != ind2 && bin != ind3 && qr[q].match >= 0) {
        ts[qr[q].match].x = -1;
        qr[q].match = -1;
      }
    }
    __syncwarp();
  }
  if (copy_to) {
    for (int i = lane; i < P.nt; i += 32) copy_to[i] = ts[i];
  }
  int n = 0;
  for (int q = lane; q < P.nq; q += 32) n += qr[q].match >= 0;
  n = __reduce_add_sync(0xffffffffu, n);
  if (lane == 0) { P.nmatches = n; a.nmatches[p] = n; }
}

// commit_tail for a whole CTA (commit_par_kernel): the same steps with the query loops spread over all threads; ts is the
// global target state.  s_hist32: 32 ints, s_misc: 4 ints of shared memory.
__device__ void commit_tail_cta(const SearchArgs& a, Problem& P, int p, QRec* qr, int2* ts, int* s_hist32, int* s_misc,
                                bool use_hist) {
  const int tid = threadIdx.x, lane = tid & 31, nq = P.nq;
  if (tid < 32) s_hist32[tid] = 0;
  if (tid < 4) s_misc[tid] = tid == 3 ? 0 : -1;          // ind1, ind2, ind3, match count
  __syncthreads();
  if (use_hist) {
    for (int q = tid; q < nq; q += blockDim.x) {
      const int t = qr[q].bin;
      if (t >= 0) {
        const int bin = rot_bin(a.src.angle[P.qsrc + qr[q].src], a.s2.angle[P.tbase + t]);
        qr[q].bin = bin;
        atomicAdd(&s_hist32[bin], 1);
      }
    }
    __syncthreads();
  }
  const bool remove = a.check_ori && use_hist;
  if (remove) {
    if (tid < 32) {                                       // ComputeThreeMaxima, src/ORBmatcher.cc:1748-1789
      const int hist = s_hist32[lane];
      int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
      for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
        const int sz = __shfl_sync(0xffffffffu, hist, i);
        if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
        else if (sz > max3) { max3 = sz; ind3 = i; }
      }
      if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
      else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
      if (lane == 0) { s_misc[0] = ind1; s_misc[1] = ind2; s_misc[2] = ind3; }
    }
    __syncthreads();
  }
  const int ind1 = s_misc[0], ind2 = s_misc[1], ind3 = s_misc[2];
  int n = 0;
  for (int q = tid; q < nq; q += blockDim.x) {
    int m = qr[q].match;
    if (remove && m >= 0) {
      const int bin = qr[q].bin;
      if (bin >= 0 && bin != ind1 && bin != ind2 && bin != ind3) {
        ts[m].x = -1;
        qr[q].match = -1;
        m = -1;
      }
    }
    n += m >= 0;
  }
  n = __reduce_add_sync(0xffffffffu, n);
  if (lane == 0 && n) atomicAdd(&s_misc[3], n);
  __syncthreads();
  if (tid == 0) { P.nmatches = s_misc[3]; a.nmatches[p] = s_misc[3]; }
}

// ------------------------------------------------------------------------------------------------------------------
// commit, PARALLEL form: one CTA per problem finds the fixed point of the
// reference's sequential query loop instead of replaying it.
// The loop-carried state is tiny: a query's decision depends on the earlier queries only through which of ITS candidate
// targets they accepted (and, for SearchForInitialization, at which distance: `vMatchedDistance[i2] <= dist` skips).
// Let A(q) be the target query q accepts.  Round 0 evaluates every query against the state on entry; round r + 1
// evaluates every query q against "state on entry + the accepts A_r(q') of the queries q' < q".  A(q) depends only on
// A(q' < q), so after round r the first r queries are final and the iteration reaches the sequential result; in
// practice conflicts are sparse and it stops after a handful of rounds (queries below the first one that changed are
// final and are not evaluated again).  Each round is fully parallel: thread = query.
//   taken kinds   taker[t] = min{q : A(q) = t} (-1: taken on entry); q skips t  <=>  taker[t] < q
//   K_INIT        the accepts on t form a list (head[t], next[q]); q skips (t, d)  <=>  min{D(q') : q' < q on t} <= d;
//                 a later accept steals the target: the owner is the LAST accepter, earlier ones lose their match
//                 (:656-663) but keep their rotation-histogram entry.
// A query's decision under ANY state is "the first two candidates, in (distance, list position) order, that the state
// does not skip", so the candidate lists are walked ONCE: a first pass keeps every query's kParTop smallest keys (and
// their targets) in shared memory and the rounds only look at those; a query whose cached keys are all skipped while
// its list holds more (rare) walks the list again.  A round then costs a few shared-memory reads per query, so even a
// greedy assignment with chains of hundreds of rounds stays cheap.
// Shared memory: A, D double-buffered [nq], next [nq], top keys / targets [kParTop][nq], flag [nq]; taker / head [nt],
// base [nt], the round's
// rescan list [nq]  ->  (25 + 6 kParTop) nq + 5 nt bytes.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kParTop = 4;
__host__ __device__ inline size_t commit_par_smem(int max_nq, int max_nt) {
  return (size_t)((max_nq + 3) & ~3) * (25 + 6 * kParTop) + (size_t)((max_nt + 3) & ~3) * 5 + 64;
}

__global__ void __launch_bounds__(1024) commit_par_kernel(SearchArgs a, int max_nq, int max_nt) {
  ORBM_PDL_ENTRY();
  extern __shared__ __align__(16) uint8_t par_smem[];
  __shared__ int s_hist[32], s_misc[4];
  __shared__ int s_first_changed, s_nres;
  const int p = blockIdx.x;
  Problem& P = a.prob[p];
  if (P.cand_base < 0) return;  // arena overflow: the host grows the arena and reruns
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int nq = P.nq, nt = P.nt;
  const int qpad = (max_nq + 3) & ~3, tpad = (max_nt + 3) & ~3;
  int* s_A0 = reinterpret_cast<int*>(par_smem);
  int* s_A1 = s_A0 + qpad;
  int* s_D0 = s_A1 + qpad;
  int* s_D1 = s_D0 + qpad;
  int* s_next = s_D1 + qpad;
  int* s_res = s_next + qpad;                  // queries of this round whose cached keys were all skipped
  uint32_t* s_key = reinterpret_cast<uint32_t*>(s_res + qpad);  // [kParTop][qpad] smallest keys of each query, ascending
  int* s_link = reinterpret_cast<int*>(s_key + kParTop * qpad);   // taker[t] (taken kinds) or head[t] (K_INIT)
  uint16_t* s_tgt = reinterpret_cast<uint16_t*>(s_link + tpad);   // [kParTop][qpad] their targets
  uint8_t* s_more = reinterpret_cast<uint8_t*>(s_tgt + kParTop * qpad);   // [qpad] the list holds more than kParTop eligible
  uint8_t* s_base = s_more + qpad;                                 // [tpad] 1 = taken on entry
  QRec* qr = a.q + P.qstart;
  int2* gts = a.tstate + P.tstart;
  const uint32_t* cand = a.cand + P.cand_base;
  const int kind = a.kind;
  const float ratio = a.ratio;
  const float kIntMaxF = 2147483648.0f;  // (float)INT_MAX
  const bool use_hist = kind == K_WINDOW || (a.check_ori && kind != K_PROJ_FF && kind != K_PROJ_LOCAL && kind != K_PROJ_INDEP);
  const int kFree = 0x7fffffff;
  float F[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (kind == K_TRIANG)
    for (int k = 0; k < 9; k++) F[k] = a.F12[(int64_t)p * 9 + k];

  for (int t = tid; t < nt; t += blockDim.x) {
    const bool taken = gts[t].x != -1;
    s_base[t] = taken;
    s_link[t] = kind == K_INIT ? -1 : (taken ? -1 : kFree);
  }
  for (int q = tid; q < nq; q += blockDim.x) { s_A0[q] = -1; s_D0[q] = 0; }
  if (tid == 0) { s_first_changed = kFree; s_nres = 0; }
  __syncthreads();

  // ---- first pass: every query's kParTop smallest keys among the candidates that are not taken on entry -------------
  // (a target taken on entry never becomes free; for K_INIT nothing is excluded on entry)
  for (int q = tid; q < nq; q += blockDim.x) {
    const int cnt = qr[q].cnt;
    const uint32_t* cl = cand + qr[q].off;
    uint32_t k0 = kNone, k1 = kNone, k2 = kNone, k3 = kNone;
    int t0 = 0, t1 = 0, t2 = 0, t3 = 0, eligible = 0;
    auto put = [&](uint32_t e, int j) {
      const int t = (int)(e & kCandMask);
      if (kind != K_INIT && s_base[t]) return;
      if (kind == K_TRIANG && (int)(e >> kCandShift) > ORBM_TH_LOW) return;        // :918-921
      eligible++;
      // (dist << kCandShift) | list position; SearchForTriangulation sorts by (dist, idx2), which the entry itself is
      const uint32_t key = kind == K_TRIANG ? e : ((e & ~(uint32_t)kCandMask) | (uint32_t)j);
      if (key < k3) {                                                   // insertion into the ascending quadruple
        if (key < k2) {
          k3 = k2; t3 = t2;
          if (key < k1) {
            k2 = k1; t2 = t1;
            if (key < k0) { k1 = k0; t1 = t0; k0 = key; t0 = t; } else { k1 = key; t1 = t; }
          } else { k2 = key; t2 = t; }
        } else { k3 = key; t3 = t; }
      }
    };
    int j = 0;
    for (; j + 4 <= cnt; j += 4) {
      const uint32_t e0 = cl[j], e1 = cl[j + 1], e2 = cl[j + 2], e3 = cl[j + 3];
      put(e0, j); put(e1, j + 1); put(e2, j + 2); put(e3, j + 3);
    }
    for (; j < cnt; j++) put(cl[j], j);
    s_key[q] = k0; s_key[qpad + q] = k1; s_key[2 * qpad + q] = k2; s_key[3 * qpad + q] = k3;
    s_tgt[q] = (uint16_t)t0; s_tgt[qpad + q] = (uint16_t)t1; s_tgt[2 * qpad + q] = (uint16_t)t2; s_tgt[3 * qpad + q] = (uint16_t)t3;
    uint32_t flags = eligible > kParTop ? 1u : 0u;
    if (kind == K_TRIANG) {                    // the epipolar test does not depend on the state: once per cached key
      const int src = qr[q].src;
      const float x1 = a.src.x[P.qsrc + src], y1 = a.src.y[P.qsrc + src];
      const uint32_t ks[kParTop] = {k0, k1, k2, k3};
#pragma unroll
      for (int k = 0; k < kParTop; k++) {
        if (ks[k] == kNone) break;
        const int gt = P.tbase + (int)(ks[k] & kCandMask);
        const int lvl = min(max(a.s2.octave[gt], 0), a.nlevels - 1);
        if (epipolar_ok(x1, y1, a.s2.x[gt], a.s2.y[gt], F, a.sigma2[lvl])) flags |= 2u << k;
      }
    }
    s_more[q] = (uint8_t)flags;               // bit 0: the list holds more eligible entries; bits 1..: epipolar test passed
  }
  __syncthreads();

  int* Aprev = s_A0; int* Dprev = s_D0; int* Acur = s_A1; int* Dcur = s_D1;
  int first = 0;                                // queries below `first` are final
  for (;;) {
    // ---- evaluate queries first .. nq-1 against the accepts of the previous round ---------------------------------
    // thread = query; the cached keys are in (distance, list position) order, so the first two that the state does not
    // skip are the loop's best and second best
    auto skipped = [&](int q, int t, int d) -> bool {
      if (kind == K_INIT) {                                     // :637  vMatchedDistance[i2] <= dist
        int dmin = INT_MAX;
        for (int ev = s_link[t]; ev >= 0; ev = s_next[ev])
          if (ev < q) dmin = min(dmin, Dprev[ev]);
        return dmin <= d;
      }
      if (kind == K_PROJ_INDEP) return false;                   // Fuse / SearchBySim3 scan every candidate
      return s_link[t] < q;                                     // `if (taken) continue`
    };
    auto decide = [&](int q, uint32_t b, uint32_t sk, int tb, int tsec) -> bool {
      int accept_t = -1, accept_d = 0;
      if (b != kNone) {
        const int bd = (int)(b >> kCandShift), bt = tb;
        const float sdf = sk == kNone ? kIntMaxF : (float)(int)(sk >> kCandShift);
        bool ok = false;
        switch (kind) {
          case K_INIT: ok = bd <= ORBM_TH_LOW && (float)bd < __fmul_rn(sdf, ratio); break;            // :652-654
          case K_WINDOW: case K_PROJ_FF: ok = (float)bd <= __fmul_rn(sdf, ratio) && bd <= ORBM_TH_HIGH; break;  // :476, :585
          case K_PROJ_LOCAL:                                                                            // :112-117
            if (bd <= ORBM_TH_HIGH) {
              const int l1 = a.s2.octave[P.tbase + bt];
              const int l2 = sk == kNone ? -1 : a.s2.octave[P.tbase + tsec];
              ok = !(l1 == l2 && (float)bd > __fmul_rn(ratio, sdf));
            }
            break;
          case K_PROJ_BEST: case K_PROJ_INDEP: ok = bd <= a.th_dist; break;                            // :1576, :1702, :398; :1116, :1245, :1395, :1480
          case K_BOW_KF_F: ok = bd <= ORBM_TH_LOW && (float)bd < __fmul_rn(ratio, sdf); break;          // :224-226
          default: ok = bd < ORBM_TH_LOW && (float)bd < __fmul_rn(ratio, sdf); break;                   // :791-793
        }
        if (ok) { accept_t = bt; accept_d = bd; }
      }
      Acur[q] = accept_t; Dcur[q] = accept_d;
      return accept_t != Aprev[q] || accept_d != Dprev[q];
    };
    auto store = [&](int q, int accept_t, int accept_d) -> bool {
      Acur[q] = accept_t; Dcur[q] = accept_d;
      return accept_t != Aprev[q] || accept_d != Dprev[q];
    };
    // (whole warps walk the query range together so that the two shared counters get ONE atomic per warp: a thousand
    // threads hitting the same shared word cost more than the evaluation itself)
    for (int qb = first + (tid - lane); qb < nq; qb += blockDim.x) {
      const int q = qb + lane;
      bool rescan = false, changed_q = false;
      if (q < nq && kind == K_TRIANG) {
        // SearchForTriangulation :896-958: the free candidates (dist <= TH_LOW) in (dist, idx2) order; the first free one
        // fixes the bound 2 * dist; the first one within the bound that passes the epipolar test wins
        const uint32_t flags = s_more[q];
        int accept_t = -1, accept_d = 0, dist_th = -1;
        bool open = true;                      // undecided after the cached keys
#pragma unroll
        for (int k = 0; k < kParTop; k++) {
          const uint32_t key = s_key[k * qpad + q];
          if (key == kNone) { open = false; break; }             // the list itself ended: no match
          const int t = (int)(key & kCandMask), d = (int)(key >> kCandShift);
          if (s_link[t] < q) continue;                            // vbMatched2 / already a MapPoint
          if (dist_th < 0) dist_th = 2 * d;
          if (d > dist_th) { open = false; break; }
          if (flags & (2u << k)) { accept_t = t; accept_d = d; open = false; break; }
        }
        rescan = open && (flags & 1u);
        if (!rescan) changed_q = store(q, accept_t, accept_d);
      } else if (q < nq) {
        uint32_t b = kNone, sk = kNone;        // best / second packed (dist, list position) keys
        int tb = 0, tsec = 0;                  // their targets
        bool exhausted = true;                 // ran through the cache without finding two
#pragma unroll
        for (int k = 0; k < kParTop; k++) {
          const uint32_t key = s_key[k * qpad + q];
          if (key == kNone) { exhausted = false; break; }        // the list itself ended
          const int t = s_tgt[k * qpad + q];
          if (!skipped(q, t, (int)(key >> kCandShift))) {
            if (b == kNone) { b = key; tb = t; }
            else { sk = key; tsec = t; exhausted = false; break; }
          }
        }
        rescan = exhausted && (s_more[q] & 1);   // walk the whole list below, a warp per query
        if (!rescan) changed_q = decide(q, b, sk, tb, tsec);
      }
      const unsigned rm = __ballot_sync(0xffffffffu, rescan);
      if (rm) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&s_nres, __popc(rm));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (rescan) s_res[base + __popc(rm & ((1u << lane) - 1))] = q;
      }
      const int fc = __reduce_min_sync(0xffffffffu, changed_q ? q : kFree);
      if (lane == 0 && fc != kFree) atomicMin(&s_first_changed, fc);
    }
    __syncthreads();
    // the (few) queries whose cached keys did not settle the matter: lanes share the list, packed-key reduce
    const int nres = s_nres;
    for (int r = warp; r < nres; r += nwarps) {
      const int q = s_res[r];
      const int cnt = qr[q].cnt;
      const uint32_t* cl = cand + qr[q].off;
      if (kind == K_TRIANG) {                  // the ordered walk of commit_kernel, against the round's taker view
        int accept_t = -1, accept_d = 0, dist_th = 0;
        uint32_t last = 0;
        bool first_free = true;
        const int src = qr[q].src;
        const float x1 = a.src.x[P.qsrc + src], y1 = a.src.y[P.qsrc + src];
        while (true) {
          uint32_t mn = kNone;
          for (int j = lane; j < cnt; j += 32) {
            const uint32_t e = cl[j];
            const int t = (int)(e & kCandMask), d = (int)(e >> kCandShift);
            if (d <= ORBM_TH_LOW && !(s_link[t] < q) && (first_free || e > last)) mn = min(mn, e);
          }
          mn = __reduce_min_sync(0xffffffffu, mn);
          if (mn == kNone) break;
          const int d = (int)(mn >> kCandShift), t = (int)(mn & kCandMask);
          if (first_free) { dist_th = 2 * d; first_free = false; }
          if (d > dist_th) break;
          const int gt = P.tbase + t;
          const int lvl = min(max(a.s2.octave[gt], 0), a.nlevels - 1);
          if (epipolar_ok(x1, y1, a.s2.x[gt], a.s2.y[gt], F, a.sigma2[lvl])) { accept_t = t; accept_d = d; break; }
          last = mn;
        }
        if (lane == 0 && store(q, accept_t, accept_d)) atomicMin(&s_first_changed, q);
        continue;
      }
      uint32_t b = kNone, sk = kNone;
      int tb = 0, tsec = 0;
      for (int j = lane; j < cnt; j += 32) {
        const uint32_t e = cl[j];
        const int t = (int)(e & kCandMask), d = (int)(e >> kCandShift);
        if (!skipped(q, t, d)) {
          const uint32_t key = ((uint32_t)d << kCandShift) | (uint32_t)j;
          if (key < b) { sk = b; tsec = tb; b = key; tb = t; }
          else if (key < sk) { sk = key; tsec = t; }
        }
      }
      const uint32_t gb = __reduce_min_sync(0xffffffffu, b);
      const uint32_t gs = __reduce_min_sync(0xffffffffu, b == gb ? sk : b);
      const int bt = __shfl_sync(0xffffffffu, tb, (int)(gb & 31u));          // list position j lives on lane j % 32
      const int t2 = __shfl_sync(0xffffffffu, b == gs ? tb : tsec, (int)(gs & 31u));
      if (lane == 0 && decide(q, gb, gs, bt, t2)) atomicMin(&s_first_changed, q);
    }
    __syncthreads();
    const int changed = s_first_changed;
    // queries below `first` were not evaluated: carry them over
    for (int q = tid; q < first; q += blockDim.x) { Acur[q] = Aprev[q]; Dcur[q] = Dprev[q]; }
    __syncthreads();
    if (changed == kFree) break;                 // fixed point: Acur == Aprev everywhere
    // ---- rebuild the per-target view of the accepts ---------------------------------------------------------------
    for (int t = tid; t < nt; t += blockDim.x) s_link[t] = kind == K_INIT ? -1 : (s_base[t] ? -1 : kFree);
    __syncthreads();
    for (int q = tid; q < nq; q += blockDim.x) {
      const int t = Acur[q];
      if (t >= 0) {
        if (kind == K_INIT) s_next[q] = atomicExch(&s_link[t], q);
        else atomicMin(&s_link[t], q);
      }
    }
    __syncthreads();
    first = changed;                             // A(q) for q < changed did not change and cannot change any more
    if (tid == 0) { s_first_changed = kFree; s_nres = 0; }   // all threads read them barriers ago; next use is after the barrier below
    int* tA = Aprev; Aprev = Acur; Acur = tA;
    int* tD = Dprev; Dprev = Dcur; Dcur = tD;
    __syncthreads();
  }
  // ---- the final state, written the way the sequential loop leaves it --------------------------------------------------
  // (s_link / s_next describe Aprev == Acur)
  for (int q = tid; q < nq; q += blockDim.x) {
    const int t = Acur[q];
    int match = t;
    if (t >= 0) {
      if (kind == K_INIT) {                      // the last accepter owns the target; earlier ones were robbed
        int owner = q;
        for (int ev = s_link[t]; ev >= 0; ev = s_next[ev]) owner = max(owner, ev);
        if (owner == q) gts[t] = make_int2(qr[q].src, Dcur[q]);
        else match = -1;
      } else if (kind != K_PROJ_INDEP) {       // (independent points leave the targets free)
        gts[t] = make_int2(qr[q].src, Dcur[q]);
      }
    }
    qr[q].match = match;
    qr[q].bin = (use_hist && t >= 0) ? t : -1;
  }
  __threadfence_block();
  __syncthreads();
  commit_tail_cta(a, P, p, qr, gts, s_hist, s_misc, use_hist);
}

// ------------------------------------------------------------------------------------------------------------------
// commit: one warp per problem replays the reference's query loop IN ORDER.  For one query the 32 lanes share the
// candidate list: each evaluates the skip rule against the current state, keeps its own (best, second) packed
// (dist << 22 | list position) keys and a butterfly merges them.  Packed-key order = (distance, position) order,
// which is what `if (dist < best) {second = best; best = dist} else if (dist < second) second = dist` computes.
// The replay is a serial chain, so everything on it is kept off global memory: the per-target state lives in shared
// memory (copied in / out once), query records are fetched 32 at a time and broadcast by shuffles, the next query's
// candidates are prefetched while the current one is resolved, the winners' targets travel with the keys (shuffle
// from the owning lane instead of a reload), and the rotation bins are computed in parallel after the loop (the
// histogram counts accept EVENTS, a later steal does not remove the entry: :656-663 leaves rotHist alone).
// ------------------------------------------------------------------------------------------------------------------
constexpr int kCommitWarps = 2, kCommitSmemTargets = 2048;
constexpr int kCommitSmemBytes = kCommitWarps * (kCommitSmemTargets * 8 + 2 * 32 * 32 * 4);

__global__ void __launch_bounds__(32 * kCommitWarps) commit_kernel(SearchArgs a) {
  ORBM_PDL_ENTRY();
  extern __shared__ __align__(16) uint8_t commit_smem[];
  __shared__ int s_hist[kCommitWarps][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = blockIdx.x * kCommitWarps + warp;
  if (p >= a.nproblems) return;
  Problem& P = a.prob[p];
  if (P.cand_base < 0) return;  // arena overflow: the host grows the arena and reruns
  QRec* qr = a.q + P.qstart;
  int2* gts = a.tstate + P.tstart;
  const bool ts_in_smem = P.nt <= kCommitSmemTargets;
  int2* ts = ts_in_smem ? reinterpret_cast<int2*>(commit_smem) + warp * kCommitSmemTargets : gts;
  // first 32 candidates of each query of a 32-query batch, double-buffered: [2][query in batch][lane]
  uint32_t* s_cb = reinterpret_cast<uint32_t*>(commit_smem + kCommitWarps * kCommitSmemTargets * 8) + warp * (2 * 32 * 32);
  const uint32_t s_cb_addr = (uint32_t)__cvta_generic_to_shared(s_cb);
  const uint32_t* cand = a.cand + P.cand_base;
  const int kind = a.kind;
  const float ratio = a.ratio;
  const float kIntMaxF = 2147483648.0f;  // (float)INT_MAX
  const bool use_hist = kind == K_WINDOW || (a.check_ori && kind != K_PROJ_FF && kind != K_PROJ_LOCAL && kind != K_PROJ_INDEP);
  float F[9];
  if (kind == K_TRIANG)
    for (int k = 0; k < 9; k++) F[k] = a.F12[(int64_t)p * 9 + k];

  // asynchronous copy of the first 32 candidates of every non-empty query of one batch (lane i holds query i's record)
  auto stage_batch = [&](int cnt_i, int off_i, int buf) {
    unsigned m = kind == K_TRIANG ? 0u : __ballot_sync(0xffffffffu, cnt_i != 0);
    while (m) {
      const int i = __ffs(m) - 1;
      m &= m - 1;
      const int c = __shfl_sync(0xffffffffu, cnt_i, i), o = __shfl_sync(0xffffffffu, off_i, i);
      if (lane < c)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s_cb_addr + 4u * (uint32_t)((buf * 32 + i) * 32 + lane)),
                     "l"(__cvta_generic_to_global(cand + o + lane)) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  // records: batch k in my_*, batch k+1 in nx_* (loaded one batch ahead), candidates of batch k+1 staged during batch k
  int my_cnt = 0, my_off = 0, my_src = 0, nx_cnt = 0, nx_off = 0, nx_src = 0;
  if (lane < P.nq) { my_cnt = qr[lane].cnt; my_off = qr[lane].off; my_src = qr[lane].src; }
  if (32 + lane < P.nq) { nx_cnt = qr[32 + lane].cnt; nx_off = qr[32 + lane].off; nx_src = qr[32 + lane].src; }
  stage_batch(my_cnt, my_off, 0);
  if (ts_in_smem) {                                            // four loads in flight per lane
    for (int i = lane; i < P.nt; i += 128) {
      int2 v[4];
#pragma unroll
      for (int k = 0; k < 4; k++) if (i + 32 * k < P.nt) v[k] = gts[i + 32 * k];
#pragma unroll
      for (int k = 0; k < 4; k++) if (i + 32 * k < P.nt) ts[i + 32 * k] = v[k];
    }
  }
  int buf = 0;
  for (int q0 = 0; q0 < P.nq; q0 += 32, buf ^= 1) {
    stage_batch(nx_cnt, nx_off, buf ^ 1);                      // reads of buf ^ 1 (batch k - 1) ended with a __syncwarp
    int nn_cnt = 0, nn_off = 0, nn_src = 0;
    if (q0 + 64 + lane < P.nq) { nn_cnt = qr[q0 + 64 + lane].cnt; nn_off = qr[q0 + 64 + lane].off; nn_src = qr[q0 + 64 + lane].src; }
    asm volatile("cp.async.wait_group 1;" ::: "memory");       // this batch's candidates have landed
    __syncwarp();
    const uint32_t* cb = s_cb + buf * (32 * 32);
    unsigned todo = __ballot_sync(0xffffffffu, my_cnt != 0);  // inactive query / empty window / empty node list: skip
    while (todo) {
      const int i = __ffs(todo) - 1;
      todo &= todo - 1;
      const int q = q0 + i;
      const int cnt = __shfl_sync(0xffffffffu, my_cnt, i), off = __shfl_sync(0xffffffffu, my_off, i);
      const int src = __shfl_sync(0xffffffffu, my_src, i);
      int accept_t = -1, accept_d = 0;
      if (kind != K_TRIANG) {
        uint32_t b = kNone, s = kNone;
        int tb = 0, tsec = 0;                                  // targets of this lane's best / second keys
        for (int j = lane; j < cnt; j += 32) {
          const uint32_t e = j < 32 ? cb[i * 32 + lane] : cand[off + j];
          const int t = (int)(e & kCandMask), d = (int)(e >> kCandShift);
          const int2 st = ts[t];
          const bool skip = kind == K_INIT ? st.y <= d : (kind == K_PROJ_INDEP ? false : st.x != -1);  // :637 / `if (taken) continue`
          if (!skip) {
            const uint32_t key = ((uint32_t)d << kCandShift) | (uint32_t)j;
            if (key < b) { s = b; tsec = tb; b = key; tb = t; }
            else if (key < s) { s = key; tsec = t; }
          }
        }
        // warp-wide (best, second) with two REDUX: the second is the smallest key that is not the best
        const uint32_t gb = __reduce_min_sync(0xffffffffu, b);
        const uint32_t gs = __reduce_min_sync(0xffffffffu, b == gb ? s : b);
        if (gb != kNone) {
          const int bd = (int)(gb >> kCandShift);
          const int bt = __shfl_sync(0xffffffffu, tb, (int)(gb & 31u));   // list position j lives on lane j % 32
          const float sdf = gs == kNone ? kIntMaxF : (float)(int)(gs >> kCandShift);
          bool ok = false;
          switch (kind) {
            case K_INIT: ok = bd <= ORBM_TH_LOW && (float)bd < __fmul_rn(sdf, ratio); break;            // :652-654
            case K_WINDOW: case K_PROJ_FF: ok = (float)bd <= __fmul_rn(sdf, ratio) && bd <= ORBM_TH_HIGH; break;  // :476, :585
            case K_PROJ_LOCAL: {                                                                          // :112-117
              // the second-best entry is either the best or the second key of the lane that owns its position
              const int t2 = __shfl_sync(0xffffffffu, b == gs ? tb : tsec, (int)(gs & 31u));
              if (bd <= ORBM_TH_HIGH) {
                const int l1 = a.s2.octave[P.tbase + bt];
                const int l2 = gs == kNone ? -1 : a.s2.octave[P.tbase + t2];
                ok = !(l1 == l2 && (float)bd > __fmul_rn(ratio, sdf));
              }
              break;
            }
            case K_PROJ_BEST: case K_PROJ_INDEP: ok = bd <= a.th_dist; break;                            // :1576, :1702, :398; :1116, :1245, :1395, :1480
            case K_BOW_KF_F: ok = bd <= ORBM_TH_LOW && (float)bd < __fmul_rn(ratio, sdf); break;          // :224-226
            default: ok = bd < ORBM_TH_LOW && (float)bd < __fmul_rn(ratio, sdf); break;                   // :791-793
          }
          if (ok) { accept_t = bt; accept_d = bd; }
        }
      } else {
        // SearchForTriangulation :896-958: candidates with dist <= TH_LOW sorted by (dist, idx2); walk while
        // dist <= 2 * best; the first one passing the epipolar test wins.  Entries already are (dist << 22) | idx2.
        uint32_t last = 0;
        bool first = true;
        int dist_th = 0;
        const float x1 = a.src.x[P.qsrc + src], y1 = a.src.y[P.qsrc + src];
        while (true) {
          uint32_t m = kNone;
          for (int j = lane; j < cnt; j += 32) {
            const uint32_t e = cand[off + j];
            const int t = (int)(e & kCandMask), d = (int)(e >> kCandShift);
            if (d <= ORBM_TH_LOW && ts[t].x == -1 && (first || e > last)) m = min(m, e);
          }
          m = __reduce_min_sync(0xffffffffu, m);
          if (m == kNone) break;
          const int d = (int)(m >> kCandShift), t = (int)(m & kCandMask);
          if (first) { dist_th = 2 * d; first = false; }
          if (d > dist_th) break;
          const int gt = P.tbase + t;
          int lvl = a.s2.octave[gt];
          lvl = min(max(lvl, 0), a.nlevels - 1);
          if (epipolar_ok(x1, y1, a.s2.x[gt], a.s2.y[gt], F, a.sigma2[lvl])) { accept_t = t; accept_d = d; break; }
          last = m;
        }
      }
      if (accept_t >= 0 && lane == 0) {
        if (kind == K_INIT) {  // steal :656-663
          const int prev = ts[accept_t].x;
          if (prev >= 0) qr[prev].match = -1;
        }
        if (kind != K_PROJ_INDEP) ts[accept_t] = make_int2(src, accept_d);
        qr[q].match = accept_t;
        qr[q].bin = use_hist ? accept_t : -1;   // the accepted target; turned into its rotation bin after the loop
      }
      __syncwarp();
    }
    my_cnt = nx_cnt; my_off = nx_off; my_src = nx_src;
    nx_cnt = nn_cnt; nx_off = nn_off; nx_src = nn_src;
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");

  commit_tail(a, P, p, qr, ts, ts_in_smem ? gts : nullptr, s_hist[warp], use_hist, lane);
}

// output rows (include/orbm.h): per-source for K_INIT / K_BOW_KF_KF / K_TRIANG / K_PROJ_INDEP, per-target otherwise
__global__ void __launch_bounds__(128) emit_kernel(SearchArgs a) {
  ORBM_PDL_ENTRY();
  const int p = blockIdx.x;
  const Problem& P = a.prob[p];
  int32_t* out = a.out + (int64_t)p * a.stride;
  const QRec* qr = a.q + P.qstart;
  const int2* ts = a.tstate + P.tstart;
  const bool per_source = a.kind == K_INIT || a.kind == K_BOW_KF_KF || a.kind == K_TRIANG || a.kind == K_PROJ_INDEP;
  if (per_source) {
    for (int i = threadIdx.x; i < a.stride; i += blockDim.x) out[i] = -1;
    __syncthreads();
    for (int q = threadIdx.x; q < P.nq; q += blockDim.x) {
      const int m = qr[q].match;
      if (m >= 0) {
        out[qr[q].src] = m;
        if (a.kind == K_INIT) {  // update of vbPrevMatched, :708-710
          a.out_xy[((int64_t)p * a.stride + q) * 2] = a.s2.x[P.tbase + m];
          a.out_xy[((int64_t)p * a.stride + q) * 2 + 1] = a.s2.y[P.tbase + m];
        }
      }
    }
  } else {
    for (int i = threadIdx.x; i < a.stride; i += blockDim.x) out[i] = i < P.nt ? max(ts[i].x, -1) : -1;
  }
}

}  // namespace
