// Search routines of the matcher (M2, M4, M5, M6, M7, M9): C-ABI entry points + the batched pipeline driver.
// Kernels: orbm_match_kernels.inl.  Public surface: include/orbm.h.
#include <cstdlib>
#include <utility>
#include <algorithm>
#include <cstring>
#include <vector>

#include "orbm_match.cuh"

using namespace orbm;

#include "orbm_match_kernels.inl"

namespace {

struct HostJob {
  int kind = 0;
  const orbm_frameset* set1 = nullptr;  // query frames (null for point problems)
  const orbm_frameset* set2 = nullptr;  // target frames
  const int32_t* idx1 = nullptr;
  const int32_t* idx2 = nullptr;
  const orbm_points* points = nullptr;
  int nproblems = 0;
  float* xy_inout = nullptr;   // K_INIT prev_xy (in/out)
  const float* xy_in = nullptr;  // K_PROJ_FF uv
  int stride = 0;
  int window = 0, min_level = 0, max_level = INT_MAX, check_ori = 0, th_dist = 0;
  float ratio = 0.f;
  const float* F12 = nullptr;
  const float* sigma2 = nullptr;
  int nlevels = 0;
  int32_t* out = nullptr;
  int32_t* nmatches = nullptr;
};

template <class T>
T* carve(uint8_t*& p, size_t n) {
  T* r = reinterpret_cast<T*>(p);
  p += (n * sizeof(T) + 255) / 256 * 256;
  return r;
}
// Kernel launch with (optionally) the programmatic-stream-serialization attribute: the chain of six small kernels of a
// search over a few problems is launch-latency bound, and every kernel starts with ORBM_PDL_ENTRY().
template <class... KArgs, class... Args>
void launch_k(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

size_t padded(size_t bytes) { return (bytes + 255) / 256 * 256; }

int run_job(orbm_handle* h, const HostJob& J) {
  if (!h) return ORBM_ERR_ARG;
  const int np = J.nproblems;
  if (np < 0 || !J.set2 || !J.idx2 || !J.out || !J.nmatches || J.stride < 1) { h->err = "bad search arguments"; return ORBM_ERR_ARG; }
  if (np == 0) return ORBM_OK;
  const bool points = J.kind == K_PROJ_LOCAL || J.kind == K_PROJ_BEST || J.kind == K_PROJ_INDEP;
  const bool bow = J.kind == K_BOW_KF_F || J.kind == K_BOW_KF_KF || J.kind == K_TRIANG;
  if (!points && (!J.set1 || !J.idx1)) { h->err = "missing source frame set"; return ORBM_ERR_ARG; }
  if (points && !J.points) { h->err = "missing points"; return ORBM_ERR_ARG; }
  if (bow && (!J.set1->has_fv || !J.set2->has_fv)) { h->err = "frame sets were created without feature vectors"; return ORBM_ERR_ARG; }
  if ((J.set1 && J.set1->h != h) || J.set2->h != h) { h->err = "frame set belongs to another handle"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;

  // ---- problem table ------------------------------------------------------------------------------------------
  std::vector<Problem> prob(np);
  int64_t qtot = 0, ttot = 0, ptot = 0;
  int max_nq = 0, max_nt = 0;   // nq is an upper bound for the BoW kinds (bow_prep writes the exact count)
  const bool per_source = J.kind == K_INIT || J.kind == K_BOW_KF_KF || J.kind == K_TRIANG || J.kind == K_PROJ_INDEP;
  for (int p = 0; p < np; p++) {
    Problem& P = prob[p];
    memset(&P, 0, sizeof(P));
    const int f2 = J.idx2[p];
    if (f2 < 0 || f2 >= J.set2->nframes) { h->err = "target frame index out of range"; return ORBM_ERR_ARG; }
    P.tframe = f2; P.tbase = J.set2->frame_off[f2]; P.nt = J.set2->frame_off[f2 + 1] - P.tbase;
    if (points) {
      const orbm_points& pt = J.points[p];
      if (pt.n < 0 || (pt.n > 0 && (!pt.u || !pt.v || !pt.radius || !pt.min_level || !pt.max_level || !pt.desc)) ||
          (pt.n > 0 && J.kind == K_PROJ_BEST && J.check_ori && !pt.angle)) {
        h->err = "malformed points";
        return ORBM_ERR_ARG;
      }
      P.f1 = -1; P.nq = P.nsrc = pt.n; P.qsrc = (int)ptot;
      ptot += pt.n;
    } else {
      const int f1 = J.idx1[p];
      if (f1 < 0 || f1 >= J.set1->nframes) { h->err = "source frame index out of range"; return ORBM_ERR_ARG; }
      P.f1 = f1; P.qsrc = J.set1->frame_off[f1]; P.nsrc = J.set1->frame_off[f1 + 1] - P.qsrc;
      P.nq = bow ? J.set1->fvitem_off[f1 + 1] - J.set1->fvitem_off[f1] : P.nsrc;
    }
    if ((per_source ? P.nsrc : P.nt) > J.stride) { h->err = "stride smaller than a frame"; return ORBM_ERR_ARG; }
    if ((J.kind == K_INIT || J.kind == K_PROJ_FF) && P.nsrc > J.stride) { h->err = "stride smaller than a frame"; return ORBM_ERR_ARG; }
    P.qstart = (int)qtot; P.tstart = (int)ttot; P.cand_base = -1;
    qtot += P.nq; ttot += P.nt;
    max_nq = std::max(max_nq, P.nq); max_nt = std::max(max_nt, P.nt);
    if (qtot > INT_MAX / 2 || ttot > INT_MAX / 2) { h->err = "batch too large"; return ORBM_ERR_UNSUPPORTED; }
  }

  // ---- device workspaces ----------------------------------------------------------------------------------------
  ORBM_CK(h->ws_problem.reserve(sizeof(Problem) * np));
  ORBM_CK(h->ws_query.reserve(sizeof(QRec) * std::max<int64_t>(qtot, 1)));
  ORBM_CK(h->ws_target.reserve(sizeof(int2) * std::max<int64_t>(ttot, 1)));
  const size_t xy_bytes = (J.kind == K_INIT || J.kind == K_PROJ_FF) ? (size_t)np * J.stride * 8 : 0;
  const size_t out_bytes = (size_t)np * J.stride * 4;
  size_t io = padded(out_bytes) + padded((size_t)np * 4) + padded(xy_bytes) + padded((size_t)np * 36) + padded(64 * 4);
  if (points) io += padded(ptot * 32) + 6 * padded(ptot * 4) + padded(ptot);
  ORBM_CK(h->ws_io.reserve(io + 1024));
  uint8_t* cur = h->ws_io.as<uint8_t>();
  int32_t* d_out = carve<int32_t>(cur, (size_t)np * J.stride);
  int32_t* d_nm = carve<int32_t>(cur, np);
  float* d_xy = xy_bytes ? carve<float>(cur, (size_t)np * J.stride * 2) : nullptr;
  float* d_F12 = carve<float>(cur, (size_t)np * 9);
  float* d_sigma = carve<float>(cur, 64);
  // Small jobs: every host input (problem table, prev_xy / uv rows, F12, sigma2, the point arrays) is gathered in pinned
  // staging that mirrors the device layout and goes up in TWO copies instead of up to a dozen pageable ones; the
  // first 256 KB of the staging area receive the results (see the read-back below).
  constexpr size_t kStageOut = 256u << 10, kStageProb = 32u << 10, kStageTotal = 1u << 20;
  uint8_t* const in_base = reinterpret_cast<uint8_t*>(d_xy ? d_xy : d_F12);
  const size_t in_span = (h->ws_io.as<uint8_t>() + io) - in_base;
  const bool staged = in_span <= kStageTotal - kStageOut - kStageProb && sizeof(Problem) * (size_t)np <= kStageProb;
  static_assert(kStageTotal == kOrbmPinnedBytes, "staging layout");
  ORBM_CK(orbm_ensure_pinned(h));
  uint8_t* const stg_prob = static_cast<uint8_t*>(h->pinned) + kStageOut;
  uint8_t* const stg_in = stg_prob + kStageProb;
  auto h2d = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
    if (!bytes) return cudaSuccess;
    if (staged) { memcpy(stg_in + (static_cast<uint8_t*>(dst) - in_base), src, bytes); return cudaSuccess; }
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
  };

  SearchArgs a;
  memset(&a, 0, sizeof(a));
  a.kind = J.kind; a.nproblems = np;
  a.prob = h->ws_problem.as<Problem>(); a.q = h->ws_query.as<QRec>(); a.tstate = h->ws_target.as<int2>();
  a.cursor = h->ws_misc.as<unsigned long long>() + 2;  // ws_misc[0] is the knn error flag
  a.s2 = view_of(J.set2);
  if (J.set1) a.s1 = view_of(J.set1);
  a.stride = J.stride; a.window = J.window; a.min_level = J.min_level; a.max_level = J.max_level;
  a.check_ori = J.check_ori; a.th_dist = J.th_dist; a.ratio = J.ratio; a.nlevels = J.nlevels;
  a.out = d_out; a.nmatches = d_nm; a.qxy = d_xy; a.out_xy = d_xy; a.F12 = d_F12; a.sigma2 = d_sigma;

  if (staged) {
    memcpy(stg_prob, prob.data(), sizeof(Problem) * np);
    ORBM_CK(cudaMemcpyAsync(a.prob, stg_prob, sizeof(Problem) * np, cudaMemcpyHostToDevice, st));
  } else {
    ORBM_CK(cudaMemcpyAsync(a.prob, prob.data(), sizeof(Problem) * np, cudaMemcpyHostToDevice, st));
  }
  if (J.kind == K_PROJ_FF) ORBM_CK(h2d(d_xy, J.xy_in, xy_bytes));
  if (J.kind == K_INIT) ORBM_CK(h2d(d_xy, J.xy_inout, xy_bytes));
  if (J.kind == K_TRIANG) {
    if (!J.F12 || !J.sigma2 || J.nlevels < 1 || J.nlevels > 64) { h->err = "bad F12 / sigma2"; return ORBM_ERR_ARG; }
    ORBM_CK(h2d(d_F12, J.F12, (size_t)np * 36));
    ORBM_CK(h2d(d_sigma, J.sigma2, (size_t)J.nlevels * 4));
  }
  std::vector<uint8_t> pvalid;  // must outlive the async copies below (synchronised before return)
  if (points) {
    uint8_t* d_desc = carve<uint8_t>(cur, (size_t)ptot * 32);
    float* d_u = carve<float>(cur, ptot); float* d_v = carve<float>(cur, ptot); float* d_r = carve<float>(cur, ptot);
    float* d_ang = carve<float>(cur, ptot);
    int32_t* d_minl = carve<int32_t>(cur, ptot); int32_t* d_maxl = carve<int32_t>(cur, ptot);
    uint8_t* d_valid = carve<uint8_t>(cur, ptot);
    pvalid.assign(std::max<int64_t>(ptot, 1), 1);
    // A large batch (thousands of problems, seven caller arrays each) is gathered on the host into one pinned block
    // that mirrors the device layout and goes up in ONE copy instead of ~7 pageable copies per problem.
    uint8_t* const pts_base = d_desc;
    const size_t pts_span = (size_t)(cur - pts_base);
    uint8_t* gather = nullptr;
    if (!staged && ptot) {
      if (h->pinned_big_cap < pts_span) {
        if (h->pinned_big) cudaFreeHost(h->pinned_big);
        h->pinned_big = nullptr; h->pinned_big_cap = 0;
        const size_t want = pts_span + pts_span / 2;
        ORBM_CK(cudaHostAlloc(&h->pinned_big, want, cudaHostAllocDefault));
        h->pinned_big_cap = want;
      }
      gather = static_cast<uint8_t*>(h->pinned_big);
      memset(gather, 0, pts_span);
    }
    auto put = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
      if (gather) { memcpy(gather + (static_cast<uint8_t*>(dst) - pts_base), src, bytes); return cudaSuccess; }
      return h2d(dst, src, bytes);
    };
    for (int p = 0; p < np; p++) {
      const orbm_points& pt = J.points[p];
      const size_t o = prob[p].qsrc, n = pt.n;
      if (!n) continue;
      ORBM_CK(put(d_desc + o * 32, pt.desc, n * 32));
      ORBM_CK(put(d_u + o, pt.u, n * 4));
      ORBM_CK(put(d_v + o, pt.v, n * 4));
      ORBM_CK(put(d_r + o, pt.radius, n * 4));
      ORBM_CK(put(d_minl + o, pt.min_level, n * 4));
      ORBM_CK(put(d_maxl + o, pt.max_level, n * 4));
      if (pt.angle) ORBM_CK(put(d_ang + o, pt.angle, n * 4));
      if (pt.valid) memcpy(&pvalid[o], pt.valid, n);
    }
    if (ptot) ORBM_CK(put(d_valid, pvalid.data(), ptot));
    if (gather) ORBM_CK(cudaMemcpyAsync(pts_base, gather, pts_span, cudaMemcpyHostToDevice, st));
    a.src.desc = reinterpret_cast<const uint4*>(d_desc); a.src.angle = d_ang; a.src.valid = d_valid;
    a.pu = d_u; a.pv = d_v; a.pr = d_r; a.pminl = d_minl; a.pmaxl = d_maxl;
  } else {
    a.src.desc = a.s1.desc; a.src.angle = a.s1.angle; a.src.x = a.s1.x; a.src.y = a.s1.y;
    a.src.octave = a.s1.octave; a.src.valid = a.s1.valid;
  }

  if (staged) ORBM_CK(cudaMemcpyAsync(in_base, stg_in, in_span, cudaMemcpyHostToDevice, st));

  // ---- kernels (rerun with a larger arena if the candidate lists do not fit) ---------------------------------------
  if (h->ws_cand.cap == 0) ORBM_CK(h->ws_cand.reserve(std::max<size_t>((size_t)qtot * 64 * 4, 4u << 20)));
  for (int attempt = 0; attempt < 3; attempt++) {
    a.cand = h->ws_cand.as<uint32_t>();
    a.cand_cap = (int64_t)(h->ws_cand.cap / 4);
    ORBM_CK(cudaMemsetAsync(a.cursor, 0, 16, st));
    if (J.kind == K_INIT && (attempt > 0 || !staged))  // emit updates prev_xy in place: a rerun needs the caller's values again
      ORBM_CK(cudaMemcpyAsync(d_xy, staged ? static_cast<const void*>(stg_in + (reinterpret_cast<uint8_t*>(d_xy) - in_base))
                                           : static_cast<const void*>(J.xy_inout), xy_bytes, cudaMemcpyHostToDevice, st));
    ORBM_CK(cudaEventRecord(h->ev0, st));
    const bool pdl = np <= 512;   // the latency regime (see the commit choice below)
    if (ttot) launch_k(pdl, target_init_kernel, dim3((unsigned)((ttot + 255) / 256)), dim3(256), 0, st, a, (int)ttot);
    if (qtot) {
      if (bow) launch_k(pdl, bow_prep_kernel, dim3((np * 32 + 127) / 128), dim3(128), 0, st, a);
      else launch_k(pdl, grid_prep_kernel, dim3((unsigned)((qtot + 255) / 256)), dim3(256), 0, st, a, (int)qtot);
    }
    launch_k(pdl, offsets_kernel, dim3((np * 32 + 127) / 128), dim3(128), 0, st, a);
    if (qtot) {
      const unsigned blocks = (unsigned)((qtot * 32 + 255) / 256);
      if (bow) launch_k(pdl, bow_fill_kernel, dim3(blocks), dim3(256), 0, st, a, (int)qtot);
      else launch_k(pdl, grid_fill_kernel, dim3(blocks), dim3(256), 0, st, a, (int)qtot);
    }
    // commit: the parallel fixed-point kernel (CTA per problem) for small jobs whose state fits in shared memory;
    // the in-order replay (warp per problem) for large jobs and oversized problems.
    // ORBM_COMMIT=serial|parallel overrides (experiments / tests).
    const size_t par_smem = commit_par_smem(max_nq, max_nt);
    const bool par_ok = par_smem <= 200 * 1024;
    // crossover measured with SearchForInitialization on 1000-keypoint pairs: 0.27 vs 0.33 ms at 256 pairs, 0.43 vs
    // 0.46 at 512, 0.86 vs 0.73 at 1024 (parallel vs replay): the replay's chain is latency, the fixed point is work
    const bool want_par = h->commit_mode ? h->commit_mode == 2 : np <= 512;
    if (par_ok && want_par) {
      const int threads = np >= 148 * 4 ? 256 : (np >= 148 ? 512 : 1024);
      cudaFuncSetAttribute(commit_par_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)par_smem);
      launch_k(pdl, commit_par_kernel, dim3(np), dim3(threads), par_smem, st, a, max_nq, max_nt);
    } else {
      cudaFuncSetAttribute(commit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kCommitSmemBytes);
      launch_k(pdl, commit_kernel, dim3((np + kCommitWarps - 1) / kCommitWarps), dim3(32 * kCommitWarps), (size_t)kCommitSmemBytes, st, a);
    }
    launch_k(pdl, emit_kernel, dim3(np), dim3(128), 0, st, a);
    ORBM_CK(cudaEventRecord(h->ev1, st));
    h->launches += 3 + (ttot ? 1 : 0) + (qtot ? 2 : 0);
    ORBM_CK(cudaGetLastError());
    // Small jobs (the per-frame calls of a tracker) are latency bound: results and the overflow cursor travel
    // together into pinned staging and ONE synchronisation ends the call; on overflow the results are simply not
    // used.  Large jobs keep the two-step form (no extra host copy of megabytes of rows).
    const size_t small_bytes = 64 + padded(out_bytes) + padded((size_t)np * 4) + padded(xy_bytes);
    const bool small = small_bytes <= (256u << 10);
    if (small) {
      uint8_t* hp = static_cast<uint8_t*>(h->pinned);
      uint8_t* h_out = hp + 64; uint8_t* h_nm = h_out + padded(out_bytes); uint8_t* h_xy = h_nm + padded((size_t)np * 4);
      ORBM_CK(cudaMemcpyAsync(hp, a.cursor, 16, cudaMemcpyDeviceToHost, st));
      ORBM_CK(cudaMemcpyAsync(h_out, d_out, out_bytes, cudaMemcpyDeviceToHost, st));
      ORBM_CK(cudaMemcpyAsync(h_nm, d_nm, (size_t)np * 4, cudaMemcpyDeviceToHost, st));
      if (J.kind == K_INIT) ORBM_CK(cudaMemcpyAsync(h_xy, d_xy, xy_bytes, cudaMemcpyDeviceToHost, st));
      ORBM_CK(cudaStreamSynchronize(st));
      unsigned long long cursor[2];
      memcpy(cursor, hp, 16);
      if (!cursor[1]) {
        memcpy(J.out, h_out, out_bytes);
        memcpy(J.nmatches, h_nm, (size_t)np * 4);
        if (J.kind == K_INIT) memcpy(J.xy_inout, h_xy, xy_bytes);
        cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
        return ORBM_OK;
      }
      if (attempt == 2) { h->err = "candidate arena overflow"; return ORBM_ERR_OVERFLOW; }
      ORBM_CK(h->ws_cand.reserve((size_t)cursor[0] * 4 + 1024));
      continue;
    }
    unsigned long long cursor[2] = {0, 0};
    ORBM_CK(cudaMemcpyAsync(cursor, a.cursor, 16, cudaMemcpyDeviceToHost, st));
    ORBM_CK(cudaStreamSynchronize(st));
    if (!cursor[1]) break;
    if (attempt == 2) { h->err = "candidate arena overflow"; return ORBM_ERR_OVERFLOW; }
    ORBM_CK(h->ws_cand.reserve((size_t)cursor[0] * 4 + 1024));
  }
  cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
  ORBM_CK(cudaMemcpyAsync(J.out, d_out, out_bytes, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(J.nmatches, d_nm, (size_t)np * 4, cudaMemcpyDeviceToHost, st));
  if (J.kind == K_INIT) ORBM_CK(cudaMemcpyAsync(J.xy_inout, d_xy, xy_bytes, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaStreamSynchronize(st));
  return ORBM_OK;
}

}  // namespace

extern "C" int orbm_search_for_initialization(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2,
                                              const int32_t* idx1, const int32_t* idx2, int npairs, float* prev_xy,
                                              int stride, int window, float ratio, int check_ori, int32_t* matches12,
                                              int32_t* nmatches) {
  if (h && !prev_xy) { h->err = "prev_xy is null"; return ORBM_ERR_ARG; }
  HostJob J;
  J.kind = K_INIT; J.set1 = set1; J.set2 = set2; J.idx1 = idx1; J.idx2 = idx2; J.nproblems = npairs;
  J.xy_inout = prev_xy; J.stride = stride; J.window = window; J.ratio = ratio; J.check_ori = check_ori;
  J.out = matches12; J.nmatches = nmatches;
  return run_job(h, J);
}

extern "C" int orbm_window_search(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2,
                                  const int32_t* idx1, const int32_t* idx2, int npairs, int window, float ratio,
                                  int check_ori, int min_level, int max_level, int32_t* matches21, int stride,
                                  int32_t* nmatches) {
  HostJob J;
  J.kind = K_WINDOW; J.set1 = set1; J.set2 = set2; J.idx1 = idx1; J.idx2 = idx2; J.nproblems = npairs;
  J.stride = stride; J.window = window; J.ratio = ratio; J.check_ori = check_ori; J.min_level = min_level;
  J.max_level = max_level; J.out = matches21; J.nmatches = nmatches;
  return run_job(h, J);
}

extern "C" int orbm_search_by_projection_frames(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2,
                                                const int32_t* idx1, const int32_t* idx2, int npairs, const float* uv,
                                                int stride, int window, float ratio, int32_t* matches21,
                                                int32_t* nmatches) {
  if (h && !uv) { h->err = "uv is null"; return ORBM_ERR_ARG; }
  HostJob J;
  J.kind = K_PROJ_FF; J.set1 = set1; J.set2 = set2; J.idx1 = idx1; J.idx2 = idx2; J.nproblems = npairs;
  J.xy_in = uv; J.stride = stride; J.window = window; J.ratio = ratio; J.out = matches21; J.nmatches = nmatches;
  return run_job(h, J);
}

extern "C" int orbm_search_by_projection(orbm_handle* h, const orbm_frameset* set, const int32_t* idx,
                                         const orbm_points* points, int nproblems, int variant, int th_dist,
                                         float ratio, int check_ori, int32_t* match_out, int stride,
                                         int32_t* nmatches) {
  if (h && variant != ORBM_PROJ_LOCAL_MAP && variant != ORBM_PROJ_BEST && variant != ORBM_PROJ_BEST_INDEPENDENT) {
    h->err = "unknown variant";
    return ORBM_ERR_ARG;
  }
  HostJob J;
  J.kind = variant == ORBM_PROJ_LOCAL_MAP ? K_PROJ_LOCAL : (variant == ORBM_PROJ_BEST ? K_PROJ_BEST : K_PROJ_INDEP);
  J.set2 = set; J.idx2 = idx; J.points = points; J.nproblems = nproblems; J.th_dist = th_dist; J.ratio = ratio;
  J.check_ori = J.kind == K_PROJ_INDEP ? 0 : check_ori;   // Fuse / SearchBySim3 have no rotation check
  J.out = match_out; J.stride = stride; J.nmatches = nmatches;
  return run_job(h, J);
}

extern "C" int orbm_search_by_sim3(orbm_handle* h, const orbm_frameset* set, const int32_t* idx1, const int32_t* idx2,
                                   const orbm_points* points12, const orbm_points* points21, int npairs, int th_dist,
                                   int32_t* match12, int stride, int32_t* nfound) {
  if (!h) return ORBM_ERR_ARG;
  if (!set || !idx1 || !idx2 || !points12 || !points21 || !match12 || !nfound || npairs < 0 || stride < 1) {
    h->err = "bad SearchBySim3 arguments";
    return ORBM_ERR_ARG;
  }
  if (npairs == 0) return ORBM_OK;
  // problem 2p: keyframe 1's points scored in keyframe 2 (vnMatch1); problem 2p + 1: keyframe 2's points in keyframe 1
  std::vector<int32_t> tgt(2 * (size_t)npairs), nm(2 * (size_t)npairs);
  std::vector<orbm_points> pts(2 * (size_t)npairs);
  int row = stride;
  for (int p = 0; p < npairs; p++) {
    if (idx1[p] < 0 || idx1[p] >= set->nframes || idx2[p] < 0 || idx2[p] >= set->nframes) { h->err = "frame index out of range"; return ORBM_ERR_ARG; }
    const int n1 = set->frame_off[idx1[p] + 1] - set->frame_off[idx1[p]], n2 = set->frame_off[idx2[p] + 1] - set->frame_off[idx2[p]];
    if (points12[p].n != n1 || points21[p].n != n2) { h->err = "SearchBySim3: one point per keypoint of the source keyframe expected"; return ORBM_ERR_ARG; }
    if (n1 > stride) { h->err = "stride smaller than a frame"; return ORBM_ERR_ARG; }
    row = std::max(row, std::max(n1, n2));
    tgt[2 * p] = idx2[p]; tgt[2 * p + 1] = idx1[p];
    pts[2 * p] = points12[p]; pts[2 * p + 1] = points21[p];
  }
  std::vector<int32_t> best((size_t)2 * npairs * row);
  const int rc = orbm_search_by_projection(h, set, tgt.data(), pts.data(), 2 * npairs, ORBM_PROJ_BEST_INDEPENDENT, th_dist,
                                           0.f, 0, best.data(), row, nm.data());
  if (rc != ORBM_OK) return rc;
  for (int p = 0; p < npairs; p++) {          // agreement, :1487-1502
    const int32_t* m1 = &best[(size_t)(2 * p) * row];
    const int32_t* m2 = &best[(size_t)(2 * p + 1) * row];
    int32_t* out = match12 + (size_t)p * stride;
    const int n1 = points12[p].n;
    int found = 0;
    for (int i1 = 0; i1 < stride; i1++) {
      int32_t r = -1;
      if (i1 < n1 && m1[i1] >= 0 && m2[m1[i1]] == i1) { r = m1[i1]; found++; }
      out[i1] = r;
    }
    nfound[p] = found;
  }
  return ORBM_OK;
}

extern "C" int orbm_search_by_bow(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2,
                                  const int32_t* idx1, const int32_t* idx2, int npairs, int variant, float ratio,
                                  int check_ori, int32_t* match_out, int stride, int32_t* nmatches) {
  if (h && variant != ORBM_BOW_KF_F && variant != ORBM_BOW_KF_KF) { h->err = "unknown variant"; return ORBM_ERR_ARG; }
  HostJob J;
  J.kind = variant == ORBM_BOW_KF_F ? K_BOW_KF_F : K_BOW_KF_KF;
  J.set1 = set1; J.set2 = set2; J.idx1 = idx1; J.idx2 = idx2; J.nproblems = npairs; J.ratio = ratio;
  J.check_ori = check_ori; J.out = match_out; J.stride = stride; J.nmatches = nmatches;
  return run_job(h, J);
}

extern "C" int orbm_search_for_triangulation(orbm_handle* h, const orbm_frameset* set1, const orbm_frameset* set2,
                                             const int32_t* idx1, const int32_t* idx2, int npairs, const float* F12,
                                             const float* sigma2, int nlevels, int check_ori, int32_t* match12,
                                             int stride, int32_t* nmatches) {
  HostJob J;
  J.kind = K_TRIANG; J.set1 = set1; J.set2 = set2; J.idx1 = idx1; J.idx2 = idx2; J.nproblems = npairs;
  J.F12 = F12; J.sigma2 = sigma2; J.nlevels = nlevels; J.check_ori = check_ori; J.out = match12; J.stride = stride;
  J.nmatches = nmatches;
  return run_job(h, J);
}
