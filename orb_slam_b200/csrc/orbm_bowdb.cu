// N4, second half: the candidate scoring of KeyFrameDatabase::DetectRelocalisationCandidates
// (reference src/KeyFrameDatabase.cc:198-308) and ::DetectLoopCandidates (:75-196) for a batch of query frames against
// a resident set of keyframes, both with BowVectors from orbm_frameset_compute_bow:
//   inverted file   word -> keyframes holding it (KeyFrameDatabase::add, :40-46), built once per keyframe set as a CSR by
//                   a counting sort over word ids
//   shared words    mnRelocWords / mnLoopWords (:211-224, :86-104): for every word of the query, every keyframe of its
//                   inverted list gets +1 (integer atomics: exact); keyframes the caller excludes (the query's
//                   covisibility neighbours, :95) are not candidates
//   cut             maxCommonWords over the candidates, minCommonWords = (int)(maxCommonWords * 0.8f) (:231-238)
//   score           mpVoc->score(query, keyframe) for candidates with words > minCommonWords (:247-255): L1Scoring::score,
//                   the double sum in ascending word order (bit-exact), returned as the float the reference stores
// What follows in the reference - accumulating scores over GetBestCovisibilityKeyFrames(10) and the 0.75 * best cut
// (:262-305) - walks the covisibility graph (std::map / std::list of KeyFrame*) and stays with the host.
#include <algorithm>
#include <cstring>

#include "orbm_internal.cuh"

using namespace orbm;

struct orbm_bowdb {
  orbm_handle* h = nullptr;
  const orbm_frameset* set = nullptr;
  int nkf = 0, nwords = 0;          // keyframes; largest word id + 1
  int64_t total = 0;                // (word, keyframe) pairs
  int32_t* d_cnt = nullptr;         // [nkf] words per keyframe
  int32_t* d_start = nullptr;       // [nwords + 1] inverted-file CSR
  int32_t* d_kf = nullptr;          // [total]
};

namespace {

__global__ void __launch_bounds__(256) bowdb_maxword_kernel(const int32_t* frame_off, const int32_t* cnt, const int32_t* word,
                                                            int nkf, int* out) {
  const int f = blockIdx.x;
  const int n = cnt[f];
  if (threadIdx.x == 0 && n > 0) atomicMax(out, word[frame_off[f] + n - 1]);   // ascending ids: the last is the largest
}

__global__ void __launch_bounds__(256) bowdb_hist_kernel(const int32_t* frame_off, const int32_t* cnt, const int32_t* word,
                                                         int32_t* hist) {
  const int f = blockIdx.x;
  const int32_t* w = word + frame_off[f];
  for (int i = threadIdx.x; i < cnt[f]; i += blockDim.x) atomicAdd(&hist[w[i]], 1);
}

// exclusive scan of n + 1 entries in place by one CTA (n ~ 10^6 words: a few microseconds of work, run once per database)
__global__ void __launch_bounds__(1024) bowdb_scan_kernel(int32_t* a, int n1) {
  __shared__ int s_part[1024];
  const int tid = threadIdx.x, per = (n1 + 1023) / 1024;
  const int lo = min(tid * per, n1), hi = min(lo + per, n1);
  int sum = 0;
  for (int i = lo; i < hi; i++) sum += a[i];
  s_part[tid] = sum;
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int i = 0; i < 1024; i++) { const int v = s_part[i]; s_part[i] = run; run += v; }
  }
  __syncthreads();
  int run = s_part[tid];
  for (int i = lo; i < hi; i++) { const int v = a[i]; a[i] = run; run += v; }
}

__global__ void __launch_bounds__(256) bowdb_fill_kernel(const int32_t* frame_off, const int32_t* cnt, const int32_t* word,
                                                         const int32_t* start, int32_t* cursor, int32_t* kf) {
  const int f = blockIdx.x;
  const int32_t* w = word + frame_off[f];
  for (int i = threadIdx.x; i < cnt[f]; i += blockDim.x) kf[start[w[i]] + atomicAdd(&cursor[w[i]], 1)] = f;
}

// +1 for every (query word, keyframe of its inverted list)
__global__ void __launch_bounds__(256) bowdb_count_kernel(const int32_t* q_off, const int32_t* q_cnt, const int32_t* q_word,
                                                          const int32_t* qidx, const int32_t* start, const int32_t* kf,
                                                          int nwords, int nkf, int32_t* common) {
  const int q = blockIdx.y, f = qidx[q];
  const int32_t* w = q_word + q_off[f];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= q_cnt[f]) return;
  const int word = w[i];
  if (word >= nwords) return;                                  // a word no keyframe holds
  int32_t* row = common + (int64_t)q * nkf;
  for (int k = start[word]; k < start[word + 1]; k++) atomicAdd(&row[kf[k]], 1);
}

__global__ void __launch_bounds__(256) bowdb_cut_kernel(const int32_t* common, const uint8_t* exclude, int nkf,
                                                        int32_t* max_common, int32_t* min_common) {
  __shared__ int s_max;
  const int q = blockIdx.x;
  if (threadIdx.x == 0) s_max = 0;
  __syncthreads();
  int m = 0;
  for (int k = threadIdx.x; k < nkf; k += blockDim.x)
    if (!exclude || !exclude[(int64_t)q * nkf + k]) m = max(m, common[(int64_t)q * nkf + k]);
  m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0) atomicMax(&s_max, m);
  __syncthreads();
  if (threadIdx.x == 0) {
    max_common[q] = s_max;
    min_common[q] = (int)__fmul_rn((float)s_max, 0.8f);       // int minCommonWords = maxCommonWords*0.8f;
  }
}

// L1Scoring::score for the candidates above the cut; one thread per (query, keyframe), the sum in ascending word order
__global__ void __launch_bounds__(128) bowdb_score_kernel(const int32_t* q_off, const int32_t* q_cnt, const int32_t* q_word,
                                                          const double* q_val, const int32_t* qidx, const int32_t* d_off,
                                                          const int32_t* d_cnt, const int32_t* d_word, const double* d_val,
                                                          const int32_t* common, const uint8_t* exclude,
                                                          const int32_t* min_common, int nkf, float* score) {
  const int q = blockIdx.y, k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nkf) return;
  const int64_t o = (int64_t)q * nkf + k;
  if ((exclude && exclude[o]) || common[o] <= min_common[q]) { score[o] = -1.f; return; }
  const int f = qidx[q];
  const int32_t *a = q_word + q_off[f], *b = d_word + d_off[k];
  const double *va = q_val + q_off[f], *vb = d_val + d_off[k];
  const int n1 = q_cnt[f], n2 = d_cnt[k];
  double s = 0.0;
  int i = 0, j = 0;
  while (i < n1 && j < n2) {
    const int x = a[i], y = b[j];
    if (x == y) {
      const double vi = va[i], wi = vb[j];
      s = __dadd_rn(s, __dsub_rn(__dsub_rn(fabs(__dsub_rn(vi, wi)), fabs(vi)), fabs(wi)));
      i++; j++;
    } else if (x < y) i++;
    else j++;
  }
  score[o] = (float)__ddiv_rn(-s, 2.0);                       // float si = mpVoc->score(...)
}

}  // namespace

extern "C" int orbm_bowdb_create(orbm_handle* h, const orbm_frameset* keyframes, orbm_bowdb** out) {
  if (!h || !out) return ORBM_ERR_ARG;
  *out = nullptr;
  if (!keyframes || keyframes->h != h || keyframes->bow_cnt.empty()) {
    h->err = "bowdb: the keyframe set has no BowVectors (orbm_frameset_compute_bow) or belongs to another handle";
    return ORBM_ERR_ARG;
  }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  orbm_bowdb* d = new orbm_bowdb();
  d->h = h; d->set = keyframes; d->nkf = keyframes->nframes;
  for (int c : keyframes->bow_cnt) d->total += c;
  auto fail = [&](int rc) { cudaFree(d->d_cnt); cudaFree(d->d_start); cudaFree(d->d_kf); delete d; return rc; };
  auto ck = [&](cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    h->err = std::string(what) + ": " + cudaGetErrorString(e);
    return false;
  };
  int* d_max = nullptr;
  if (!ck(cudaMalloc(&d->d_cnt, (size_t)d->nkf * 4 + 16), "cudaMalloc")) return fail(ORBM_ERR_CUDA);
  if (!ck(cudaMemcpyAsync(d->d_cnt, keyframes->bow_cnt.data(), (size_t)d->nkf * 4, cudaMemcpyHostToDevice, st), "copy")) return fail(ORBM_ERR_CUDA);
  if (!ck(cudaMalloc(&d_max, 4), "cudaMalloc")) return fail(ORBM_ERR_CUDA);
  cudaMemsetAsync(d_max, 0xFF, 4, st);                         // -1
  if (d->nkf) bowdb_maxword_kernel<<<d->nkf, 32, 0, st>>>(keyframes->d_frame_off, d->d_cnt, keyframes->d_bow_word, d->nkf, d_max);
  int maxw = -1;
  cudaMemcpyAsync(&maxw, d_max, 4, cudaMemcpyDeviceToHost, st);
  if (!ck(cudaStreamSynchronize(st), "bowdb max word")) { cudaFree(d_max); return fail(ORBM_ERR_CUDA); }
  cudaFree(d_max);
  d->nwords = maxw + 1;
  int32_t* d_cursor = nullptr;
  if (!ck(cudaMalloc(&d->d_start, ((size_t)d->nwords + 2) * 4), "cudaMalloc") ||
      !ck(cudaMalloc(&d->d_kf, (size_t)std::max<int64_t>(d->total, 1) * 4), "cudaMalloc") ||
      !ck(cudaMalloc(&d_cursor, ((size_t)d->nwords + 2) * 4), "cudaMalloc")) { cudaFree(d_cursor); return fail(ORBM_ERR_CUDA); }
  cudaMemsetAsync(d->d_start, 0, ((size_t)d->nwords + 2) * 4, st);
  cudaMemsetAsync(d_cursor, 0, ((size_t)d->nwords + 2) * 4, st);
  if (d->nkf && d->total) {
    bowdb_hist_kernel<<<d->nkf, 256, 0, st>>>(keyframes->d_frame_off, d->d_cnt, keyframes->d_bow_word, d->d_start);
    bowdb_scan_kernel<<<1, 1024, 0, st>>>(d->d_start, d->nwords + 1);
    bowdb_fill_kernel<<<d->nkf, 256, 0, st>>>(keyframes->d_frame_off, d->d_cnt, keyframes->d_bow_word, d->d_start, d_cursor, d->d_kf);
    h->launches += 4;
  }
  const bool ok = ck(cudaGetLastError(), "bowdb kernels") && ck(cudaStreamSynchronize(st), "bowdb build");
  cudaFree(d_cursor);
  if (!ok) return fail(ORBM_ERR_CUDA);
  *out = d;
  return ORBM_OK;
}

extern "C" void orbm_bowdb_destroy(orbm_bowdb* d) {
  if (!d) return;
  cudaSetDevice(d->h->device);
  cudaStreamSynchronize(d->h->stream);
  cudaFree(d->d_cnt); cudaFree(d->d_start); cudaFree(d->d_kf);
  delete d;
}

extern "C" int orbm_bowdb_query(orbm_bowdb* d, const orbm_frameset* queries, const int32_t* qidx, int nq,
                                const uint8_t* exclude, int32_t* common, int32_t* max_common, int32_t* min_common,
                                float* score) {
  if (!d) return ORBM_ERR_ARG;
  orbm_handle* h = d->h;
  if (!queries || queries->h != h || queries->bow_cnt.empty() || !qidx || nq < 0 || !common || !max_common || !min_common || !score) {
    h->err = "bowdb query: null buffer, or the query set has no BowVectors / belongs to another handle";
    return ORBM_ERR_ARG;
  }
  if (nq == 0) return ORBM_OK;
  int max_words = 0;
  for (int q = 0; q < nq; q++) {
    if (qidx[q] < 0 || qidx[q] >= queries->nframes) { h->err = "bowdb query: frame index out of range"; return ORBM_ERR_ARG; }
    max_words = std::max(max_words, queries->bow_cnt[qidx[q]]);
  }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  const int nkf = d->nkf;
  const size_t cells = (size_t)nq * std::max(nkf, 1);
  auto pad = [](size_t b) { return (b + 255) / 256 * 256; };
  const size_t o_common = 0, o_score = o_common + pad(cells * 4), o_excl = o_score + pad(cells * 4),
               o_qidx = o_excl + pad(exclude ? cells : 0), o_qcnt = o_qidx + pad((size_t)nq * 4),
               o_max = o_qcnt + pad((size_t)queries->nframes * 4), o_min = o_max + pad((size_t)nq * 4),
               total = o_min + pad((size_t)nq * 4);
  ORBM_CK(h->ws_io.reserve(total + 256));
  uint8_t* sb = h->ws_io.as<uint8_t>();
  int32_t* d_common = reinterpret_cast<int32_t*>(sb + o_common);
  float* d_score = reinterpret_cast<float*>(sb + o_score);
  uint8_t* d_excl = exclude ? sb + o_excl : nullptr;
  int32_t* d_qidx = reinterpret_cast<int32_t*>(sb + o_qidx);
  int32_t* d_qcnt = reinterpret_cast<int32_t*>(sb + o_qcnt);
  int32_t* d_max = reinterpret_cast<int32_t*>(sb + o_max);
  int32_t* d_min = reinterpret_cast<int32_t*>(sb + o_min);
  ORBM_CK(cudaMemsetAsync(d_common, 0, cells * 4, st));
  if (exclude) ORBM_CK(cudaMemcpyAsync(d_excl, exclude, cells, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(d_qidx, qidx, (size_t)nq * 4, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(d_qcnt, queries->bow_cnt.data(), (size_t)queries->nframes * 4, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaEventRecord(h->ev0, st));
  if (nkf && max_words && d->total)
    bowdb_count_kernel<<<dim3((max_words + 255) / 256, nq), 256, 0, st>>>(queries->d_frame_off, d_qcnt, queries->d_bow_word, d_qidx,
                                                                        d->d_start, d->d_kf, d->nwords, nkf, d_common);
  bowdb_cut_kernel<<<nq, 256, 0, st>>>(d_common, d_excl, nkf, d_max, d_min);
  if (nkf)
    bowdb_score_kernel<<<dim3((nkf + 127) / 128, nq), 128, 0, st>>>(queries->d_frame_off, d_qcnt, queries->d_bow_word,
                                                                   queries->d_bow_val, d_qidx, d->set->d_frame_off, d->d_cnt,
                                                                   d->set->d_bow_word, d->set->d_bow_val, d_common, d_excl, d_min,
                                                                   nkf, d_score);
  h->launches += 3;
  ORBM_CK(cudaGetLastError());
  ORBM_CK(cudaEventRecord(h->ev1, st));
  if (nkf) {
    ORBM_CK(cudaMemcpyAsync(common, d_common, cells * 4, cudaMemcpyDeviceToHost, st));
    ORBM_CK(cudaMemcpyAsync(score, d_score, cells * 4, cudaMemcpyDeviceToHost, st));
  }
  ORBM_CK(cudaMemcpyAsync(max_common, d_max, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(min_common, d_min, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaStreamSynchronize(st));
  cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
  return ORBM_OK;
}
