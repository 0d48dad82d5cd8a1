// N2: the BoW transform on the device.  include/orbm.h.
// Reference: Frame::ComputeBoW src/Frame.cc:279-286 -> TemplatedVocabulary::transform(features, BowVector,
// FeatureVector, levelsup) Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1194, the per-feature tree descent
// :1217-1259, FORB::distance FORB.cpp:81-101, BowVector::addWeight / addIfNotExist / normalize BowVector.cpp:34-85,
// FeatureVector::addFeature FeatureVector.cpp:31-47, L1Scoring::score ScoringObject.cpp:23-67, and the ORBvoc.txt
// loader TemplatedVocabulary.h:1337-1424.
//
// Layout in HBM.  The tree is renumbered breadth first so that the children of a node are CONSECUTIVE: one int2
// (first child, child count) per node and the children's 32-byte descriptors as one contiguous, 32-byte aligned run
// (k = 10 -> 320 B).  ORBvoc (k = 10, L = 6, 1.1 M nodes) is 36 MB of descriptors + 9 MB of node records: it lives in
// the 126 MB L2 after the first touch.  The reference's node / word ids are kept in side tables (orig, word) and only
// read at the end of a descent.
//
// Kernels:  bow_descend (one thread per descriptor, all frames of a set in one launch: L levels x k distances of
// 8 XOR + 8 POPC each, POPC-pipe bound), bow_frame (one CTA per frame: two shared-memory bitonic sorts of
// (id << 32 | feature) keys give the std::map orders, run heads become map entries, weights are summed in the
// reference's order with individually rounded double operations), bow_scan + bow_compact (exact-size CSR of the
// feature vectors, the layout orbm_search_by_bow reads).
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>

#include "orbm_internal.cuh"

using namespace orbm;

struct orbm_vocabulary {
  orbm_handle* h = nullptr;
  int k = 0, L = 0, scoring = 0, weighting = 0, nnodes = 0, nwords = 0;
  int2* d_info = nullptr;       // [nnodes] breadth-first numbering: (first child, number of children); 0 children = leaf
  uint4* d_desc = nullptr;      // [nnodes][2]
  int32_t* d_orig = nullptr;    // [nnodes] the reference's NodeId
  int32_t* d_word = nullptr;    // [nnodes] WordId of a leaf, -1 otherwise
  double* d_weight = nullptr;   // [nnodes]
};

namespace {

constexpr int kSortSmemKeys = 4096;   // 32 KB of keys; larger frames sort in a global scratch slice
constexpr int kFrameThreads = 256;

struct VocView {
  const int2* info;
  const uint4* desc;
  const int32_t *orig, *word;
  const double* weight;
};

// ------------------------------------------------------------------------------------------------------------------
// bow_descend: TemplatedVocabulary::transform(feature, word, weight, nid, levelsup) for every descriptor.
// kp_word = -1 for a stopped word (weight <= 0: the `if(w > 0)` of :1161 / :1186).
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bow_descend_kernel(const uint4* __restrict__ desc, int n, VocView v,
                                                          int nid_level, int32_t* __restrict__ kp_word,
                                                          int32_t* __restrict__ kp_node,
                                                          double* __restrict__ kp_weight) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 q0 = __ldg(desc + 2 * (int64_t)i), q1 = __ldg(desc + 2 * (int64_t)i + 1);
  int cur = 0, at = nid_level <= 0 ? 0 : -1, level = 0;
  for (;;) {
    const int2 nd = __ldg(v.info + cur);
    if (nd.y == 0) break;
    ++level;
    const uint4* c = v.desc + 2 * (int64_t)nd.x;
    int best = INT_MAX, bi = 0;
    int j = 0;
    for (; j + 2 <= nd.y; j += 2) {   // two children per round: four independent 16-byte loads in flight
      const uint4 a0 = __ldg(c + 2 * j), a1 = __ldg(c + 2 * j + 1), b0 = __ldg(c + 2 * j + 2), b1 = __ldg(c + 2 * j + 3);
      const int da = hamming256(q0, q1, a0, a1), db = hamming256(q0, q1, b0, b1);
      if (da < best) { best = da; bi = j; }
      if (db < best) { best = db; bi = j + 1; }
    }
    if (j < nd.y) {
      const int da = hamming256(q0, q1, __ldg(c + 2 * j), __ldg(c + 2 * j + 1));
      if (da < best) { best = da; bi = j; }
    }
    cur = nd.x + bi;
    if (level == nid_level) at = cur;
  }
  if (at < 0) at = cur;   // ragged tree: leaf above nid_level
  const double w = __ldg(v.weight + cur);
  kp_word[i] = w > 0.0 ? __ldg(v.word + cur) : -1;
  kp_node[i] = __ldg(v.orig + at);
  kp_weight[i] = w;
}

// ------------------------------------------------------------------------------------------------------------------
// bow_frame: one CTA per frame.
// ------------------------------------------------------------------------------------------------------------------
struct BowFrameArgs {
  const int32_t* frame_off;
  const int32_t *kp_word, *kp_node;
  const double* kp_weight;
  int accumulate;    // TF_IDF / TF: addWeight;  IDF / BINARY: addIfNotExist
  int norm;          // 0: none (then TF* divides by the number of words), 1: L1, 2: L2
  // feature vector, not yet compacted: nodes of frame f at frame_off[f], starts at frame_off[f] + f
  int32_t *t_node_id, *t_node_start, *t_items, *cnt_nodes, *cnt_items;
  // bow vector: frame_off[f]
  int32_t *bow_word, *bow_cnt;
  double* bow_val;
  unsigned long long* scratch;     // for frames with more than kSortSmemKeys keys (padded)
  const long long* scratch_off;    // [nframes], -1 = sort in shared memory
};

__device__ __forceinline__ void bitonic_sort(unsigned long long* keys, int npad) {
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < npad; i += blockDim.x) {
        const int p = i ^ j;
        if (p > i) {
          const unsigned long long a = keys[i], b = keys[p];
          if ((a > b) == ((i & k) == 0)) { keys[i] = b; keys[p] = a; }
        }
      }
      __syncthreads();
    }
  }
}

// rank of every run head among the first m sorted keys (block-wide, in index order); calls emit(i, rank) for heads
// and returns the number of runs
template <class F>
__device__ __forceinline__ int for_each_run_head(const unsigned long long* keys, int m, int* s_warp, F emit) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  int base = 0;
  for (int i0 = 0; i0 < m; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const bool head = i < m && (i == 0 || (uint32_t)(keys[i] >> 32) != (uint32_t)(keys[i - 1] >> 32));
    const unsigned bal = __ballot_sync(0xffffffffu, head);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int off = 0, tot = 0;
    for (int w = 0; w < nwarp; w++) {
      const int c = s_warp[w];
      if (w < warp) off += c;
      tot += c;
    }
    if (head) emit(i, base + off + __popc(bal & ((1u << lane) - 1)));
    base += tot;
    __syncthreads();
  }
  return base;
}

__global__ void __launch_bounds__(kFrameThreads) bow_frame_kernel(BowFrameArgs a) {
  extern __shared__ unsigned long long s_keys[];
  __shared__ int s_warp[kFrameThreads / 32];
  __shared__ int s_m;
  const int f = blockIdx.x, tid = threadIdx.x;
  const int base = a.frame_off[f], n = a.frame_off[f + 1] - base;
  int npad = 1;
  while (npad < n) npad <<= 1;
  unsigned long long* keys = a.scratch_off[f] < 0 ? s_keys : a.scratch + a.scratch_off[f];
  const int32_t* word = a.kp_word + base;
  const int32_t* node = a.kp_node + base;

  // ---- FeatureVector: sort by (node, feature); stopped words are left out -------------------------------------
  if (tid == 0) s_m = 0;
  __syncthreads();
  int mine = 0;
  for (int i = tid; i < npad; i += blockDim.x) {
    unsigned long long key = ~0ull;
    if (i < n && word[i] >= 0) { key = ((unsigned long long)(uint32_t)node[i] << 32) | (uint32_t)i; mine++; }
    keys[i] = key;
  }
  if (mine) atomicAdd(&s_m, mine);
  __syncthreads();
  const int m = s_m;
  bitonic_sort(keys, npad);
  int32_t* t_id = a.t_node_id + base;
  int32_t* t_start = a.t_node_start + base + f;
  int32_t* t_items = a.t_items + base;
  for (int i = tid; i < m; i += blockDim.x) t_items[i] = (int32_t)(uint32_t)keys[i];
  const int nnodes = for_each_run_head(keys, m, s_warp, [&](int i, int r) {
    t_id[r] = (int32_t)(keys[i] >> 32);
    t_start[r] = i;
  });
  if (tid == 0) { t_start[nnodes] = m; a.cnt_nodes[f] = nnodes; a.cnt_items[f] = m; }
  __syncthreads();

  // ---- BowVector: sort by (word, feature); one entry per distinct word --------------------------------------------
  for (int i = tid; i < npad; i += blockDim.x) {
    unsigned long long key = ~0ull;
    if (i < n && word[i] >= 0) key = ((unsigned long long)(uint32_t)word[i] << 32) | (uint32_t)i;
    keys[i] = key;
  }
  __syncthreads();
  bitonic_sort(keys, npad);
  int32_t* bw = a.bow_word + base;
  double* bv = a.bow_val + base;
  const double* kw = a.kp_weight + base;
  const int accumulate = a.accumulate;
  const int nb = for_each_run_head(keys, m, s_warp, [&](int i, int r) {
    const uint32_t wd = (uint32_t)(keys[i] >> 32);
    const double w = kw[(uint32_t)keys[i]];
    double v = w;   // first occurrence inserts (id, w); each further one does `+= w` (addWeight) or nothing
    if (accumulate)
      for (int j = i + 1; j < m && (uint32_t)(keys[j] >> 32) == wd; j++) v = __dadd_rn(v, w);
    bw[r] = (int32_t)wd;
    bv[r] = v;
  });
  __syncthreads();
  if (tid == 0) a.bow_cnt[f] = nb;
  if (nb == 0) return;
  // TemplatedVocabulary.h:1167-1173: without normalisation, TF* weights are divided by the number of words
  __shared__ double s_norm;
  if (a.norm == 0) {
    if (!accumulate) return;
    if (tid == 0) s_norm = (double)nb;
  } else if (tid == 0) {
    // BowVector::normalize: a running sum in map order.  One thread, nb dependent adds (a few microseconds).
    double norm = 0.0;
    if (a.norm == 1) { for (int i = 0; i < nb; i++) norm = __dadd_rn(norm, fabs(bv[i])); }
    else { for (int i = 0; i < nb; i++) norm = __dadd_rn(norm, __dmul_rn(bv[i], bv[i])); norm = __dsqrt_rn(norm); }
    s_norm = norm;
  }
  __syncthreads();
  const double norm = s_norm;
  if (norm > 0.0)
    for (int i = tid; i < nb; i += blockDim.x) bv[i] = __ddiv_rn(bv[i], norm);
}

// exclusive sums of the per-frame node / item counts (one CTA; frames in chunks of blockDim)
__global__ void __launch_bounds__(1024) bow_scan_kernel(const int32_t* cnt_nodes, const int32_t* cnt_items, int nframes,
                                                        int32_t* node_off, int32_t* item_off) {
  __shared__ int s_a[32], s_b[32];
  __shared__ int s_base[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { s_base[0] = 0; s_base[1] = 0; }
  __syncthreads();
  for (int i0 = 0; i0 < nframes; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const int va = i < nframes ? cnt_nodes[i] : 0, vb = i < nframes ? cnt_items[i] : 0;
    int ia = va, ib = vb;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int oa = __shfl_up_sync(0xffffffffu, ia, d), ob = __shfl_up_sync(0xffffffffu, ib, d);
      if (lane >= d) { ia += oa; ib += ob; }
    }
    if (lane == 31) { s_a[warp] = ia; s_b[warp] = ib; }
    __syncthreads();
    int wa = 0, wb = 0, ta = 0, tb = 0;
    for (int w = 0; w < 32; w++) {
      if (w < warp) { wa += s_a[w]; wb += s_b[w]; }
      ta += s_a[w]; tb += s_b[w];
    }
    const int ba = s_base[0], bb = s_base[1];
    if (i < nframes) { node_off[i] = ba + wa + ia - va; item_off[i] = bb + wb + ib - vb; }
    __syncthreads();
    if (threadIdx.x == 0) { s_base[0] = ba + ta; s_base[1] = bb + tb; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { node_off[nframes] = s_base[0]; item_off[nframes] = s_base[1]; }
}

__global__ void __launch_bounds__(128) bow_compact_kernel(const int32_t* frame_off, const int32_t* t_node_id,
                                                          const int32_t* t_node_start, const int32_t* t_items,
                                                          const int32_t* node_off, const int32_t* item_off,
                                                          int32_t* node_id, int32_t* node_start, int32_t* items) {
  const int f = blockIdx.x;
  const int base = frame_off[f], no = node_off[f], nn = node_off[f + 1] - no, io = item_off[f],
            ni = item_off[f + 1] - io;
  for (int i = threadIdx.x; i < nn; i += blockDim.x) node_id[no + i] = t_node_id[base + i];
  for (int i = threadIdx.x; i <= nn; i += blockDim.x) node_start[no + f + i] = t_node_start[base + f + i];
  for (int i = threadIdx.x; i < ni; i += blockDim.x) items[io + i] = t_items[base + i];
}

// L1Scoring::score: one thread per pair walks the two ascending word lists; the sum runs in word order
__global__ void __launch_bounds__(128) bow_score_l1_kernel(const int32_t* off1, const int32_t* cnt1,
                                                           const int32_t* w1, const double* v1, const int32_t* off2,
                                                           const int32_t* cnt2, const int32_t* w2, const double* v2,
                                                           const int32_t* idx1, const int32_t* idx2, int npairs,
                                                           double* out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npairs) return;
  const int f1 = idx1[p], f2 = idx2[p];
  const int32_t *a = w1 + off1[f1], *b = w2 + off2[f2];
  const double *va = v1 + off1[f1], *vb = v2 + off2[f2];
  const int n1 = cnt1[f1], n2 = cnt2[f2];
  double score = 0.0;
  int i = 0, j = 0;
  while (i < n1 && j < n2) {
    const int x = a[i], y = b[j];
    if (x == y) {
      const double vi = va[i], wi = vb[j];
      score = __dadd_rn(score, __dsub_rn(__dsub_rn(fabs(__dsub_rn(vi, wi)), fabs(vi)), fabs(wi)));
      i++; j++;
    } else if (x < y) i++;
    else j++;
  }
  out[p] = __ddiv_rn(-score, 2.0);
}

// ------------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------------
template <class T>
cudaError_t put(T** dptr, const std::vector<T>& v, cudaStream_t s) {
  cudaError_t e = cudaMalloc((void**)dptr, std::max<size_t>(v.size(), 1) * sizeof(T) + 64);
  if (e != cudaSuccess) return e;
  return v.empty() ? cudaSuccess
                   : cudaMemcpyAsync(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s);
}

struct BowTargets {   // where the results go (device)
  int32_t *node_off, *fvitem_off, *node_id, *node_start, *fv_items, *bow_word;
  double* bow_val;
};

inline size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

// descend + per-frame assembly + compaction for frames laid out as a frame set (packed descriptors + frame_off)
int bow_run(orbm_handle* h, const orbm_vocabulary* v, const uint4* d_desc, const int32_t* d_frame_off,
            const std::vector<int32_t>& frame_off, int levelsup, const BowTargets& o,
            std::vector<int32_t>* h_node_off, std::vector<int32_t>* h_item_off, std::vector<int32_t>* h_bow_cnt) {
  const int nframes = (int)frame_off.size() - 1;
  const int64_t N = frame_off[nframes];
  cudaStream_t st = h->stream;
  // scratch slices for frames too large for the shared-memory sort
  std::vector<long long> soff(nframes, -1);
  long long sneed = 0;
  for (int f = 0; f < nframes; f++) {
    const int n = frame_off[f + 1] - frame_off[f];
    int npad = 1;
    while (npad < n) npad <<= 1;
    if (npad > kSortSmemKeys) { soff[f] = sneed; sneed += npad; }
  }
  const size_t n1 = (size_t)std::max<int64_t>(N, 1);
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t at = off; off = align_up(off + bytes); return at; };
  const size_t o_weight = take(n1 * 8), o_scratch = take((size_t)std::max<long long>(sneed, 1) * 8),
               o_soff = take((size_t)nframes * 8), o_word = take(n1 * 4), o_node = take(n1 * 4),
               o_tid = take(n1 * 4), o_tstart = take((n1 + nframes) * 4), o_titems = take(n1 * 4),
               o_cn = take((size_t)nframes * 4), o_ci = take((size_t)nframes * 4), o_bc = take((size_t)nframes * 4);
  ORBM_CK(h->ws_bow.reserve(off));
  char* ws = h->ws_bow.as<char>();
  double* kp_weight = reinterpret_cast<double*>(ws + o_weight);
  int32_t* kp_word = reinterpret_cast<int32_t*>(ws + o_word);
  int32_t* kp_node = reinterpret_cast<int32_t*>(ws + o_node);
  ORBM_CK(cudaMemcpyAsync(ws + o_soff, soff.data(), (size_t)nframes * 8, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaEventRecord(h->ev0, st));
  VocView vv{v->d_info, v->d_desc, v->d_orig, v->d_word, v->d_weight};
  if (N > 0) {
    bow_descend_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(d_desc, (int)N, vv, v->L - levelsup, kp_word,
                                                                     kp_node, kp_weight);
    h->launches++;
  }
  BowFrameArgs a;
  a.frame_off = d_frame_off; a.kp_word = kp_word; a.kp_node = kp_node; a.kp_weight = kp_weight;
  a.accumulate = (v->weighting == 0 || v->weighting == 1) ? 1 : 0;          // TF_IDF, TF
  a.norm = v->scoring == 5 ? 0 : (v->scoring == 1 ? 2 : 1);                 // ScoringObject.h:72-87
  a.t_node_id = reinterpret_cast<int32_t*>(ws + o_tid);
  a.t_node_start = reinterpret_cast<int32_t*>(ws + o_tstart);
  a.t_items = reinterpret_cast<int32_t*>(ws + o_titems);
  a.cnt_nodes = reinterpret_cast<int32_t*>(ws + o_cn);
  a.cnt_items = reinterpret_cast<int32_t*>(ws + o_ci);
  a.bow_word = o.bow_word; a.bow_val = o.bow_val;
  a.bow_cnt = reinterpret_cast<int32_t*>(ws + o_bc);
  a.scratch = reinterpret_cast<unsigned long long*>(ws + o_scratch);
  a.scratch_off = reinterpret_cast<const long long*>(ws + o_soff);
  int max_pad = 1;
  for (int f = 0; f < nframes; f++) {
    const int n = frame_off[f + 1] - frame_off[f];
    if (soff[f] < 0) while (max_pad < n) max_pad <<= 1;
  }
  bow_frame_kernel<<<nframes, kFrameThreads, (size_t)max_pad * 8, st>>>(a);
  bow_scan_kernel<<<1, 1024, 0, st>>>(a.cnt_nodes, a.cnt_items, nframes, o.node_off, o.fvitem_off);
  bow_compact_kernel<<<nframes, 128, 0, st>>>(d_frame_off, a.t_node_id, a.t_node_start, a.t_items, o.node_off,
                                              o.fvitem_off, o.node_id, o.node_start, o.fv_items);
  h->launches += 3;
  ORBM_CK(cudaGetLastError());
  ORBM_CK(cudaEventRecord(h->ev1, st));
  h_node_off->resize(nframes + 1); h_item_off->resize(nframes + 1); h_bow_cnt->resize(nframes);
  ORBM_CK(cudaMemcpyAsync(h_node_off->data(), o.node_off, (size_t)(nframes + 1) * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(h_item_off->data(), o.fvitem_off, (size_t)(nframes + 1) * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(h_bow_cnt->data(), a.bow_cnt, (size_t)nframes * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaStreamSynchronize(st));
  ORBM_CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
  return ORBM_OK;
}

int build_vocabulary(orbm_handle* h, int k, int L, int scoring, int weighting, int nnodes, const int32_t* parent,
                     const uint8_t* desc, const double* weight, const uint8_t* is_leaf, orbm_vocabulary** out) {
  if (!h) return ORBM_ERR_ARG;
  if (!out || !parent || !desc || !weight || !is_leaf || nnodes < 2 || L < 1 || scoring < 0 || scoring > 5 ||
      weighting < 0 || weighting > 3) {
    h->err = "vocabulary: null arrays or parameters out of range";
    return ORBM_ERR_ARG;
  }
  *out = nullptr;
  // children in ascending id order = the order loadFromTextFile pushes them (:1392)
  std::vector<int32_t> cstart(nnodes + 1, 0), child(nnodes - 1);
  for (int i = 1; i < nnodes; i++) {
    if (parent[i] < 0 || parent[i] >= i) { h->err = "vocabulary: a parent must precede its children"; return ORBM_ERR_ARG; }
    cstart[parent[i] + 1]++;
  }
  for (int i = 0; i < nnodes; i++) cstart[i + 1] += cstart[i];
  {
    std::vector<int32_t> cur(cstart.begin(), cstart.end() - 1);
    for (int i = 1; i < nnodes; i++) child[cur[parent[i]]++] = i;
  }
  std::vector<int32_t> wid(nnodes, -1);
  int nwords = 0;
  for (int i = 1; i < nnodes; i++) {
    const bool none = cstart[i + 1] == cstart[i];
    if ((is_leaf[i] != 0) != none) { h->err = "vocabulary: leaf flag disagrees with the child lists"; return ORBM_ERR_ARG; }
    if (none) wid[i] = nwords++;
  }
  // breadth-first renumbering: the children of a node become consecutive
  std::vector<int32_t> order;
  order.reserve(nnodes);
  order.push_back(0);
  std::vector<int2> info(nnodes);
  for (size_t q = 0; q < order.size(); q++) {
    const int node = order[q];
    info[q] = make_int2((int)order.size(), cstart[node + 1] - cstart[node]);
    for (int c = cstart[node]; c < cstart[node + 1]; c++) order.push_back(child[c]);
  }
  if ((int)order.size() != nnodes) { h->err = "vocabulary: unreachable nodes"; return ORBM_ERR_ARG; }
  std::vector<uint4> bdesc((size_t)nnodes * 2);
  std::vector<int32_t> bword(nnodes);
  std::vector<double> bweight(nnodes);
  for (int q = 0; q < nnodes; q++) {
    const int node = order[q];
    memcpy(&bdesc[(size_t)q * 2], desc + (size_t)node * 32, 32);
    bword[q] = wid[node];
    bweight[q] = weight[node];
  }
  ORBM_CK(cudaSetDevice(h->device));
  orbm_vocabulary* v = new orbm_vocabulary();
  v->h = h; v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting; v->nnodes = nnodes; v->nwords = nwords;
  cudaStream_t st = h->stream;
  cudaError_t e;
  if ((e = put(&v->d_info, info, st)) != cudaSuccess || (e = put(&v->d_desc, bdesc, st)) != cudaSuccess ||
      (e = put(&v->d_orig, order, st)) != cudaSuccess || (e = put(&v->d_word, bword, st)) != cudaSuccess ||
      (e = put(&v->d_weight, bweight, st)) != cudaSuccess || (e = cudaStreamSynchronize(st)) != cudaSuccess) {
    h->err = std::string("vocabulary upload: ") + cudaGetErrorString(e);
    orbm_vocabulary_destroy(v);
    return ORBM_ERR_CUDA;
  }
  *out = v;
  return ORBM_OK;
}

}  // namespace

extern "C" int orbm_vocabulary_create(orbm_handle* h, int k, int L, int scoring, int weighting, int nnodes,
                                      const int32_t* parent, const uint8_t* desc, const double* weight,
                                      const uint8_t* is_leaf, orbm_vocabulary** out) {
  return build_vocabulary(h, k, L, scoring, weighting, nnodes, parent, desc, weight, is_leaf, out);
}

// ORBVocabulary::loadFromTextFile.  Header "k L scoring weighting" with the reference's range checks (:1358), then one
// node per line: parent id, leaf flag, 32 descriptor bytes in decimal, weight.
extern "C" int orbm_vocabulary_load_text(orbm_handle* h, const char* path, orbm_vocabulary** out) {
  if (!h) return ORBM_ERR_ARG;
  if (!path || !out) { h->err = "vocabulary: null path"; return ORBM_ERR_ARG; }
  std::ifstream f(path);
  std::string line;
  if (!f || !std::getline(f, line)) { h->err = std::string("vocabulary: cannot read ") + path; return ORBM_ERR_ARG; }
  int k, L, n1, n2;
  if (sscanf(line.c_str(), "%d %d %d %d", &k, &L, &n1, &n2) != 4 || k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 ||
      n1 > 5 || n2 < 0 || n2 > 3) {
    h->err = "vocabulary: not a vocabulary text file (header out of range)";
    return ORBM_ERR_ARG;
  }
  std::vector<int32_t> parent(1, -1);
  std::vector<uint8_t> desc(32, 0), leaf(1, 0);
  std::vector<double> weight(1, 0.0);
  while (std::getline(f, line)) {
    const char* p = line.c_str();
    char* e;
    const long pid = strtol(p, &e, 10);
    if (e == p) break;   // blank line ends the file
    p = e;
    const long lf = strtol(p, &e, 10);
    if (e == p) { h->err = "vocabulary: truncated node line"; return ORBM_ERR_ARG; }
    p = e;
    parent.push_back((int32_t)pid);
    leaf.push_back(lf > 0);
    for (int i = 0; i < 32; i++) {
      const long b = strtol(p, &e, 10);
      if (e == p) { h->err = "vocabulary: truncated node line"; return ORBM_ERR_ARG; }
      p = e;
      desc.push_back((uint8_t)b);
    }
    weight.push_back(strtod(p, &e));
    if (e == p) { h->err = "vocabulary: truncated node line"; return ORBM_ERR_ARG; }
  }
  return build_vocabulary(h, k, L, n1, n2, (int)parent.size(), parent.data(), desc.data(), weight.data(), leaf.data(),
                          out);
}

extern "C" void orbm_vocabulary_destroy(orbm_vocabulary* v) {
  if (!v) return;
  cudaSetDevice(v->h->device);
  cudaStreamSynchronize(v->h->stream);
  void* ptrs[] = {v->d_info, v->d_desc, v->d_orig, v->d_word, v->d_weight};
  for (void* p : ptrs) if (p) cudaFree(p);
  delete v;
}

extern "C" int orbm_vocabulary_info(const orbm_vocabulary* v, int32_t info[6]) {
  if (!v || !info) return ORBM_ERR_ARG;
  info[0] = v->k; info[1] = v->L; info[2] = v->scoring; info[3] = v->weighting; info[4] = v->nwords;
  info[5] = v->nnodes;
  return ORBM_OK;
}

extern "C" int orbm_vocabulary_transform(orbm_vocabulary* v, const uint8_t* desc, int n, int levelsup,
                                         int32_t* bow_word, double* bow_weight, int32_t* n_bow, int32_t* fv_node,
                                         int32_t* fv_start, int32_t* fv_items, int32_t* n_nodes) {
  if (!v) return ORBM_ERR_ARG;
  orbm_handle* h = v->h;
  if (n < 0 || (n > 0 && !desc) || !bow_word || !bow_weight || !n_bow || !fv_node || !fv_start || !fv_items ||
      !n_nodes) {
    h->err = "transform: null buffers";
    return ORBM_ERR_ARG;
  }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  const size_t n1 = (size_t)std::max(n, 1);
  // staging: descriptors | frame_off[2] | node_off[2] | item_off[2] | node_id | node_start | items | bow_word | bow_val
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t at = off; off = align_up(off + bytes); return at; };
  const size_t o_val = take(n1 * 8), o_desc = take(n1 * 32), o_foff = take(8), o_noff = take(8), o_ioff = take(8),
               o_nid = take(n1 * 4), o_nst = take((n1 + 1) * 4), o_it = take(n1 * 4), o_bw = take(n1 * 4);
  ORBM_CK(h->stage_a.reserve(off));
  char* sa = h->stage_a.as<char>();
  const std::vector<int32_t> frame_off = {0, n};
  if (n) ORBM_CK(cudaMemcpyAsync(sa + o_desc, desc, (size_t)n * 32, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(sa + o_foff, frame_off.data(), 8, cudaMemcpyHostToDevice, st));
  BowTargets o;
  o.node_off = reinterpret_cast<int32_t*>(sa + o_noff); o.fvitem_off = reinterpret_cast<int32_t*>(sa + o_ioff);
  o.node_id = reinterpret_cast<int32_t*>(sa + o_nid); o.node_start = reinterpret_cast<int32_t*>(sa + o_nst);
  o.fv_items = reinterpret_cast<int32_t*>(sa + o_it); o.bow_word = reinterpret_cast<int32_t*>(sa + o_bw);
  o.bow_val = reinterpret_cast<double*>(sa + o_val);
  std::vector<int32_t> noff, ioff, bcnt;
  const int rc = bow_run(h, v, reinterpret_cast<const uint4*>(sa + o_desc), reinterpret_cast<int32_t*>(sa + o_foff),
                         frame_off, levelsup, o, &noff, &ioff, &bcnt);
  if (rc != ORBM_OK) return rc;
  *n_bow = bcnt[0];
  *n_nodes = noff[1];
  if (bcnt[0]) {
    ORBM_CK(cudaMemcpyAsync(bow_word, o.bow_word, (size_t)bcnt[0] * 4, cudaMemcpyDeviceToHost, st));
    ORBM_CK(cudaMemcpyAsync(bow_weight, o.bow_val, (size_t)bcnt[0] * 8, cudaMemcpyDeviceToHost, st));
  }
  if (noff[1]) ORBM_CK(cudaMemcpyAsync(fv_node, o.node_id, (size_t)noff[1] * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaMemcpyAsync(fv_start, o.node_start, (size_t)(noff[1] + 1) * 4, cudaMemcpyDeviceToHost, st));
  if (ioff[1]) ORBM_CK(cudaMemcpyAsync(fv_items, o.fv_items, (size_t)ioff[1] * 4, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaStreamSynchronize(st));
  return ORBM_OK;
}

extern "C" int orbm_frameset_compute_bow(orbm_frameset* s, const orbm_vocabulary* v, int levelsup) {
  if (!s) return ORBM_ERR_ARG;
  orbm_handle* h = s->h;
  if (!v || v->h != h) { h->err = "compute_bow: vocabulary missing or created on another handle"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  ORBM_CK(cudaStreamSynchronize(h->stream));
  // (re)allocate the upper-bound sized targets: a frame has at most n nodes, n items and n words
  const size_t n1 = (size_t)std::max<int64_t>(s->total_kp, 1), nf = (size_t)s->nframes;
  s->has_fv = false;
  s->bow_cnt.clear();
  if (!s->bow_sized) {   // first ComputeBoW on this set (or feature vectors uploaded at their exact size): reallocate
    void** old[] = {(void**)&s->d_node_off, (void**)&s->d_fvitem_off, (void**)&s->d_node_id, (void**)&s->d_node_start,
                    (void**)&s->d_fv_items, (void**)&s->d_bow_word, (void**)&s->d_bow_val};
    for (void** p : old) { orbm_free(h, *p); *p = nullptr; }
    ORBM_CK(orbm_alloc(h, &s->d_node_off, (nf + 1) * 4));
    ORBM_CK(orbm_alloc(h, &s->d_fvitem_off, (nf + 1) * 4));
    ORBM_CK(orbm_alloc(h, &s->d_node_id, n1 * 4));
    ORBM_CK(orbm_alloc(h, &s->d_node_start, (n1 + nf) * 4));
    ORBM_CK(orbm_alloc(h, &s->d_fv_items, n1 * 4));
    ORBM_CK(orbm_alloc(h, &s->d_bow_word, n1 * 4));
    ORBM_CK(orbm_alloc(h, &s->d_bow_val, n1 * 8));
    s->bow_sized = true;
  }
  BowTargets o{s->d_node_off, s->d_fvitem_off, s->d_node_id, s->d_node_start, s->d_fv_items, s->d_bow_word,
               s->d_bow_val};
  const int rc = bow_run(h, v, reinterpret_cast<const uint4*>(s->d_desc), s->d_frame_off, s->frame_off, levelsup, o,
                         &s->node_off, &s->fvitem_off, &s->bow_cnt);
  if (rc != ORBM_OK) return rc;
  s->total_nodes = s->node_off[s->nframes];
  s->total_fv_items = s->fvitem_off[s->nframes];
  s->has_fv = true;
  return ORBM_OK;
}

extern "C" int orbm_frameset_bow_sizes(orbm_frameset* s, int32_t* n_words, int32_t* n_nodes, int32_t* n_items) {
  if (!s) return ORBM_ERR_ARG;
  for (int f = 0; f < s->nframes; f++) {
    if (n_words) n_words[f] = s->bow_cnt.empty() ? 0 : s->bow_cnt[f];
    if (n_nodes) n_nodes[f] = s->has_fv ? s->node_off[f + 1] - s->node_off[f] : 0;
    if (n_items) n_items[f] = s->has_fv ? s->fvitem_off[f + 1] - s->fvitem_off[f] : 0;
  }
  return ORBM_OK;
}

extern "C" int orbm_frameset_bowvec(orbm_frameset* s, int frame, int32_t* word, double* weight) {
  if (!s) return ORBM_ERR_ARG;
  orbm_handle* h = s->h;
  if (frame < 0 || frame >= s->nframes || !word || !weight) { h->err = "bowvec: bad frame / null buffers"; return ORBM_ERR_ARG; }
  if (s->bow_cnt.empty()) { h->err = "bowvec: orbm_frameset_compute_bow has not run on this set"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  const size_t n = s->bow_cnt[frame], o = s->frame_off[frame];
  if (!n) return ORBM_OK;
  ORBM_CK(cudaMemcpy(word, s->d_bow_word + o, n * 4, cudaMemcpyDeviceToHost));
  ORBM_CK(cudaMemcpy(weight, s->d_bow_val + o, n * 8, cudaMemcpyDeviceToHost));
  return ORBM_OK;
}

extern "C" int orbm_frameset_featvec(orbm_frameset* s, int frame, int32_t* node_id, int32_t* start, int32_t* items) {
  if (!s) return ORBM_ERR_ARG;
  orbm_handle* h = s->h;
  if (frame < 0 || frame >= s->nframes || !node_id || !start || !items) { h->err = "featvec: bad frame / null buffers"; return ORBM_ERR_ARG; }
  if (!s->has_fv) { h->err = "featvec: the set holds no feature vectors"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  const size_t no = s->node_off[frame], nn = s->node_off[frame + 1] - no, io = s->fvitem_off[frame],
               ni = s->fvitem_off[frame + 1] - io;
  if (nn) ORBM_CK(cudaMemcpy(node_id, s->d_node_id + no, nn * 4, cudaMemcpyDeviceToHost));
  ORBM_CK(cudaMemcpy(start, s->d_node_start + no + frame, (nn + 1) * 4, cudaMemcpyDeviceToHost));
  if (ni) ORBM_CK(cudaMemcpy(items, s->d_fv_items + io, ni * 4, cudaMemcpyDeviceToHost));
  return ORBM_OK;
}

extern "C" int orbm_bow_score_l1(orbm_frameset* s1, orbm_frameset* s2, const int32_t* idx1, const int32_t* idx2,
                                 int npairs, double* score) {
  if (!s1 || !s2) return ORBM_ERR_ARG;
  orbm_handle* h = s1->h;
  if (s2->h != h || !idx1 || !idx2 || !score || npairs < 0) { h->err = "score: bad arguments"; return ORBM_ERR_ARG; }
  if (s1->bow_cnt.empty() || s2->bow_cnt.empty()) { h->err = "score: compute_bow has not run on both sets"; return ORBM_ERR_ARG; }
  if (!npairs) return ORBM_OK;
  for (int p = 0; p < npairs; p++)
    if (idx1[p] < 0 || idx1[p] >= s1->nframes || idx2[p] < 0 || idx2[p] >= s2->nframes) { h->err = "score: frame index out of range"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  // staging: score | idx1 | idx2 | cnt1 | cnt2
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t at = off; off = align_up(off + bytes); return at; };
  const size_t o_sc = take((size_t)npairs * 8), o_i1 = take((size_t)npairs * 4), o_i2 = take((size_t)npairs * 4),
               o_c1 = take((size_t)s1->nframes * 4), o_c2 = take((size_t)s2->nframes * 4);
  ORBM_CK(h->stage_b.reserve(off));
  char* sb = h->stage_b.as<char>();
  ORBM_CK(cudaMemcpyAsync(sb + o_i1, idx1, (size_t)npairs * 4, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(sb + o_i2, idx2, (size_t)npairs * 4, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(sb + o_c1, s1->bow_cnt.data(), (size_t)s1->nframes * 4, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(sb + o_c2, s2->bow_cnt.data(), (size_t)s2->nframes * 4, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaEventRecord(h->ev0, st));
  bow_score_l1_kernel<<<(npairs + 127) / 128, 128, 0, st>>>(
      s1->d_frame_off, reinterpret_cast<int32_t*>(sb + o_c1), s1->d_bow_word, s1->d_bow_val, s2->d_frame_off,
      reinterpret_cast<int32_t*>(sb + o_c2), s2->d_bow_word, s2->d_bow_val, reinterpret_cast<int32_t*>(sb + o_i1),
      reinterpret_cast<int32_t*>(sb + o_i2), npairs, reinterpret_cast<double*>(sb + o_sc));
  h->launches++;
  ORBM_CK(cudaGetLastError());
  ORBM_CK(cudaEventRecord(h->ev1, st));
  ORBM_CK(cudaMemcpyAsync(score, sb + o_sc, (size_t)npairs * 8, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaStreamSynchronize(st));
  ORBM_CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
  return ORBM_OK;
}
