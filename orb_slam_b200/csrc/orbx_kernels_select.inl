// Included by orbx_kernels_detect.inl.  E5 quota redistribution and E6 KeyPointsFilter::retainBest.
namespace orbx {

// ------------------------------------------------------------------------------------------------------------------
// Functional port of libstdc++ 13 std::nth_element (bits/stl_algo.h __introselect / __unguarded_partition_pivot /
// __move_median_to_first / __heap_select / __insertion_sort, bits/stl_heap.h __make_heap / __pop_heap /
// __adjust_heap / __push_heap) on 8-byte Cand records with the comparator of OpenCV's KeypointResponseGreater
// (a.response > b.response).  The reference truncates retainBest's result to n entries IN THE ORDER nth_element
// left them (:683-685, :697-701), so the element order — not just the set — is part of the output.
// tests/test_gpu_select.py checks this port (orbx_debug_retain_best) against the real std::nth_element on random,
// tie-heavy, sorted and depth-limit lists.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool cgt(const Cand& a, const Cand& b) { return a.resp > b.resp; }
__device__ __forceinline__ void cswap(Cand* v, int i, int j) { Cand t = v[i]; v[i] = v[j]; v[j] = t; }

__device__ void nth_push_heap(Cand* first, int hole, int top, Cand value) {
  int parent = (hole - 1) / 2;
  while (hole > top && cgt(first[parent], value)) {
    first[hole] = first[parent];
    hole = parent;
    parent = (hole - 1) / 2;
  }
  first[hole] = value;
}

__device__ void nth_adjust_heap(Cand* first, int hole, int len, Cand value) {
  const int top = hole;
  int child = hole;
  while (child < (len - 1) / 2) {
    child = 2 * (child + 1);
    if (cgt(first[child], first[child - 1])) child--;
    first[hole] = first[child];
    hole = child;
  }
  if ((len & 1) == 0 && child == (len - 2) / 2) {
    child = 2 * (child + 1);
    first[hole] = first[child - 1];
    hole = child - 1;
  }
  nth_push_heap(first, hole, top, value);
}

__device__ void nth_heap_select(Cand* first, int middle, int last) {  // [0,middle) heap, scan [middle,last)
  const int len = middle;
  if (len >= 2) {  // __make_heap
    int parent = (len - 2) / 2;
    while (true) {
      Cand value = first[parent];
      nth_adjust_heap(first, parent, len, value);
      if (parent == 0) break;
      parent--;
    }
  }
  for (int i = middle; i < last; ++i)
    if (cgt(first[i], first[0])) {  // __pop_heap(first, middle, i)
      Cand value = first[i];
      first[i] = first[0];
      nth_adjust_heap(first, 0, len, value);
    }
}

// kBatch: the two scan loops of the partition read four consecutive keys per trip (independent loads, one latency) and
// then test them in order — same result, a quarter of the dependent round trips.  The scans may read up to 3 records
// beyond either end of [first, last): only legal where the caller pads the array (the shared-memory kernel does).
template <bool kBatch>
__device__ void nth_element_port(Cand* v, int nth, int n) {
  if (n == 0 || nth == n) return;
  int first = 0, last = n;
  int depth = 2 * (31 - __clz(n));  // std::__lg(n) * 2
  while (last - first > 3) {
    if (depth == 0) {
      nth_heap_select(v + first, nth + 1 - first, last - first);
      cswap(v, first, nth);
      return;
    }
    --depth;
    // __unguarded_partition_pivot
    const int mid = first + (last - first) / 2;
    {  // __move_median_to_first(first, first+1, mid, last-1)
      const int a = first + 1, b = mid, c = last - 1;
      if (cgt(v[a], v[b])) {
        if (cgt(v[b], v[c])) cswap(v, first, b);
        else if (cgt(v[a], v[c])) cswap(v, first, c);
        else cswap(v, first, a);
      } else if (cgt(v[a], v[c])) cswap(v, first, a);
      else if (cgt(v[b], v[c])) cswap(v, first, c);
      else cswap(v, first, b);
    }
    int lo = first + 1, hi = last;
    const float pivot = v[first].resp;  // the pivot element itself never moves during the partition
    while (true) {
      if (kBatch) {
        for (;;) {
          const float a0 = v[lo].resp, a1 = v[lo + 1].resp, a2 = v[lo + 2].resp, a3 = v[lo + 3].resp;
          if (!(a0 > pivot)) break;
          if (!(a1 > pivot)) { lo += 1; break; }
          if (!(a2 > pivot)) { lo += 2; break; }
          if (!(a3 > pivot)) { lo += 3; break; }
          lo += 4;
        }
        --hi;
        for (;;) {
          const float a0 = v[hi].resp, a1 = v[hi - 1].resp, a2 = v[hi - 2].resp, a3 = v[hi - 3].resp;
          if (!(pivot > a0)) break;
          if (!(pivot > a1)) { hi -= 1; break; }
          if (!(pivot > a2)) { hi -= 2; break; }
          if (!(pivot > a3)) { hi -= 3; break; }
          hi -= 4;
        }
      } else {
        while (v[lo].resp > pivot) ++lo;
        --hi;
        while (pivot > v[hi].resp) --hi;
      }
      if (!(lo < hi)) break;
      cswap(v, lo, hi);
      ++lo;
    }
    const int cut = lo;
    if (cut <= nth) first = cut;
    else last = cut;
  }
  // __insertion_sort(first, last)
  for (int i = first + 1; i < last; ++i) {
    Cand val = v[i];
    if (cgt(val, v[first])) {
      for (int k = i; k > first; --k) v[k] = v[k - 1];
      v[first] = val;
    } else {
      int k = i;
      while (cgt(val, v[k - 1])) { v[k] = v[k - 1]; --k; }
      v[k] = val;
    }
  }
}

// retainBest(keys, n) followed by resize(n): the first min(n, size) entries survive            (:683-685)
template <bool kBatch = false>
__device__ __forceinline__ void retain_best_port(Cand* v, int size, int n, int retain_mode) {
  if (n > 0 && size > n) nth_element_port<kBatch>(v, retain_mode == ORBX_RETAIN_CV24 ? n : n - 1, size);
}

// ------------------------------------------------------------------------------------------------------------------
// One warp per (level, frame):
//   lane 0   quota per cell + redistribution loop                                  (reference :622-636, :644-670)
//   lanes    per-cell retainBest in place in the cell's candidate list             (:678-685)
//   lanes    copy the survivors, cells in row-major order, into the level list     (:687-694)
//   lane 0   level cap retainBest(keypoints, nDesired) + resize                     (:697-701)
//   lanes    publish the level's keypoints into the frame's slots [kp_base, kp_base + count)
// Serial sections are a few hundred steps; thousands of (level, frame) warps run concurrently.
// ------------------------------------------------------------------------------------------------------------------
// kSmemList: the LEVEL list (survivors of the cells, <= list_cap records) lives in shared memory, `stride` records per
// warp.  The level cap is one lane's chain of several hundred dependent steps (nth_element over ~240 records): ~30
// cycles per step from shared memory instead of an L2 round trip, and it is the longest chain of the kernel.
template <bool kSmemList>
__global__ void __launch_bounds__(128) select_kernel(ChunkBuffers b, int nwarps_total, int stride) {
  ORBX_PDL_ENTRY();
  extern __shared__ __align__(16) uint8_t sel_list_smem[];
  const DevicePlan* P = b.plan;
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (gw >= nwarps_total) return;
  const int level = gw % P->nlevels, frame = gw / P->nlevels;
  const LevelPlan& L = P->lv[level];
  const int nCells = L.nCells, q = L.quota;
  const int32_t* nTotal = b.n_total + (int64_t)frame * P->total_cells + L.cell_base;
  int32_t* nToRetain = b.n_retain + (int64_t)frame * P->total_cells + L.cell_base;
  Cand* cand = b.cand + (int64_t)frame * P->cand_per_frame;
  Cand* list = kSmemList ? reinterpret_cast<Cand*>(sel_list_smem) + (threadIdx.x >> 5) * stride + 4   // 4 pad records either side
                         : b.list + (int64_t)frame * P->list_per_frame + L.list_base;

  if (lane == 0) {
    // bNoMore is kept in the sign bit of nToRetain while the loop runs (-(v)-1 encodes "no more")
    int nNoMore = 0, nToDistribute = 0;
    for (int c = 0; c < nCells; c++) {
      const CellPlan& cp = b.cells[L.cell_base + c];
      const int nKeys = nTotal[c];
      // a cell the reference `continue`s over (:571, :595) keeps nToRetain = 0 / bNoMore = false and is NOT
      // counted in nNoMore; it still takes part in the redistribution below with nTotal = 0
      if (cp.skipped & 1) { nToRetain[c] = 0; continue; }
      if (nKeys > q) nToRetain[c] = q;
      else { nToRetain[c] = -nKeys - 1; nToDistribute += q - nKeys; nNoMore++; }
    }
    while (nToDistribute > 0 && nNoMore < nCells) {
      const int nNew = q + (int)ceilf(__fdiv_rn((float)nToDistribute, (float)(nCells - nNoMore)));
      nToDistribute = 0;
      for (int c = 0; c < nCells; c++) {
        if (nToRetain[c] >= 0) {
          if (nTotal[c] > nNew) nToRetain[c] = nNew;
          else { nToRetain[c] = -nTotal[c] - 1; nToDistribute += nNew - nTotal[c]; nNoMore++; }
        }
      }
    }
    for (int c = 0; c < nCells; c++)
      if (nToRetain[c] < 0) nToRetain[c] = -nToRetain[c] - 1;
  }
  __syncwarp();
  for (int c = lane; c < nCells; c += 32) {
    const CellPlan& cp = b.cells[L.cell_base + c];
    // batched scans (four independent loads per trip): the over-read of <= 3 records lands in the neighbouring cell's
    // block or in the pad records around the array; whatever it reads is ignored
    retain_best_port<true>(cand + cp.cand_off, nTotal[c], nToRetain[c], P->retain_mode);
  }
  __syncwarp();
  // exclusive prefix of nToRetain over the cells (warp scan in chunks of 32)
  int run = 0;
  for (int c0 = 0; c0 < nCells; c0 += 32) {
    const int c = c0 + lane;
    const int n = c < nCells ? min(nToRetain[c], nTotal[c]) : 0;
    int incl = n;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    const int off = run + incl - n;
    if (c < nCells && off + n <= L.list_cap) {
      const Cand* src = cand + b.cells[L.cell_base + c].cand_off;
      for (int k = 0; k < n; k++) list[off + k] = src[k];
    }
    run += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (run > L.list_cap) {
    if (lane == 0) { atomicExch(b.err_flag, 2); b.kp_count[frame * kMaxLevels + level] = 0; }
    return;
  }
  __syncwarp();
  int count = run;
  if (run > L.nDesired) {
    if (lane == 0) retain_best_port<kSmemList>(list, run, L.nDesired, P->retain_mode);   // (padded: batched scans allowed)
    count = L.nDesired;
  }
  __syncwarp();
  Cand* kp = b.kp + (int64_t)frame * P->nfeatures + L.kp_base;
  for (int k = lane; k < count; k += 32) kp[k] = list[k];
  if (lane == 0) b.kp_count[frame * kMaxLevels + level] = count;
}

// ------------------------------------------------------------------------------------------------------------------
// The same selection with every serial chain in SHARED memory: one CTA per (level, frame), one warp per cell.
// std::nth_element's element order is part of the output, so the partitions stay sequential (one lane per cell, lane 0
// for the level cap) — what can change is the cost of one dependent step: ~25 cycles in shared memory instead of an
// L1/L2 round trip.  Single-frame latency of the stage: 90 -> ~20 us; per 10k frames 5.5 -> ~1.5 ms.
//   s_list  [list_cap]        the level list (survivors of the cells, row-major cell order)
//   s_slab  [32][kSelStride]  one row per lane: the candidate list of the lane's cell while it is partitioned
//   s_tot / s_ret / s_off [nCells]
// A cell with more than kSelCap candidates is partitioned in place in global memory as before (never seen at the
// headline configurations: ~35 candidates per cell).
// ------------------------------------------------------------------------------------------------------------------
#ifndef ORBX_SEL_CAP
#define ORBX_SEL_CAP 128
#endif
constexpr int kSelCap = ORBX_SEL_CAP, kSelStride = kSelCap + 5;   // row padding: the batched scans over-read <= 3 records
constexpr int kSelPad = 4;

__host__ __device__ inline int select_smem_bytes(int list_cap, int ncells) {
  return (kSelPad + list_cap + kSelPad + 32 * kSelStride + kSelPad) * (int)sizeof(Cand) +
         3 * ((ncells + 3) & ~3) * (int)sizeof(int32_t);
}

__global__ void __launch_bounds__(1024) select_smem_kernel(ChunkBuffers b, int nwarps_total, int max_list_cap,
                                                          int max_cells) {
  ORBX_PDL_ENTRY();
  extern __shared__ __align__(16) uint8_t sel_smem[];
  const DevicePlan* P = b.plan;
  const int gw = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  if (gw >= nwarps_total) return;
  const int level = gw % P->nlevels, frame = gw / P->nlevels;
  const LevelPlan& L = P->lv[level];
  const int nCells = L.nCells, q = L.quota, retain_mode = P->retain_mode;
  Cand* s_list = reinterpret_cast<Cand*>(sel_smem) + kSelPad;
  Cand* s_slab = s_list + max_list_cap + kSelPad;
  int32_t* s_tot = reinterpret_cast<int32_t*>(s_slab + 32 * kSelStride + kSelPad);
  int32_t* s_ret = s_tot + ((max_cells + 3) & ~3);
  int32_t* s_off = s_ret + ((max_cells + 3) & ~3);
  __shared__ int s_run;
  const int32_t* nTotal = b.n_total + (int64_t)frame * P->total_cells + L.cell_base;
  int32_t* nToRetain = b.n_retain + (int64_t)frame * P->total_cells + L.cell_base;
  Cand* cand = b.cand + (int64_t)frame * P->cand_per_frame;
  const CellPlan* cells = b.cells + L.cell_base;

  for (int c = tid; c < nCells; c += blockDim.x) {
    s_tot[c] = nTotal[c];
    s_off[c] = cells[c].skipped & 1;                       // borrowed until the prefix below
  }
  __syncthreads();
  if (tid == 0) {                                      // quota + redistribution, :622-636 / :644-670 (see select_kernel)
    int nNoMore = 0, nToDistribute = 0;
    for (int c = 0; c < nCells; c++) {
      const int nKeys = s_tot[c];
      if (s_off[c]) { s_ret[c] = 0; continue; }
      if (nKeys > q) s_ret[c] = q;
      else { s_ret[c] = -nKeys - 1; nToDistribute += q - nKeys; nNoMore++; }
    }
    while (nToDistribute > 0 && nNoMore < nCells) {
      const int nNew = q + (int)ceilf(__fdiv_rn((float)nToDistribute, (float)(nCells - nNoMore)));
      nToDistribute = 0;
      for (int c = 0; c < nCells; c++) {
        if (s_ret[c] >= 0) {
          if (s_tot[c] > nNew) s_ret[c] = nNew;
          else { s_ret[c] = -s_tot[c] - 1; nToDistribute += nNew - s_tot[c]; nNoMore++; }
        }
      }
    }
    for (int c = 0; c < nCells; c++)
      if (s_ret[c] < 0) s_ret[c] = -s_ret[c] - 1;
  }
  __syncthreads();
  // exclusive prefix of the survivors per cell -> offsets into the level list (warp 0)
  if (warp == 0) {
    int run = 0;
    for (int c0 = 0; c0 < nCells; c0 += 32) {
      const int c = c0 + lane;
      const int n = c < nCells ? min(s_ret[c], s_tot[c]) : 0;
      if (c < nCells) nToRetain[c] = s_ret[c];
      int incl = n;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
      }
      if (c < nCells) s_off[c] = run + incl - n;
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) s_run = run;
  }
  __syncthreads();
  const int run = s_run;
  if (run > L.list_cap) {
    if (tid == 0) { atomicExch(b.err_flag, 2); b.kp_count[frame * kMaxLevels + level] = 0; }
    return;
  }
  // per-cell retainBest (:678-685): ONE WARP per cell - the lanes copy the list in and the survivors out, lane 0 runs
  // the sequential partition.  (A lane per cell made the 30 partitions of a level diverge inside one warp and
  // serialised them: 33k cycles for this phase; a warp per cell runs them side by side on the SM's schedulers.)
  Cand* row = s_slab + warp * kSelStride;
  for (int c = warp; c < nCells; c += nw) {
    const int tot = s_tot[c], ret = s_ret[c], n = min(ret, tot);
    Cand* src = cand + cells[c].cand_off;
    const bool reorder = ret > 0 && tot > ret;
    if (reorder && tot <= kSelCap) {
      for (int k = lane; k < tot; k += 32) row[k] = src[k];
      __syncwarp();
      if (lane == 0) nth_element_port<true>(row, retain_mode == ORBX_RETAIN_CV24 ? ret : ret - 1, tot);
      __syncwarp();
      src = row;
    } else if (reorder) {
      if (lane == 0) nth_element_port<false>(src, retain_mode == ORBX_RETAIN_CV24 ? ret : ret - 1, tot);
      __syncwarp();
    }
    Cand* dst = s_list + s_off[c];
    for (int k = lane; k < n; k += 32) dst[k] = src[k];
    __syncwarp();
  }
  __syncthreads();
  int count = run;
  if (run > L.nDesired) {                              // level cap retainBest + resize (:697-701)
    if (tid == 0) retain_best_port<true>(s_list, run, L.nDesired, retain_mode);
    count = L.nDesired;
    __syncthreads();
  }
  Cand* kp = b.kp + (int64_t)frame * P->nfeatures + L.kp_base;
  for (int k = tid; k < count; k += blockDim.x) kp[k] = s_list[k];
  if (tid == 0) b.kp_count[frame * kMaxLevels + level] = count;
}

// Which one: a chunk of a few frames is a handful of warps whose serial chains ARE the stage's latency -> shared memory
// (43 instead of 90 us for one frame).  A full chunk runs thousands of such warps beside blur7_kernel, hidden behind it;
// there the global version wins because it leaves the SMs' shared memory to the blur CTAs (A/B on one box, 10k frames:
// 56.00 ms with the global kernel, 57.39 with this one at kSelCap 128, 56.27 at kSelCap 64).
constexpr int kSelectSmemMaxFrames = 32;

void launch_select(const ChunkBuffers& b, const DevicePlan& hp, int nframes, cudaStream_t s, int force) {
  const int nwarps = hp.nlevels * nframes;
  int max_list = 0, max_cells = 0;
  for (int l = 0; l < hp.nlevels; l++) { max_list = std::max(max_list, hp.lv[l].list_cap); max_cells = std::max(max_cells, hp.lv[l].nCells); }
  max_list = (max_list + 1) & ~1;
  const int smem = select_smem_bytes(max_list, max_cells);
  // force: 0 = by chunk size, 1 = global, 2 = shared memory (ORBX_SELECT, read when the handle is created)
  const bool want_smem = force ? force == 2 : nframes <= kSelectSmemMaxFrames;
  if (want_smem && smem <= 160 * 1024) {
    cudaFuncSetAttribute(select_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int warps = std::min(32, std::max(1, max_cells));
    launch_k(select_smem_kernel, dim3(nwarps), dim3(32 * warps), (size_t)smem, s, b, nwarps, max_list, max_cells);
    return;
  }
  const int stride = max_list + 8;                               // 4 pad records on either side
  const size_t list_smem = (size_t)4 * stride * sizeof(Cand);
  if (list_smem <= 96 * 1024) {
    cudaFuncSetAttribute(select_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)list_smem);
    launch_k(select_kernel<true>, dim3((nwarps * 32 + 127) / 128), dim3(128), list_smem, s, b, nwarps, stride);
  } else {
    launch_k(select_kernel<false>, dim3((nwarps * 32 + 127) / 128), dim3(128), 0, s, b, nwarps, 0);
  }
}

// Unit-test hook (orbx_debug_retain_best): retain_best_port on caller-made lists, one thread per list.  Lists lie back
// to back behind 4 pad records on either side, so the kBatch scans' over-reads stay inside the buffer.
template <bool kBatch>
__global__ void retain_best_debug_kernel(Cand* v, const int32_t* start, const int32_t* keep, int nlists, int retain_mode) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlists) return;
  retain_best_port<kBatch>(v + start[i], start[i + 1] - start[i], keep[i], retain_mode);
}

void launch_retain_best_debug(Cand* v, const int32_t* start, const int32_t* keep, int nlists, int retain_mode, int batch,
                              cudaStream_t s) {
  if (batch) retain_best_debug_kernel<true><<<(nlists + 63) / 64, 64, 0, s>>>(v, start, keep, nlists, retain_mode);
  else retain_best_debug_kernel<false><<<(nlists + 63) / 64, 64, 0, s>>>(v, start, keep, nlists, retain_mode);
}

}  // namespace orbx
