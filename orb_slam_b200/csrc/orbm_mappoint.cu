// N4: MapPoint::ComputeDistinctiveDescriptors batched (reference src/MapPoint.cc:185-250).  include/orbm.h.
// For every map point: all-pairs Hamming distances between its observed descriptors, the median of every row
// (`vDists[0.5*(N-1)]` after std::sort, the 0 self-distance included) and the first row with the least median.
// A sort is not needed: the k-th smallest of a row of integers in [0, 256] is read off a 257-bin histogram.
//
// One warp per map point.  Row i: lanes take columns j = lane, lane + 32, ... (the row descriptor is a warp-wide
// broadcast load, column descriptors stay in L1), bump the warp's shared-memory histogram, then each lane sums its 9
// bins, a shuffle scan locates the bin holding rank k.  Work per point: N^2 distances (the reference computes
// N(N-1)/2 and mirrors them; recomputing is cheaper here than a N x N scratch matrix).
#include "orbm_internal.cuh"

using namespace orbm;

namespace {

constexpr int kBins = 288;            // 257 distance values, padded to 9 per lane
constexpr int kWarpsPerCta = 4;

__global__ void __launch_bounds__(kWarpsPerCta * 32) distinctive_kernel(const uint4* __restrict__ desc,
                                                                        const int32_t* __restrict__ start,
                                                                        int npoints, int32_t* __restrict__ best,
                                                                        uint4* __restrict__ out_desc) {
  __shared__ int s_hist[kWarpsPerCta][kBins];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int p = blockIdx.x * kWarpsPerCta + warp;
  if (p >= npoints) return;
  const int s0 = start[p], n = start[p + 1] - s0;
  if (n <= 0) {
    if (lane == 0) best[p] = -1;
    return;
  }
  int* hist = s_hist[warp];
  const uint4* d = desc + 2 * (int64_t)s0;
  const int k = (n - 1) >> 1;         // (int)(0.5 * (N - 1))
  int best_median = INT_MAX, best_idx = 0;
  for (int i = 0; i < n; i++) {
#pragma unroll
    for (int b = 0; b < 9; b++) hist[lane * 9 + b] = 0;
    __syncwarp();
    const uint4 q0 = __ldg(d + 2 * i), q1 = __ldg(d + 2 * i + 1);
    for (int j = lane; j < n; j += 32)
      atomicAdd(&hist[hamming256(q0, q1, __ldg(d + 2 * j), __ldg(d + 2 * j + 1))], 1);
    __syncwarp();
    int c[9], sum = 0;
#pragma unroll
    for (int b = 0; b < 9; b++) { c[b] = hist[lane * 9 + b]; sum += c[b]; }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    // the lane whose bins cover rank k: exclusive prefix <= k < inclusive prefix
    int median = 0;
    int run = incl - sum;
    if (run <= k && k < incl) {
#pragma unroll
      for (int b = 0; b < 9; b++) {
        if (run <= k && k < run + c[b]) median = lane * 9 + b;
        run += c[b];
      }
    }
    median = __reduce_max_sync(0xffffffffu, median);   // exactly one lane holds a value (others 0)
    if (median < best_median) { best_median = median; best_idx = i; }
    __syncwarp();
  }
  if (lane == 0) best[p] = best_idx;
  if (out_desc && lane < 2) out_desc[2 * (int64_t)p + lane] = __ldg(d + 2 * best_idx + lane);
}

}  // namespace

extern "C" int orbm_distinctive_descriptors(orbm_handle* h, const uint8_t* desc, const int32_t* start, int npoints,
                                            int32_t* best, uint8_t* out_desc) {
  if (!h) return ORBM_ERR_ARG;
  if (!start || !best || npoints < 0) { h->err = "distinctive: null buffers"; return ORBM_ERR_ARG; }
  if (npoints == 0) return ORBM_OK;
  if (start[0] != 0) { h->err = "distinctive: start[0] must be 0"; return ORBM_ERR_ARG; }
  for (int p = 0; p < npoints; p++)
    if (start[p + 1] < start[p]) { h->err = "distinctive: starts must be monotone"; return ORBM_ERR_ARG; }
  const int64_t total = start[npoints];
  if (total > 0 && !desc) { h->err = "distinctive: null descriptors"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  ORBM_CK(h->stage_a.reserve((size_t)std::max<int64_t>(total, 1) * 32));
  ORBM_CK(h->stage_b.reserve((size_t)(npoints + 1) * 4));
  ORBM_CK(h->stage_c.reserve((size_t)npoints * 4 + 256 + (size_t)npoints * 32));
  int32_t* d_best = h->stage_c.as<int32_t>();
  uint4* d_out = reinterpret_cast<uint4*>(h->stage_c.as<char>() + (((size_t)npoints * 4 + 255) & ~(size_t)255));
  if (total) ORBM_CK(cudaMemcpyAsync(h->stage_a.p, desc, (size_t)total * 32, cudaMemcpyHostToDevice, st));
  ORBM_CK(cudaMemcpyAsync(h->stage_b.p, start, (size_t)(npoints + 1) * 4, cudaMemcpyHostToDevice, st));
  if (out_desc) ORBM_CK(cudaMemsetAsync(d_out, 0, (size_t)npoints * 32, st));
  ORBM_CK(cudaEventRecord(h->ev0, st));
  distinctive_kernel<<<(npoints + kWarpsPerCta - 1) / kWarpsPerCta, kWarpsPerCta * 32, 0, st>>>(
      h->stage_a.as<uint4>(), h->stage_b.as<int32_t>(), npoints, d_best, out_desc ? d_out : nullptr);
  h->launches++;
  ORBM_CK(cudaGetLastError());
  ORBM_CK(cudaEventRecord(h->ev1, st));
  ORBM_CK(cudaMemcpyAsync(best, d_best, (size_t)npoints * 4, cudaMemcpyDeviceToHost, st));
  if (out_desc) ORBM_CK(cudaMemcpyAsync(out_desc, d_out, (size_t)npoints * 32, cudaMemcpyDeviceToHost, st));
  ORBM_CK(cudaStreamSynchronize(st));
  ORBM_CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
  return ORBM_OK;
}
