// Stand-alone probe of knn2_tc_kernel against knn2_kernel<4,2> (same part/merge format).  argv: nq ndb [chunks] [reps]
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -I orb_slam_b200/csrc tools/scratch/tc_probe.cu -o gpurun_out/tc_probe
#include <climits>
#include <cuda_fp16.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <vector>
#include "orbm_internal.cuh"
using namespace orbm;
#include "orbm_knn_kernels.inl"
#include "orbm_knn_tc.inl"
#ifndef TC_DBG
#define TC_DBG 0
#endif
#ifndef TC_CTAS
#define TC_CTAS 2
#endif

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

int main(int argc, char** argv) {
  const int nq = argc > 1 ? atoi(argv[1]) : 1000;
  const int64_t ndb = argc > 2 ? atoll(argv[2]) : 100000;
  int want_chunks = argc > 3 ? atoi(argv[3]) : 0;
  const int reps = argc > 4 ? atoi(argv[4]) : 3;
  std::vector<uint32_t> hq((size_t)nq * 8), hdb((size_t)ndb * 8);
  uint64_t st = 88172645463325252ull;
  auto rnd = [&]() { st ^= st << 13; st ^= st >> 7; st ^= st << 17; return (uint32_t)(st >> 16); };
  for (auto& v : hdb) v = rnd();
  for (int i = 0; i < nq; i++)
    for (int k = 0; k < 8; k++) {
      uint32_t v = hdb[(size_t)((int64_t)i * 7919 % ndb) * 8 + k];
      if (i % 3 != 0) for (int b = 0; b < 3; b++) v ^= 1u << (rnd() & 31);   // planted near-duplicates; every third query is an exact copy
      hq[(size_t)i * 8 + k] = v;
    }
  uint32_t *dq, *ddb, *part, *out_a, *out_b;
  int* derr;
  CK(cudaMalloc(&dq, hq.size() * 4)); CK(cudaMalloc(&ddb, hdb.size() * 4));
  CK(cudaMemcpy(dq, hq.data(), hq.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(ddb, hdb.data(), hdb.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&derr, 64)); CK(cudaMemset(derr, 0, 64));
  CK(cudaMalloc(&out_a, (size_t)nq * 8)); CK(cudaMalloc(&out_b, (size_t)nq * 8));
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);

  // ---- reference: knn2_kernel<4,2>
  {
    const int qpt = 4, qtiles = (nq + kKnnThreads * qpt - 1) / (kKnnThreads * qpt);
    int64_t want = std::max<int64_t>(1, (2 * (int64_t)sms + qtiles - 1) / qtiles);
    int64_t cl = (ndb + want - 1) / want;
    cl = std::max<int64_t>((cl + kKnnTile - 1) / kKnnTile * kKnnTile, 8 * kKnnTile);
    const int nch = (int)((ndb + cl - 1) / cl);
    CK(cudaMalloc(&part, (size_t)nch * nq * 8 + 16));
    float best_ms = 1e9f;
    for (int r = 0; r < reps; r++) {
      cudaEventRecord(e0);
      knn2_kernel<4, 2><<<dim3(nch, qtiles), kKnnThreads>>>((const uint4*)dq, nq, (const uint4*)ddb, ndb, (int)cl, 0u, part, derr);
      merge_top2_kernel<<<(nq + 255) / 256, 256>>>((const uint2*)part, nch, nq, 0, 0.f, (uint2*)out_a, nullptr);
      cudaEventRecord(e1);
      CK(cudaDeviceSynchronize());
      float ms; cudaEventElapsedTime(&ms, e0, e1); best_ms = std::min(best_ms, ms);
    }
    printf("popc  kernel: %.3f ms  %.1f Gpairs/s (chunks %d x qtiles %d)\n", best_ms, (double)nq * ndb / best_ms / 1e6, nch, qtiles);
    CK(cudaFree(part));
  }
  // ---- tensor-core kernel
  {
    const int npairs = ((nq + kTcQ - 1) / kTcQ + 1) / 2;
    const int ntiles = (int)((ndb + kTcN - 1) / kTcN);
    const int64_t total = (int64_t)npairs * ntiles;
    int ncl = want_chunks > 0 ? want_chunks : sms / 2;
    ncl = (int)std::max<int64_t>(1, std::min<int64_t>(ncl, total));
    const int64_t per = (total + ncl - 1) / ncl;
    ncl = (int)((total + per - 1) / per);
    const int pieces = (int)((ntiles + per - 1) / per) + 1;
    CK(cudaMalloc(&part, (size_t)pieces * nq * 8 + 16));
    CK(cudaFuncSetAttribute(knn2_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
    float best_ms = 1e9f;
    for (int r = 0; r < reps; r++) {
      cudaEventRecord(e0);
      CK(cudaMemsetAsync(part, 0xFF, (size_t)pieces * nq * 8));
      {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * ncl); cfg.blockDim = dim3(kTcThreads); cfg.dynamicSmemBytes = kTcSmemBytes; cfg.stream = 0;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        CK(cudaLaunchKernelEx(&cfg, knn2_tc_kernel, (const uint4*)dq, nq, (const uint4*)ddb, ndb, ntiles, per, total, 0u, part, derr));
      }
      merge_top2_kernel<<<(nq + 255) / 256, 256>>>((const uint2*)part, pieces, nq, 0, 0.f, (uint2*)out_b, nullptr);
      cudaEventRecord(e1);
      CK(cudaGetLastError());
      CK(cudaDeviceSynchronize());
      float ms; cudaEventElapsedTime(&ms, e0, e1); best_ms = std::min(best_ms, ms);
    }
    printf("tcgen05 kernel: %.3f ms  %.1f Gpairs/s (%d clusters x %lld tile steps, %d pairs x %d tiles, %d pieces)\n", best_ms,
           (double)nq * ndb / best_ms / 1e6, ncl, (long long)per, npairs, ntiles, pieces);
  }
  int herr = 0; CK(cudaMemcpy(&herr, derr, 4, cudaMemcpyDeviceToHost));
  std::vector<uint32_t> a((size_t)nq * 2), b((size_t)nq * 2);
  CK(cudaMemcpy(a.data(), out_a, a.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(b.data(), out_b, b.size() * 4, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int i = 0; i < nq; i++)
    if (a[2 * i] != b[2 * i] || a[2 * i + 1] != b[2 * i + 1]) {
      if (bad < 8) printf("  q %d: popc (%u,%u | %u,%u)  tc (%u,%u | %u,%u)\n", i, a[2 * i] >> 23, a[2 * i] & 0x7fffff, a[2 * i + 1] >> 23, a[2 * i + 1] & 0x7fffff,
                          b[2 * i] >> 23, b[2 * i] & 0x7fffff, b[2 * i + 1] >> 23, b[2 * i + 1] & 0x7fffff);
      bad++;
    }
  printf("err flag %d, mismatching queries %d of %d\n", herr, bad, nq);
  return bad || herr ? 2 : 0;
}
