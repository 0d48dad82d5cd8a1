// Probe 2: matrix of tensor-map shapes.  argv: rank bw bh x y
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
template <int RANK>
__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tm, int bytes, int x, int y, int f, uint8_t* out) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"((uint32_t)bytes) : "memory");
    if (RANK == 2)
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                   ::"r"((uint32_t)__cvta_generic_to_shared(sm)), "l"(reinterpret_cast<uint64_t>(&tm)), "r"(x), "r"(y), "r"(b) : "memory");
    else
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                   ::"r"((uint32_t)__cvta_generic_to_shared(sm)), "l"(reinterpret_cast<uint64_t>(&tm)), "r"(x), "r"(y), "r"(f), "r"(b) : "memory");
  }
  uint32_t ok = 0;
  for (int spin = 0; spin < (1 << 22) && !ok; ++spin)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(b) : "memory");
  for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = ok ? sm[i] : 0xEE;
}
int main(int argc, char** argv) {
  const int rank = atoi(argv[1]), bw = atoi(argv[2]), bh = atoi(argv[3]), x = atoi(argv[4]), y = atoi(argv[5]);
  const int W = 640, H = 480, F = 3, P = 640, f = 1;
  std::vector<uint8_t> img((size_t)P * H * F);
  for (size_t i = 0; i < img.size(); i++) img[i] = (uint8_t)((i * 2654435761u) >> 24);
  uint8_t *d, *o;
  cudaMalloc(&d, img.size()); cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice);
  cudaMalloc(&o, bw * bh);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  alignas(64) CUtensorMap m;
  cuuint64_t dims[3] = {W, H, F}, st[2] = {P, (cuuint64_t)P * H};
  cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
  CUresult r = ((Enc)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, rank == 2 ? d + (size_t)f * P * H : d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(probe<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  if (rank == 2) probe<2><<<1, 128, bw * bh + 128>>>(m, bw * bh, x, y, f, o);
  else probe<3><<<1, 128, bw * bh + 128>>>(m, bw * bh, x, y, f, o);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<uint8_t> h(bw * bh);
  cudaMemcpy(h.data(), o, h.size(), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int rr = 0; rr < bh; rr++) for (int c = 0; c < bw; c++) {
    const int gx = x + c, gy = y + rr;
    const uint8_t exp = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? img[(size_t)f * P * H + (size_t)gy * P + gx] : 0;
    bad += h[rr * bw + c] != exp;
  }
  printf("rank %d box %dx%d at (%d,%d): encode %d, %s, mismatches %d\n", rank, bw, bh, x, y, (int)r, cudaGetErrorString(e), bad);
  return 0;
}
