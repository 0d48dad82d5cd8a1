// Probe: 3-D byte tensor TMA load with the descriptor passed as a __grid_constant__ struct member (dynamic index).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cstdlib>
struct Maps { CUtensorMap map[16]; int bw[16], bh[16]; int smem; };
template <int V>
__global__ void __launch_bounds__(256) probe(const __grid_constant__ Maps tm, const __grid_constant__ CUtensorMap one, const CUtensorMap* gmap, const int* lvl, int x, int y, int f, uint8_t* out) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ __align__(8) uint64_t bar;
  const int level = lvl[0];
  const int bw = tm.bw[level], bh = tm.bh[level];
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"((uint32_t)(bw * bh)) : "memory");
    const uint64_t desc = V == 2 ? reinterpret_cast<uint64_t>(&one) : (V == 3 ? reinterpret_cast<uint64_t>(gmap) : reinterpret_cast<uint64_t>(&tm.map[level]));
    if (V == 1)
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"((uint32_t)__cvta_generic_to_shared(sm)), "l"(desc), "r"(x), "r"(y), "r"(f), "r"(b) : "memory");
    else
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"((uint32_t)__cvta_generic_to_shared(sm)), "l"(desc), "r"(x), "r"(y), "r"(f), "r"(b) : "memory");
  }
  __syncthreads();
  uint32_t ok = 0;
  for (int spin = 0; spin < (1 << 22) && !ok; ++spin)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(b) : "memory");
  for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = ok ? sm[i] : 0xEE;
}
int main(int argc, char** argv) {
  const int V = argc > 1 ? atoi(argv[1]) : 0;
  const int W = 640, H = 480, F = 3, P = 640;
  std::vector<uint8_t> img((size_t)P * H * F);
  for (size_t i = 0; i < img.size(); i++) img[i] = (uint8_t)((i * 2654435761u) >> 24);
  uint8_t *d, *o; int* dl;
  cudaMalloc(&d, img.size()); cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice);
  const int bw = 144, bh = 81;
  cudaMalloc(&o, bw * bh); cudaMalloc(&dl, 4);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  Maps* m = new Maps();
  cuuint64_t dims[3] = {W, H, F}, st[2] = {P, (cuuint64_t)P * H};
  cuuint32_t box[3] = {bw, bh, 1}, es[3] = {1, 1, 1};
  for (int level = 0; level < 3; level++) {
    CUresult r = ((Enc)fn)(&m->map[level], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode level %d -> %d (q %d)\n", level, (int)r, (int)q);
    m->bw[level] = bw; m->bh[level] = bh;
  }
  int lv = 2; cudaMemcpy(dl, &lv, 4, cudaMemcpyHostToDevice);
  CUtensorMap* gm; cudaMalloc(&gm, sizeof(CUtensorMap)); cudaMemcpy(gm, &m->map[2], sizeof(CUtensorMap), cudaMemcpyHostToDevice);
  for (int t = 0; t < 3; t++) {
    const int x = t == 0 ? 11 : (t == 1 ? 560 : -5), y = t == 0 ? 13 : (t == 1 ? 430 : -2), f = t;
    if (V == 0) probe<0><<<1, 256, bw * bh + 128>>>(*m, m->map[2], gm, dl, x, y, f, o);
    if (V == 1) probe<1><<<1, 256, bw * bh + 128>>>(*m, m->map[2], gm, dl, x, y, f, o);
    if (V == 2) probe<2><<<1, 256, bw * bh + 128>>>(*m, m->map[2], gm, dl, x, y, f, o);
    if (V == 3) probe<3><<<1, 256, bw * bh + 128>>>(*m, m->map[2], gm, dl, x, y, f, o);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<uint8_t> h(bw * bh);
    cudaMemcpy(h.data(), o, h.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r = 0; r < bh; r++) for (int c = 0; c < bw; c++) {
      const int gx = x + c, gy = y + r;
      const uint8_t exp = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? img[(size_t)f * P * H + (size_t)gy * P + gx] : 0;
      bad += h[r * bw + c] != exp;
    }
    printf("case %d: %s, mismatches %d\n", t, cudaGetErrorString(e), bad);
  }
  return 0;
}
