// Register-only issue-rate microbenchmark of the integer / byte-SIMD instructions the extractor kernels lean on
// (DP4A, DP2A, IMAD, VIMNMX3.U16x2, VABSDIFF4, PRMT, SHF funnel, FLO, BREV, LOP3).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o int_ops_mb int_ops_microbench.cu
// Prints warp-instructions per clock per SM (4.0 = one per scheduler per clock).
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__device__ __forceinline__ unsigned op(unsigned a, unsigned b, unsigned c) {
  if (OP == 0) return __dp4a(a, b, c);
  if (OP == 1) return __dp2a_lo(a, b, c);
  if (OP == 2) return a * b + c;
  if (OP == 3) return __vimax3_u16x2(a, b, c);
  if (OP == 4) return __vabsdiffu4(a, b) ;
  if (OP == 5) return __byte_perm(a, b, 0x5432);
  if (OP == 6) return __funnelshift_r(a, b, 8);
  if (OP == 7) return (31 - __clz(a)) ^ b;   // FLO + LOP3
  if (OP == 8) return __brev(a) ^ b;         // BREV + LOP3
  if (OP == 9) { unsigned r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
  if (OP == 10) return __vminu2(a, b);
  if (OP == 11) return a + b + c;
  if (OP == 12) return __vsadu4(a, b) + c;
  return a;
}

template <int OP>
__global__ void __launch_bounds__(1024) k(unsigned* out, int iters, unsigned seed) {
  unsigned a[8];
#pragma unroll
  for (int u = 0; u < 8; u++) a[u] = (threadIdx.x + seed) * (2 * u + 3);
  const unsigned b = seed * 0x01010101u + 0x01020304u, c = seed + 7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
#pragma unroll
      for (int u = 0; u < 8; u++) a[u] = op<OP>(a[u], a[(u + 1) & 7], c);
    }
  }
  unsigned s = 0;
#pragma unroll
  for (int u = 0; u < 8; u++) s += a[u];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char* name, int sms, double mhz) {
  unsigned* out;
  cudaMalloc(&out, sms * 2 * 1024 * 4);
  const int iters = 2048;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<OP><<<sms * 2, 1024>>>(out, 64, 1);
  cudaEventRecord(e0);
  k<OP><<<sms * 2, 1024>>>(out, iters, 1);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double warp_inst = (double)sms * 2 * 32 * iters * 64;
  printf("%-18s %8.3f ms  %5.2f warp-inst/clk/SM at %.0f MHz nominal\n", name, ms,
         warp_inst / (ms * 1e-3) / sms / (mhz * 1e6), mhz);
  cudaFree(out);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double mhz = khz / 1e3;
  printf("%s, %d SMs, nominal %.0f MHz\n", p.name, p.multiProcessorCount, mhz);
  const int s = p.multiProcessorCount;
  run<0>("IDP.4A (dp4a)", s, mhz);
  run<1>("IDP.2A (dp2a)", s, mhz);
  run<2>("IMAD", s, mhz);
  run<3>("VIMNMX3.U16x2", s, mhz);
  run<4>("VABSDIFF4", s, mhz);
  run<5>("PRMT", s, mhz);
  run<6>("SHF.R funnel", s, mhz);
  run<7>("FLO+LOP3 (x2 inst)", s, mhz);
  run<8>("BREV+LOP3 (x2 inst)", s, mhz);
  run<9>("LOP3", s, mhz);
  run<10>("VIMNMX.U16x2", s, mhz);
  run<11>("IADD3", s, mhz);
  run<12>("VABSDIFF4.ACC", s, mhz);
  return 0;
}
