// Register-only issue-rate microbenchmark for the integer instructions of the Hamming kernels (the
// per-SM POPC rate of sm_100 must be measured, not assumed).  Build: nvcc -arch=sm_100a -O3 -o popc_mb popc_microbench.cu
// Prints warp-instructions per clock per SM for POPC, LOP3 (xor), IADD3, VIMNMX and a POPC+LOP3 mix.
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(1024) k(unsigned* out, int iters, unsigned seed) {
  unsigned a0 = threadIdx.x + seed, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, a4 = a0 * 11, a5 = a0 * 13, a6 = a0 * 17, a7 = a0 * 19;
  for (int i = 0; i < iters; i++) {
#pragma unroll 8
    for (int u = 0; u < 8; u++) {
      if (OP == 0) { a0 = __popc(a0) + 1u; a1 = __popc(a1) + 1u; a2 = __popc(a2) + 1u; a3 = __popc(a3) + 1u;
                     a4 = __popc(a4) + 1u; a5 = __popc(a5) + 1u; a6 = __popc(a6) + 1u; a7 = __popc(a7) + 1u; }
      if (OP == 1) { a0 ^= a1; a1 ^= a2; a2 ^= a3; a3 ^= a4; a4 ^= a5; a5 ^= a6; a6 ^= a7; a7 ^= a0; }
      if (OP == 2) { a0 += a1 + a2; a1 += a2 + a3; a2 += a3 + a4; a3 += a4 + a5; a4 += a5 + a6; a5 += a6 + a7; a6 += a7 + a0; a7 += a0 + a1; }
      if (OP == 3) { a0 = min(a0, a1) + 0; a1 = max(a1, a2); a2 = min(a2, a3); a3 = max(a3, a4); a4 = min(a4, a5); a5 = max(a5, a6); a6 = min(a6, a7); a7 = max(a7, a0); }
      if (OP == 4) { a0 = __popc(a0 ^ a4); a1 = __popc(a1 ^ a5); a2 = __popc(a2 ^ a6); a3 = __popc(a3 ^ a7);
                     a4 ^= a0; a5 ^= a1; a6 ^= a2; a7 ^= a3; }
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

template <int OP>
void run(const char* name, double inst_per_iter, int sms, double mhz) {
  unsigned* out;
  cudaMalloc(&out, sms * 2 * 1024 * 4);
  const int iters = 4096;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<OP><<<sms * 2, 1024>>>(out, 64, 1);
  cudaEventRecord(e0);
  k<OP><<<sms * 2, 1024>>>(out, iters, 1);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double warp_inst = (double)sms * 2 * 32 * iters * 8 * inst_per_iter;
  printf("%-22s %8.3f ms  %7.2f warp-inst/us/SM  = %5.2f warp-inst/clk/SM at %.0f MHz (thread-ops/clk/SM %.1f)\n", name, ms,
         warp_inst / (ms * 1e3) / sms, warp_inst / (ms * 1e-3) / sms / (mhz * 1e6), mhz,
         32 * warp_inst / (ms * 1e-3) / sms / (mhz * 1e6));
  cudaFree(out);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double mhz = khz / 1e3;
  printf("%s, %d SMs, nominal %.0f MHz (rates below assume the nominal clock; see clocks in the bench line)\n", p.name,
         p.multiProcessorCount, mhz);
  run<0>("POPC(+IADD)", 16, p.multiProcessorCount, mhz);
  run<1>("LOP3 xor", 8, p.multiProcessorCount, mhz);
  run<2>("IADD3", 8, p.multiProcessorCount, mhz);
  run<3>("VIMNMX", 8, p.multiProcessorCount, mhz);
  run<4>("POPC+2xLOP3 mix", 12, p.multiProcessorCount, mhz);
  return 0;
}
