#include <cstdio>
#include <cstdint>
#include <vector>
extern "C" void orb_oracle_synth(uint8_t*, int, int, long, uint64_t);
__device__ __noinline__ int score_simple(const uint8_t* p, int pitch, bool pr) {
  const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
  const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  int d[25]; int v = p[0];
  for (int k = 0; k < 16; k++) d[k] = v - p[dy[k] * pitch + dx[k]];
  for (int k = 16; k < 25; k++) d[k] = d[k - 16];
  int best = -1000;
  for (int k = 0; k < 16; k++) { int mn = d[k], mx = d[k]; for (int j = 1; j < 9; j++) { mn = min(mn, d[k + j]); mx = max(mx, d[k + j]); }
    if (pr) printf("k=%d mn=%d mx=%d\n", k, mn, mx);
    best = max(best, max(mn, -mx)); }
  return max(best - 1, 0);
}
__global__ void k(const uint8_t* img, int w, int* out) {
  out[threadIdx.x] = score_simple(img + 3 * w + 191, w, threadIdx.x == 0);
}
int main() {
  const int w = 640, h = 480;
  std::vector<uint8_t> img(w * h);
  orb_oracle_synth(img.data(), w, h, w, 0);
  uint8_t* d; cudaMalloc(&d, w * h); int* o; cudaMalloc(&o, 128);
  cudaMemcpy(d, img.data(), w * h, cudaMemcpyHostToDevice);
  k<<<1, 2>>>(d, w, o); cudaDeviceSynchronize();
  int ho[2]; cudaMemcpy(ho, o, 8, cudaMemcpyDeviceToHost); printf("score %d %d\n", ho[0], ho[1]);
  return 0;
}
