// ORACLE — TEST INFRASTRUCTURE ONLY (see orb_oracle.cc header).
// CPU twin of the product's device-side synthetic frame generator (orb_slam_b200/csrc/synth.cu); the two must be
// byte-identical (tests/test_synth.py).  Integer-only restatement of SURVEY.md §8(d)'s generator: grey canvas,
// w*h/768 random rectangles painted in order, a sigma~1 binomial blur, +-2 uniform noise.
#include <cstdint>
#include <cstring>
#include <vector>

namespace {
inline uint64_t mix64(uint64_t seed, uint64_t i) {
  uint64_t z = seed * 0xD1342543DE82EF95ull + (i + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
inline int refl(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * (n - 1) - i : i); }
}  // namespace

extern "C" void orb_oracle_synth(uint8_t* out, int w, int h, long stride, uint64_t seed) {
  const int nrect = w * h / 768;
  std::vector<uint8_t> canvas((size_t)w * h, 128);
  for (int r = 0; r < nrect; r++) {
    uint64_t u = mix64(seed, (uint64_t)r);
    int x0 = (int)(((u & 0xFFFF) * (uint64_t)w) >> 16);
    int y0 = (int)((((u >> 16) & 0xFFFF) * (uint64_t)h) >> 16);
    int ww = 8 + (int)((((u >> 32) & 0xFFFF) * 88ull) >> 16);
    int hh = 8 + (int)((((u >> 48) & 0xFFFF) * 88ull) >> 16);
    uint8_t val = (uint8_t)(mix64(seed, 0x10000ull + r) & 255);
    int x1 = x0 + ww < w ? x0 + ww : w, y1 = y0 + hh < h ? y0 + hh : h;
    for (int y = y0; y < y1; y++) memset(&canvas[(size_t)y * w + x0], val, x1 - x0);
  }
  static const int k[5] = {1, 4, 6, 4, 1};
  std::vector<int> hb((size_t)w * h);
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) {
      int s = 0;
      for (int i = 0; i < 5; i++) s += k[i] * canvas[(size_t)y * w + refl(x + i - 2, w)];
      hb[(size_t)y * w + x] = s;
    }
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) {
      int s = 0;
      for (int j = 0; j < 5; j++) s += k[j] * hb[(size_t)refl(y + j - 2, h) * w + x];
      int v = (s + 128) >> 8;
      int n = (int)(mix64(seed, 0x1000000ull + (uint64_t)y * w + x) % 5) - 2;
      v += n;
      out[(size_t)y * stride + x] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
    }
}
