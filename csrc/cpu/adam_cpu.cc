// tf.train.AdamOptimizer's update (HVD.py:282-284) on the host, one pass over the data instead of the nine tensor
// ops of the eager formula: what the parameter servers run for every pushed gradient, and the CPU trainer for its own.
//     m <- b1 m + (1 - b1) g;   v <- b2 v + (1 - b2) g^2;   p <- p - lr_t m / (sqrt(v) + eps)
// lr_t = lr sqrt(1 - b2^t) / (1 - b1^t) is computed by the caller (utils/optimizers.py).
#include "cpu_common.h"

using dfcpu::parallel_ranges;

namespace {

DFM_CLONES
void adam_span(float* p, float* m, float* v, const float* g, int64_t n, float lr_t, float b1, float b2, float eps) {
  const float c1 = 1.f - b1, c2 = 1.f - b2;
  for (int64_t i = 0; i < n; ++i) {
    const float gi = g[i];
    const float mi = b1 * m[i] + c1 * gi;
    const float vi = b2 * v[i] + c2 * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= lr_t * mi / (__builtin_sqrtf(vi) + eps);
  }
}

}  // namespace

extern "C" {

void dfcpu_adam_dense(float* p, float* m, float* v, const float* g, int64_t n, float lr_t, float b1, float b2, float eps,
                      int threads) {
  parallel_ranges(n, threads < 8 ? threads : 8, 1 << 16, [=](int64_t lo, int64_t hi) {
    adam_span(p + lo, m + lo, v + lo, g + lo, hi - lo, lr_t, b1, b2, eps);
  });
}

// lazy sparse update: rows[] are DISTINCT row indices of p / m / v [*, K]; g is [n_rows, K]
void dfcpu_adam_rows(float* p, float* m, float* v, const int64_t* rows, const float* g, int64_t n_rows, int64_t K,
                     float lr_t, float b1, float b2, float eps, int threads) {
  parallel_ranges(n_rows, threads < 8 ? threads : 8, 2048, [=](int64_t lo, int64_t hi) {
    for (int64_t r = lo; r < hi; ++r) {
      const int64_t o = rows[r] * K;
      adam_span(p + o, m + o, v + o, g + r * K, K, lr_t, b1, b2, eps);
    }
  });
}

}  // extern "C"
