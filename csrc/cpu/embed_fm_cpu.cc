// Host-side fused embedding gather + first-order term + FM second-order interaction, forward and backward
// (HVD.py:196-206), for the CPU trainers: the parameter-server workers (parallel/ps.py) and the eager back-end on a
// host without a GPU.  The GPU engine has its own kernels (csrc/cuda/embed_fm.cu); this is the same fusion for the
// mode the reference runs on CPU instances (1-ps-cpu), where stock PyTorch spends more than half of a step in a
// dozen element-wise passes over the [B, F, K] embedding block and in embedding_dense_backward.
//
// Inputs are the batch's DISTINCT table rows (w_rows[U], v_rows[U, K]) and inverse[B, F] indexing into them, so the
// backward pass yields the de-duplicated sparse gradient directly.
//
//   e[b, f, :] = v_rows[inverse[b, f], :] * x[b, f]
//   s[b, :]    = sum_f e[b, f, :]
//   y_v[b]     = 0.5 * sum_k (s[b, k]^2 - sum_f e[b, f, k]^2)          y_w[b] = sum_f w_rows[inverse[b, f]] * x[b, f]
//
//   de[b, f, k]                 = g_e[b, f, k] + g_yv[b] * (s[b, k] - e[b, f, k])
//   g_v_rows[inverse[b, f], k] += de[b, f, k] * x[b, f]                g_w_rows[inverse[b, f]] += g_yw[b] * x[b, f]
//
// One pass over the block each way; the K loops vectorise (the hot functions are compiled for AVX2 + FMA as well as
// for the baseline ISA and picked at load time).
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#if defined(__GNUC__) && defined(__x86_64__)
#define DFM_CLONES __attribute__((target_clones("arch=haswell", "default")))
#else
#define DFM_CLONES
#endif

namespace {

// rows [b0, b1) of the batch
DFM_CLONES
void fwd_rows(const float* w_rows, const float* v_rows, const int64_t* inverse, const float* x, int64_t b0, int64_t b1,
              int64_t F, int64_t K, float* y_w, float* y_v, float* e, float* s) {
  for (int64_t b = b0; b < b1; ++b) {
    float* sb = s + b * K;
    float qsum = 0.f, yw = 0.f;
    for (int64_t k = 0; k < K; ++k) sb[k] = 0.f;
    for (int64_t f = 0; f < F; ++f) {
      const int64_t id = inverse[b * F + f];
      const float xv = x[b * F + f];
      const float* row = v_rows + id * K;
      float* eb = e + (b * F + f) * K;
      float q = 0.f;
      for (int64_t k = 0; k < K; ++k) {
        const float ev = row[k] * xv;
        eb[k] = ev;
        sb[k] += ev;
        q += ev * ev;
      }
      qsum += q;
      yw += w_rows[id] * xv;
    }
    float ss = 0.f;
    for (int64_t k = 0; k < K; ++k) ss += sb[k] * sb[k];
    y_v[b] = 0.5f * (ss - qsum);
    y_w[b] = yw;
  }
}

// the distinct rows [id_lo, id_hi): every thread scans all (b, f) pairs and accumulates only into its own rows, so the
// scatter-add needs no atomics and the sums are formed in batch order whatever the thread count
DFM_CLONES
void bwd_ids(const int64_t* inverse, const float* x, const float* e, const float* s, const float* g_yw,
             const float* g_yv, const float* g_e, int64_t B, int64_t F, int64_t K, int64_t id_lo, int64_t id_hi,
             float* g_w_rows, float* g_v_rows) {
  for (int64_t b = 0; b < B; ++b) {
    const float gw = g_yw ? g_yw[b] : 0.f;
    const float gv = g_yv ? g_yv[b] : 0.f;
    const float* sb = s + b * K;
    for (int64_t f = 0; f < F; ++f) {
      const int64_t id = inverse[b * F + f];
      if (id < id_lo || id >= id_hi) continue;
      const float xv = x[b * F + f];
      const float* eb = e + (b * F + f) * K;
      float* out = g_v_rows + id * K;
      if (g_e) {
        const float* ge = g_e + (b * F + f) * K;
        for (int64_t k = 0; k < K; ++k) out[k] += (ge[k] + gv * (sb[k] - eb[k])) * xv;
      } else {
        for (int64_t k = 0; k < K; ++k) out[k] += gv * (sb[k] - eb[k]) * xv;
      }
      g_w_rows[id] += gw * xv;
    }
  }
}

template <typename Fn>
void parallel_ranges(int64_t n, int threads, int64_t min_per_thread, Fn fn) {
  int t = threads < 1 ? 1 : threads;
  if (n / t < min_per_thread) t = static_cast<int>(n / min_per_thread);
  if (t <= 1) {
    fn(int64_t(0), n);
    return;
  }
  std::vector<std::thread> pool;
  const int64_t step = (n + t - 1) / t;
  for (int i = 1; i < t; ++i) {
    const int64_t lo = i * step, hi = lo + step < n ? lo + step : n;
    if (lo < hi) pool.emplace_back([=] { fn(lo, hi); });
  }
  fn(int64_t(0), step < n ? step : n);
  for (auto& th : pool) th.join();
}

}  // namespace

extern "C" {

// `threads`: how many threads may be used (the caller passes torch.get_num_threads(); small problems stay on one)
void dfcpu_embed_fm_fwd(const float* w_rows, const float* v_rows, const int64_t* inverse, const float* x, int64_t B,
                        int64_t F, int64_t K, float* y_w, float* y_v, float* e, float* s, int threads) {
  parallel_ranges(B, threads < 8 ? threads : 8, 256, [=](int64_t b0, int64_t b1) {
    fwd_rows(w_rows, v_rows, inverse, x, b0, b1, F, K, y_w, y_v, e, s);
  });
}

// g_w_rows[U] and g_v_rows[U, K] must be zero-filled by the caller; g_yw / g_yv / g_e may be null (no gradient)
void dfcpu_embed_fm_bwd(const int64_t* inverse, const float* x, const float* e, const float* s, const float* g_yw,
                        const float* g_yv, const float* g_e, int64_t B, int64_t F, int64_t K, int64_t U, float* g_w_rows,
                        float* g_v_rows, int threads) {
  // every thread re-reads all (b, f) pairs, so more than a few threads only add scanning (measured: 4.6 / 3.7 / 2.9 ms
  // with 1 / 2 / 4 threads at B = 2048, slower again with 8)
  const int t = B * F >= 32768 ? (threads < 4 ? threads : 4) : 1;
  parallel_ranges(U, t, 64, [=](int64_t lo, int64_t hi) {
    bwd_ids(inverse, x, e, s, g_yw, g_yv, g_e, B, F, K, lo, hi, g_w_rows, g_v_rows);
  });
}

int dfcpu_version() { return 2; }

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// The fused engine's dropout stream on the host (csrc/cuda/common.cuh: philox4x32, dropout_keep_bits8): keep decision
// of activation (row, column) after hidden layer `layer` at 1-based optimizer step `step`, one Philox4x32-10 block
// per 8 columns.  out[rows * cols] = 1 (kept) or 0.  PyTorch's own CPU dropout draws one scalar Mersenne-twister
// Bernoulli per element (3.4 ms for a [1024, 128] activation); this takes a tenth of that and gives the CPU
// trainers the very masks the GPU engine uses for the same seed and step.
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

// ---------------------------------------------------------------------------------------------------------------
// tf.train.AdamOptimizer's update (HVD.py:282-284) on the host, one pass over the data instead of the nine tensor
// ops of the eager formula: what the parameter servers run for every pushed gradient, and the CPU trainer for its own.
//     m <- b1 m + (1 - b1) g;   v <- b2 v + (1 - b2) g^2;   p <- p - lr_t m / (sqrt(v) + eps)
// lr_t = lr sqrt(1 - b2^t) / (1 - b1^t) is computed by the caller (utils/optimizers.py).
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
