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
#include "cpu_common.h"

using dfcpu::parallel_ranges;

namespace {

constexpr int64_t kMaxK = 512;     // embedding sizes above this take the PyTorch expression (ops/embed_fm_cpu.py)

// rows [b0, b1) of the batch
DFM_CLONES
void fwd_rows(const float* w_rows, const float* v_rows, const int64_t* inverse, const float* x, int64_t b0, int64_t b1,
              int64_t F, int64_t K, float* y_w, float* y_v, float* e, float* s) {
  for (int64_t b = b0; b < b1; ++b) {
    float* sb = s + b * K;
    float yw = 0.f;
    // sum of squares per component, like s: a scalar running sum would be a loop-carried dependence the compiler may
    // not reorder, and the whole k loop would stay scalar
    float qk[kMaxK];
    for (int64_t k = 0; k < K; ++k) sb[k] = 0.f;
    for (int64_t k = 0; k < K; ++k) qk[k] = 0.f;
    for (int64_t f = 0; f < F; ++f) {
      const int64_t id = inverse[b * F + f];
      const float xv = x[b * F + f];
      const float* row = v_rows + id * K;
      float* eb = e + (b * F + f) * K;
      for (int64_t k = 0; k < K; ++k) {
        const float ev = row[k] * xv;
        eb[k] = ev;
        sb[k] += ev;
        qk[k] += ev * ev;
      }
      yw += w_rows[id] * xv;
    }
    float acc = 0.f;
    for (int64_t k = 0; k < K; ++k) acc += sb[k] * sb[k] - qk[k];
    y_v[b] = 0.5f * acc;
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
