// Optimizer kernels (tf.train.{Adam,Adagrad,Momentum,Ftrl,GradientDescent}Optimizer,
// HVD.py:281-294) for the two kinds of state:
//
//   dense_opt_kernel     all MLP / out-layer / fm_bias parameters as one flat fp32 vector:
//                        apply, refresh the bf16 tensor-core copy, zero the gradient.
//   sparse_apply_kernel  "lazy" update of fm_v / fm_w: one warp per distinct id touched by the
//                        batch; reads the reduced row gradient from the slot buffer, adds the
//                        L2 term of the touched row, applies, clears the slot bookkeeping.
//   table_dense_kernel   "dense_exact" update: every row every step with g = sparse + l2*row,
//                        which is the math the reference's l2_loss(whole table) implies
//                        (HVD.py:266-267).
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

// `step` is the number of COMPLETED optimizer steps; the first kernel of a training step bumps it, so
// every later kernel of that step reads the 1-based index of the update in flight.
__global__ void set_state_kernel(StepState* st, int count, float lr, int train, int step) {
  st->count = count;
  st->lr = lr;
  st->train = train;
  if (step >= 0) st->step = step;
}

// n = number of flat dense parameters; element n-1 is fm_bias whose gradient equals that of
// b_out (element n-2): y = fm_bias + ... + b_out.
__global__ void __launch_bounds__(256)
dense_opt_kernel(float* __restrict__ P, float* __restrict__ s0, float* __restrict__ s1,
                 const float* __restrict__ part, int ksplit, size_t part_stride,
                 __nv_bfloat16* __restrict__ Wb, int n,
                 const StepState* __restrict__ st, const OptHyper hyper, float grad_scale, int wb_skip_lo,
                 int wb_skip_hi) {
  pdl_wait();
  const OptStepConsts c = make_opt_consts(hyper, st);
  kstamp(st, kStampDenseOpt);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n - 1; i += gridDim.x * blockDim.x) {
    float g = 0.f;
    for (int k = 0; k < ksplit; ++k) g += __ldg(part + size_t(k) * part_stride + i);   // fixed order
    g *= grad_scale;
    float p = P[i], a = s0[i], b = s1[i];
    opt_apply(c, g, p, a, b);
    P[i] = p; s0[i] = a; s1[i] = b;
    // [wb_skip_lo, wb_skip_hi) = W_0, whose bf16 copy embed_bwd may still be reading: embed_fm_fwd refreshes it
    if (i < wb_skip_lo || i >= wb_skip_hi) Wb[i] = __float2bfloat16(p);
    if (i == n - 2) {   // fm_bias shares b_out's gradient
      float p2 = P[n - 1], a2 = s0[n - 1], b2 = s1[n - 1];
      opt_apply(c, g, p2, a2, b2);
      P[n - 1] = p2; s0[n - 1] = a2; s1[n - 1] = b2;
    }
  }
  kstamp(st, kStampDenseEnd);
}

__global__ void cast_weights_kernel(const float* __restrict__ P, __nv_bfloat16* __restrict__ Wb, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    Wb[i] = __float2bfloat16(P[i]);
}


// Ke/4 lanes per touched row (four columns each, float4 accesses), so a warp covers 32/(Ke/4) rows per instruction
// and, at <= 51 registers, an SM keeps 160 rows in flight -- the warp-per-row version it replaces needed 76 registers
// for two rows per warp (54 rows per SM, 2.7 waves of CTAs for one Criteo batch).  A row is a dependent chain
// (id -> table row), so what counts is how many chains are in flight.
__global__ void __launch_bounds__(256, 5)
sparse_apply_kernel(const SparseArgs a, const OptHyper hyper) {
  pdl_wait();                       // launched behind embed_bwd with programmatic serialization
  const OptStepConsts c = make_opt_consts(hyper, a.st);
  const int n = a.st->n_unique;
  kstamp(a.st, kStampSparse);
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const int Ke = a.Ke, GP = g_pitch(a.Ke);
  const int LPR = Ke >> 2, G = 32 / LPR;
  const int g = lane / LPR, sub = lane - g * LPR;
  if (g < G) {
    for (int s = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * G + g; s < n; s += warps * G) {
      const int id = a.uniq_ids[s];
      float* grow = a.G + size_t(s) * GP;
      const size_t pi = size_t(id) * Ke + 4 * sub;
      const float4 gr = *reinterpret_cast<const float4*>(grow + 4 * sub);
      float4 p = *reinterpret_cast<const float4*>(a.fm_v + pi);
      float4 m = *reinterpret_cast<const float4*>(a.v_s0 + pi);
      float4 v = *reinterpret_cast<const float4*>(a.v_s1 + pi);
      float gw = 0.f, pw = 0.f, mw = 0.f, vw = 0.f;
      if (sub == 0) { gw = grow[Ke]; pw = a.fm_w[id]; mw = a.w_s0[id]; vw = a.w_s1[id]; }
      *reinterpret_cast<float4*>(grow + 4 * sub) = make_float4(0.f, 0.f, 0.f, 0.f);
      opt_apply(c, gr.x * a.grad_scale + hyper.l2_reg * p.x, p.x, m.x, v.x);
      opt_apply(c, gr.y * a.grad_scale + hyper.l2_reg * p.y, p.y, m.y, v.y);
      opt_apply(c, gr.z * a.grad_scale + hyper.l2_reg * p.z, p.z, m.z, v.z);
      opt_apply(c, gr.w * a.grad_scale + hyper.l2_reg * p.w, p.w, m.w, v.w);
      *reinterpret_cast<float4*>(a.fm_v + pi) = p;
      *reinterpret_cast<float4*>(a.v_s0 + pi) = m;
      *reinterpret_cast<float4*>(a.v_s1 + pi) = v;
      if (sub == 0) {
        grow[Ke] = 0.f;
        opt_apply(c, gw * a.grad_scale + hyper.l2_reg * pw, pw, mw, vw);
        a.fm_w[id] = pw; a.w_s0[id] = mw; a.w_s1[id] = vw;
        a.slot_of[id] = -1;
      }
    }
  }
  kstamp(a.st, kStampSparseEnd);
}

// dense_exact: one warp per table row, all rows.
__global__ void __launch_bounds__(256)
table_dense_kernel(const SparseArgs a, const OptHyper hyper) {
  const OptStepConsts c = make_opt_consts(hyper, a.st);
  const int lane = threadIdx.x & 31;
  const long long warps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  for (long long id = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; id < a.V; id += warps) {
    const int s = slot_index(a.slot_of[id]);   // -1 when the id is not in the batch (slot_index(-1) < 0)
    for (int k = lane; k < a.Ke; k += 32) {
      const size_t pi = size_t(id) * a.Ke + k;
      float p = a.fm_v[pi], m = a.v_s0[pi], v = a.v_s1[pi];
      float g = hyper.l2_reg * p;
      if (s >= 0) {
        g += a.G[size_t(s) * g_pitch(a.Ke) + k] * a.grad_scale;
        a.G[size_t(s) * g_pitch(a.Ke) + k] = 0.f;
      }
      opt_apply(c, g, p, m, v);
      a.fm_v[pi] = p; a.v_s0[pi] = m; a.v_s1[pi] = v;
    }
    __syncwarp();
    if (lane == 0) {
      float p = a.fm_w[id], m = a.w_s0[id], v = a.w_s1[id];
      float g = hyper.l2_reg * p;
      if (s >= 0) {
        g += a.G[size_t(s) * g_pitch(a.Ke) + a.Ke] * a.grad_scale;
        a.G[size_t(s) * g_pitch(a.Ke) + a.Ke] = 0.f;
        a.slot_of[id] = -1;
      }
      opt_apply(c, g, p, m, v);
      a.fm_w[id] = p; a.w_s0[id] = m; a.w_s1[id] = v;
    }
  }
}

// dense_exact on several GPUs: the owner of each touched row has published the row's gradient summed over all ranks
// into every replica's `gsum` table ([V, g_pitch], zero elsewhere); every replica then sweeps its whole table with
// g = gsum / N + l2 * row -- the reference's math (HVD.py:266-267) -- and clears what it consumed.
__global__ void __launch_bounds__(256)
table_dense_from_kernel(const SparseArgs a, const OptHyper hyper, float* __restrict__ gsum, float inv_n) {
  const OptStepConsts c = make_opt_consts(hyper, a.st);
  const int lane = threadIdx.x & 31;
  const int gp = g_pitch(a.Ke);
  const long long warps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  for (long long id = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; id < a.V; id += warps) {
    float* grow = gsum + size_t(id) * gp;
    for (int k = lane; k < a.Ke; k += 32) {
      const size_t pi = size_t(id) * a.Ke + k;
      float p = a.fm_v[pi], m = a.v_s0[pi], v = a.v_s1[pi];
      const float gs = grow[k];
      if (gs != 0.f) grow[k] = 0.f;
      opt_apply(c, hyper.l2_reg * p + gs * inv_n, p, m, v);
      a.fm_v[pi] = p; a.v_s0[pi] = m; a.v_s1[pi] = v;
    }
    if (lane == 0) {
      float p = a.fm_w[id], m = a.w_s0[id], v = a.w_s1[id];
      const float gs = grow[a.Ke];
      if (gs != 0.f) grow[a.Ke] = 0.f;
      opt_apply(c, hyper.l2_reg * p + gs * inv_n, p, m, v);
      a.fm_w[id] = p; a.w_s0[id] = m; a.w_s1[id] = v;
    }
  }
}

// ---- streaming AUC accumulators, tf.metrics.auc(num_thresholds=200) (HVD.py:271) ----
// bucket(p) = number of thresholds t_j with p > t_j, thresholds = {-eps, 1/199 .. 198/199, 1+eps}.
// hist[0][k] counts positives, hist[1][k] negatives; TP_j = sum_{k>j} hist[0][k].
__global__ void __launch_bounds__(256)
auc_hist_kernel(const float* __restrict__ prob, const float* __restrict__ label, int n,
                unsigned long long* __restrict__ hist, int num_thresholds) {
  extern __shared__ unsigned int sh[];   // 2 * (num_thresholds + 1)
  const int nb = num_thresholds + 1;
  for (int i = threadIdx.x; i < 2 * nb; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const float eps = 1e-7f;
  const float scale = float(num_thresholds - 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float pr = prob[i];
    // thresholds: t_0 = -eps, t_j = j/(T-1) for 1 <= j <= T-2, t_{T-1} = 1+eps
    int k;
    if (!(pr > -eps)) k = 0;
    else if (pr > 1.f + eps) k = num_thresholds;
    else {
      int j = int(ceilf(pr * scale)) - 1;          // largest j with j/(T-1) < pr, approximately
      j = max(0, min(num_thresholds - 2, j));
      while (j + 1 <= num_thresholds - 2 && pr > float(j + 1) / scale) ++j;
      while (j >= 1 && !(pr > float(j) / scale)) --j;
      k = j + 1;                                   // t_0 .. t_j are below pr
    }
    atomicAdd(&sh[(label[i] > 0.5f ? 0 : nb) + k], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * nb; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], static_cast<unsigned long long>(sh[i]));
}

// ---------------------------------------------------------------------------------------------
int launch_set_state(StepState* st, int count, float lr, int train, int step, cudaStream_t stream) {
  set_state_kernel<<<1, 1, 0, stream>>>(st, count, lr, train, step);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_dense_opt(float* P, float* s0, float* s1, const float* part, int ksplit, size_t part_stride,
                     __nv_bfloat16* Wb, int n, const StepState* st, const OptHyper& h, float grad_scale,
                     int wb_skip_lo, int wb_skip_hi, cudaStream_t stream) {
  const int grid = std::min(148 * 8, ceil_div(n, 256));
  DFM_CUDA_CHECK(launch_pdl(kPdlDense, dense_opt_kernel, dim3(grid), dim3(256), 0, stream, P, s0, s1, part, ksplit, part_stride, Wb,
                            n, st, h, grad_scale, wb_skip_lo, wb_skip_hi));
  return DFM_OK;
}

int launch_cast_weights(const float* P, __nv_bfloat16* Wb, int n, cudaStream_t stream) {
  cast_weights_kernel<<<std::min(148 * 4, ceil_div(n, 256)), 256, 0, stream>>>(P, Wb, n);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_table_dense_from(const SparseArgs& a, const OptHyper& h, float* gsum, float inv_n, cudaStream_t stream) {
  DFM_REQUIRE(gsum != nullptr, "dense_exact on several GPUs: no summed-gradient table");
  const int grid = int(std::min<long long>(148 * 8, (a.V + 7) / 8));
  table_dense_from_kernel<<<grid, 256, 0, stream>>>(a, h, gsum, inv_n);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_sparse_apply(const SparseArgs& a, const OptHyper& h, int max_rows, bool dense_exact, cudaStream_t stream) {
  if (dense_exact) {
    const long long warps_needed = a.V;
    const int grid = int(std::min<long long>(148 * 8, (warps_needed + 7) / 8));
    table_dense_kernel<<<grid, 256, 0, stream>>>(a, h);
  } else {
    // one CTA covers 8 warps x 32/(Ke/4) rows per pass; 5 CTAs fit an SM
    const int grid = std::max(1, std::min(148 * 5, ceil_div(max_rows, 8 * (32 / (a.Ke / 4)))));
    DFM_CUDA_CHECK(launch_pdl(kPdlSparse, sparse_apply_kernel, dim3(grid), dim3(256), 0, stream, a, h));
  }
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm

extern "C" int dfm_auc_hist(const float* prob, const float* label, int n, unsigned long long* hist,
                            int num_thresholds, dfm_stream_t stream) {
  using namespace dfm;
  DFM_REQUIRE(num_thresholds >= 3 && num_thresholds <= 4096, "dfm_auc_hist: num_thresholds %d out of range", num_thresholds);
  if (n <= 0) return DFM_OK;
  const int grid = std::min(148, ceil_div(n, 256));
  auc_hist_kernel<<<grid, 256, 2 * (num_thresholds + 1) * sizeof(unsigned int), static_cast<cudaStream_t>(stream)>>>(
      prob, label, n, hist, num_thresholds);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}
