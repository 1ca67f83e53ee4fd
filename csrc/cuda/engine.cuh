// Shared definitions of the DeepFM step engine: the device-side state block, the
// per-step buffer table, and the launch prototypes of the fused kernels.
//
// Layout conventions (all row-major):
//   K[0] = field_size * embedding_size, N[l] = deep_layers[l], K[l+1] = N[l]
//   act[l]  bf16 [B, pitch[l]]  input of layer l (act[0] = e = v[id]*x), pitch[l] = round_up(K[l]+1, 16);
//           column K[l] is a constant 1.0 so that the weight-gradient GEMM also yields the bias row
//   dz[l]   bf16 [B, N[l]]      gradient w.r.t. the pre-activation of layer l
//   dz[L]   bf16 [B, 16]        column 0 = dlogit (out layer has one output; padded to a legal UMMA N)
//   P[l]    fp32 [K[l]+1, N[l]] weights with the bias as last row (l = L: out layer, N = 1)
//   Wb[l]   bf16 copy of P[l] (same shape) used as the tensor-core operand
#pragma once

#include "common.cuh"

namespace dfm {

constexpr int kMaxLayers = 6;       // hidden layers
constexpr int kTileM = 128;         // batch rows per CTA
constexpr int kChunkK = 64;         // K elements per pipeline stage (128 B of bf16)

// Scalars that change every step live in device memory so a captured CUDA graph can be
// replayed without re-recording kernel arguments.
struct StepState {
  int count;          // valid rows of the current batch
  int step;           // optimizer steps started so far; inside a training step: 1-based index of this update
  float lr;           // learning rate (already multiplied by world size, HVD.py:171)
  float loss_sum;     // sum of per-example BCE over the batch (this rank)
  int n_unique;       // distinct table rows touched by the current batch
  int bad_ids;        // feature ids outside [0, feature_size) seen so far (they are clamped; the host reports them)
  int train;          // 1: dropout active, gradients produced
  int reserved1;
  int comm_epoch;     // multi-GPU: completed collective rounds (flag value of the next round = +1)
  float adam_b1, adam_b2;   // > 0 when the optimizer is Adam: the step's first kernel derives lr_t from them
  float lr_t;               // Adam: lr * sqrt(1 - b2^t) / (1 - b1^t) of the update in flight (one powf pair per step,
                            // not one per warp of every optimizer kernel)
  // %globaltimer when CTA 0 of each kernel of the last step started (slots: kStamp*); read by the
  // timeline probe (scripts/gpu_timeline.py) to see launch gaps inside the replayed graph
  unsigned long long kstamp[24];
  // multi-GPU: distinct ids of the current batch owned by rank o (id % world == o); dedupe hands out the compact
  // slots of owner o from segment o of the slot space, so the records that go to one owner are contiguous
  int n_seg[8];
};

enum : int { kStampEmbedFwd = 0, kStampDedupe = 1, kStampFwd0 = 2, kStampLoss = 8, kStampBwd0 = 9, kStampEmbedBwd = 15,
             kStampWgrad = 16, kStampDenseOpt = 17, kStampSparse = 18, kStampSparseEnd = 19, kStampEmbedBwdEnd = 20,
             kStampWgradEnd = 21, kStampDenseEnd = 22, kStampDedupeEnd = 23 };

#ifdef __CUDACC__
// any one thread of CTA (0,0) chosen by the caller
__device__ __forceinline__ void kstamp_any(const StepState* st, int slot) {
  if (blockIdx.x == 0 && blockIdx.y == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    const_cast<StepState*>(st)->kstamp[slot] = t;
  }
}
__device__ __forceinline__ void kstamp(const StepState* st, int slot) {
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    const_cast<StepState*>(st)->kstamp[slot] = t;
  }
}
#endif

struct OptHyper {
  int kind;            // DFM_OPT_*
  float beta1, beta2, eps;       // Adam
  float momentum;                // Momentum
  float ftrl_lr_power, ftrl_l1, ftrl_l2;
  float l2_reg;                  // applied to fm_w / fm_v only (HVD.py:266-267)
};

// Row pitch (floats) of the compact sparse-gradient buffer G: Ke factor gradients, then the first-order
// gradient at column Ke, padded to a whole number of 32-byte sectors so no two rows share a sector
// (4-byte read-modify-writes of one sector from different warps serialise badly on this part).
__host__ __device__ inline int g_pitch(int Ke) { return Ke + 8; }

// slot_of[id] encoding: -1 = id not in the batch; otherwise bits 0..29 = slot + 1 (0 while the claimer is
// still allocating), bit 30 = the id occurs more than once in the batch (its gradient rows must be
// combined with atomics; an id that occurs once gets a plain store).
constexpr int kSlotMulti = 1 << 30;
constexpr int kSlotMask = kSlotMulti - 1;
__host__ __device__ inline int slot_index(int v) { return v < 0 ? -1 : (v & kSlotMask) - 1; }
__host__ __device__ inline bool slot_multi(int v) { return (v & kSlotMulti) != 0; }

struct MlpShape {
  int L;                       // hidden layers
  int K[kMaxLayers + 1];       // K[l] input width of layer l (l = L: out layer)
  int N[kMaxLayers + 1];       // N[l] output width (N[L] = 1)
  int pitch[kMaxLayers + 1];   // act[l] row pitch in elements
  float keep[kMaxLayers];      // dropout keep probability after layer l (1 = off)
};

// -------- one uniform apply rule for all five optimizers (TF1 formulas) ----------
//   Adam      m,v slots;   lr_t = lr*sqrt(1-b2^t)/(1-b1^t); p -= lr_t*m/(sqrt(v)+eps)      HVD.py:282-284
//   Adagrad   acc slot (init 1e-8); p -= lr*g/sqrt(acc)                                     HVD.py:286-288
//   Momentum  acc = mom*acc + g; p -= lr*acc                                                HVD.py:290-292
//   Ftrl      accum (init 0.1), linear; TF FtrlOptimizer defaults                            HVD.py:294
//   GD        p -= lr*g
struct OptStepConsts {
  float lr, lr_t, b1, b2, eps, mom, lr_power, l1, l2;
  int kind;
};

__device__ __forceinline__ OptStepConsts make_opt_consts(const OptHyper& h, const StepState* st) {
  OptStepConsts c;
  c.kind = h.kind; c.lr = st->lr; c.b1 = h.beta1; c.b2 = h.beta2; c.eps = h.eps; c.mom = h.momentum;
  c.lr_power = h.ftrl_lr_power; c.l1 = h.ftrl_l1; c.l2 = h.ftrl_l2;
  c.lr_t = h.kind == DFM_OPT_ADAM ? st->lr_t : st->lr;
  return c;
}

// Called by the one thread that opens a training step (after it bumped st->step).
__device__ __forceinline__ void open_step_consts(StepState* st) {
  if (st->adam_b1 > 0.f) {
    const float t = static_cast<float>(st->step);
    st->lr_t = st->lr * sqrtf(1.f - powf(st->adam_b2, t)) / (1.f - powf(st->adam_b1, t));
  }
}

__device__ __forceinline__ void opt_apply(const OptStepConsts& c, float g, float& p, float& s0, float& s1) {
  switch (c.kind) {
    case DFM_OPT_ADAM: {
      s0 = c.b1 * s0 + (1.f - c.b1) * g;
      s1 = c.b2 * s1 + (1.f - c.b2) * g * g;
      p -= c.lr_t * s0 / (sqrtf(s1) + c.eps);
    } break;
    case DFM_OPT_ADAGRAD: {
      s0 += g * g;
      p -= c.lr * g * rsqrtf(s0);
    } break;
    case DFM_OPT_MOMENTUM: {
      s0 = c.mom * s0 + g;
      p -= c.lr * s0;
    } break;
    case DFM_OPT_FTRL: {
      // s0 = accum, s1 = linear
      const float acc_new = s0 + g * g;
      float sigma, quad;
      if (c.lr_power == -0.5f) {
        sigma = (sqrtf(acc_new) - sqrtf(s0)) / c.lr;
        quad = sqrtf(acc_new) / c.lr + 2.f * c.l2;
      } else {
        sigma = (powf(acc_new, -c.lr_power) - powf(s0, -c.lr_power)) / c.lr;
        quad = powf(acc_new, -c.lr_power) / c.lr + 2.f * c.l2;
      }
      s1 += g - sigma * p;
      const float lin = s1;
      p = fabsf(lin) > c.l1 ? (copysignf(c.l1, lin) - lin) / quad : 0.f;
      s0 = acc_new;
    } break;
    default: {  // GD
      p -= c.lr * g;
    } break;
  }
}

// ---------------------------------------------------------------------------
// Kernel parameter blocks (passed as __grid_constant__) and launchers
// ---------------------------------------------------------------------------
struct FwdMaps {
  CUtensorMap act[kMaxLayers];    // act[l] as [B, K_l], box {64, 128}   (A operand, forward)
  CUtensorMap w[kMaxLayers];      // Wb[l] as [K_l, N_l] (N contiguous), box {64, 64}
                                  //   forward: one MN-major atom at (n0, k0); backward: K-major rows at (k0, n0)
  CUtensorMap dz[kMaxLayers];     // dz[l] as [B, N_l], box {64, 128}    (A operand, backward)
};

struct BwdParams {
  MlpShape sh;
  const __nv_bfloat16* act[kMaxLayers + 1];
  __nv_bfloat16* dz[kMaxLayers + 1];
  const float* w_out;          // P[L]
  const float* dlogit;
  StepState* st;
  int stages;
  int b_slot_bytes;
  int a_region_chunks;
  // embed_bwd only
  const int* ids;
  const float* vals;
  const float* s_sum;
  const int* slot_of;
  float* G;                    // [slots, Ke]
  float* Gw;                   // [slots]
  int F, Ke;
  int V;                       // feature_size: ids are clamped into [0, V)
  unsigned long long* dbg;     // optional phase timestamps (globaltimer ns) written by CTA (0,0)
};

struct WgradItem {
  int layer;      // 0..L (L = out layer)
  int m0;         // first row of dP_l handled by this CTA (multiple of 128)
  int n0;         // first column (multiple of 256)
  int n_here;     // columns (multiple of 16, <= 256)
  int k_begin;    // batch range [k_begin, k_end), multiples of 64
  int k_end;
  int ks;         // split-K slice this CTA writes
};

struct WgradMaps {
  CUtensorMap act[kMaxLayers + 1];   // [B, K_l+1], box {64, 64}
  CUtensorMap dz[kMaxLayers + 1];    // [B, N_l],   box {64, 64}
};

struct WgradParams {
  MlpShape sh;
  float* dP[kMaxLayers + 1];     // layer bases inside slice 0 of the partial buffer
  size_t part_stride;            // elements between split-K slices
  const WgradItem* items;
  StepState* st;
};

struct SparseArgs {
  float* fm_v; float* v_s0; float* v_s1;
  float* fm_w; float* w_s0; float* w_s1;
  float* G; float* Gw;
  int* slot_of; const int* uniq_ids;
  StepState* st;
  int Ke;
  long long V;
  float grad_scale;
};

// One MLP layer launch (mlp_layer.cu)
struct LayerParams {
  StepState* st;
  const float* bias;              // forward: b_l [N_l]
  const __nv_bfloat16* mask;      // backward: h_l (input of layer l), row pitch mask_pitch
  __nv_bfloat16* out;             // forward: h_{l+1} ; backward: dz_{l-1}
  int out_pitch, mask_pitch;
  int n_out;                      // output columns (N_l forward, K_l backward)
  int k_red;                      // reduction length (K_l forward, N_l backward)
  int b_box_rows;                 // rows of the B operand's TMA box (64)
  int layer;
  float keep;                     // forward: keep_l ; backward: keep_{l-1}
  unsigned long long seed;
  int bn;                         // backward with batch norm: the output is masked by the dropout bits of layer-1 only
                                  // (recomputed), the ReLU / BN part of the chain is bn_bwd's
  unsigned long long* dbg;        // optional: 8 phase timestamps of CTA (0,0)
};

// Batch normalisation after the ReLU of a hidden layer (HVD.py:230-233,306-328), one launch per layer and direction.
struct BnParams {
  StepState* st;
  const __nv_bfloat16* pre;       // a = relu(h W + b)  [rows, n], row pitch n
  __nv_bfloat16* out;             // forward: dropout(BN(a)), row pitch out_pitch
  __nv_bfloat16* dy;              // backward: in: gradient w.r.t. BN output (dropout already applied); out: dz of the layer
  const float* gamma; const float* beta;
  float* moving;                  // [mean n | variance n]
  float* save;                    // [mean n | rstd n] of the current batch (forward writes, backward reads)
  float* dgamma;                  // backward: [dgamma n | dbeta n] inside slice 0 of the dense gradient buffer
  int n, out_pitch, layer;
  float keep, decay, eps;
  unsigned long long seed;
};
int launch_bn_fwd(const BnParams&, cudaStream_t);
int launch_bn_bwd(const BnParams&, int rows_pad, cudaStream_t);

struct LossParams {
  StepState* st;
  int loss_type;                  // DFM_LOSS_*
  int bn;                         // batch-norm tower: dz_last is the gradient w.r.t. the BN output (no ReLU mask here)
  unsigned long long seed; int last_layer;   // dropout stream of the last hidden layer (bn: keep bits are recomputed)
  const __nv_bfloat16* h_last; int h_pitch; int n_last;
  const float* w_out;             // P_out: [n_last] weights then b_out
  const float* fm_part; const float* labels;
  float* logit; float* prob; float* dlogit;
  __nv_bfloat16* dz_last;         // [rows, n_last] or null (inference)
  __nv_bfloat16* dz_out;          // [rows, 16]
  float keep_last;
  int rows_pad;
};

// Multi-GPU view: every rank's symmetric block and embedding tables are addressable from every rank (peer mappings over
// NVLink; optionally also through one NVLS multicast address that stores to / reduces over all of them).
// Identical on all ranks except `rank`.
constexpr int kCommMaxRanks = 8;
constexpr int kCommMaxCtas = 160;
struct CommView {
  uint8_t* base[kCommMaxRanks];      // symmetric block of rank r
  uint8_t* mc_base;                  // multicast alias of all symmetric blocks (null: no NVLS; peers are looped over)
  float* tab_v[kCommMaxRanks];       // fm_v of rank r
  float* tab_w[kCommMaxRanks];       // fm_w of rank r
  float* mc_v; float* mc_w;          // multicast aliases of the tables (null: none)
  int world, rank;
  int cap;            // inbox rows per sender and slots per owner segment (= batch_max * field_size)
  int ctas;           // CTAs per rank of the sparse exchange kernel (all co-resident)
  int dense_ctas;     // CTAs per rank of the dense all-reduce kernel
  // symmetric block: dense gradient vector | 4 flag arrays [world][kCommMaxCtas] (dense round 1, 2; sparse round 2; spare)
  // | inbox counts [world][kCommMaxCtas] | inbox gradients [world][cap][Ke] fp32 | inbox tails [world][cap]{gw, id}
  // | (id, sender) index | arrival counter
  size_t off_dP, off_flags, off_in_count, off_in_grads, off_in_tails;
  size_t off_entries;                // [V + world] u64 {epoch, slot + 1}: the record sender s pushed for an id this rank owns,
                                     // at index id - rank + s; written by the SENDER, stale epochs are ignored, never cleared
  size_t off_counter;                // u64 arrivals of the sparse exchange's first barrier (all CTAs of all ranks)
  unsigned int epoch_base;           // StepState::comm_epoch when the views were attached (arrivals are counted from there)
  size_t flags_bytes;                // one flag array
  long long watchdog;                // clock64 ticks a cross-GPU or grid wait may take before the kernel traps
  int exact;                         // dense_exact table update: the owner publishes summed gradient rows, not parameters
  float* tab_g[kCommMaxRanks];       // dense_exact: per-rank [V, g_pitch] summed-gradient table
  float* mc_g;
  unsigned long long* dbg;           // 32 phase timestamps (globaltimer ns) written by CTA 0 of the collective kernels
};

struct DenseArgs {
  float* P; float* s0; float* s1;        // flat dense parameters and optimizer slots
  const float* part; int ksplit; size_t part_stride;   // split-K slices of the dense gradient
  __nv_bfloat16* Wb;
  int n;
  int wb_skip_lo, wb_skip_hi;            // bf16 copy of W_0 is left to the next step's first kernel (see W0Cast)
};

// The bf16 copy of W_0 is refreshed by the FIRST kernel of the next pass (embed_fm_fwd), not by the dense optimizer
// kernel: that one runs while embed_bwd's GEMM may still be reading W_0 through TMA.
struct W0Cast { const float* src; __nv_bfloat16* dst; int n; };
// Everything the two collective kernels need.  A real rank passes its own block by value; the single-GPU
// emulation of N ranks (one cooperative launch over all ranks' data) passes an array indexed by virtual rank.
struct XchgArgs {
  CommView cv;
  SparseArgs a;
  DenseArgs d;
  unsigned int* bar_local;               // {arrival counter, gate} x 2: this rank's side of the sparse kernel's two barriers
  OptHyper hyper;
};
int launch_dense_sync(const XchgArgs&, const XchgArgs* emu, int emu_ranks, cudaStream_t);
int launch_exchange_apply(const XchgArgs&, const XchgArgs* emu, int emu_ranks, cudaStream_t);
int launch_embed_fm_fwd(const int*, const float*, const float*, const float*, const float*, __nv_bfloat16*, int,
                        float*, float*, int, int, int, StepState*, int, W0Cast, long long V, cudaStream_t);
int launch_fill_ones_column(__nv_bfloat16*, int, int, int, cudaStream_t);
int launch_dedupe(const int*, int, StepState*, int*, int*, int, int world, int cap, long long V, cudaStream_t);
int mlp_layer_set_smem();
int launch_mlp_layer(bool, const CUtensorMap&, const CUtensorMap&, const LayerParams&, int, cudaStream_t);
int launch_loss_init(const LossParams&, int, cudaStream_t);
int embed_bwd_set_smem(int);
int launch_embed_bwd(const FwdMaps&, const BwdParams&, int, cudaStream_t);
int wgrad_set_smem();
int launch_wgrad(const WgradMaps&, const WgradParams&, int, cudaStream_t);
int launch_set_state(StepState*, int, float, int, int, cudaStream_t);
int launch_dense_opt(float*, float*, float*, const float*, int, size_t, __nv_bfloat16*, int, const StepState*,
                     const OptHyper&, float, int wb_skip_lo, int wb_skip_hi, cudaStream_t);
int launch_cast_weights(const float*, __nv_bfloat16*, int, cudaStream_t);
int launch_sparse_apply(const SparseArgs&, const OptHyper&, int, bool, cudaStream_t);
int launch_table_dense_from(const SparseArgs&, const OptHyper&, float* gsum, float inv_n, cudaStream_t);

}  // namespace dfm
