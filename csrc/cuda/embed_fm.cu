// deepfm_embed_fm_fwd: one pass over the batch that replaces the reference's
//   embedding_lookup(fm_w) * x, reduce_sum                      (HVD.py:196-198)
//   embedding_lookup(fm_v) * x, (sum e)^2, sum e^2, 0.5*sum     (HVD.py:200-206)
//   reshape to [B, F*K]                                         (HVD.py:219-221)
// and, when training, also starts the sparse-gradient bookkeeping: every distinct id of
// the batch claims a compact "slot" (slot_of[id]) and is appended to uniq_ids, so the
// backward pass can reduce row gradients into a dense [n_unique, K] buffer and the
// optimizer touches each table row exactly once (lazy update; dedupe that TF's
// IndexedSlices path never did).
//
// One warp per example; lane k owns embedding component(s) k, k+32, ...
// Outputs: e (bf16, MLP input, row pitch `pitch`), s_sum [B,K] fp32 = sum_f e,
//          fm_part [B] = fm_bias + y_w + y_v.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kEmbedWarps = 8;

__device__ __forceinline__ void cast_w0(const W0Cast& wc) {
  const int n2 = wc.n >> 1;                          // src / dst are 8- / 4-byte aligned (checked by the launcher)
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += gridDim.x * blockDim.x) {
    const float2 v = __ldg(reinterpret_cast<const float2*>(wc.src) + i);
    reinterpret_cast<uint32_t*>(wc.dst)[i] = pack_bf16x2(v.x, v.y);
  }
  if ((wc.n & 1) && blockIdx.x == 0 && threadIdx.x == 0) wc.dst[wc.n - 1] = __float2bfloat16(wc.src[wc.n - 1]);
}

// Ke/4 lanes per table row (float4 each), so one load instruction of a warp gathers 32/(Ke/4) rows and U of them
// put U*32/(Ke/4) rows in flight: all 39 fields of a Criteo example in ONE round trip at Ke = 32 (a lane-per-column
// version needed five dependent rounds of eight).
// Feature ids outside [0, V) are clamped into the table (every kernel that touches ids does the same, so the step
// stays memory-safe) and counted in StepState::bad_ids for the host to report.
template <int U>
__global__ void __launch_bounds__(kEmbedWarps * 32)
embed_fm_fwd_v4_kernel(const int* __restrict__ ids, const float* __restrict__ vals,
                       const float* __restrict__ fm_w, const float* __restrict__ fm_v,
                       const float* __restrict__ fm_bias, __nv_bfloat16* __restrict__ e, int pitch,
                       float* __restrict__ s_sum, float* __restrict__ fm_part, int F, int Ke,
                       StepState* __restrict__ st, int dedupe, const W0Cast wc, int V) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * kEmbedWarps + warp;
  const int count = st->count;
  pdl_launch_dependents();     // the first MLP layer may be scheduled now; it waits for this grid before reading
  kstamp(st, kStampEmbedFwd);
  cast_w0(wc);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->loss_sum = 0.f;            // accumulated by the loss kernel of this step
    if (dedupe) {                  // == training step: open the optimizer step (see StepState::step)
      st->step += 1;
      st->comm_epoch += 1;
      open_step_consts(st);
      st->n_unique = 0;            // counted by dedupe, which is queued behind this kernel
#pragma unroll
      for (int o = 0; o < 8; ++o) st->n_seg[o] = 0;
    }
  }
  __shared__ __align__(16) float part_out[kEmbedWarps];
  const int LPR = Ke >> 2;                 // lanes per row
  const int G = 32 / LPR;                  // rows per load instruction
  const int g = lane / LPR, sub = lane - g * LPR;
  const bool lane_on = g < G;
  if (b < count) {
    const int* my_ids = ids + size_t(b) * F;
    const float* my_vals = vals + size_t(b) * F;
    float4 S = make_float4(0.f, 0.f, 0.f, 0.f);
    float acc = 0.f;                       // yw - 0.5 * sum e^2, this lane's share
    for (int f0 = 0; f0 < F; f0 += G * U) {
      int id[U];
      float x[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int f = f0 + u * G + g;
        const bool on = lane_on && f < F;
        id[u] = -1;
        if (on) {
          const int raw = __ldg(my_ids + f);
          id[u] = min(max(raw, 0), V - 1);
          if (raw != id[u] && sub == 0) atomicAdd(&st->bad_ids, 1);
        }
        x[u] = on ? __ldg(my_vals + f) : 0.f;
      }
      float4 row[U];
      float w[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        row[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        w[u] = 0.f;
        if (id[u] >= 0) {
          row[u] = __ldg(reinterpret_cast<const float4*>(fm_v + size_t(id[u]) * Ke) + sub);
          if (sub == 0) w[u] = __ldg(fm_w + id[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (id[u] < 0) continue;
        const int f = f0 + u * G + g;
        const float4 ev = make_float4(row[u].x * x[u], row[u].y * x[u], row[u].z * x[u], row[u].w * x[u]);
        S.x += ev.x; S.y += ev.y; S.z += ev.z; S.w += ev.w;
        acc += w[u] * x[u] - 0.5f * (ev.x * ev.x + ev.y * ev.y + ev.z * ev.z + ev.w * ev.w);
        uint2 pk;
        pk.x = pack_bf16x2(ev.x, ev.y);
        pk.y = pack_bf16x2(ev.z, ev.w);
        *reinterpret_cast<uint2*>(e + size_t(b) * pitch + size_t(f) * Ke + 4 * sub) = pk;
      }
    }
    // sum S over the G row groups; every lane of group 0 ends with the full sum of its four columns
    float4 T = S;
    for (int k = 1; k < G; ++k) {
      const int src = lane_on ? ((g + k) % G) * LPR + sub : lane;
      T.x += __shfl_sync(0xffffffffu, S.x, src);
      T.y += __shfl_sync(0xffffffffu, S.y, src);
      T.z += __shfl_sync(0xffffffffu, S.z, src);
      T.w += __shfl_sync(0xffffffffu, S.w, src);
    }
    if (lane_on && g == 0) {
      *reinterpret_cast<float4*>(s_sum + size_t(b) * Ke + 4 * sub) = T;
      acc += 0.5f * (T.x * T.x + T.y * T.y + T.z * T.z + T.w * T.w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) part_out[warp] = acc + __ldg(fm_bias);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int b0 = blockIdx.x * kEmbedWarps;
    if (b0 + kEmbedWarps <= count) {
      float4* dst = reinterpret_cast<float4*>(fm_part + b0);
      dst[0] = make_float4(part_out[0], part_out[1], part_out[2], part_out[3]);
      dst[1] = make_float4(part_out[4], part_out[5], part_out[6], part_out[7]);
    } else {
      for (int i = 0; b0 + i < count; ++i) fm_part[b0 + i] = part_out[i];
    }
  }
}

// Sparse-gradient bookkeeping, off the critical path (runs on the side stream while the forward pass
// computes): every distinct id of the batch claims a compact slot.  One thread per (example, field).
//
// Same-address atomics serialise in L2 at ~50 ns each and, while they queue, hold up unrelated traffic
// through that slice (the co-running layer's TMA ring slowed 3x).  Criteo ids are Zipf-distributed and the 13
// numeric fields carry one id for the whole batch, so the kernel first de-duplicates inside the CTA with a
// shared-memory hash set: only the first thread to insert an id goes to global memory, which bounds the
// traffic on a hot id to one CAS per CTA.
constexpr int kDedupeHash = 2048;                    // >= 2x threads per CTA, power of two

__global__ void __launch_bounds__(1024)
dedupe_kernel(const int* __restrict__ ids, int F, StepState* __restrict__ st, int* __restrict__ slot_of,
              int* __restrict__ uniq_ids, int world, int cap, int V) {
  const int total = st->count * F;
  kstamp(st, kStampDedupe);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ int tab_id[kDedupeHash];              // -1 = empty
  __shared__ unsigned char tab_multi[kDedupeHash]; // the id was inserted more than once in this CTA
  __shared__ int warp_base[32];
  __shared__ int cta_base;
  __shared__ int seg_cnt[8], seg_base[8];          // multi-GPU: winners of this CTA per owner rank
  if (threadIdx.x < 8) seg_cnt[threadIdx.x] = 0;
  for (int t = threadIdx.x; t < kDedupeHash; t += blockDim.x) { tab_id[t] = -1; tab_multi[t] = 0; }
  __syncthreads();

  int id = 0, h = 0;
  bool leader = false;
  if (i < total) {
    id = min(max(__ldg(ids + i), 0), V - 1);       // out-of-range ids: clamped, like everywhere else in the step
    h = int((unsigned(id) * 2654435761u) >> 21) & (kDedupeHash - 1);
    while (true) {
      const int prev = atomicCAS(&tab_id[h], -1, id);
      if (prev == -1) { leader = true; break; }
      if (prev == id) { tab_multi[h] = 1; break; }
      h = (h + 1) & (kDedupeHash - 1);
    }
  }
  __syncthreads();

  int won = 0;
  if (leader) {
    const int pre = tab_multi[h] ? kSlotMulti : 0;   // known duplicate: claim with the flag set
    int v = slot_of[id];
    if (v == -1) v = atomicCAS(&slot_of[id], -1, pre);
    if (v == -1) won = 1;
    else if (!slot_multi(v)) atomicOr(&slot_of[id], kSlotMulti);   // another CTA holds it: the id repeats
  }

  if (world > 1) {
    // Multi-GPU: the slot space has one segment per owner rank (owner = id % world, segment o = slots
    // [o * cap, (o + 1) * cap)), so the records the exchange kernel streams to one owner are contiguous.
    // Winners count per owner in shared memory; one returning global atomic per owner per CTA.
    const int o = id % world;
    int pos = 0;
    if (won) pos = atomicAdd(&seg_cnt[o], 1);
    __syncthreads();
    if (threadIdx.x < world)
      seg_base[threadIdx.x] = seg_cnt[threadIdx.x] ? atomicAdd(&st->n_seg[threadIdx.x], seg_cnt[threadIdx.x]) : 0;
    __syncthreads();
    if (won) {
      const int s = o * cap + seg_base[o] + pos;
      uniq_ids[s] = id;
      atomicOr(&slot_of[id], s + 1);     // keeps a multi flag that arrived while the slot was being allocated
    }
    return;
  }

  // one counter bump per CTA: returning atomics on a single address serialise at ~50 ns each
  const unsigned winners = __ballot_sync(0xffffffffu, won);
  if (lane == 0) warp_base[warp] = __popc(winners);
  __syncthreads();
  if (warp == 0) {
    int c = warp_base[lane];
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    warp_base[lane] = incl - c;
    if (lane == 31) cta_base = incl ? atomicAdd(&st->n_unique, incl) : 0;
  }
  __syncthreads();
  if (won) {
    const int s = cta_base + warp_base[warp] + __popc(winners & ((1u << lane) - 1u));
    uniq_ids[s] = id;
    atomicOr(&slot_of[id], s + 1);       // keeps a multi flag that arrived while the slot was being allocated
  }
}

int launch_dedupe(const int* ids, int F, StepState* st, int* slot_of, int* uniq_ids, int batch_max, int world, int cap,
                  long long V, cudaStream_t stream) {
  DFM_REQUIRE(V > 0 && V <= 0x7fffffffLL, "dedupe: feature_size %lld does not fit int32 ids", V);
  DFM_REQUIRE(world >= 1 && world <= 8 && (world == 1 || (long long)world * cap < kSlotMulti),
              "dedupe: %d ranks x %d slots exceed the slot encoding", world, cap);
  dedupe_kernel<<<ceil_div(batch_max * F, 1024), 1024, 0, stream>>>(ids, F, st, slot_of, uniq_ids, world, cap, int(V));
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

// Writes the constant ones column (and zeroes the pad) of an activation buffer.
__global__ void fill_ones_column_kernel(__nv_bfloat16* act, int rows, int pitch, int k_valid) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  for (int c = k_valid; c < pitch; ++c) act[size_t(r) * pitch + c] = __float2bfloat16(c == k_valid ? 1.f : 0.f);
}

int launch_embed_fm_fwd(const int* ids, const float* vals, const float* fm_w, const float* fm_v,
                        const float* fm_bias, __nv_bfloat16* e, int pitch, float* s_sum, float* fm_part,
                        int F, int Ke, int batch_max, StepState* st, int dedupe, W0Cast wc, long long V,
                        cudaStream_t stream) {
  DFM_REQUIRE((reinterpret_cast<uintptr_t>(wc.src) & 7) == 0 && (reinterpret_cast<uintptr_t>(wc.dst) & 3) == 0,
              "W_0 cast: misaligned buffers");
  DFM_REQUIRE(Ke % 4 == 0 && Ke >= 8 && Ke <= 128 && pitch % 4 == 0, "embedding_size %d is not supported", Ke);
  DFM_REQUIRE(V > 0 && V <= 0x7fffffffLL, "feature_size %lld does not fit int32 ids", V);
  const int grid = ceil_div(batch_max, kEmbedWarps);
  // rows in flight per warp = U * (32 / (Ke/4)); enough to cover F in one round when it fits in registers
  const int G = 32 / (Ke / 4);
  if (ceil_div(F, G) <= 10)
    embed_fm_fwd_v4_kernel<10><<<grid, kEmbedWarps * 32, 0, stream>>>(ids, vals, fm_w, fm_v, fm_bias, e, pitch, s_sum,
                                                                      fm_part, F, Ke, st, dedupe, wc, int(V));
  else
    embed_fm_fwd_v4_kernel<16><<<grid, kEmbedWarps * 32, 0, stream>>>(ids, vals, fm_w, fm_v, fm_bias, e, pitch, s_sum,
                                                                      fm_part, F, Ke, st, dedupe, wc, int(V));
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_fill_ones_column(__nv_bfloat16* act, int rows, int pitch, int k_valid, cudaStream_t stream) {
  fill_ones_column_kernel<<<ceil_div(rows, 256), 256, 0, stream>>>(act, rows, pitch, k_valid);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
