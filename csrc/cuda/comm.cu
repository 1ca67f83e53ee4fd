// In-kernel collectives over NVLink / NVSwitch peer memory: no NCCL call on the training path.
//
// Every rank owns one "symmetric block" (same layout on all ranks) that its peers map through
// CUDA IPC.  Kernels address peer memory directly with ld/st (P2P over NVLink 5); cross-GPU
// ordering uses per-CTA flag words with release/acquire at system scope and a monotonically
// increasing epoch kept in device memory, so the same code replays inside a CUDA graph.
//
//   dense_allreduce_opt_kernel   replaces Horovod's fused-buffer ncclAllReduce + ApplyAdam for the
//       MLP / out-layer / fm_bias gradients (SURVEY X3 + K14): two-shot all-reduce (each rank
//       reduces 1/N of the vector from all peers, writes the average back to all peers) fused with
//       the 1/N scale, the optimizer update, the bf16 weight refresh and the gradient clear.
//   sparse_push_claim_kernel +   replace Horovod's ncclAllGather of IndexedSlices + the sparse
//   sparse_merge_apply_kernel    apply on the tables (SURVEY X4 + K13 + K14): every rank stores its
//       de-duplicated (id, gradient row) list into each peer's inbox, the receivers merge the N
//       lists by id in rank order (bit-identical on every replica), scale by 1/N and run the lazy
//       optimizer on the union of touched rows.
//
// Deadlock safety: CTA i only ever waits for CTA i of the other ranks, and all ranks launch the same
// grid in the same stream order; spins are bounded by a watchdog that traps instead of hanging.
#include "engine.cuh"

namespace dfm {

constexpr int kCommThreads = 512;      // dense all-reduce kernel
constexpr int kOneShotMaxRanks = 4;    // up to here every rank reduces the whole dense gradient itself
constexpr int kSparseThreads = 1024;   // sparse exchange kernels: one warp per row, rows are latency-bound
constexpr long long kCommWatchdog = 10000000000LL;   // ~5 s

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_volatile_f(const float* p) {
  float v;
  asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_volatile_i(const int* p) {
  int v;
  asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void stamp(const CommView& cv, int slot) {
  if (blockIdx.x == 0 && threadIdx.x == 0 && cv.dbg) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    cv.dbg[slot] = t;
  }
}

// Barrier among the CTAs with index blockIdx.x of all ranks.  `slot` selects one of the flag arrays.
// Everything the calling CTA wrote before (to any GPU) is visible to its peers' matching CTA after.
__device__ __forceinline__ unsigned int ld_relaxed_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void cta_barrier_all_ranks(const CommView& cv, int slot, unsigned int epoch) {
  // bar.sync orders every thread's earlier writes before the signalling threads, and their release
  // store at system scope is cumulative over those writes: no separate membar.sys is needed (measured:
  // a system fence costs ~4 us here, and st.release.sys contains one).
  __syncthreads();
  if (slot == 0) stamp(cv, 16);
  const int t = threadIdx.x;
  if (t < cv.world) {
    unsigned int* peer_flags = reinterpret_cast<unsigned int*>(cv.base[t] + cv.off_flags);
    if (slot == 0) stamp(cv, 17);
    st_release_sys(peer_flags + (size_t(slot) * cv.world + cv.rank) * kCommMaxCtas + blockIdx.x, epoch);
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(cv.base[cv.rank] + cv.off_flags) +
                               (size_t(slot) * cv.world + t) * kCommMaxCtas + blockIdx.x;
    if (slot == 0) stamp(cv, 18);
    const long long t0 = clock64();
    while (ld_relaxed_sys(mine) < epoch) {
      if (clock64() - t0 > kCommWatchdog) {
        printf("[dfm comm watchdog] rank %d cta %d slot %d waiting for rank %d epoch %u\n", cv.rank, blockIdx.x,
               slot, t, epoch);
        __trap();
      }
    }
    if (slot == 0) stamp(cv, 19);
    (void)ld_acquire_sys(mine);        // one acquire once the flag is up: peers' writes are visible from here
    if (slot == 0) stamp(cv, 20);
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCommThreads)
dense_allreduce_opt_kernel(const CommView cv, float* __restrict__ P, float* __restrict__ s0,
                           float* __restrict__ s1, const float* __restrict__ part, int ksplit,
                           size_t part_stride, __nv_bfloat16* __restrict__ Wb, int n,
                           const StepState* __restrict__ st, const OptHyper hyper) {
  const OptStepConsts c = make_opt_consts(hyper, st->lr, st->step);
  const unsigned int epoch = unsigned(st->comm_epoch);       // bumped by the first kernel of the step
  const int N = cv.world;
  const float inv_n = 1.f / float(N);
  const int n_red = n - 1;                            // element n-1 (fm_bias) shares the gradient of n-2
  const int n4 = (n_red + 3) / 4;                     // float4 groups (buffers are padded)
  const int shard4 = (n4 + N - 1) / N;                // groups per rank
  const int sub4 = (shard4 + gridDim.x - 1) / gridDim.x;   // groups per (rank, CTA)

  float* my_dP = reinterpret_cast<float*>(cv.base[cv.rank] + cv.off_dP);
  float* my_avg = reinterpret_cast<float*>(cv.base[cv.rank] + cv.off_avg);

  stamp(cv, 0);
  // local: add the split-K slices of the weight-gradient GEMM into the symmetric gradient buffer.
  // CTA b sums exactly the groups that the peers' CTA b will read, so the per-CTA barrier below is enough.
  for (int r = 0; r < N; ++r) {
    const int g0 = r * shard4 + blockIdx.x * sub4;
    const int g1 = min(min(g0 + sub4, (r + 1) * shard4), n4);
    for (int g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = 0; k < ksplit; ++k) {
        const float4 v = *reinterpret_cast<const float4*>(part + size_t(k) * part_stride + 4 * size_t(g));
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      *reinterpret_cast<float4*>(my_dP + 4 * size_t(g)) = acc;
    }
  }

  // A: every rank's local gradients are complete -> safe to read them remotely
  stamp(cv, 1);
  cta_barrier_all_ranks(cv, 0, epoch);
  stamp(cv, 2);

  if (N <= kOneShotMaxRanks) {
    // one-shot: every rank reads every peer's gradient itself and applies the update -- one barrier
    // instead of two; the extra NVLink reads (N-1 x 1.4 MB) are cheaper than a second system fence.
    for (int r = 0; r < N; ++r) {
      const int g0 = r * shard4 + blockIdx.x * sub4;
      const int g1 = min(min(g0 + sub4, (r + 1) * shard4), n4);
      for (int g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int q = 0; q < N; ++q) {               // rank order: identical sum on every replica
          const float4 v = ld_volatile_f4(reinterpret_cast<const float*>(cv.base[q] + cv.off_dP) + 4 * size_t(g));
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        const float gr[4] = {acc.x * inv_n, acc.y * inv_n, acc.z * inv_n, acc.w * inv_n};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int i = 4 * g + j;
          if (i >= n_red) break;
          float p = P[i], a = s0[i], b = s1[i];
          opt_apply(c, gr[j], p, a, b);
          P[i] = p; s0[i] = a; s1[i] = b;
          Wb[i] = __float2bfloat16(p);
          if (i == n - 2) {
            float p2 = P[n - 1], a2 = s0[n - 1], b2 = s1[n - 1];
            opt_apply(c, gr[j], p2, a2, b2);
            P[n - 1] = p2; s0[n - 1] = a2; s1[n - 1] = b2;
          }
        }
      }
    }
    stamp(cv, 5);
    return;
  }

  // reduce-scatter my shard, fused with the 1/N scale; all-gather by storing to every peer
  {
    const int g0 = cv.rank * shard4 + blockIdx.x * sub4;
    const int g1 = min(min(g0 + sub4, (cv.rank + 1) * shard4), n4);
    for (int g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < N; ++r) {
        const float4 v = ld_volatile_f4(reinterpret_cast<const float*>(cv.base[r] + cv.off_dP) + 4 * size_t(g));
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      acc.x *= inv_n; acc.y *= inv_n; acc.z *= inv_n; acc.w *= inv_n;
      for (int r = 0; r < N; ++r)
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(cv.base[r] + cv.off_avg) + 4 * size_t(g)) = acc;
    }
  }

  // B: the averaged slices written by CTA blockIdx.x of every rank have landed here, and every peer
  //    has finished reading the matching slices of my gradient buffer
  stamp(cv, 3);
  cta_barrier_all_ranks(cv, 1, epoch);
  stamp(cv, 4);

  for (int r = 0; r < N; ++r) {
    const int g0 = r * shard4 + blockIdx.x * sub4;
    const int g1 = min(min(g0 + sub4, (r + 1) * shard4), n4);
    for (int g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
      const float4 a4 = ld_volatile_f4(my_avg + 4 * size_t(g));
      const float gr[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = 4 * g + j;
        if (i >= n_red) break;
        float p = P[i], a = s0[i], b = s1[i];
        opt_apply(c, gr[j], p, a, b);
        P[i] = p; s0[i] = a; s1[i] = b;
        Wb[i] = __float2bfloat16(p);
        if (i == n - 2) {
          float p2 = P[n - 1], a2 = s0[n - 1], b2 = s1[n - 1];
          opt_apply(c, gr[j], p2, a2, b2);
          P[n - 1] = p2; s0[n - 1] = a2; s1[n - 1] = b2;
        }
      }
    }
  }
  stamp(cv, 5);
}

// ------------------------------------------------------------------------------------------------
// Inbox layout inside the symmetric block, per sender rank r:
//   count   int[world]                    at off_in_count
//   ids     int  [world][cap]             at off_in_ids
//   rows    float[world][cap][row_pitch]  at off_in_rows     row = Ke gradient values, then gw
__global__ void __launch_bounds__(kSparseThreads)
sparse_push_claim_kernel(const CommView cv, const SparseArgs a, int* __restrict__ entry_of,
                         int* __restrict__ touched, int* __restrict__ union_ids) {
  const unsigned int epoch = unsigned(a.st->comm_epoch);
  const int N = cv.world, Ke = a.Ke, pitch = cv.row_pitch;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const int warp_g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int n = a.st->n_unique;

  stamp(cv, 8);
  // ---- push: my (row gradient, first-order gradient, id) records into inbox[cv.rank] of every rank ----
  // record = g_pitch(Ke) floats: [0,Ke) factor gradients, [Ke] first-order gradient, [Ke+1] id (bit pattern)
  for (int s = warp_g; s < n; s += warps) {
    const int id = a.uniq_ids[s];
    float* grow = a.G + size_t(s) * pitch;
    for (int k = lane; k < pitch; k += 32) {
      float v = 0.f;
      if (k <= Ke) { v = grow[k]; grow[k] = 0.f; }
      else if (k == Ke + 1) v = __int_as_float(id);
      for (int r = 0; r < N; ++r) {
        float* rows = reinterpret_cast<float*>(cv.base[r] + cv.off_in_rows) + (size_t(cv.rank) * cv.cap + s) * pitch;
        rows[k] = v;
      }
    }
    if (lane == 0) a.slot_of[id] = -1;
  }
  if (threadIdx.x < N)
    reinterpret_cast<int*>(cv.base[threadIdx.x] + cv.off_in_count)[cv.rank * kCommMaxCtas + blockIdx.x] = n;

  stamp(cv, 9);
  cta_barrier_all_ranks(cv, 2, epoch);
  stamp(cv, 10);

  // ---- claim: index every received entry by (id, sender) and build the union id list ----
  // One thread per entry.  The rows pushed by CTA blockIdx.x of rank r are those with
  // (s % warps) in this CTA's warp range, i.e. s = first_warp + (t % wpc) + (t / wpc) * warps.
  const int* in_count = reinterpret_cast<const int*>(cv.base[cv.rank] + cv.off_in_count);
  const float* in_rows_c = reinterpret_cast<const float*>(cv.base[cv.rank] + cv.off_in_rows);
  const int wpc = blockDim.x >> 5;
  const int first_warp = blockIdx.x * wpc;
  __shared__ int wsum[kSparseThreads / 32];
  __shared__ int cta_base;
  for (int r = 0; r < N; ++r) {
    const int nr = ld_volatile_i(in_count + r * kCommMaxCtas + blockIdx.x);
    const int per_cta = (nr > first_warp) ? ((nr - first_warp + warps - 1) / warps) * wpc : 0;   // upper bound
    for (int t0 = 0; t0 < per_cta; t0 += blockDim.x) {       // uniform trip count for the whole CTA
      const int t = t0 + threadIdx.x;
      const int s = first_warp + (t % wpc) + (t / wpc) * warps;
      int id = 0, won = 0;
      if (t < per_cta && s < nr) {
        id = __float_as_int(ld_volatile_f(in_rows_c + (size_t(r) * cv.cap + s) * pitch + Ke + 1));
        entry_of[size_t(id) * N + r] = s + 1;
        won = atomicCAS(&touched[id], 0, 1) == 0;
      }
      // one n_union bump per CTA round (returning atomics on one address serialise)
      const unsigned winners = __ballot_sync(0xffffffffu, won);
      if (lane == 0) wsum[threadIdx.x >> 5] = __popc(winners);
      __syncthreads();
      if (threadIdx.x == 0) {
        int tot = 0;
        for (int w = 0; w < wpc; ++w) { const int c = wsum[w]; wsum[w] = tot; tot += c; }
        cta_base = tot ? atomicAdd(&a.st->n_union, tot) : 0;
      }
      __syncthreads();
      if (won) union_ids[cta_base + wsum[threadIdx.x >> 5] + __popc(winners & ((1u << lane) - 1u))] = id;
      __syncthreads();
    }
  }
  stamp(cv, 11);
}

__global__ void __launch_bounds__(kSparseThreads)
sparse_merge_apply_kernel(const CommView cv, const SparseArgs a, int* __restrict__ entry_of,
                          int* __restrict__ touched, const int* __restrict__ union_ids, const OptHyper hyper) {
  const OptStepConsts c = make_opt_consts(hyper, a.st->lr, a.st->step);
  const int N = cv.world, Ke = a.Ke, pitch = cv.row_pitch;
  const float inv_n = 1.f / float(N);
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const int n = a.st->n_union;
  // the inbox was written by peers before the push barrier of the previous kernel; this is a new launch
  // (L1 starts empty), so plain L2 loads see it
  const float* in_rows = reinterpret_cast<const float*>(cv.base[cv.rank] + cv.off_in_rows);
  stamp(cv, 12);

  // The work per row is a chain of dependent loads (id -> entries -> inbox rows -> table row), so each
  // warp keeps kRows rows in flight: every stage is issued for all of them before any result is used.
  constexpr int kRows = 2;
  constexpr int kKPL = 2;                            // embedding components per lane (Ke <= 64)
  for (int u0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * kRows; u0 < n; u0 += warps * kRows) {
    int id[kRows], ent[kRows];
    bool live[kRows];
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      live[j] = u0 + j < n;
      id[j] = live[j] ? __ldcg(union_ids + u0 + j) : 0;
    }
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      ent[j] = 0;
      if (live[j] && lane < N) {
        ent[j] = __ldcg(entry_of + size_t(id[j]) * N + lane);
        entry_of[size_t(id[j]) * N + lane] = 0;
      }
    }
    // lane l accumulates columns l, l+32 (factor gradients) and, for l == 0, column Ke (first-order gradient)
    float acc[kRows][kKPL + 1];
    float p[kRows][kKPL + 1], m[kRows][kKPL + 1], v[kRows][kKPL + 1];
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
#pragma unroll
      for (int q = 0; q <= kKPL; ++q) acc[j][q] = 0.f;
#pragma unroll
      for (int r = 0; r < kCommMaxRanks; ++r) {     // rank order: identical sum on every replica
        const int e = __shfl_sync(0xffffffffu, ent[j], r);
        if (e) {
          const float* row = in_rows + (size_t(r) * cv.cap + (e - 1)) * pitch;
#pragma unroll
          for (int q = 0; q < kKPL; ++q)
            if (lane + 32 * q < Ke) acc[j][q] += __ldcg(row + lane + 32 * q);
          if (lane == 0) acc[j][kKPL] += __ldcg(row + Ke);
        }
      }
      if (live[j]) {
#pragma unroll
        for (int q = 0; q < kKPL; ++q)
          if (lane + 32 * q < Ke) {
            const size_t pi = size_t(id[j]) * Ke + lane + 32 * q;
            p[j][q] = a.fm_v[pi]; m[j][q] = a.v_s0[pi]; v[j][q] = a.v_s1[pi];
          }
        if (lane == 0) { p[j][kKPL] = a.fm_w[id[j]]; m[j][kKPL] = a.w_s0[id[j]]; v[j][kKPL] = a.w_s1[id[j]]; }
      }
    }
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      if (!live[j]) continue;
#pragma unroll
      for (int q = 0; q < kKPL; ++q)
        if (lane + 32 * q < Ke) {
          const size_t pi = size_t(id[j]) * Ke + lane + 32 * q;
          opt_apply(c, acc[j][q] * inv_n + hyper.l2_reg * p[j][q], p[j][q], m[j][q], v[j][q]);
          a.fm_v[pi] = p[j][q]; a.v_s0[pi] = m[j][q]; a.v_s1[pi] = v[j][q];
        }
      if (lane == 0) {
        opt_apply(c, acc[j][kKPL] * inv_n + hyper.l2_reg * p[j][kKPL], p[j][kKPL], m[j][kKPL], v[j][kKPL]);
        a.fm_w[id[j]] = p[j][kKPL]; a.w_s0[id[j]] = m[j][kKPL]; a.w_s1[id[j]] = v[j][kKPL];
        touched[id[j]] = 0;
      }
    }
  }
  stamp(cv, 13);
}

// ------------------------------------------------------------------------------------------------
int launch_dense_allreduce_opt(const CommView& cv, float* P, float* s0, float* s1, const float* part, int ksplit,
                               size_t part_stride, __nv_bfloat16* Wb, int n, const StepState* st, const OptHyper& h,
                               cudaStream_t stream) {
  dense_allreduce_opt_kernel<<<cv.dense_ctas, kCommThreads, 0, stream>>>(cv, P, s0, s1, part, ksplit, part_stride,
                                                                      Wb, n, st, h);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_sparse_exchange_apply(const CommView& cv, const SparseArgs& a, int* entry_of, int* touched, int* union_ids,
                                 const OptHyper& h, cudaStream_t stream) {
  sparse_push_claim_kernel<<<cv.sparse_ctas, kSparseThreads, 0, stream>>>(cv, a, entry_of, touched, union_ids);
  DFM_LAUNCH_CHECK();
  sparse_merge_apply_kernel<<<cv.sparse_ctas, kSparseThreads, 0, stream>>>(cv, a, entry_of, touched, union_ids, h);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
