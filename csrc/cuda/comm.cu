// In-kernel collectives over NVLink / NVSwitch peer memory: no NCCL call on the training path.
//
// Every rank owns one "symmetric block" (same layout on all ranks) that its peers map through CUDA
// IPC.  The kernel addresses peer memory directly with ld/st (P2P over NVLink 5); cross-GPU ordering
// uses per-CTA flag words with release/acquire at system scope and an epoch kept in device memory,
// so the same code replays inside a CUDA graph.
//
// exchange_apply_kernel is the whole multi-GPU tail of a training step in ONE launch with ONE
// cross-GPU barrier.  It replaces what Horovod's DistributedOptimizer (HVD.py:296) and the optimizer
// apply ops do with a fusion-buffer ncclAllReduce, two ncclAllGathers and ~20 Apply kernels
// (SURVEY X3, X4, K13, K14):
//
//   phase 0  local: add the split-K slices of the weight-gradient GEMM into my dense gradient vector
//            (symmetric memory); store my de-duplicated sparse records [K grads | w grad | id] into
//            the inbox every rank keeps for me (plain stores to peer pointers)
//   barrier  CTA b of every rank waits for CTA b of every other rank (one system-scope release each)
//   phase 1  dense: every rank reads every peer's gradient slice itself (one-shot all-reduce: the
//            extra NVLink reads are cheaper than a second barrier), averages in rank order, applies
//            the optimizer and refreshes the bf16 weights;
//            sparse: index every received record by (id, sender)
//   grid barrier (this GPU only)
//   phase 2  sparse: for each received record, the lowest-ranked sender of that id "owns" the row:
//            its warp sums the row over all senders in rank order (bit-identical on every replica,
//            so the replicated tables never need a broadcast), scales by 1/N, adds the L2 term of
//            the row and applies the lazy optimizer update.
//
// The (id, sender) index is stamped with the step's epoch instead of being cleared, so nothing has
// to be reset afterwards.  Deadlock safety: the grid is at most one CTA per SM (all co-resident),
// CTA b only ever waits for CTA b of the other ranks, every rank launches the same grid in the same
// stream order, and every spin is bounded by a watchdog that traps instead of hanging.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kXThreads = 1024;
constexpr long long kCommWatchdog = 10000000000LL;   // ~5 s

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int ld_relaxed_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint2 ld_volatile_u2(const uint32_t* p) {
  uint2 v;
  asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_volatile_i(const int* p) {
  int v;
  asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// caller picks the thread (CTA 0 only)
__device__ __forceinline__ void stamp_any(const CommView& cv, int slot) {
  if (blockIdx.x == 0 && cv.dbg) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    cv.dbg[slot] = t;
  }
}
__device__ __forceinline__ void stamp(const CommView& cv, int slot) {
  if (blockIdx.x == 0 && threadIdx.x == 0 && cv.dbg) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    cv.dbg[slot] = t;
  }
}

static __device__ __noinline__ void comm_watchdog_trip(int rank, int what, int peer, unsigned int epoch) {
  printf("[dfm comm watchdog] rank %d cta %d: %s barrier, waiting for %d, epoch %u\n", rank, blockIdx.x,
         what ? "grid" : "cross-GPU", peer, epoch);
  __trap();
}

// Barrier among the CTAs with index blockIdx.x of all ranks, split into signal and wait.  bar.sync orders every
// thread's earlier writes before the signalling threads, and their release store at system scope is cumulative
// over those writes: no separate membar.sys is needed (measured: a system fence costs ~4 us here, and
// st.release.sys contains one).  `off` selects the flag array (two rounds per step use two arrays).
__device__ __forceinline__ void cta_signal_all_ranks(const CommView& cv, size_t off, unsigned int epoch) {
  __syncthreads();
  if (threadIdx.x < cv.world) {
    unsigned int* peer_flags = reinterpret_cast<unsigned int*>(cv.base[threadIdx.x] + off);
    st_release_sys(peer_flags + size_t(cv.rank) * kCommMaxCtas + blockIdx.x, epoch);
  }
}
__device__ __forceinline__ void cta_wait_all_ranks(const CommView& cv, size_t off, unsigned int epoch) {
  const int t = threadIdx.x;
  if (t < cv.world) {
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(cv.base[cv.rank] + off) +
                               size_t(t) * kCommMaxCtas + blockIdx.x;
    const long long t0 = clock64();
    while (ld_relaxed_sys(mine) < epoch)
      if (clock64() - t0 > kCommWatchdog) comm_watchdog_trip(cv.rank, 0, t, epoch);
    (void)ld_acquire_sys(mine);        // one acquire once the flag is up: peers' writes are visible from here
  }
  __syncthreads();
}
__device__ __forceinline__ void cta_barrier_all_ranks(const CommView& cv, unsigned int epoch) {
  cta_signal_all_ranks(cv, cv.off_flags, epoch);
  cta_wait_all_ranks(cv, cv.off_flags, epoch);
}

__device__ __forceinline__ void st_volatile_f4(float* p, const float4& v) {
  asm volatile("st.volatile.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Barrier among all CTAs of this GPU's grid (they are co-resident), split into arrive and wait so that work which
// does not depend on the other CTAs can sit between the two.  `bar` only ever grows.
__device__ __forceinline__ void grid_arrive(unsigned int* bar) {
  __syncthreads();
  if (threadIdx.x == 0) red_release_gpu_add(bar, 1u);
}
__device__ __forceinline__ void grid_wait(const CommView& cv, unsigned int* bar, unsigned int epoch) {
  if (threadIdx.x == 0) {
    const unsigned int target = epoch * gridDim.x;
    const long long t0 = clock64();
    while (ld_acquire_gpu(bar) < target)
      if (clock64() - t0 > kCommWatchdog) comm_watchdog_trip(cv.rank, 1, -1, epoch);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kXThreads, 1)
exchange_apply_kernel(const __grid_constant__ CommView cv, const SparseArgs a, const DenseArgs d,
                      unsigned long long* __restrict__ entry_of, unsigned int* __restrict__ grid_bar,
                      const OptHyper hyper) {
  const OptStepConsts c = make_opt_consts(hyper, a.st);
  const unsigned int epoch = unsigned(a.st->comm_epoch);       // bumped by the first kernel of the step
  const int N = cv.world, Ke = a.Ke;
  const int hw = Ke / 2;                               // 32-bit words of a record's bf16 gradient pairs
  const size_t tail_off = size_t(N) * cv.cap * hw;     // words: the {first-order gradient fp32, id} pairs follow all gradients
  const int gp = g_pitch(Ke);                          // floats per row of the local gradient buffer G
  const float inv_n = 1.f / float(N);
  const int lane = threadIdx.x & 31;
  // Two warp groups (cv.split): the first 8 warps own the dense side (local split-K sum, then all-reduce + optimizer),
  // the other 24 the sparse side (push, index, merge).  The two sides are independent after the cross-GPU barrier
  // -- one is bound by NVLink reads, the other by random HBM rows -- so they run side by side instead of one after
  // the other.  Without the split every thread does both jobs in turn (needed by the two-shot dense variant).
  const bool split = cv.split != 0;
  const int s0 = split ? 256 : 0;                      // first thread of the sparse group
  const int ts = int(threadIdx.x) - s0;                // index inside the sparse group (< 0: dense-only thread)
  const bool in_s = ts >= 0;
  const int s_threads = int(blockDim.x) - s0;
  const int d_threads = split ? 256 : int(blockDim.x);
  const bool in_d = int(threadIdx.x) < d_threads;
  const int wpc = s_threads >> 5;                      // sparse-side warps per CTA
  const int warps = gridDim.x * wpc;
  const int warp_g = blockIdx.x * wpc + (max(ts, 0) >> 5);
  const int n_mine = a.st->n_unique;

  // dense work split: CTA b owns float4 groups [g_lo, g_hi) of the gradient vector on every rank
  const int n_red = d.n - 1;                           // element n-1 (fm_bias) shares the gradient of n-2
  const int n4 = (n_red + 3) / 4;
  const int per = (n4 + gridDim.x - 1) / gridDim.x;
  const int g_lo = blockIdx.x * per, g_hi = min(g_lo + per, n4);
  const size_t par = epoch & 1u;
  const size_t off_dP = cv.off_dP + par * cv.dP_bytes;
  const size_t off_in_count = cv.off_in_count + par * cv.in_count_bytes;
  const size_t off_in_rows = cv.off_in_rows + par * cv.in_rows_bytes;
  float* my_dP = reinterpret_cast<float*>(cv.base[cv.rank] + off_dP);

  stamp(cv, 0);
  kstamp(a.st, kStampSparse);
  // ---------------- phase 0a: local sum of the split-K slices ----------------
  // A CTA sums exactly the groups its namesakes on the other ranks read after the barrier (the per-CTA flags only
  // cover what the same-numbered CTA wrote): [g_lo, g_hi) for the one-shot all-reduce, sub-slice b of every
  // rank's chunk for the two-shot one.
  const int c4 = (n4 + N - 1) / N;                     // groups per rank chunk
  const int per2 = (c4 + gridDim.x - 1) / gridDim.x;   // groups per CTA sub-slice
  auto local_sum = [&](int gq) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < d.ksplit; ++k) {
      const float4 v = *reinterpret_cast<const float4*>(d.part + size_t(k) * d.part_stride + 4 * size_t(gq));
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    if (cv.dense_bf16)   // wire format of the one-shot all-reduce: 4 bf16 = 8 bytes per group (half the NVLink bytes)
      *reinterpret_cast<uint2*>(my_dP + 2 * size_t(gq)) = make_uint2(pack_bf16x2(acc.x, acc.y), pack_bf16x2(acc.z, acc.w));
    else
      *reinterpret_cast<float4*>(my_dP + 4 * size_t(gq)) = acc;
  };
  if (cv.two_shot) {
    for (int q = 0; q < N; ++q) {
      const int lo = q * c4 + blockIdx.x * per2, hi = min(min(lo + per2, (q + 1) * c4), n4);
      for (int gq = lo + threadIdx.x; gq < hi; gq += blockDim.x) local_sum(gq);
    }
  } else if (in_d) {
    for (int gq = g_lo + threadIdx.x; gq < g_hi; gq += d_threads) local_sum(gq);
  }
  // ---------------- phase 0b: push my sparse records to every rank ----------------
  // wire format, structure of arrays: grads[sender][slot][hw words of bf16 pairs] then tails[sender][slot]{first-order
  // gradient fp32, id}.  The G records a warp moves per instruction are consecutive slots, so it writes 32*8
  // contiguous bytes of gradients and G*8 of tails per destination -- whole sectors on the NVLink.  bf16 on the wire
  // halves the all-gather's bytes (the accumulation on the receiving side is fp32, in rank order).
  // LPR = Ke/4 lanes serve one record (four columns each), so a warp moves G = 32/LPR records per instruction and
  // keeps kRows * G of them in flight.  Record s belongs to group s / G; group gi is handled by warp gi % warps
  // on EVERY rank and in every phase -- so the per-CTA barrier below covers exactly what a CTA reads afterwards.
  constexpr int kRows = 2;
  const int LPR = Ke >> 2, G = 32 / LPR;
  const int g = lane / LPR, sub = lane - g * LPR;
  const bool lane_on = g < G;
  if (in_s) {
    const int n_groups = (n_mine + G - 1) / G;
    for (int gi0 = warp_g; gi0 < n_groups; gi0 += warps * kRows) {
      int sidx[kRows], id[kRows];
      float4 x[kRows];
      float gw[kRows];
#pragma unroll
      for (int j = 0; j < kRows; ++j) {
        sidx[j] = (gi0 + j * warps) * G + g;
        const bool on = lane_on && sidx[j] < n_mine;
        id[j] = on ? a.uniq_ids[sidx[j]] : -1;
        x[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        gw[j] = 0.f;
        if (on) {
          const float* grow = a.G + size_t(sidx[j]) * gp;
          x[j] = *reinterpret_cast<const float4*>(grow + 4 * sub);
          if (sub == 0) gw[j] = grow[Ke];
        }
      }
#pragma unroll
      for (int j = 0; j < kRows; ++j) {
        if (id[j] < 0) continue;
        float* grow = a.G + size_t(sidx[j]) * gp;
        *reinterpret_cast<float4*>(grow + 4 * sub) = make_float4(0.f, 0.f, 0.f, 0.f);
        const uint2 w2 = make_uint2(pack_bf16x2(x[j].x, x[j].y), pack_bf16x2(x[j].z, x[j].w));
        const uint2 tail = make_uint2(__float_as_uint(gw[j]), uint32_t(id[j]));
        const size_t rec = size_t(cv.rank) * cv.cap + sidx[j];
        for (int r = 0; r < N; ++r) {
          uint32_t* dst = reinterpret_cast<uint32_t*>(cv.base[r] + off_in_rows);
          *reinterpret_cast<uint2*>(dst + rec * hw + 2 * sub) = w2;
          if (sub == 0) *reinterpret_cast<uint2*>(dst + tail_off + rec * 2) = tail;
        }
        if (sub == 0) { grow[Ke] = 0.f; a.slot_of[id[j]] = -1; }
      }
    }
  }
  if (in_s && ts < N)
    reinterpret_cast<int*>(cv.base[ts] + off_in_count)[cv.rank * kCommMaxCtas + blockIdx.x] = n_mine;
  if (ts == 0) stamp_any(cv, 1);
  cta_barrier_all_ranks(cv, epoch);
  stamp(cv, 2);

  // ---------------- phase 1a: two-shot dense all-reduce, first half ----------------
  // Rank r owns chunk r of the gradient vector; CTA b a sub-slice of it: it sums the N copies in rank order
  // (identical on every replica by construction: only one rank computes each sum) and stores the sum back into
  // every rank's copy, in place -- nobody else touches that sub-slice.  NVLink moves 2 (N-1)/N of the vector per
  // rank instead of the one-shot's N-1 copies; the second cross-GPU round is only signalled here and waited for
  // after the index phase.
  if (cv.two_shot) {
    const int lo = cv.rank * c4 + blockIdx.x * per2, hi = min(min(lo + per2, (cv.rank + 1) * c4), n4);
    for (int gq = lo + threadIdx.x; gq < hi; gq += blockDim.x) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
      for (int q = 0; q < N; ++q) {
        const float4 v = ld_volatile_f4(reinterpret_cast<const float*>(cv.base[q] + off_dP) + 4 * size_t(gq));
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      for (int q = 0; q < N; ++q)
        st_volatile_f4(reinterpret_cast<float*>(cv.base[q] + off_dP) + 4 * size_t(gq), acc);
    }
    cta_signal_all_ranks(cv, cv.off_flags2, epoch);
  }

  // ---------------- phase 1b: index the records I received by (id, sender) ----------------
  // CTA b reads the groups its namesakes on the other ranks pushed (gi % warps in [b*wpc, (b+1)*wpc)).  One thread
  // per (sender, warp slot, record of the group); entries carry the epoch, so stale ones are ignored.
  const int* in_count = reinterpret_cast<const int*>(cv.base[cv.rank] + off_in_count);
  const uint32_t* in_rows = reinterpret_cast<const uint32_t*>(cv.base[cv.rank] + off_in_rows);
  const int first_warp = blockIdx.x * wpc;
  auto index_phase = [&]() {
  for (int t = ts; t < N * wpc * G; t += s_threads) {
    const int r = t / (wpc * G), u = t - r * (wpc * G);
    const int wi = u / G, gg = u - wi * G;
    const int nr = ld_volatile_i(in_count + r * kCommMaxCtas + blockIdx.x);
    for (int gi = first_warp + wi; gi * G + gg < nr; gi += warps) {
      const int sr = gi * G + gg;
      const int id = int(__ldcg(in_rows + tail_off + (size_t(r) * cv.cap + sr) * 2 + 1));
      entry_of[size_t(id) * N + r] = (static_cast<unsigned long long>(epoch) << 32) | unsigned(sr + 1);
    }
  }
  };

  // ---------------- phase 1c: dense optimizer on the averaged gradient ----------------
  auto apply_group = [&](int gq, const float4& sum) {
    const float gr[4] = {sum.x * inv_n, sum.y * inv_n, sum.z * inv_n, sum.w * inv_n};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = 4 * gq + j;
      if (i >= n_red) break;
      float p = d.P[i], m = d.s0[i], v = d.s1[i];
      opt_apply(c, gr[j], p, m, v);
      d.P[i] = p; d.s0[i] = m; d.s1[i] = v;
      d.Wb[i] = __float2bfloat16(p);
      if (i == d.n - 2) {                              // fm_bias shares b_out's gradient
        float p2 = d.P[d.n - 1], m2 = d.s0[d.n - 1], v2 = d.s1[d.n - 1];
        opt_apply(c, gr[j], p2, m2, v2);
        d.P[d.n - 1] = p2; d.s0[d.n - 1] = m2; d.s1[d.n - 1] = v2;
      }
    }
  };
  auto dense_phase = [&]() {
  if (cv.two_shot) {
    // the sums of my namesake CTAs' sub-slices (one per rank's chunk) have landed in my dP once their flags are up
    cta_wait_all_ranks(cv, cv.off_flags2, epoch);
    for (int q = 0; q < N; ++q) {
      const int lo = q * c4 + blockIdx.x * per2, hi = min(min(lo + per2, (q + 1) * c4), n4);
      for (int gq = lo + threadIdx.x; gq < hi; gq += blockDim.x)
        apply_group(gq, ld_volatile_f4(my_dP + 4 * size_t(gq)));
    }
  } else {
    // one-shot all-reduce: every rank reads all N copies of its groups
    for (int gq = g_lo + threadIdx.x; gq < g_hi; gq += d_threads) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
      for (int q = 0; q < N; ++q) {                    // rank order: identical sum on every replica
        const float* src = reinterpret_cast<const float*>(cv.base[q] + off_dP);
        if (cv.dense_bf16) {                           // every rank, the owner included, adds the rounded values
          const uint2 w = ld_volatile_u2(reinterpret_cast<const uint32_t*>(src + 2 * size_t(gq)));
          const __nv_bfloat162 b0 = *reinterpret_cast<const __nv_bfloat162*>(&w.x);
          const __nv_bfloat162 b1 = *reinterpret_cast<const __nv_bfloat162*>(&w.y);
          acc.x += __low2float(b0); acc.y += __high2float(b0); acc.z += __low2float(b1); acc.w += __high2float(b1);
        } else {
          const float4 v = ld_volatile_f4(src + 4 * size_t(gq));
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      }
      apply_group(gq, acc);
    }
  }
  };

  if (split) {
    if (!in_s) {                                       // dense group: done after its phase (no CTA-wide barrier follows)
      dense_phase();
      stamp(cv, 4);
      return;
    }
    index_phase();
    if (ts == 0) stamp_any(cv, 3);
    // grid barrier among the sparse groups of this GPU's CTAs (named barrier: the dense warps are not part of it)
    asm volatile("bar.sync 1, 768;" ::: "memory");
    if (ts == 0) {
      red_release_gpu_add(grid_bar, 1u);
      const unsigned int target = epoch * gridDim.x;
      const long long t0 = clock64();
      while (ld_acquire_gpu(grid_bar) < target)
        if (clock64() - t0 > kCommWatchdog) comm_watchdog_trip(cv.rank, 1, -1, epoch);
    }
    asm volatile("bar.sync 1, 768;" ::: "memory");
    if (ts == 0) stamp_any(cv, 5);
  } else {
    index_phase();
    stamp(cv, 3);
    grid_arrive(grid_bar);
    dense_phase();     // sits between the grid barrier's arrive and wait: it hides the barrier and the index imbalance
    stamp(cv, 4);
    grid_wait(cv, grid_bar, epoch);
    stamp(cv, 5);
  }

  // ---------------- phase 2: owners merge and apply ----------------
  // Work items of a warp: (sender r, round k) -> group k*warps + warp_g of sender r; kRows items in flight.  Each
  // record is a chain id -> {index row, table row (speculative: needed if this record owns the id)} -> the
  // senders' records.  The lowest-ranked sender of an id owns it; it sums all senders in rank order.
  int nr_l = lane < N ? ld_volatile_i(in_count + lane * kCommMaxCtas + blockIdx.x) : 0;
  int nr_max = nr_l;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) nr_max = max(nr_max, __shfl_xor_sync(0xffffffffu, nr_max, o));
  const int k_rounds = ((nr_max + G - 1) / G + warps - 1) / warps;
  const int n_items = k_rounds * N;
  const unsigned grp_mask = LPR >= 32 ? 0xffffffffu : ((1u << LPR) - 1u);
  for (int it0 = 0; it0 < n_items; it0 += kRows) {
    int id[kRows], rr[kRows], ent[kRows];
    float4 p[kRows], m[kRows], v[kRows];
    float pw[kRows], mw[kRows], vw[kRows];
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      const int it = it0 + j;
      const int k = it / N;
      rr[j] = it - k * N;
      const int nr = __shfl_sync(0xffffffffu, nr_l, rr[j] & 31);
      const int sr = (k * warps + warp_g) * G + g;
      const bool on = lane_on && it < n_items && sr < nr;
      id[j] = on ? int(__ldcg(in_rows + tail_off + (size_t(rr[j]) * cv.cap + sr) * 2 + 1)) : -1;
    }
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      ent[j] = 0;
      if (id[j] >= 0) {
        // lane sub reads the entry of sender sub (and sub + LPR when there are more ranks than lanes per record)
        if (sub < N) {
          const unsigned long long e = __ldcg(entry_of + size_t(id[j]) * N + sub);
          if (unsigned(e >> 32) == epoch) ent[j] = int(unsigned(e));
        }
        const size_t pi = size_t(id[j]) * Ke + 4 * sub;
        p[j] = *reinterpret_cast<const float4*>(a.fm_v + pi);
        m[j] = *reinterpret_cast<const float4*>(a.v_s0 + pi);
        v[j] = *reinterpret_cast<const float4*>(a.v_s1 + pi);
        if (sub == 0) { pw[j] = a.fm_w[id[j]]; mw[j] = a.w_s0[id[j]]; vw[j] = a.w_s1[id[j]]; }
      }
    }
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      const unsigned have = (__ballot_sync(0xffffffffu, ent[j] != 0) >> (lane_on ? g * LPR : 0)) & grp_mask;
      const bool own = id[j] >= 0 && (__ffs(have) - 1) == rr[j];
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float accw = 0.f;
#pragma unroll
      for (int q = 0; q < kCommMaxRanks; ++q) {             // rank order: identical sum on every replica
        const int e = __shfl_sync(0xffffffffu, ent[j], lane_on ? g * LPR + (q < LPR ? q : 0) : lane);
        if (own && q < N && e) {
          const size_t rec = size_t(q) * cv.cap + (e - 1);
          const uint32_t* row = in_rows + rec * hw;
          const uint2 w2 = __ldcg(reinterpret_cast<const uint2*>(row + 2 * sub));
          const __nv_bfloat162 b0 = *reinterpret_cast<const __nv_bfloat162*>(&w2.x);
          const __nv_bfloat162 b1 = *reinterpret_cast<const __nv_bfloat162*>(&w2.y);
          acc.x += __low2float(b0); acc.y += __high2float(b0);
          acc.z += __low2float(b1); acc.w += __high2float(b1);
          if (sub == 0) accw += __uint_as_float(__ldcg(in_rows + tail_off + rec * 2));
        }
      }
      if (own) {
        const size_t pi = size_t(id[j]) * Ke + 4 * sub;
        opt_apply(c, acc.x * inv_n + hyper.l2_reg * p[j].x, p[j].x, m[j].x, v[j].x);
        opt_apply(c, acc.y * inv_n + hyper.l2_reg * p[j].y, p[j].y, m[j].y, v[j].y);
        opt_apply(c, acc.z * inv_n + hyper.l2_reg * p[j].z, p[j].z, m[j].z, v[j].z);
        opt_apply(c, acc.w * inv_n + hyper.l2_reg * p[j].w, p[j].w, m[j].w, v[j].w);
        *reinterpret_cast<float4*>(a.fm_v + pi) = p[j];
        *reinterpret_cast<float4*>(a.v_s0 + pi) = m[j];
        *reinterpret_cast<float4*>(a.v_s1 + pi) = v[j];
        if (sub == 0) {
          opt_apply(c, accw * inv_n + hyper.l2_reg * pw[j], pw[j], mw[j], vw[j]);
          a.fm_w[id[j]] = pw[j]; a.w_s0[id[j]] = mw[j]; a.w_s1[id[j]] = vw[j];
        }
      }
    }
  }
  if (ts == 0) { stamp_any(cv, 6); kstamp_any(a.st, kStampSparseEnd); }
}

// ------------------------------------------------------------------------------------------------
int launch_exchange_apply(const CommView& cv, const SparseArgs& a, const DenseArgs& d, unsigned long long* entry_of,
                          unsigned int* grid_bar, const OptHyper& h, cudaStream_t stream) {
  exchange_apply_kernel<<<cv.ctas, kXThreads, 0, stream>>>(cv, a, d, entry_of, grid_bar, h);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
