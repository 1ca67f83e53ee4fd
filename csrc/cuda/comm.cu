// In-kernel collectives over NVLink / NVSwitch peer memory: no NCCL call on the training path.
//
// Every rank owns one "symmetric block" (same layout on all ranks) and its replica of the embedding tables; all of
// them are mapped into every rank (CUDA IPC or cuMem handles), and -- where the fabric offers NVLS -- also bound to
// one multicast address per buffer, so a single `multimem.st` lands in every replica and a single
// `multimem.ld_reduce` returns the sum over all replicas, formed inside the switch.  Cross-GPU ordering uses per-CTA
// flag words with release/acquire at system scope and an epoch kept in device memory, so the same code replays
// inside a CUDA graph.  Together the two kernels replace what Horovod's DistributedOptimizer (HVD.py:296) and the
// optimizer apply ops do with a fusion-buffer ncclAllReduce, two ncclAllGathers and ~20 Apply kernels (SURVEY X3,
// X4, K13, K14).  Peer memory is only ever touched with ld / st / multimem instructions issued by threads.
//
// dense_sync_kernel (side stream, right behind the weight-gradient GEMM, so it runs under embed_bwd):
//   phase 0  add the split-K slices of the weight-gradient GEMM into my dense gradient vector (symmetric memory)
//   barrier  CTA b of every rank waits for CTA b of every other rank
//   phase 1  rank r owns chunk r of the vector: it fetches the sum over all ranks (multimem.ld_reduce, or N peer
//            loads added in rank order) and stores it into every rank's copy (multimem.st, or N peer stores).  Only
//            one rank forms each sum, so every replica applies bit-identical gradients.
//   barrier
//   phase 2  every rank scales by 1/N, applies the optimizer to all dense parameters, refreshes the bf16 weights.
//
// exchange_apply_kernel (main stream, behind embed_bwd) -- owner-sharded sparse update:
//   phase 0  the row with id belongs to rank id % N.  dedupe already grouped my distinct ids by owner (segment o of
//            the slot space), so I stream segment o -- gradient rows in fp32, {first-order gradient, id} tails -- into
//            the inbox rank o keeps for me: 1/N of an all-gather's bytes, each record crosses NVLink once.
//            The sender also writes the owner's (id, sender) index entry, stamped with the step's epoch.
//   barrier  all CTAs of all ranks, hierarchical: the CTAs of a rank meet at a local counter, CTA 0 alone signals the
//            peers (one system-scope release, cumulative over what the local CTAs wrote) and opens a local gate
//   phase 2  for every received record whose sender is the lowest-ranked sender of its id: add the senders' rows in
//            rank order, scale by 1/N, add the row's L2 term, apply the lazy optimizer to MY copy of the slots
//            (m, v of a row live on its owner only), and store the new parameter row into every replica.
//            HBM read-modify-write per GPU is union/N rows instead of the whole union.
//   barrier  (same kind) every replica has every row before the next step's gather starts.
//
// Deadlock safety: every rank launches the same grids in the same stream order, the dense kernel's CTAs only wait for
// their namesakes on the other ranks, the sparse kernel's grid is at most one CTA per SM (its barriers need them
// co-resident), and every spin is bounded by a watchdog that traps instead of hanging.
//
// With kEmu the same kernels run N virtual ranks on ONE GPU as a single cooperative launch (blockIdx.x / ctas = rank):
// that is how the multi-rank protocol is tested and profiled on a one-GPU box.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kXThreads = 1024;

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
__device__ __forceinline__ void red_release_sys_add_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("red.release.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_f4(float* p, const float4& v) {
  asm volatile("st.volatile.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ int ld_volatile_i(const int* p) {
  int v;
  asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// NVLS: one instruction addresses every replica bound to the multicast object
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_f1(float* mc, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}

__device__ __forceinline__ void stamp_any(const CommView& cv, int cta, int slot) {
  if (cta == 0 && cv.dbg) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    cv.dbg[slot] = t;
  }
}
__device__ __forceinline__ void stamp(const CommView& cv, int cta, int slot) {
  if (threadIdx.x == 0) stamp_any(cv, cta, slot);
}

static __device__ __noinline__ void comm_watchdog_trip(int rank, int cta, int what, int peer, unsigned int epoch) {
  printf("[dfm comm watchdog] rank %d cta %d: %s wait timed out (peer %d, epoch %u): a peer died, stalled on the host "
         "for longer than DFM_COMM_TIMEOUT_S, or did not launch this step\n",
         rank, cta, what == 1 ? "grid-barrier" : what == 2 ? "dense all-reduce" : "sparse exchange", peer, epoch);
  __trap();
}

// Barrier among the CTAs with index `cta` of all ranks, split into signal and wait.  bar.sync orders every thread's
// earlier writes before the signalling threads, and their release store at system scope is cumulative over those
// writes: no separate membar.sys is needed (measured: a system fence costs ~4 us here, and st.release.sys contains
// one).  `which` selects one of the four flag arrays.
__device__ __forceinline__ void cta_signal_all_ranks(const CommView& cv, int cta, int which, unsigned int epoch) {
  __syncthreads();
  if (threadIdx.x < cv.world) {
    unsigned int* peer_flags = reinterpret_cast<unsigned int*>(cv.base[threadIdx.x] + cv.off_flags + which * cv.flags_bytes);
    st_release_sys(peer_flags + size_t(cv.rank) * kCommMaxCtas + cta, epoch);
  }
}
__device__ __forceinline__ void cta_wait_all_ranks(const CommView& cv, int cta, int which, unsigned int epoch) {
  const int t = threadIdx.x;
  if (t < cv.world) {
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(cv.base[cv.rank] + cv.off_flags + which * cv.flags_bytes) +
                               size_t(t) * kCommMaxCtas + cta;
    const long long t0 = clock64();
    while (ld_relaxed_sys(mine) < epoch)
      if (clock64() - t0 > cv.watchdog) comm_watchdog_trip(cv.rank, cta, which < 2 ? 2 : 0, t, epoch);
    (void)ld_acquire_sys(mine);        // one acquire once the flag is up: peers' writes are visible from here
  }
  __syncthreads();
}

// Barrier among ALL CTAs of ALL ranks, hierarchical: the CTAs of a rank meet at a local counter (gpu scope, a
// non-returning reduction each), CTA 0 alone talks to the other GPUs -- one system-scope release per peer, whose fence
// is cumulative over everything the local CTAs wrote before they arrived -- and then opens a local gate.  Measured
// here: a system-scope release from every one of 116 CTAs costs 10-12 us per barrier, from 32 CTAs 5-8 us; the fence is
// the expensive part, so only one CTA pays it.  `local` = {arrival counter, gate} of this barrier in local memory.
__device__ __forceinline__ void hier_barrier_all(const CommView& cv, int cta, int which, unsigned int epoch,
                                                 unsigned int* local) {
  __syncthreads();
  const unsigned int gen = epoch - cv.epoch_base;                 // barriers passed so far, this one included
  if (cta != 0) {
    if (threadIdx.x == 0) {
      red_release_gpu_add(local, 1u);
      const long long t0 = clock64();
      while (ld_acquire_gpu(local + 1) < gen)
        if (clock64() - t0 > cv.watchdog) comm_watchdog_trip(cv.rank, cta, 1, -1, epoch);
    }
  } else {
    if (threadIdx.x == 0) {
      const unsigned int target = gen * unsigned(cv.ctas - 1);
      const long long t0 = clock64();
      while (ld_acquire_gpu(local) < target)
        if (clock64() - t0 > cv.watchdog) comm_watchdog_trip(cv.rank, cta, 1, -1, epoch);
    }
    __syncthreads();
    if (threadIdx.x < cv.world) {
      unsigned int* peer_flags = reinterpret_cast<unsigned int*>(cv.base[threadIdx.x] + cv.off_flags + which * cv.flags_bytes);
      st_release_sys(peer_flags + size_t(cv.rank) * kCommMaxCtas, epoch);
      const unsigned int* mine = reinterpret_cast<const unsigned int*>(cv.base[cv.rank] + cv.off_flags + which * cv.flags_bytes) +
                                 size_t(threadIdx.x) * kCommMaxCtas;
      const long long t0 = clock64();
      while (ld_relaxed_sys(mine) < epoch)
        if (clock64() - t0 > cv.watchdog) comm_watchdog_trip(cv.rank, cta, 0, threadIdx.x, epoch);
      (void)ld_acquire_sys(mine);
    }
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(local + 1), "r"(gen) : "memory");
  }
  __syncthreads();
}

// this CTA's argument block: the kernel parameter (a real rank) or virtual rank blockIdx.x / ctas of the emulation
template <bool kEmu>
__device__ __forceinline__ const XchgArgs& select_args(const XchgArgs& pa, const XchgArgs* emu, XchgArgs* smem_copy,
                                                       bool dense, int& cta) {
  cta = blockIdx.x;
  if (!kEmu) return pa;
  const int cpr = dense ? emu[0].cv.dense_ctas : emu[0].cv.ctas;
  const int vr = blockIdx.x / cpr;
  cta = blockIdx.x - vr * cpr;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(emu + vr);
  uint32_t* dst = reinterpret_cast<uint32_t*>(smem_copy);
  for (int i = threadIdx.x; i < int(sizeof(XchgArgs) / 4); i += blockDim.x) dst[i] = src[i];
  __syncthreads();
  return *smem_copy;
}

// =================================================================================================
// dense all-reduce + optimizer
// =================================================================================================
template <bool kEmu>
__global__ void __launch_bounds__(kXThreads, 1)
dense_sync_kernel(const __grid_constant__ XchgArgs pa, const XchgArgs* __restrict__ emu) {
  __shared__ XchgArgs sx;
  int cta;
  const XchgArgs& x = select_args<kEmu>(pa, emu, &sx, true, cta);
  pdl_wait();                                          // launched behind the weight-gradient GEMM
  const CommView& cv = x.cv;
  const DenseArgs& d = x.d;
  const OptStepConsts c = make_opt_consts(x.hyper, x.a.st);
  const unsigned int epoch = unsigned(x.a.st->comm_epoch);     // bumped by the first kernel of the step
  const int N = cv.world, ctas = cv.dense_ctas;
  const float inv_n = 1.f / float(N);
  const int n_red = d.n - 1;                           // element n-1 (fm_bias) shares the gradient of n-2
  const int n4 = (n_red + 3) / 4;
  const int c4 = (n4 + N - 1) / N;                     // float4 groups per rank chunk
  const int per2 = (c4 + ctas - 1) / ctas;             // groups per CTA sub-slice of a chunk
  float* my_dP = reinterpret_cast<float*>(cv.base[cv.rank] + cv.off_dP);
  if (threadIdx.x == 0) kstamp_any(x.a.st, kStampDenseOpt);
  stamp(cv, cta, 8);

  // phase 0: CTA b sums sub-slice b of EVERY chunk -- exactly what its namesakes read after the barrier (the
  // per-CTA flags only cover what the same-numbered CTA wrote)
  for (int q = 0; q < N; ++q) {
    const int lo = q * c4 + cta * per2, hi = min(min(lo + per2, (q + 1) * c4), n4);
    for (int gq = lo + threadIdx.x; gq < hi; gq += 2 * blockDim.x) {
      const int g1 = min(gq + int(blockDim.x), hi - 1);       // second group of the pair (clamped: a harmless repeat)
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f), acc1 = acc;
      for (int k = 0; k < d.ksplit; ++k) {
        const float4 v = *reinterpret_cast<const float4*>(d.part + size_t(k) * d.part_stride + 4 * size_t(gq));
        const float4 w = *reinterpret_cast<const float4*>(d.part + size_t(k) * d.part_stride + 4 * size_t(g1));
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        acc1.x += w.x; acc1.y += w.y; acc1.z += w.z; acc1.w += w.w;
      }
      *reinterpret_cast<float4*>(my_dP + 4 * size_t(gq)) = acc;
      *reinterpret_cast<float4*>(my_dP + 4 * size_t(g1)) = acc1;
    }
  }
  stamp(cv, cta, 12);
  cta_signal_all_ranks(cv, cta, 0, epoch);
  cta_wait_all_ranks(cv, cta, 0, epoch);
  stamp(cv, cta, 9);

  // phase 1: reduce my chunk over all ranks and publish the sum into every rank's copy, in place
  {
    const int lo = cv.rank * c4 + cta * per2, hi = min(min(lo + per2, (cv.rank + 1) * c4), n4);
    float* mc = cv.mc_base ? reinterpret_cast<float*>(cv.mc_base + cv.off_dP) : nullptr;
    for (int gq = lo + threadIdx.x; gq < hi; gq += blockDim.x) {
      if (mc) {
        const float4 s = multimem_ld_reduce_f4(mc + 4 * size_t(gq));     // summed inside the switch
        multimem_st_f4(mc + 4 * size_t(gq), s);
      } else {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int q = 0; q < N; ++q) {                  // rank order
          const float4 v = ld_volatile_f4(reinterpret_cast<const float*>(cv.base[q] + cv.off_dP) + 4 * size_t(gq));
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        for (int q = 0; q < N; ++q)
          st_volatile_f4(reinterpret_cast<float*>(cv.base[q] + cv.off_dP) + 4 * size_t(gq), acc);
      }
    }
  }
  stamp(cv, cta, 13);
  cta_signal_all_ranks(cv, cta, 1, epoch);
  cta_wait_all_ranks(cv, cta, 1, epoch);
  stamp(cv, cta, 10);

  // phase 2: the sums of my namesake CTAs' sub-slices (one per chunk) have landed in my copy.  Two float4 groups per
  // thread in flight (the loop is a chain of dependent loads), parameters and slots moved as float4.
  auto apply_group = [&](int gq, const float4& sum, float4& p4, float4& m4, float4& v4) {
    float* pp = &p4.x; float* mm = &m4.x; float* vv = &v4.x;
    const float gr[4] = {sum.x * inv_n, sum.y * inv_n, sum.z * inv_n, sum.w * inv_n};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = 4 * gq + j;
      if (i >= n_red) break;
      opt_apply(c, gr[j], pp[j], mm[j], vv[j]);
      if (i == d.n - 2) {                              // fm_bias shares b_out's gradient
        float p2 = d.P[d.n - 1], m2 = d.s0[d.n - 1], v2 = d.s1[d.n - 1];
        opt_apply(c, gr[j], p2, m2, v2);
        d.P[d.n - 1] = p2; d.s0[d.n - 1] = m2; d.s1[d.n - 1] = v2;
      }
    }
  };
  auto store_group = [&](int gq, const float4& p4, const float4& m4, const float4& v4) {
    const int i0 = 4 * gq;
    if (i0 + 4 <= n_red) {
      *reinterpret_cast<float4*>(d.P + i0) = p4;
      *reinterpret_cast<float4*>(d.s0 + i0) = m4;
      *reinterpret_cast<float4*>(d.s1 + i0) = v4;
      if (i0 + 4 <= d.wb_skip_lo || i0 >= d.wb_skip_hi) {
        *reinterpret_cast<uint2*>(d.Wb + i0) = make_uint2(pack_bf16x2(p4.x, p4.y), pack_bf16x2(p4.z, p4.w));
        return;
      }
    }
    const float* pp = &p4.x; const float* mm = &m4.x; const float* vv = &v4.x;
    for (int j = 0; j < 4 && i0 + j < n_red; ++j) {    // ragged end of the vector, or a group straddling W_0's edge
      const int i = i0 + j;
      d.P[i] = pp[j]; d.s0[i] = mm[j]; d.s1[i] = vv[j];
      if (i < d.wb_skip_lo || i >= d.wb_skip_hi) d.Wb[i] = __float2bfloat16(pp[j]);
    }
  };
  auto load_group = [&](int gq, float4& p4, float4& m4, float4& v4) {
    const int i0 = 4 * gq;
    if (i0 + 4 <= d.n) {                               // (n_red = n - 1: the vectors hold one more element)
      p4 = *reinterpret_cast<const float4*>(d.P + i0);
      m4 = *reinterpret_cast<const float4*>(d.s0 + i0);
      v4 = *reinterpret_cast<const float4*>(d.s1 + i0);
    } else {
      float* pp = &p4.x; float* mm = &m4.x; float* vv = &v4.x;
      for (int j = 0; j < 4; ++j) {
        const bool in = i0 + j < d.n;
        pp[j] = in ? d.P[i0 + j] : 0.f; mm[j] = in ? d.s0[i0 + j] : 0.f; vv[j] = in ? d.s1[i0 + j] : 0.f;
      }
    }
  };
  for (int q = 0; q < N; ++q) {
    const int lo = q * c4 + cta * per2, hi = min(min(lo + per2, (q + 1) * c4), n4);
    for (int gq = lo + threadIdx.x; gq < hi; gq += 2 * blockDim.x) {
      const int g1 = gq + blockDim.x;
      const bool two = g1 < hi;
      float4 sa = ld_volatile_f4(my_dP + 4 * size_t(gq)), sb = make_float4(0.f, 0.f, 0.f, 0.f);
      float4 pa4, ma4, va4, pb4, mb4, vb4;
      load_group(gq, pa4, ma4, va4);
      if (two) { sb = ld_volatile_f4(my_dP + 4 * size_t(g1)); load_group(g1, pb4, mb4, vb4); }
      apply_group(gq, sa, pa4, ma4, va4);
      store_group(gq, pa4, ma4, va4);
      if (two) { apply_group(g1, sb, pb4, mb4, vb4); store_group(g1, pb4, mb4, vb4); }
    }
  }
  stamp(cv, cta, 11);
  if (threadIdx.x == 0) kstamp_any(x.a.st, kStampDenseEnd);
}

// =================================================================================================
// sparse exchange + owner-sharded lazy update
// =================================================================================================
// Lane layout: LPR = Ke/4 lanes serve one record (four columns each), so a warp moves G = 32/LPR records per
// instruction.  Record s of the stream sender q -> owner o belongs to group s / G; group gi of that stream is
// handled by warp (gi + shift(q, o)) % warps on BOTH ranks, shift = ((q + o) % N) * (warps / N): a sender's N
// outgoing streams and an owner's N incoming streams each tile the warps evenly, and the per-CTA flags cover
// exactly what a CTA reads after the barrier.
template <bool kEmu>
__global__ void __launch_bounds__(kXThreads, 1)
exchange_apply_kernel(const __grid_constant__ XchgArgs pa, const XchgArgs* __restrict__ emu) {
  __shared__ XchgArgs sx;
  int cta;
  const XchgArgs& x = select_args<kEmu>(pa, emu, &sx, false, cta);
  pdl_wait();                                          // launched behind embed_bwd
  const CommView& cv = x.cv;
  const SparseArgs& a = x.a;
  const OptHyper& hyper = x.hyper;
  const OptStepConsts c = make_opt_consts(hyper, a.st);
  const unsigned int epoch = unsigned(a.st->comm_epoch);
  const int N = cv.world, Ke = a.Ke, me = cv.rank, cap = cv.cap;
  const int gp = g_pitch(Ke);                          // floats per row of the local gradient buffer G
  const float inv_n = 1.f / float(N);
  const int lane = threadIdx.x & 31;
  const int wpc = blockDim.x >> 5;
  const int warps = cv.ctas * wpc;
  const int warp_g = cta * wpc + (threadIdx.x >> 5);
  const int stripe = warps / N;
  const int LPR = Ke >> 2, G = 32 / LPR;
  const int g = lane / LPR, sub = lane - g * LPR;
  const bool lane_on = g < G;
  constexpr int kRows = 2;                             // record groups in flight per warp

  stamp(cv, cta, 0);
  if (threadIdx.x == 0) kstamp_any(a.st, kStampSparse);

  // ---------------- phase 0: stream segment o of my records to its owner ----------------
  // item = (round k, owner) -> one group of G records of that stream; kPush items in flight per warp, so the
  // dependent loads (id, gradient row) of all owners' groups overlap instead of queueing owner after owner
  constexpr int kPush = 4;
  {
    const int ns_l = lane < N ? a.st->n_seg[lane] : 0;   // my records for owner `lane`
    int ng_p = (ns_l + G - 1) / G;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ng_p = max(ng_p, __shfl_xor_sync(0xffffffffu, ng_p, o));
    const int n_pitems = ((ng_p + warps - 1) / warps) * N;
    for (int it0 = 0; it0 < n_pitems; it0 += kPush) {
      int sidx[kPush], id[kPush], ow[kPush];
      float4 v[kPush];
      float gw[kPush];
      bool any = false;
#pragma unroll
      for (int j = 0; j < kPush; ++j) {
        const int it = min(it0 + j, n_pitems - 1);
        const int k = it / N;
        ow[j] = (me + 1 + (it - k * N)) % N;             // peers first, my own shard last
        const int n_o = __shfl_sync(0xffffffffu, ns_l, ow[j]);
        const int first = (warp_g - ((me + ow[j]) % N) * stripe + warps) % warps;
        sidx[j] = (k * warps + first) * G + g;
        const bool on = lane_on && it0 + j < n_pitems && sidx[j] < n_o;
        const size_t slot = size_t(ow[j]) * cap + sidx[j];
        id[j] = on ? a.uniq_ids[slot] : -1;
        v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        gw[j] = 0.f;
        if (on) {
          const float* grow = a.G + slot * gp;
          v[j] = *reinterpret_cast<const float4*>(grow + 4 * sub);
          if (sub == 0) gw[j] = grow[Ke];
        }
        any |= on;
      }
      if (!__any_sync(0xffffffffu, any)) continue;
#pragma unroll
      for (int j = 0; j < kPush; ++j) {
        if (id[j] < 0) continue;
        const int o = ow[j];
        float* grow = a.G + (size_t(o) * cap + sidx[j]) * gp;
        float* dst_g = reinterpret_cast<float*>(cv.base[o] + cv.off_in_grads) + (size_t(me) * cap + sidx[j]) * Ke;
        *reinterpret_cast<float4*>(dst_g + 4 * sub) = v[j];
        *reinterpret_cast<float4*>(grow + 4 * sub) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (sub == 0) {
          reinterpret_cast<uint2*>(cv.base[o] + cv.off_in_tails)[size_t(me) * cap + sidx[j]] =
              make_uint2(__float_as_uint(gw[j]), uint32_t(id[j]));
          // the owner's (id, sender) index entry, written from here: the owner needs no index pass of its own
          reinterpret_cast<unsigned long long*>(cv.base[o] + cv.off_entries)[size_t(id[j]) - o + me] =
              (static_cast<unsigned long long>(epoch) << 32) | unsigned(sidx[j] + 1);
          grow[Ke] = 0.f;
          a.slot_of[id[j]] = -1;
        }
      }
    }
  }
  if (threadIdx.x < N)
    reinterpret_cast<int*>(cv.base[threadIdx.x] + cv.off_in_count)[me * kCommMaxCtas + cta] = a.st->n_seg[threadIdx.x];
  stamp(cv, cta, 1);
  // ---------------- barrier: ALL CTAs of ALL ranks have pushed ----------------
  // a cross-GPU barrier and a grid barrier in one, so the merge below may read any record and any index entry
  hier_barrier_all(cv, cta, 2, epoch, x.bar_local);
  stamp(cv, cta, 2);
  stamp(cv, cta, 3);
  stamp(cv, cta, 4);

  const int* in_count = reinterpret_cast<const int*>(cv.base[me] + cv.off_in_count);
  const float* in_grads = reinterpret_cast<const float*>(cv.base[me] + cv.off_in_grads);
  const uint2* in_tails = reinterpret_cast<const uint2*>(cv.base[me] + cv.off_in_tails);
  const unsigned long long* entry_of = reinterpret_cast<const unsigned long long*>(cv.base[me] + cv.off_entries);
  int nr_l = lane < N ? ld_volatile_i(in_count + lane * kCommMaxCtas + cta) : 0;   // records sender `lane` pushed
  int ng_max = (nr_l + G - 1) / G;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ng_max = max(ng_max, __shfl_xor_sync(0xffffffffu, ng_max, o));
  const int k_rounds = (ng_max + warps - 1) / warps;
  const int n_items = k_rounds * N;                    // item = (round k, sender r) -> one group of G records
  auto item_record = [&](int it, int& r, int& nr) -> int {
    const int k = it / N;
    r = it - k * N;
    nr = __shfl_sync(0xffffffffu, nr_l, r);
    const int first = (warp_g - ((r + me) % N) * stripe + warps) % warps;
    return (k * warps + first) * G + g;
  };

  // ---------------- phase 2: owners merge, apply, publish ----------------
  // Each record is a chain: id -> entries of all senders -> {senders' rows, my table row + slots} -> update ->
  // stores.  The lowest-ranked sender's record carries the row; the other records of the id do nothing.
  const unsigned grp_mask = LPR >= 32 ? 0xffffffffu : ((1u << LPR) - 1u);
  unsigned int n_own = 0;                              // rows this warp published (counted by lane 0)
  for (int it0 = 0; it0 < n_items; it0 += kRows) {
    int id[kRows], rr[kRows], ent[kRows];
    bool any = false;
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      int nr;
      const int s = item_record(min(it0 + j, n_items - 1), rr[j], nr);
      const bool on = lane_on && it0 + j < n_items && s < nr;
      id[j] = on ? int(__ldcg(reinterpret_cast<const unsigned int*>(in_tails + size_t(rr[j]) * cap + s) + 1)) : -1;
      any |= on;
    }
    if (!__any_sync(0xffffffffu, any)) continue;
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      ent[j] = 0;
      if (id[j] >= 0 && sub < N) {                     // lane sub reads the entry of sender sub
        const unsigned long long e = __ldcg(entry_of + size_t(id[j]) - me + sub);
        if (unsigned(e >> 32) == epoch) ent[j] = int(unsigned(e));
      }
    }
    bool own[kRows];
    float4 p[kRows], m[kRows], v[kRows], acc[kRows];
    float pw[kRows], mw[kRows], vw[kRows], accw[kRows];
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      const unsigned have = (__ballot_sync(0xffffffffu, ent[j] != 0) >> (lane_on ? g * LPR : 0)) & grp_mask;
      own[j] = id[j] >= 0 && (__ffs(have) - 1) == rr[j];
      acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      accw[j] = 0.f;
      if (own[j] && !cv.exact) {
        const size_t pi = size_t(id[j]) * Ke + 4 * sub;
        p[j] = *reinterpret_cast<const float4*>(a.fm_v + pi);
        m[j] = *reinterpret_cast<const float4*>(a.v_s0 + pi);
        v[j] = *reinterpret_cast<const float4*>(a.v_s1 + pi);
        if (sub == 0) { pw[j] = a.fm_w[id[j]]; mw[j] = a.w_s0[id[j]]; vw[j] = a.w_s1[id[j]]; }
      }
#pragma unroll
      for (int q = 0; q < kCommMaxRanks; ++q) {        // rank order: the same sum whatever the arrival order
        const int e = __shfl_sync(0xffffffffu, ent[j], lane_on ? g * LPR + (q < LPR ? q : 0) : lane);
        if (own[j] && q < N && e) {
          const size_t rec = size_t(q) * cap + (e - 1);
          const float4 w4 = __ldcg(reinterpret_cast<const float4*>(in_grads + rec * Ke + 4 * sub));
          acc[j].x += w4.x; acc[j].y += w4.y; acc[j].z += w4.z; acc[j].w += w4.w;
          if (sub == 0) accw[j] += __uint_as_float(__ldcg(reinterpret_cast<const unsigned int*>(in_tails + rec)));
        }
      }
    }
#pragma unroll
    for (int j = 0; j < kRows; ++j) {
      n_own += __popc(__ballot_sync(0xffffffffu, own[j] && sub == 0));
      if (!own[j]) continue;
      if (cv.exact) {
        // dense_exact: publish the summed gradient row; every replica sweeps its whole table afterwards
        const size_t gi = size_t(id[j]) * gp + 4 * sub;
        if (cv.mc_g) {
          multimem_st_f4(cv.mc_g + gi, acc[j]);
          if (sub == 0) multimem_st_f1(cv.mc_g + size_t(id[j]) * gp + Ke, accw[j]);
        } else {
          for (int r = 0; r < N; ++r) {
            *reinterpret_cast<float4*>(cv.tab_g[r] + gi) = acc[j];
            if (sub == 0) cv.tab_g[r][size_t(id[j]) * gp + Ke] = accw[j];
          }
        }
        continue;
      }
      const size_t pi = size_t(id[j]) * Ke + 4 * sub;
      opt_apply(c, acc[j].x * inv_n + hyper.l2_reg * p[j].x, p[j].x, m[j].x, v[j].x);
      opt_apply(c, acc[j].y * inv_n + hyper.l2_reg * p[j].y, p[j].y, m[j].y, v[j].y);
      opt_apply(c, acc[j].z * inv_n + hyper.l2_reg * p[j].z, p[j].z, m[j].z, v[j].z);
      opt_apply(c, acc[j].w * inv_n + hyper.l2_reg * p[j].w, p[j].w, m[j].w, v[j].w);
      *reinterpret_cast<float4*>(a.v_s0 + pi) = m[j];
      *reinterpret_cast<float4*>(a.v_s1 + pi) = v[j];
      if (sub == 0) {
        opt_apply(c, accw[j] * inv_n + hyper.l2_reg * pw[j], pw[j], mw[j], vw[j]);
        a.w_s0[id[j]] = mw[j]; a.w_s1[id[j]] = vw[j];
      }
      if (cv.mc_v) {                                   // one store, replicated by the switch into every table
        multimem_st_f4(cv.mc_v + pi, p[j]);
        if (sub == 0) multimem_st_f1(cv.mc_w + id[j], pw[j]);
      } else {
        for (int rq = 1; rq <= N; ++rq) {              // peers first, my own replica last
          const int r = (me + rq) % N;
          *reinterpret_cast<float4*>(cv.tab_v[r] + pi) = p[j];
          if (sub == 0) cv.tab_w[r][id[j]] = pw[j];
        }
      }
    }
  }
  // traffic counters (cumulative since attach; read by bench.py around its timed region): one atomic per CTA
  {
    __shared__ unsigned int rows_cta;
    if (threadIdx.x == 0) rows_cta = 0;
    __syncthreads();
    if (lane == 0 && n_own) atomicAdd(&rows_cta, n_own);
    __syncthreads();
    if (threadIdx.x == 0 && cv.dbg) {
      if (rows_cta) atomicAdd(cv.dbg + 20, static_cast<unsigned long long>(rows_cta));       // rows I published
      if (cta == 0) {
        unsigned long long sent = 0;
        for (int o = 0; o < N; ++o) if (o != me) sent += unsigned(a.st->n_seg[o]);
        cv.dbg[21] += sent;                                                                // records I pushed to peers
        cv.dbg[22] += 1;                                                                   // steps
      }
    }
  }
  stamp(cv, cta, 5);
  // every replica must hold every row before the next step gathers from the tables
  hier_barrier_all(cv, cta, 3, epoch, x.bar_local + 2);
  stamp(cv, cta, 6);
  if (threadIdx.x == 0) kstamp_any(a.st, kStampSparseEnd);
}

// =================================================================================================
// stand-alone all-reduce over a symmetric buffer (the dense half of the step without the optimizer; used by the
// collective sweep of bench.py and available to callers that keep other tensors in symmetric memory)
// =================================================================================================
// buf[r]: rank r's copy of n floats (n % 4 == 0); flags: 2 arrays [world][kCommMaxCtas] per rank behind the data.
// Sum lands in every copy.  Same protocol as dense_sync_kernel: barrier, rank r reduces chunk r and publishes it,
// barrier.  `epoch` must increase by one per call (kept by the caller on the host: this kernel is not graph-replayed).
struct AllReduceArgs {
  float* buf[kCommMaxRanks];
  float* mc;                       // multicast alias of the copies or null
  unsigned int* flags[kCommMaxRanks];
  int world, rank, ctas;
  long long n4;                    // float4 groups
  unsigned int epoch;
  long long watchdog;
};

__global__ void __launch_bounds__(kXThreads, 1)
allreduce_kernel(const __grid_constant__ AllReduceArgs a) {
  const int N = a.world, cta = blockIdx.x;
  auto barrier = [&](int which) {
    __syncthreads();
    if (threadIdx.x < N)
      st_release_sys(a.flags[threadIdx.x] + size_t(which * N + a.rank) * kCommMaxCtas + cta, a.epoch);
    if (threadIdx.x < N) {
      const unsigned int* mine = a.flags[a.rank] + size_t(which * N + threadIdx.x) * kCommMaxCtas + cta;
      const long long t0 = clock64();
      while (ld_relaxed_sys(mine) < a.epoch)
        if (clock64() - t0 > a.watchdog) comm_watchdog_trip(a.rank, cta, 2, threadIdx.x, a.epoch);
      (void)ld_acquire_sys(mine);
    }
    __syncthreads();
  };
  const long long c4 = (a.n4 + N - 1) / N;              // groups per rank chunk
  const long long per = (c4 + a.ctas - 1) / a.ctas;     // groups per CTA sub-slice
  barrier(0);                                           // every rank's input is in place
  const long long lo = a.rank * c4 + cta * per, hi = min(min(lo + per, (a.rank + 1) * c4), a.n4);
  constexpr int kU = 4;                                 // independent groups in flight per thread (bandwidth-bound sizes)
  for (long long g = lo + threadIdx.x; g < hi; g += (long long)blockDim.x * (a.mc ? kU : 1)) {
    if (a.mc) {
      float4 s[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (g + (long long)u * blockDim.x < hi) s[u] = multimem_ld_reduce_f4(a.mc + 4 * (g + (long long)u * blockDim.x));
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (g + (long long)u * blockDim.x < hi) multimem_st_f4(a.mc + 4 * (g + (long long)u * blockDim.x), s[u]);
    } else {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int q = 0; q < N; ++q) {
        const float4 v = ld_volatile_f4(a.buf[q] + 4 * g);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      for (int q = 0; q < N; ++q) st_volatile_f4(a.buf[q] + 4 * g, acc);
    }
  }
  barrier(1);                                           // every chunk has been published everywhere
}

// ------------------------------------------------------------------------------------------------
template <typename K>
static int launch_collective(K real_kernel, K emu_kernel, int ctas, const XchgArgs& x, const XchgArgs* emu,
                             int emu_ranks, cudaStream_t stream) {
  if (emu_ranks <= 0) {
    // programmatic dependent launch: the grid is scheduled while its predecessor in the stream drains, and waits
    // (griddepcontrol.wait) before it reads anything that kernel wrote
    const XchgArgs* none = nullptr;
    DFM_CUDA_CHECK(launch_pdl(kPdlSparse, real_kernel, dim3(ctas), dim3(kXThreads), 0, stream, x, none));
  } else {
    // all virtual ranks' CTAs wait on one another: a cooperative launch guarantees (or refuses) co-residency
    const XchgArgs* ep = emu;
    void* args[] = {const_cast<XchgArgs*>(&x), &ep};
    DFM_CUDA_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(emu_kernel), dim3(ctas * emu_ranks),
                                               dim3(kXThreads), args, 0, stream));
  }
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_dense_sync(const XchgArgs& x, const XchgArgs* emu, int emu_ranks, cudaStream_t stream) {
  return launch_collective(dense_sync_kernel<false>, dense_sync_kernel<true>, x.cv.dense_ctas, x, emu, emu_ranks, stream);
}

int launch_exchange_apply(const XchgArgs& x, const XchgArgs* emu, int emu_ranks, cudaStream_t stream) {
  return launch_collective(exchange_apply_kernel<false>, exchange_apply_kernel<true>, x.cv.ctas, x, emu, emu_ranks, stream);
}

}  // namespace dfm

using namespace dfm;

// data_ptrs / flag_ptrs: `world` device pointers each (peer-mapped); flags = 2 * world * 160 zero-initialised uint32 per
// rank; epoch: 1, 2, 3, ... per call on every rank.
extern "C" int dfm_allreduce_f32(int world, int rank, void* const* data_ptrs, void* data_mc, void* const* flag_ptrs,
                                 long long n, unsigned int epoch, dfm_stream_t stream) {
  DFM_REQUIRE(world >= 2 && world <= kCommMaxRanks && rank >= 0 && rank < world, "dfm_allreduce_f32: world %d rank %d", world, rank);
  DFM_REQUIRE(n > 0 && n % 4 == 0 && epoch > 0, "dfm_allreduce_f32: n must be a positive multiple of 4, epoch >= 1");
  AllReduceArgs a{};
  for (int r = 0; r < world; ++r) {
    a.buf[r] = static_cast<float*>(data_ptrs[r]);
    a.flags[r] = static_cast<unsigned int*>(flag_ptrs[r]);
  }
  a.mc = static_cast<float*>(data_mc);
  a.world = world; a.rank = rank; a.n4 = n / 4; a.epoch = epoch;
  // latency-bound below ~1 MB: few CTAs keep the barrier cheap; bandwidth-bound above: every SM moves data
  const long long per_rank = a.n4 / world;
  a.ctas = int(std::max<long long>(1, std::min<long long>(kCommMaxCtas - 12, (per_rank + kXThreads - 1) / kXThreads)));
  a.watchdog = (long long)(120.0 * 1.9e9);
  allreduce_kernel<<<a.ctas, kXThreads, 0, static_cast<cudaStream_t>(stream)>>>(a);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}
