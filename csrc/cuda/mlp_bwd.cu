// embed_bwd_kernel: gradient of the embedding input, fused with the FM backward and the sparse scatter.
// Per (128-example tile, 256-column slab of the F*K embedding gradient):
//     de = dz_0 W_0^T + dlogit * (sum_f e - e)        (deep path + FM second order, HVD.py:200-206)
//     g_v[id] += de * x ;  g_w[id] += dlogit * x      (reduced into the compact slot buffer)
// The scatter is fused into the tcgen05 GEMM's epilogue: accumulator rows are read from TMEM and
// red.global.add'ed into G[slot_of[id]]; a field whose 32 lanes all hit the same table row (the 13
// numeric Criteo fields) is first summed across the warp so the hot rows see one atomic per warp
// instead of 32.  Replaces the gather-gradient (UnsortedSegmentSum / IndexedSlices) of TF autodiff
// for HVD.py:197,201.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kBwdThreads = 320;   // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue (two per TMEM lane quarter)


__device__ __forceinline__ int cdiv(int a, int b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------
constexpr int kEbStages = 2;                       // K' = N_0 is short (4 chunks for 256 units)
constexpr uint32_t kEbStage = 16384u + 32768u;     // A [128x64] + B [256x64]
constexpr int kEbRowBytes = 512 + 16;              // one example's 256-column slab of e (bf16) + pad
constexpr int kEbMaxFields = 16;                   // fields per slab = 256 / Ke  (Ke >= 16)

__host__ __device__ inline int eb_s_row_bytes(int Ke) { return Ke * 4 + 16; }

__device__ __forceinline__ void eb_stamp(const BwdParams& p, int slot) {
  if (p.dbg && (threadIdx.x & 31) == 0 && (blockIdx.y == 0 || blockIdx.y == 3)) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    if (blockIdx.x == 0 && blockIdx.y == 0) p.dbg[slot] = t;
    if (blockIdx.y == 3) p.dbg[200 + blockIdx.x * 8 + slot] = t;
  }
}

__global__ void __launch_bounds__(kBwdThreads, 1)
embed_bwd_kernel(const __grid_constant__ FwdMaps maps, const __grid_constant__ BwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* e_tile = smem + kEbStages * kEbStage;                       // [128][kEbRowBytes]
  uint8_t* s_tile = e_tile + kTileM * kEbRowBytes;                     // [128][Ke*4 + 16]
  const int Ke = p.Ke, F = p.F;
  const int GP = g_pitch(Ke);
  const int s_row = eb_s_row_bytes(Ke);
  int2* meta = reinterpret_cast<int2*>(s_tile + kTileM * s_row);       // [128][16] {slot, x}
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(meta) + kTileM * kEbMaxFields * 8);
  uint64_t* empty_bar = full_bar + kEbStages;
  uint64_t* acc_bar = empty_bar + kEbStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) eb_stamp(p, 0);
  kstamp(p.st, kStampEmbedBwd);
  if (threadIdx.x == 0 && p.dbg) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); p.dbg[8 + 2 * (blockIdx.y * gridDim.x + blockIdx.x)] = t; }
  const int m0 = blockIdx.x * kTileM;
  // A slab is four 64-column boxes of de [B, K0] taken with stride gridDim.y, not four adjacent ones:
  // the numeric Criteo fields come first and need ~1/3 of the atomics of a categorical field, so
  // adjacent boxes would give a few CTAs all the cheap fields and the rest all the expensive ones.
  // Local column c (0..255) is global column gcol(c).
  const int n_slabs = gridDim.y;
  auto gcol = [&](int c) { return ((int(blockIdx.y) + n_slabs * (c >> 6)) << 6) + (c & 63); };
  const int count = p.st->count;
  if (m0 >= count) return;
  const int K0 = p.sh.K[0], N0 = p.sh.N[0];
  const int n_here = 256;                          // boxes past K0 are zero-filled by TMA and skipped below
  const int num_kc = cdiv(N0, kChunkK);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.dz[0]);
    tma_prefetch_desc(&maps.w[0]);
    for (int s = 0; s < kEbStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(acc_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) eb_stamp(p, 1);

  if (warp == 0) {
    if (lane == 0) {
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kEbStages;
        mbar_wait(&empty_bar[s], ((kc / kEbStages) & 1) ^ 1, 100);
        uint8_t* a_dst = smem + size_t(s) * kEbStage;
        // B = 256 rows of W_0 as four 64-row boxes (rows past K0 are zero-filled but still counted)
        mbar_arrive_expect_tx(&full_bar[s], kEbStage);
        tma_load_2d(a_dst, &maps.dz[0], &full_bar[s], kc * kChunkK, m0);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          tma_load_2d(a_dst + 16384 + j * 8192, &maps.w[0], &full_bar[s], kc * kChunkK, gcol(64 * j));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kTileM, n_here, 0, 0);
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kEbStages;
        mbar_wait(&full_bar[s], (kc / kEbStages) & 1, 200);
        tc_fence_after_sync();
        const uint32_t a_addr = smem_u32(smem + size_t(s) * kEbStage);
        const uint32_t b_addr = a_addr + 16384u;
        const int kk_n = min(4, (N0 - kc * kChunkK + 15) / 16);
        for (int kk = 0; kk < kk_n; ++kk)
          umma_bf16(tmem_base, umma_smem_desc(a_addr + kk * 32, 1024, 0),
                    umma_smem_desc(b_addr + kk * 32, 1024, 0), idesc, (kc | kk) != 0);
        umma_commit(&empty_bar[s]);
      }
      umma_commit(acc_bar);
      eb_stamp(p, 2);
    }
  } else {
    // Two warps share each TMEM lane quarter (a warp may only read lanes 32*(warp%4)..+31): warp pair (q, half)
    // handles rows 32q..32q+31, local columns [128*half, 128*half+128).  The scatter is bound by how many REDs /
    // stores the SM has in flight, so eight issuing warps instead of four shorten it.
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const int b = m0 + row;
    const bool valid = b < count;
    const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16);
    const float dl = valid ? __ldg(p.dlogit + b) : 0.f;
    const bool all_valid = __all_sync(0xffffffffu, valid);
    const int fpb = 64 / Ke;                        // fields per box (Ke is 16, 32 or 64)
    const int nf = 4 * fpb;                         // metadata slot j <-> field of local column j * Ke

    // ---- while the MMAs run: stage everything the epilogue needs, with all loads in flight at once ----
    // (1) per-(example, field) metadata: slot of the id, feature value.  Three passes so that the loads of
    //     all fields are in flight together (ids -> slots is a dependent chain; one field at a time costs
    //     two memory latencies per field).
    {
      int idv[kEbMaxFields / 2];
      float xv[kEbMaxFields / 2];
      const int j_lo = half * (nf >> 1);            // the fields of this warp's column half
#pragma unroll
      for (int jj = 0; jj < kEbMaxFields / 2; ++jj) {
        const int j = j_lo + jj;
        const int f = gcol(j * Ke) / Ke;
        const bool on = valid && jj < (nf >> 1) && f < F;
        idv[jj] = on ? min(max(__ldg(p.ids + size_t(b) * F + f), 0), p.V - 1) : -1;
        xv[jj] = on ? __ldg(p.vals + size_t(b) * F + f) : 0.f;
      }
      int sl[kEbMaxFields / 2];
#pragma unroll
      for (int jj = 0; jj < kEbMaxFields / 2; ++jj) sl[jj] = idv[jj] >= 0 ? p.slot_of[idv[jj]] : 0;   // encoded: see slot_index()
#pragma unroll
      for (int jj = 0; jj < kEbMaxFields / 2; ++jj)
        if (jj < (nf >> 1)) meta[row * kEbMaxFields + j_lo + jj] = make_int2(sl[jj], __float_as_int(xv[jj]));
    }
    // (2) this warp's 32 rows of the e slab: lane j fetches 16-byte granule j of each row (coalesced)
    {
      // lane j fetches 16-byte granule j of the 512-byte local row = 8 columns of box j / 8
      const int gc = gcol(lane * 8);
      const __nv_bfloat16* e_base = p.act[0] + size_t(m0 + q * 32) * p.sh.pitch[0] + gc - lane * 8;
      const bool g_ok = gc < K0;
#pragma unroll
      for (int r0 = 16 * half; r0 < 16 * half + 16; r0 += 8) {     // this warp stages 16 of the quarter's 32 rows
        uint4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          v[i] = make_uint4(0u, 0u, 0u, 0u);
          if (g_ok && m0 + q * 32 + r0 + i < count)
            v[i] = __ldg(reinterpret_cast<const uint4*>(e_base + size_t(r0 + i) * p.sh.pitch[0]) + lane);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<uint4*>(e_tile + size_t(q * 32 + r0 + i) * kEbRowBytes + lane * 16) = v[i];
      }
    }
    // (3) this warp's 32 rows of sum_f e (contiguous in memory: rows are Ke floats), 8 loads in flight
    {
      const int r_first = q * 32 + 16 * half;         // 16 rows per warp
      const float* s_base = p.s_sum + size_t(m0 + r_first) * Ke;
      const int total4 = 16 * Ke / 4;                 // multiple of 64 (Ke % 16 == 0)
      for (int i0 = lane; i0 < total4; i0 += 32 * 8) {
        float4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = i0 + 32 * u;
          v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (i < total4 && m0 + r_first + (i * 4) / Ke < count) v[u] = __ldg(reinterpret_cast<const float4*>(s_base) + i);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = i0 + 32 * u;
          if (i < total4) {
            const int e0 = i * 4, r = e0 / Ke, c = e0 - r * Ke;
            *reinterpret_cast<float4*>(s_tile + size_t(r_first + r) * s_row + c * 4) = v[u];
          }
        }
      }
    }
    // the two warps of a quarter staged half of its rows each
    asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
    if (warp == 2) eb_stamp(p, 3);

    mbar_wait(acc_bar, 0, 300);
    tc_fence_after_sync();
    if (warp == 2) eb_stamp(p, 4);

    const uint8_t* my_e = e_tile + size_t(row) * kEbRowBytes;
    const float* my_s = reinterpret_cast<const float*>(s_tile + size_t(row) * s_row);
    int cur_j = -1, slot = 0;
    bool multi = true;        // this (example, field)'s id occurs elsewhere in the batch too
    float x = 0.f;
    bool uniform = false;     // every lane of the warp maps this field to the same slot
    for (int c0 = 128 * half; c0 < 128 * half + 128; c0 += 16) {
      uint32_t r[16];
      tmem_ld_x16(t_row + c0, r);
      tmem_ld_wait();
      if (gcol(c0) >= K0) continue;             // warp-uniform: an empty box (past the last field)
      // a 16-column group lies inside one field (Ke % 16 == 0, checked by the host)
      const int j = c0 / Ke;
      const int k0 = c0 - j * Ke;
      if (j != cur_j) {
        cur_j = j;
        const int2 mt = meta[row * kEbMaxFields + j];
        slot = slot_index(mt.x);
        multi = slot_multi(mt.x);
        x = __int_as_float(mt.y);
        const int s0 = __shfl_sync(0xffffffffu, slot, 0);
        uniform = all_valid && __all_sync(0xffffffffu, slot == s0);
        // first-order weight gradient, once per (example, field)
        float gw = dl * x;
        if (uniform) {
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) gw += __shfl_xor_sync(0xffffffffu, gw, o);
          if (lane == 0) atomicAdd(p.G + size_t(slot) * GP + Ke, gw);
        } else if (valid) {
          if (multi) atomicAdd(p.G + size_t(slot) * GP + Ke, gw);
          else p.G[size_t(slot) * GP + Ke] = gw;
        }
      }
      float g[16];
      {
        const uint4 e0 = *reinterpret_cast<const uint4*>(my_e + c0 * 2);
        const uint4 e1 = *reinterpret_cast<const uint4*>(my_e + c0 * 2 + 16);
        const uint32_t ev[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
        float sv[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 t = *reinterpret_cast<const float4*>(my_s + k0 + 4 * i);
          sv[4 * i] = t.x; sv[4 * i + 1] = t.y; sv[4 * i + 2] = t.z; sv[4 * i + 3] = t.w;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&ev[i]);
          g[2 * i] = (__uint_as_float(r[2 * i]) + dl * (sv[2 * i] - __low2float(v))) * x;
          g[2 * i + 1] = (__uint_as_float(r[2 * i + 1]) + dl * (sv[2 * i + 1] - __high2float(v))) * x;
        }
      }
      if (uniform) {
        // reduce-scatter 16 values over 32 lanes: lane (i & 15) ends with column i's total
#pragma unroll
        for (int off = 8; off >= 1; off >>= 1) {
          const bool upper = (lane & off) != 0;
#pragma unroll
          for (int jj = 0; jj < off; ++jj) {
            const float send = upper ? g[jj] : g[jj + off];
            const float keepv = upper ? g[jj + off] : g[jj];
            g[jj] = keepv + __shfl_xor_sync(0xffffffffu, send, off);
          }
        }
        g[0] += __shfl_xor_sync(0xffffffffu, g[0], 16);
        if (lane < 16) atomicAdd(p.G + size_t(slot) * GP + k0 + lane, g[0]);
      } else if (valid) {
        float* dst = p.G + size_t(slot) * GP + k0;
        if (multi) {
          red_add_v4(dst, g[0], g[1], g[2], g[3]);
          red_add_v4(dst + 4, g[4], g[5], g[6], g[7]);
          red_add_v4(dst + 8, g[8], g[9], g[10], g[11]);
          red_add_v4(dst + 12, g[12], g[13], g[14], g[15]);
        } else {
          // the only occurrence of this id in the batch: the row is zero (the optimizer cleared it), a
          // plain store is the sum -- and costs the LSU a fraction of what four vector REDs do
          st_global_v8f(dst, g[0], g[1], g[2], g[3], g[4], g[5], g[6], g[7]);
          st_global_v8f(dst + 8, g[8], g[9], g[10], g[11], g[12], g[13], g[14], g[15]);
        }
      }
    }
  }

  if (warp == 2) eb_stamp(p, 5);
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
  if (threadIdx.x == 0) eb_stamp(p, 6);
  if (threadIdx.x == 0 && p.dbg) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); p.dbg[9 + 2 * (blockIdx.y * gridDim.x + blockIdx.x)] = t; }
}

// ------------------------------------- host side -------------------------------------------
int embed_bwd_smem(int Ke) {
  return kEbStages * int(kEbStage) + kTileM * kEbRowBytes + kTileM * eb_s_row_bytes(Ke) + kTileM * kEbMaxFields * 8 + 512 + 1024;
}

int embed_bwd_set_smem(int Ke) {
  DFM_REQUIRE(embed_bwd_smem(Ke) <= 227 * 1024, "embedding_size %d needs too much shared memory in embed_bwd", Ke);
  DFM_CUDA_CHECK(cudaFuncSetAttribute(embed_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, embed_bwd_smem(Ke)));
  return DFM_OK;
}

int launch_embed_bwd(const FwdMaps& maps, const BwdParams& prm, int batch_max, cudaStream_t stream) {
  dim3 grid(ceil_div(batch_max, kTileM), ceil_div(prm.sh.K[0], 256));
  embed_bwd_kernel<<<grid, kBwdThreads, embed_bwd_smem(prm.Ke), stream>>>(maps, prm);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
