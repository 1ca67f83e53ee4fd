// Backward data-gradient path (what TF autodiff generates for HVD.py:196-249).
//
// mlp_bwd_chain_kernel   per 128-example tile:
//     dz[L-1] = dlogit (x) w_out, masked by h_L > 0 (ReLU and dropout share the zero pattern)
//     for l = L-1 .. 1:  dh_l = dz_l W_l^T   (tcgen05, accumulator in TMEM)
//                        dz_{l-1} = dh_l masked by h_l > 0, / keep_{l-1}
//   dz_l goes to global memory (bf16, operand of the weight-gradient GEMM) and into shared
//   memory as the next GEMM's swizzled A operand, so the chain never leaves the SM.
//
// embed_bwd_kernel       per (tile, 256-column slab of the F*K embedding gradient):
//     de = dz_0 W_0^T + dlogit * (sum_f e - e)        (deep path + FM second order)
//     g_v[id] += de * x ;  g_w[id] += dlogit * x      (reduced into the compact slot buffer)
//   The scatter is fused into the GEMM epilogue: accumulator rows are read from TMEM and
//   red.global.add'ed into G[slot_of[id]]; a field whose 32 lanes all hit the same table
//   row (the 13 numeric Criteo fields) is first summed across the warp so the hot rows
//   see one atomic per warp instead of 32.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kBwdThreads = 192;


__device__ __forceinline__ int cdiv(int a, int b) { return (a + b - 1) / b; }

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// Writes 16 consecutive bf16 (packed in pk[8]) of row `row`, starting at column c0, into a
// K-major SWIZZLE_128B operand region made of 16 KiB chunks of 64 columns.
__device__ __forceinline__ void store_a_operand(uint32_t region, int row, int c0, const uint32_t (&pk)[8]) {
  const uint32_t chunk = region + uint32_t(c0 >> 6) * 16384u;
  const uint32_t g0 = uint32_t(c0 & 63) >> 3;
  st_shared_v4(chunk + sw128_offset(row, g0), pk[0], pk[1], pk[2], pk[3]);
  st_shared_v4(chunk + sw128_offset(row, g0 + 1), pk[4], pk[5], pk[6], pk[7]);
}

__global__ void __launch_bounds__(kBwdThreads, 1)
mlp_bwd_chain_kernel(const __grid_constant__ BwdMaps maps, const __grid_constant__ BwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const int S = p.stages;
  const uint32_t stage_bytes = uint32_t(p.b_slot_bytes);
  uint8_t* a_region = smem + size_t(S) * stage_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(a_region + size_t(p.a_region_chunks) * 16384);
  uint64_t* empty_bar = full_bar + S;
  uint64_t* acc_bar = empty_bar + S;
  uint64_t* aready_bar = acc_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aready_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kTileM;
  const int L = p.sh.L;
  const int count = p.st->count;
  if (m0 >= count) return;

  int max_k = 32;
  for (int l = 1; l < L; ++l) max_k = max(max_k, p.sh.K[l]);
  uint32_t tmem_cols = 32;
  while (tmem_cols < uint32_t(max_k)) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    for (int l = 1; l < L; ++l) tma_prefetch_desc(&maps.w[l]);
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(acc_bar, 1);
    mbar_init(aready_bar, 128);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // producer: weights of layers L-1 .. 1, K-major ([K_l rows] x 64 columns of N_l per chunk)
    if (lane == 0) {
      int it = 0;
      for (int l = L - 1; l >= 1; --l) {
        const int Kl = p.sh.K[l], Nl = p.sh.N[l];
        const int n_pass = cdiv(Kl, 256);
        const int num_kc = cdiv(Nl, kChunkK);
        for (int ps = 0; ps < n_pass; ++ps) {
          const int rows = min(256, Kl - 256 * ps);
          for (int kc = 0; kc < num_kc; ++kc, ++it) {
            const int s = it % S;
            mbar_wait(&empty_bar[s], ((it / S) & 1) ^ 1, 100 + l);
            // the box is min(256, K_l) rows tall; rows past K_l are zero-filled but still counted
            mbar_arrive_expect_tx(&full_bar[s], uint32_t(min(256, Kl)) * 128u);
            tma_load_2d(smem + size_t(s) * stage_bytes, &maps.w[l], &full_bar[s], kc * kChunkK, 256 * ps);
            (void)rows;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int it = 0, gemm = 0;
      for (int l = L - 1; l >= 1; --l, ++gemm) {
        const int Kl = p.sh.K[l], Nl = p.sh.N[l];
        const int n_pass = cdiv(Kl, 256);
        const int num_kc = cdiv(Nl, kChunkK);
        mbar_wait(aready_bar, gemm & 1, 250 + l);
        tc_fence_after_sync();
        for (int ps = 0; ps < n_pass; ++ps) {
          const int n_here = min(256, Kl - 256 * ps);
          const uint32_t idesc = umma_idesc_bf16(kTileM, n_here, 0, 0);
          const uint32_t d_tmem = tmem_base + 256u * ps;
          for (int kc = 0; kc < num_kc; ++kc, ++it) {
            const int s = it % S;
            mbar_wait(&full_bar[s], (it / S) & 1, 200 + l);
            tc_fence_after_sync();
            const uint32_t a_addr = smem_u32(a_region + size_t(kc) * 16384);
            const uint32_t b_addr = smem_u32(smem + size_t(s) * stage_bytes);
            const int kk_n = min(4, (Nl - kc * kChunkK) / 16);
            for (int kk = 0; kk < kk_n; ++kk)
              umma_bf16(d_tmem, umma_smem_desc(a_addr + kk * 32, 1024, 0),
                        umma_smem_desc(b_addr + kk * 32, 1024, 0), idesc, (kc | kk) != 0);
            umma_commit(&empty_bar[s]);
          }
        }
        umma_commit(acc_bar);
      }
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int b = m0 + row;
    const bool valid = b < count;
    const int train = p.st->train;
    const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16);
    const uint32_t a_region_addr = smem_u32(a_region);
    const float dl = valid ? __ldg(p.dlogit + b) : 0.f;

    // ---- dz[L-1] from the out layer; dz[L] = [dlogit, 0, ... 0] ----
    {
      const int l = L - 1;
      const int Nl = p.sh.N[l];
      const float keep = p.sh.keep[l];
      const float scale = (train && keep < 1.f) ? 1.f / keep : 1.f;
      const __nv_bfloat16* h = p.act[L] + size_t(b) * p.sh.pitch[L];
      __nv_bfloat16* dzo = p.dz[l] + size_t(b) * Nl;
      for (int c0 = 0; c0 < Nl; c0 += 16) {
        uint32_t hv[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (valid) {
          const uint4 h0 = *reinterpret_cast<const uint4*>(h + c0);
          const uint4 h1 = *reinterpret_cast<const uint4*>(h + c0 + 8);
          hv[0] = h0.x; hv[1] = h0.y; hv[2] = h0.z; hv[3] = h0.w;
          hv[4] = h1.x; hv[5] = h1.y; hv[6] = h1.z; hv[7] = h1.w;
        }
        uint32_t pk[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&hv[i]);
          const float d0 = __low2float(v) > 0.f ? dl * __ldg(p.w_out + c0 + 2 * i) * scale : 0.f;
          const float d1 = __high2float(v) > 0.f ? dl * __ldg(p.w_out + c0 + 2 * i + 1) * scale : 0.f;
          pk[i] = pack_bf16x2(d0, d1);
        }
        uint4* g = reinterpret_cast<uint4*>(dzo + c0);
        g[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        g[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        if (L > 1) store_a_operand(a_region_addr, row, c0, pk);
      }
      uint4* go = reinterpret_cast<uint4*>(p.dz[L] + size_t(b) * 16);
      go[0] = make_uint4(pack_bf16x2(dl, 0.f), 0u, 0u, 0u);
      go[1] = make_uint4(0u, 0u, 0u, 0u);
      if (L > 1) {
        fence_proxy_async_smem();
        tc_fence_before_sync();
        mbar_arrive(aready_bar);
      }
    }

    int gemm = 0;
    for (int l = L - 1; l >= 1; --l, ++gemm) {
      const int Kl = p.sh.K[l];                 // columns of dh_l == N_{l-1}
      const float keep = p.sh.keep[l - 1];
      const float scale = (train && keep < 1.f) ? 1.f / keep : 1.f;
      const __nv_bfloat16* h = p.act[l] + size_t(b) * p.sh.pitch[l];
      __nv_bfloat16* dzo = p.dz[l - 1] + size_t(b) * Kl;
      mbar_wait(acc_bar, gemm & 1, 300 + l);
      tc_fence_after_sync();
      for (int c0 = 0; c0 < Kl; c0 += 16) {
        uint32_t r[16];
        tmem_ld_x16(t_row + c0, r);
        tmem_ld_wait();
        uint32_t hv[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (valid) {
          const uint4 h0 = *reinterpret_cast<const uint4*>(h + c0);
          const uint4 h1 = *reinterpret_cast<const uint4*>(h + c0 + 8);
          hv[0] = h0.x; hv[1] = h0.y; hv[2] = h0.z; hv[3] = h0.w;
          hv[4] = h1.x; hv[5] = h1.y; hv[6] = h1.z; hv[7] = h1.w;
        }
        uint32_t pk[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&hv[i]);
          const float d0 = __low2float(v) > 0.f ? __uint_as_float(r[2 * i]) * scale : 0.f;
          const float d1 = __high2float(v) > 0.f ? __uint_as_float(r[2 * i + 1]) * scale : 0.f;
          pk[i] = pack_bf16x2(d0, d1);
        }
        uint4* g = reinterpret_cast<uint4*>(dzo + c0);
        g[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        g[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        if (l > 1) store_a_operand(a_region_addr, row, c0, pk);
      }
      if (l > 1) {
        fence_proxy_async_smem();
        tc_fence_before_sync();
        mbar_arrive(aready_bar);
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// ---------------------------------------------------------------------------------------------
constexpr int kEbStages = 4;

__global__ void __launch_bounds__(kBwdThreads, 1)
embed_bwd_kernel(const __grid_constant__ BwdMaps maps, const __grid_constant__ BwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t kStage = 16384u + 32768u;   // A [128x64] + B [256x64]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kEbStages * kStage);
  uint64_t* empty_bar = full_bar + kEbStages;
  uint64_t* acc_bar = empty_bar + kEbStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kTileM;
  const int n0 = blockIdx.y * 256;                 // first column of this slab of de [B, K0]
  const int count = p.st->count;
  if (m0 >= count) return;
  const int K0 = p.sh.K[0], N0 = p.sh.N[0];
  const int n_here = min(256, ((K0 - n0) + 15) & ~15);
  const int num_kc = cdiv(N0, kChunkK);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.dz0);
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

  if (warp == 0) {
    if (lane == 0) {
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kEbStages;
        mbar_wait(&empty_bar[s], ((kc / kEbStages) & 1) ^ 1, 100);
        uint8_t* a_dst = smem + size_t(s) * kStage;
        // W_0's box is min(256, K0) rows tall (rows past K0 are zero-filled but counted)
        mbar_arrive_expect_tx(&full_bar[s], 16384u + uint32_t(min(256, K0)) * 128u);
        tma_load_2d(a_dst, &maps.dz0, &full_bar[s], kc * kChunkK, m0);
        tma_load_2d(a_dst + 16384, &maps.w[0], &full_bar[s], kc * kChunkK, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kTileM, n_here, 0, 0);
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kEbStages;
        mbar_wait(&full_bar[s], (kc / kEbStages) & 1, 200);
        tc_fence_after_sync();
        const uint32_t a_addr = smem_u32(smem + size_t(s) * kStage);
        const uint32_t b_addr = a_addr + 16384u;
        const int kk_n = min(4, (N0 - kc * kChunkK + 15) / 16);
        for (int kk = 0; kk < kk_n; ++kk)
          umma_bf16(tmem_base, umma_smem_desc(a_addr + kk * 32, 1024, 0),
                    umma_smem_desc(b_addr + kk * 32, 1024, 0), idesc, (kc | kk) != 0);
        umma_commit(&empty_bar[s]);
      }
      umma_commit(acc_bar);
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int b = m0 + row;
    const bool valid = b < count;
    const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16);
    const int Ke = p.Ke, F = p.F;
    const float dl = valid ? __ldg(p.dlogit + b) : 0.f;
    const __nv_bfloat16* e_row = p.act[0] + size_t(valid ? b : 0) * p.sh.pitch[0];
    const float* s_row = p.s_sum + size_t(valid ? b : 0) * Ke;
    const int* id_row = p.ids + size_t(valid ? b : 0) * F;
    const float* x_row = p.vals + size_t(valid ? b : 0) * F;
    const bool all_valid = __all_sync(0xffffffffu, valid);

    mbar_wait(acc_bar, 0, 300);
    tc_fence_after_sync();

    int cur_f = -1, slot = 0;
    float x = 0.f;
    bool uniform = false;     // every lane of the warp maps this field to the same slot
    for (int c0 = 0; c0 < n_here; c0 += 16) {
      uint32_t r[16];
      tmem_ld_x16(t_row + c0, r);
      tmem_ld_wait();
      const int c = n0 + c0;
      if (c >= K0) break;                       // warp-uniform
      // a 16-column group lies inside one field when Ke % 16 == 0 (checked by the host)
      const int f = c / Ke;
      const int k0 = c - f * Ke;
      if (f != cur_f) {
        cur_f = f;
        slot = 0; x = 0.f;
        if (valid) {
          const int id = __ldg(id_row + f);
          x = __ldg(x_row + f);
          slot = p.slot_of[id];
        }
        const int s0 = __shfl_sync(0xffffffffu, slot, 0);
        uniform = all_valid && __all_sync(0xffffffffu, slot == s0);
        // first-order weight gradient, once per (example, field)
        float gw = dl * x;
        if (uniform) {
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) gw += __shfl_xor_sync(0xffffffffu, gw, o);
          if (lane == 0) atomicAdd(p.Gw + slot, gw);
        } else if (valid) {
          atomicAdd(p.Gw + slot, gw);
        }
      }
      float g[16];
      {
        uint32_t ev[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (valid) {
          const uint4 e0 = *reinterpret_cast<const uint4*>(e_row + c);
          const uint4 e1 = *reinterpret_cast<const uint4*>(e_row + c + 8);
          ev[0] = e0.x; ev[1] = e0.y; ev[2] = e0.z; ev[3] = e0.w;
          ev[4] = e1.x; ev[5] = e1.y; ev[6] = e1.z; ev[7] = e1.w;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&ev[i]);
          const float s0v = valid ? __ldg(s_row + k0 + 2 * i) : 0.f;
          const float s1v = valid ? __ldg(s_row + k0 + 2 * i + 1) : 0.f;
          g[2 * i] = (__uint_as_float(r[2 * i]) + dl * (s0v - __low2float(v))) * x;
          g[2 * i + 1] = (__uint_as_float(r[2 * i + 1]) + dl * (s1v - __high2float(v))) * x;
        }
      }
      if (uniform) {
        // reduce-scatter 16 values over 32 lanes: lane (i & 15) ends with column i's total
#pragma unroll
        for (int off = 8; off >= 1; off >>= 1) {
          const bool upper = (lane & off) != 0;
#pragma unroll
          for (int j = 0; j < off; ++j) {
            const float send = upper ? g[j] : g[j + off];
            const float keepv = upper ? g[j + off] : g[j];
            g[j] = keepv + __shfl_xor_sync(0xffffffffu, send, off);
          }
        }
        g[0] += __shfl_xor_sync(0xffffffffu, g[0], 16);
        if (lane < 16) atomicAdd(p.G + size_t(slot) * Ke + k0 + lane, g[0]);
      } else if (valid) {
        float* dst = p.G + size_t(slot) * Ke + k0;
        red_add_v4(dst, g[0], g[1], g[2], g[3]);
        red_add_v4(dst + 4, g[4], g[5], g[6], g[7]);
        red_add_v4(dst + 8, g[8], g[9], g[10], g[11]);
        red_add_v4(dst + 12, g[12], g[13], g[14], g[15]);
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// ------------------------------------- host side -------------------------------------------
int mlp_bwd_plan(const MlpShape& sh, int* stages, int* b_slot_bytes, int* a_region_chunks) {
  int max_rows = 8, chunks = 1;
  for (int l = 1; l < sh.L; ++l) {
    max_rows = std::max(max_rows, std::min(256, sh.K[l]));
    chunks = std::max(chunks, ceil_div(sh.N[l], 64));
  }
  chunks = std::max(chunks, ceil_div(sh.N[sh.L - 1], 64));
  *b_slot_bytes = round_up(max_rows * 128, 1024);
  *a_region_chunks = chunks;
  const int fixed = chunks * 16384 + 512 + 1024;
  int s = (227 * 1024 - fixed) / *b_slot_bytes;
  if (s < 2) return -1;
  if (s > 4) s = 4;
  *stages = s;
  return fixed + s * *b_slot_bytes;
}

int embed_bwd_smem() { return kEbStages * (16384 + 32768) + 512 + 1024; }

int mlp_bwd_set_smem(int chain_bytes) {
  DFM_CUDA_CHECK(cudaFuncSetAttribute(mlp_bwd_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, chain_bytes));
  DFM_CUDA_CHECK(cudaFuncSetAttribute(embed_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, embed_bwd_smem()));
  return DFM_OK;
}

int launch_mlp_bwd_chain(const BwdMaps& maps, const BwdParams& prm, int batch_max, int smem_bytes,
                         cudaStream_t stream) {
  mlp_bwd_chain_kernel<<<ceil_div(batch_max, kTileM), kBwdThreads, smem_bytes, stream>>>(maps, prm);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_embed_bwd(const BwdMaps& maps, const BwdParams& prm, int batch_max, cudaStream_t stream) {
  dim3 grid(ceil_div(batch_max, kTileM), ceil_div(prm.sh.K[0], 256));
  embed_bwd_kernel<<<grid, kBwdThreads, embed_bwd_smem(), stream>>>(maps, prm);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
