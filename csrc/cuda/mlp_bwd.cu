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

constexpr int kBwdThreads = 192;


__device__ __forceinline__ int cdiv(int a, int b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------
constexpr int kEbStages = 4;

__global__ void __launch_bounds__(kBwdThreads, 1)
embed_bwd_kernel(const __grid_constant__ FwdMaps maps, const __grid_constant__ BwdParams p) {
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

  if (warp == 0) {
    if (lane == 0) {
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kEbStages;
        mbar_wait(&empty_bar[s], ((kc / kEbStages) & 1) ^ 1, 100);
        uint8_t* a_dst = smem + size_t(s) * kStage;
        // B = 256 rows of W_0 as four 64-row boxes (rows past K0 are zero-filled but still counted)
        mbar_arrive_expect_tx(&full_bar[s], kStage);
        tma_load_2d(a_dst, &maps.dz[0], &full_bar[s], kc * kChunkK, m0);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          tma_load_2d(a_dst + 16384 + j * 8192, &maps.w[0], &full_bar[s], kc * kChunkK, n0 + 64 * j);
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
int embed_bwd_smem() { return kEbStages * (16384 + 32768) + 512 + 1024; }

int embed_bwd_set_smem() {
  DFM_CUDA_CHECK(cudaFuncSetAttribute(embed_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, embed_bwd_smem()));
  return DFM_OK;
}

int launch_embed_bwd(const FwdMaps& maps, const BwdParams& prm, int batch_max, cudaStream_t stream) {
  dim3 grid(ceil_div(batch_max, kTileM), ceil_div(prm.sh.K[0], 256));
  embed_bwd_kernel<<<grid, kBwdThreads, embed_bwd_smem(), stream>>>(maps, prm);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
