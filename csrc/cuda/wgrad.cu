// Weight gradients of all MLP layers (and the out layer) in one launch.
//
//   dP_l [K_l+1, N_l] += act_l[:, 0..K_l]^T  dz_l          (last row = bias gradient, because
//                                                           column K_l of act_l is the constant 1)
//
// Both operands are used exactly as the forward/backward kernels left them in global memory
// (row-major [batch, features]); with the batch as the GEMM K dimension that makes them
// MN-major UMMA operands, so no transposed copies exist anywhere.  The batch is split across
// CTAs (split-K) so the launch fills the chip; every split writes its partial tile with plain
// vector stores into its own slice of the partial buffer [ksplit][n_dense] (no atomics, no
// zero-fill pass, deterministic), and the optimizer kernel adds the slices when it reads them.
// Replaces the dW GEMMs + BiasAddGrad of TF autodiff for HVD.py:224-243.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kWgThreads = 192;
constexpr int kWgStages = 4;


__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_kernel(const __grid_constant__ WgradMaps maps, const __grid_constant__ WgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t kStage = 16384u + 32768u;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kWgStages * kStage);
  uint64_t* empty_bar = full_bar + kWgStages;
  uint64_t* acc_bar = empty_bar + kWgStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);
  // no early pdl_launch_dependents() here: dense_opt's 1184 small CTAs would fill every SM's thread slots while they
  // wait and keep embed_bwd (other stream) from being scheduled; it is released when this grid's CTAs exit

  const WgradItem item = p.items[blockIdx.x];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int count = p.st->count;
  const int k_lim = min(item.k_end, (count + 63) & ~63);
  const int l = item.layer;
  if (item.k_begin >= k_lim) {
    pdl_wait();
    // this batch range is empty (short batch): the slice still has to read as zero
    const int Nl = p.sh.N[l];
    float* base = p.dP[l] + size_t(item.ks) * p.part_stride;
    for (int r = threadIdx.x; r < kTileM; r += blockDim.x) {
      const int r_out = item.m0 + r;
      if (r_out >= p.sh.K[l] + 1) break;
      for (int c = item.n0; c < min(Nl, item.n0 + item.n_here); ++c) base[size_t(r_out) * Nl + c] = 0.f;
    }
    return;
  }
  const int num_kc = (k_lim - item.k_begin) / kChunkK;
  const int b_atoms = (item.n_here + 63) / 64;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.act[l]);
    tma_prefetch_desc(&maps.dz[l]);
    for (int s = 0; s < kWgStages; ++s) {
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
  pdl_wait();                              // set-up done; the activations and dz come from the preceding kernels
  kstamp(p.st, kStampWgrad);

  if (warp == 0) {
    if (lane == 0) {
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kWgStages;
        mbar_wait(&empty_bar[s], ((kc / kWgStages) & 1) ^ 1, 100);
        uint8_t* a_dst = smem + size_t(s) * kStage;
        uint8_t* b_dst = a_dst + 16384;
        const int k0 = item.k_begin + kc * kChunkK;
        mbar_arrive_expect_tx(&full_bar[s], 16384u + uint32_t(b_atoms) * 8192u);
        tma_load_2d(a_dst, &maps.act[l], &full_bar[s], item.m0, k0);
        tma_load_2d(a_dst + 8192, &maps.act[l], &full_bar[s], item.m0 + 64, k0);
        for (int a = 0; a < b_atoms; ++a)
          tma_load_2d(b_dst + a * 8192, &maps.dz[l], &full_bar[s], item.n0 + a * 64, k0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kTileM, item.n_here, 1, 1);
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kWgStages;
        mbar_wait(&full_bar[s], (kc / kWgStages) & 1, 200);
        tc_fence_after_sync();
        const uint32_t a_addr = smem_u32(smem + size_t(s) * kStage);
        const uint32_t b_addr = a_addr + 16384u;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16(tmem_base, umma_smem_desc(a_addr + kk * 2048, 1024, 8192),
                    umma_smem_desc(b_addr + kk * 2048, 1024, 8192), idesc, (kc | kk) != 0);
        umma_commit(&empty_bar[s]);
      }
      umma_commit(acc_bar);
    }
  } else {
    const int q = warp & 3;
    const int r_out = item.m0 + q * 32 + lane;        // row of dP_l
    const int Nl = p.sh.N[l];
    const bool row_ok = r_out < p.sh.K[l] + 1;
    float* dst_row = p.dP[l] + size_t(item.ks) * p.part_stride + size_t(r_out) * Nl;
    const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16);
    mbar_wait(acc_bar, 0, 300);
    tc_fence_after_sync();
    for (int c0 = 0; c0 < item.n_here; c0 += 16) {
      uint32_t r[16];
      tmem_ld_x16(t_row + c0, r);
      tmem_ld_wait();
      if (!row_ok) continue;
      const int c = item.n0 + c0;
      if (Nl == 1) {
        if (c0 == 0) dst_row[0] = __uint_as_float(r[0]);
      } else {
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          if (c + 4 * v < Nl)
            *reinterpret_cast<float4*>(dst_row + c + 4 * v) =
                make_float4(__uint_as_float(r[4 * v]), __uint_as_float(r[4 * v + 1]),
                            __uint_as_float(r[4 * v + 2]), __uint_as_float(r[4 * v + 3]));
        }
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

int wgrad_smem() { return kWgStages * (16384 + 32768) + 512 + 1024; }

int wgrad_set_smem() {
  DFM_CUDA_CHECK(cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wgrad_smem()));
  return DFM_OK;
}

int launch_wgrad(const WgradMaps& maps, const WgradParams& prm, int n_items, cudaStream_t stream) {
  // plain launch: on its stream it follows a cross-stream event wait (not a kernel), so there is nothing to overlap
  wgrad_kernel<<<n_items, kWgThreads, wgrad_smem(), stream>>>(maps, prm);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
