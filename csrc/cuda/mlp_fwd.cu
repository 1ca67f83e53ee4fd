// deepfm_mlp_fwd: the whole deep tower for a 128-example tile in ONE kernel.
//
//   h_{l+1} = dropout_keep[l]( relu(h_l W_l + b_l) ),  l = 0..L-1        (HVD.py:223-235)
//   y_d     = h_L w_out + b_out                                           (HVD.py:237-244)
//   y       = fm_bias + y_w + y_v + y_d ; prob = sigmoid(y)               (HVD.py:247-249)
//   loss   += BCE(y, label) ; dlogit = (prob - label) / batch             (HVD.py:264-265)
//
// Every layer is a tcgen05 GEMM whose accumulator stays in TMEM; the epilogue warps read
// it back, apply bias/ReLU/dropout, and write the bf16 result (a) to global memory for the
// backward pass and (b) straight into shared memory in the 128-byte-swizzled K-major
// layout the next layer's UMMA reads as its A operand -- activations never round-trip
// through HBM between layers.  Weights stream through a TMA ring as MN-major B operands,
// i.e. in their natural TF layout [in, out].
//
// Warp roles: 0 = TMA producer, 1 = TMEM owner + UMMA issuer, 2..5 = epilogue.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kMlpThreads = 192;


__device__ __forceinline__ int ceil_div_d(int a, int b) { return (a + b - 1) / b; }

__global__ void __launch_bounds__(kMlpThreads, 1)
mlp_fwd_kernel(const __grid_constant__ FwdMaps maps, const __grid_constant__ FwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const int S = p.stages;
  const uint32_t stage_bytes = 16384u + uint32_t(p.b_slot_bytes);
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
  if (m0 >= count) return;   // whole tile past the batch (uniform per CTA, before any barrier)

  int max_n = 32;
  for (int l = 0; l < L; ++l) max_n = max(max_n, p.sh.N[l]);
  uint32_t tmem_cols = 32;
  while (tmem_cols < uint32_t(max_n)) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.e);
    for (int l = 0; l < L; ++l) tma_prefetch_desc(&maps.w[l]);
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
    // ============================== TMA producer ==============================
    if (lane == 0) {
      int it = 0;
      for (int l = 0; l < L; ++l) {
        const int n_atoms = ceil_div_d(p.sh.N[l], 64);
        const int n_pass = ceil_div_d(n_atoms, 4);
        const int num_kc = ceil_div_d(p.sh.K[l], kChunkK);
        for (int ps = 0; ps < n_pass; ++ps) {
          const int atoms = min(4, n_atoms - 4 * ps);
          for (int kc = 0; kc < num_kc; ++kc, ++it) {
            const int s = it % S;
            mbar_wait(&empty_bar[s], ((it / S) & 1) ^ 1, 100 + l);
            uint8_t* a_dst = smem + size_t(s) * stage_bytes;
            uint8_t* b_dst = a_dst + 16384;
            mbar_arrive_expect_tx(&full_bar[s], uint32_t(atoms) * 8192u + (l == 0 ? 16384u : 0u));
            if (l == 0) tma_load_2d(a_dst, &maps.e, &full_bar[s], kc * kChunkK, m0);
            for (int a = 0; a < atoms; ++a)
              tma_load_2d(b_dst + a * 8192, &maps.w[l], &full_bar[s], (4 * ps + a) * 64, kc * kChunkK);
          }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== UMMA issuer ===============================
    if (lane == 0) {
      int it = 0;
      for (int l = 0; l < L; ++l) {
        const int Nl = p.sh.N[l], Kl = p.sh.K[l];
        const int n_atoms = ceil_div_d(Nl, 64);
        const int n_pass = ceil_div_d(n_atoms, 4);
        const int num_kc = ceil_div_d(Kl, kChunkK);
        if (l > 0) {
          mbar_wait(aready_bar, (l - 1) & 1, 250 + l);   // A_l is in smem, TMEM has been drained
          tc_fence_after_sync();
        }
        for (int ps = 0; ps < n_pass; ++ps) {
          const int n_here = min(256, Nl - 256 * ps);
          const uint32_t idesc = umma_idesc_bf16(kTileM, n_here, 0, 1);
          const uint32_t d_tmem = tmem_base + 256u * ps;
          for (int kc = 0; kc < num_kc; ++kc, ++it) {
            const int s = it % S;
            mbar_wait(&full_bar[s], (it / S) & 1, 200 + l);
            tc_fence_after_sync();
            const uint32_t stage_addr = smem_u32(smem + size_t(s) * stage_bytes);
            const uint32_t a_addr = (l == 0) ? stage_addr : smem_u32(a_region + size_t(kc) * 16384);
            const uint32_t b_addr = stage_addr + 16384u;
            const int kk_n = (l == 0) ? 4 : min(4, (Kl - kc * kChunkK) / 16);
            for (int kk = 0; kk < kk_n; ++kk) {
              const uint64_t adesc = umma_smem_desc(a_addr + kk * 32, 1024, 0);
              const uint64_t bdesc = umma_smem_desc(b_addr + kk * 2048, 1024, 8192);
              umma_bf16(d_tmem, adesc, bdesc, idesc, (kc | kk) != 0);
            }
            umma_commit(&empty_bar[s]);
          }
        }
        umma_commit(acc_bar);
      }
    }
  } else {
    // ================================ epilogue =================================
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int b = m0 + row;
    const bool valid = b < count;
    const int train = p.st->train;
    const uint32_t step = uint32_t(p.st->step);
    const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16);
    const uint32_t a_region_addr = smem_u32(a_region);
    float y_d = 0.f;

    for (int l = 0; l < L; ++l) {
      const int Nl = p.sh.N[l];
      const bool last = (l == L - 1);
      const float* bias = p.P[l] + size_t(p.sh.K[l]) * Nl;
      const float keep = p.sh.keep[l];
      const bool drop = train && keep < 1.f;
      const float inv_keep = drop ? 1.f / keep : 1.f;
      const float* w_out = p.P[L];
      __nv_bfloat16* out_row = p.act[l + 1] + size_t(b) * p.sh.pitch[l + 1];

      mbar_wait(acc_bar, l & 1, 300 + l);
      tc_fence_after_sync();

      for (int c0 = 0; c0 < Nl; c0 += 16) {
        uint32_t r[16];
        tmem_ld_x16(t_row + c0, r);
        tmem_ld_wait();
        float a[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fmaxf(__uint_as_float(r[i]) + __ldg(bias + c0 + i), 0.f);
        if (drop) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint32_t rnd[4];
            philox4x32(uint32_t(p.seed), uint32_t(p.seed >> 32), uint32_t(b), uint32_t(l),
                       uint32_t((c0 >> 2) + g), step, rnd);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              a[g * 4 + i] = (u32_to_unit(rnd[i]) < keep) ? a[g * 4 + i] * inv_keep : 0.f;
          }
        }
        uint32_t pk[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) pk[i] = valid ? pack_bf16x2(a[2 * i], a[2 * i + 1]) : 0u;
        if (valid) {
          uint4* g = reinterpret_cast<uint4*>(out_row + c0);
          g[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          g[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        }
        if (!last) {
          // next layer's A operand: K-major, SWIZZLE_128B, one 16 KiB chunk per 64 columns
          const uint32_t chunk = a_region_addr + uint32_t(c0 >> 6) * 16384u;
          const uint32_t g0 = uint32_t(c0 & 63) >> 3;
          const uint32_t o0 = chunk + sw128_offset(row, g0);
          const uint32_t o1 = chunk + sw128_offset(row, g0 + 1);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(o0), "r"(pk[0]), "r"(pk[1]),
                       "r"(pk[2]), "r"(pk[3]) : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(o1), "r"(pk[4]), "r"(pk[5]),
                       "r"(pk[6]), "r"(pk[7]) : "memory");
        } else {
          // out layer (one unit) folded into the epilogue; uses the bf16-rounded activations
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&pk[i]);
            y_d += __low2float(v) * __ldg(w_out + c0 + 2 * i) + __high2float(v) * __ldg(w_out + c0 + 2 * i + 1);
          }
        }
      }
      if (!last) {
        fence_proxy_async_smem();   // generic-proxy smem writes -> visible to UMMA
        tc_fence_before_sync();     // our TMEM reads are ordered before the issuer's next MMAs
        mbar_arrive(aready_bar);
      }
    }

    // ------------------------- logit, probability, loss -------------------------
    float loss = 0.f;
    if (valid) {
      const float b_out = __ldg(p.P[L] + p.sh.K[L]);
      const float z = __ldg(p.fm_part + b) + y_d + b_out;
      const float pr = 1.f / (1.f + __expf(-z));
      p.logit[b] = z;
      p.prob[b] = pr;
      if (p.labels) {
        const float y = __ldg(p.labels + b);
        loss = fmaxf(z, 0.f) - z * y + log1pf(__expf(-fabsf(z)));
        if (p.dlogit) p.dlogit[b] = (pr - y) / float(count);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
    if (lane == 0 && p.labels) atomicAdd(&p.st->loss_sum, loss);
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// Shared-memory plan for the forward kernel; returns total dynamic smem bytes or -1.
int mlp_fwd_plan(const MlpShape& sh, int* stages, int* b_slot_bytes, int* a_region_chunks) {
  int max_atoms = 1, chunks = 1;
  for (int l = 0; l < sh.L; ++l) {
    max_atoms = std::max(max_atoms, std::min(4, ceil_div(sh.N[l], 64)));
    if (l > 0) chunks = std::max(chunks, ceil_div(sh.K[l], 64));
  }
  *b_slot_bytes = max_atoms * 8192;
  *a_region_chunks = chunks;
  const int fixed = chunks * 16384 + 512 + 1024;
  const int stage = 16384 + *b_slot_bytes;
  int s = (227 * 1024 - fixed) / stage;
  if (s < 2) return -1;
  if (s > 4) s = 4;
  *stages = s;
  return fixed + s * stage;
}

int launch_mlp_fwd(const FwdMaps& maps, const FwdParams& prm, int batch_max, int smem_bytes,
                   cudaStream_t stream) {
  mlp_fwd_kernel<<<ceil_div(batch_max, kTileM), kMlpThreads, smem_bytes, stream>>>(maps, prm);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int mlp_fwd_set_smem(int smem_bytes) {
  DFM_CUDA_CHECK(cudaFuncSetAttribute(mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  return DFM_OK;
}

}  // namespace dfm
