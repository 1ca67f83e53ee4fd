// One MLP layer per launch, tiled (128 examples) x (64 output units) so even a 2048-example batch
// spreads over 64 SMs -- at this batch size the deep tower is latency-bound, and what matters is how
// many SMs share the epilogue and the weight stream, not how long one CTA can stay busy.
//
//   forward   h_{l+1} = dropout_keep( relu(h_l W_l + b_l) )                    (HVD.py:223-235)
//             A = h_l  [B, K_l]  K-major (TMA, 128-byte swizzle)
//             B = W_l  [K_l, N_l] in its TF layout == MN-major UMMA operand (one 64-wide atom per CTA)
//   backward  dz_{l-1} = (h_l > 0) * (dz_l W_l^T) / keep_{l-1}                  (autodiff of the above)
//             A = dz_l [B, N_l]  K-major
//             B = W_l  rows n0..n0+63 of [K_l, N_l] == K-major UMMA operand
//
// Pipeline: warp 0 = TMA producer, warp 1 = TMEM owner + UMMA issuer, warps 2..5 = epilogue; 9-stage
// smem ring (24 KiB per stage); accumulator 128 lanes x 64 fp32 columns in TMEM.
// loss_init_kernel closes the forward pass (out layer, sigmoid, BCE) and opens the backward pass
// (dlogit, dz of the last hidden layer) with one warp per example.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kLayerThreads = 192;
constexpr int kLayerStages = 9;     // 9 x 24 KiB in flight: the first layer's 20-step K loop is latency-bound
constexpr int kLayerN = 64;
constexpr uint32_t kLayerStageBytes = 16384u + 8192u;

// phase stamps of CTA (0,0): 0 start, 1 set-up done, 2 first stage landed, 3 MMAs issued, 4 accumulator ready,
// 5 epilogue done, 6 end
__device__ __forceinline__ void layer_stamp(const LayerParams& p, int i) {
  if (p.dbg && blockIdx.x == 0 && blockIdx.y == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.dbg[i] = t;
  }
}

template <bool kBackward>
__global__ void __launch_bounds__(kLayerThreads, 1)
mlp_layer_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const LayerParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kLayerStages * kLayerStageBytes);
  uint64_t* empty_bar = full_bar + kLayerStages;
  uint64_t* acc_bar = empty_bar + kLayerStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kTileM;
  const int n0 = blockIdx.y * kLayerN;
  const int count = p.st->count;        // set before the step starts, never by a kernel of the step
  pdl_launch_dependents();
  if (m0 >= count) return;
  if (threadIdx.x == 0) layer_stamp(p, 0);
  const int n_here = min(kLayerN, p.n_out - n0);          // multiple of 16
  const int num_kc = (p.k_red + kChunkK - 1) / kChunkK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kLayerStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(acc_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 64);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  // everything above is CTA-local set-up; from here on the kernel reads what its predecessor wrote
  pdl_wait();
  kstamp(p.st, (kBackward ? kStampBwd0 : kStampFwd0) + p.layer);

  if (warp == 0) {
    if (lane == 0) {
      layer_stamp(p, 1);
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kLayerStages;
        mbar_wait(&empty_bar[s], ((kc / kLayerStages) & 1) ^ 1, 100);
        uint8_t* a_dst = smem + size_t(s) * kLayerStageBytes;
        mbar_arrive_expect_tx(&full_bar[s], 16384u + uint32_t(p.b_box_rows) * 128u);
        tma_load_2d(a_dst, &tmA, &full_bar[s], kc * kChunkK, m0);
        if (!kBackward) tma_load_2d(a_dst + 16384, &tmB, &full_bar[s], n0, kc * kChunkK);   // [64 k][64 n]
        else            tma_load_2d(a_dst + 16384, &tmB, &full_bar[s], kc * kChunkK, n0);   // [64 n][64 k]
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kTileM, n_here, 0, kBackward ? 0 : 1);
      for (int kc = 0; kc < num_kc; ++kc) {
        const int s = kc % kLayerStages;
        mbar_wait(&full_bar[s], (kc / kLayerStages) & 1, 200);
        if (kc == 0) layer_stamp(p, 2);
        tc_fence_after_sync();
        const uint32_t a_addr = smem_u32(smem + size_t(s) * kLayerStageBytes);
        const uint32_t b_addr = a_addr + 16384u;
        // both operands are zero-filled past the tensor edge, so a ragged last chunk needs no special case
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t adesc = umma_smem_desc(a_addr + kk * 32, 1024, 0);
          const uint64_t bdesc = kBackward ? umma_smem_desc(b_addr + kk * 32, 1024, 0)
                                           : umma_smem_desc(b_addr + kk * 2048, 1024, 8192);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (kc | kk) != 0);
        }
        umma_commit(&empty_bar[s]);
      }
      umma_commit(acc_bar);
      layer_stamp(p, 3);
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int b = m0 + row;
    const bool valid = b < count;
    const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16);
    const bool drop = p.st->train && p.keep < 1.f;
    const float inv_keep = drop ? 1.f / p.keep : 1.f;
    const uint32_t step = uint32_t(p.st->step);
    __nv_bfloat16* out_row = p.out + size_t(b) * p.out_pitch + n0;
    const __nv_bfloat16* mask_row = kBackward ? p.mask + size_t(valid ? b : 0) * p.mask_pitch + n0 : nullptr;

    // operands that do not depend on the accumulator are fetched while the MMAs run
    float bias[kLayerN];
    uint4 hmask[kLayerN / 8];
    if (!kBackward) {
#pragma unroll
      for (int i = 0; i < kLayerN / 4; ++i) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (4 * i < n_here) v = __ldg(reinterpret_cast<const float4*>(p.bias + n0) + i);
        bias[4 * i] = v.x; bias[4 * i + 1] = v.y; bias[4 * i + 2] = v.z; bias[4 * i + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < kLayerN / 8; ++i) {
        hmask[i] = make_uint4(0u, 0u, 0u, 0u);
        if (valid && !p.bn && 8 * i < n_here) hmask[i] = __ldg(reinterpret_cast<const uint4*>(mask_row) + i);
      }
    }

    // The dropout decisions do not depend on the accumulator either: one Philox4x32-10 call yields eight 16-bit
    // uniforms, so a row's 64 keep bits cost eight calls, all hidden behind the MMA main loop.
    uint32_t keep_bits[kLayerN / 32] = {0u, 0u};
    if (!kBackward && drop) {
#pragma unroll
      for (int g = 0; g < kLayerN / 8; ++g) {
        if (8 * g >= n_here) break;
        const uint32_t bits = dropout_keep_bits8(p.seed, uint32_t(b), uint32_t(p.layer), uint32_t((n0 >> 3) + g), step, p.keep);
        keep_bits[g >> 2] |= bits << (8 * (g & 3));
      }
    }
    if (kBackward && p.bn) {
      // batch-norm tower: the output is the gradient w.r.t. the BN output of layer-1, i.e. only that layer's dropout
      // applies here (bits recomputed); bn_bwd finishes the chain through BN and the ReLU
#pragma unroll
      for (int g = 0; g < kLayerN / 8; ++g) {
        if (8 * g >= n_here) break;
        const uint32_t bits = drop ? dropout_keep_bits8(p.seed, uint32_t(b), uint32_t(p.layer - 1), uint32_t((n0 >> 3) + g), step, p.keep)
                                   : 0xffu;
        keep_bits[g >> 2] |= bits << (8 * (g & 3));
      }
    }

    mbar_wait(acc_bar, 0, 300);
    if (threadIdx.x == 64) layer_stamp(p, 4);
    tc_fence_after_sync();
#pragma unroll
    for (int c0 = 0; c0 < kLayerN; c0 += 16) {
      if (c0 >= n_here) break;
      uint32_t r[16];
      tmem_ld_x16(t_row + c0, r);
      tmem_ld_wait();
      float a[16];
      if (!kBackward) {
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fmaxf(__uint_as_float(r[i]) + bias[c0 + i], 0.f);
        if (drop) {
          const uint32_t kb = keep_bits[c0 >> 5] >> (c0 & 31);
#pragma unroll
          for (int i = 0; i < 16; ++i) a[i] = ((kb >> i) & 1u) ? a[i] * inv_keep : 0.f;
        }
      } else {
        if (p.bn) {
          const uint32_t kb = keep_bits[c0 >> 5] >> (c0 & 31);
#pragma unroll
          for (int i = 0; i < 16; ++i) a[i] = ((kb >> i) & 1u) ? __uint_as_float(r[i]) * inv_keep : 0.f;
        } else {
          const uint4 h0 = hmask[c0 / 8], h1 = hmask[c0 / 8 + 1];
          const uint32_t hv[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&hv[i]);
            a[2 * i] = __low2float(v) > 0.f ? __uint_as_float(r[2 * i]) * inv_keep : 0.f;
            a[2 * i + 1] = __high2float(v) > 0.f ? __uint_as_float(r[2 * i + 1]) * inv_keep : 0.f;
          }
        }
      }
      // rows past the batch: forward leaves them untouched, backward must write zeros (they are
      // inside the weight-gradient GEMM's batch range)
      if (valid || kBackward) {
        uint32_t pk[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) pk[i] = valid ? pack_bf16x2(a[2 * i], a[2 * i + 1]) : 0u;
        st_global_v8(out_row + c0, pk);      // 16 bf16 = one whole sector
      }
    }
    if (threadIdx.x == 64) layer_stamp(p, 5);
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 64);
  if (threadIdx.x == 0) layer_stamp(p, 6);
}

// ---------------------------------------------------------------------------------------------
// Eight examples per warp (4 lanes each), so every 32-byte sector of the per-example outputs (logit,
// prob, dlogit) is written by exactly one warp in one instruction:
//   y_d = h_L . w_out + b_out ; logit ; prob ; BCE ; dlogit ;
//   dz_{L-1} = (h_L > 0) * dlogit * w_out / keep_{L-1} ; dz_L = [dlogit, 0 ...].
// LPE lanes per example (4, 8, 16 or 32, chosen from the width of the last hidden layer): with 4 lanes a warp
// covers eight examples and writes one 32-byte sector of logit / prob / dlogit per instruction, but a 400-wide
// layer then costs 13 dependent 16-byte loads per lane (17 us measured); wider layers get more lanes per example.
template <int LPE>
__global__ void __launch_bounds__(256)
loss_init_kernel(const LossParams p) {
  constexpr int EPW = 32 / LPE;                             // examples per warp
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane / LPE, part = lane % LPE;            // example within the warp, slice of its row
  const int b = (blockIdx.x * 8 + warp) * EPW + sub;
  const int count = p.st->count;
  pdl_launch_dependents();
  pdl_wait();
  kstamp(p.st, kStampLoss);
  __shared__ float block_loss[8];
  float loss = 0.f;
  const bool valid = b < count;
  const __nv_bfloat16* h = p.h_last + size_t(valid ? b : 0) * p.h_pitch;
  float yd = 0.f;
  if (valid)
    for (int c = part * 8; c < p.n_last; c += 8 * LPE) {     // 8 consecutive columns (16 bytes) per lane
      const uint4 hv = __ldg(reinterpret_cast<const uint4*>(h + c));
      const uint32_t w4[4] = {hv.x, hv.y, hv.z, hv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&w4[i]);
        if (c + 2 * i < p.n_last) yd += __low2float(v) * __ldg(p.w_out + c + 2 * i);
        if (c + 2 * i + 1 < p.n_last) yd += __high2float(v) * __ldg(p.w_out + c + 2 * i + 1);
      }
    }
#pragma unroll
  for (int o = 1; o < LPE; o <<= 1) yd += __shfl_xor_sync(0xffffffffu, yd, o);
  float dl = 0.f;
  if (valid) {
    const float z = __ldg(p.fm_part + b) + yd + __ldg(p.w_out + p.n_last);
    const float pr = 1.f / (1.f + __expf(-z));
    if (p.labels) {
      const float y = __ldg(p.labels + b);
      if (p.loss_type == DFM_LOSS_SQUARE) {            // mean (sigmoid(y) - label)^2
        if (part == 0) loss = (pr - y) * (pr - y);
        dl = 2.f * (pr - y) * pr * (1.f - pr) / float(count);
      } else {
        if (part == 0) loss = fmaxf(z, 0.f) - z * y + log1pf(__expf(-fabsf(z)));
        dl = (pr - y) / float(count);
      }
    }
    if (part == 0) {
      p.logit[b] = z;
      p.prob[b] = pr;
      if (p.dlogit) p.dlogit[b] = dl;
    }
  }
  if (p.dz_last && b < p.rows_pad) {
    // rows past the batch inside the weight-gradient GEMM's K range are written as zeros (dl == 0)
    const bool drop = p.st->train && p.keep_last < 1.f;
    const float scale = drop ? 1.f / p.keep_last : 1.f;
    __nv_bfloat16* dz = p.dz_last + size_t(b) * p.n_last;
    for (int c = part * 8; c < p.n_last; c += 8 * LPE) {
      uint4 hv = make_uint4(0u, 0u, 0u, 0u);
      if (valid && !p.bn) hv = __ldg(reinterpret_cast<const uint4*>(h + c));
      const uint32_t w4[4] = {hv.x, hv.y, hv.z, hv.w};
      // batch-norm tower: h_L = dropout(BN(a)) can be <= 0 where it was kept, so the keep bits are recomputed
      const uint32_t kb = !p.bn ? 0u : drop ? dropout_keep_bits8(p.seed, uint32_t(b), uint32_t(p.last_layer), uint32_t(c >> 3),
                                                                uint32_t(p.st->step), p.keep_last) : 0xffu;
      uint32_t o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&w4[i]);
        const bool on0 = p.bn ? ((kb >> (2 * i)) & 1u) != 0 : __low2float(v) > 0.f;
        const bool on1 = p.bn ? ((kb >> (2 * i + 1)) & 1u) != 0 : __high2float(v) > 0.f;
        const float d0 = (valid && c + 2 * i < p.n_last && on0) ? dl * __ldg(p.w_out + c + 2 * i) * scale : 0.f;
        const float d1 = (valid && c + 2 * i + 1 < p.n_last && on1) ? dl * __ldg(p.w_out + c + 2 * i + 1) * scale : 0.f;
        o[i] = pack_bf16x2(d0, d1);
      }
      *reinterpret_cast<uint4*>(dz + c) = make_uint4(o[0], o[1], o[2], o[3]);
    }
    // dz_L row: 16 bf16 = 32 bytes = two 16-byte halves, written by parts 0 and 1
    if (part < 2)
      *reinterpret_cast<uint4*>(p.dz_out + size_t(b) * 16 + part * 8) =
          make_uint4(part == 0 ? pack_bf16x2(dl, 0.f) : 0u, 0u, 0u, 0u);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
  if (lane == 0) block_loss[warp] = loss;
  __syncthreads();
  if (threadIdx.x == 0 && p.labels) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += block_loss[i];
    if (t != 0.f) atomicAdd(&p.st->loss_sum, t);
  }
}

// ---------------------------------------------------------------------------------------------
int mlp_layer_smem() { return kLayerStages * int(kLayerStageBytes) + 512 + 1024; }

int mlp_layer_set_smem() {
  DFM_CUDA_CHECK(cudaFuncSetAttribute(mlp_layer_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mlp_layer_smem()));
  DFM_CUDA_CHECK(cudaFuncSetAttribute(mlp_layer_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mlp_layer_smem()));
  return DFM_OK;
}

int launch_mlp_layer(bool backward, const CUtensorMap& tmA, const CUtensorMap& tmB, const LayerParams& p,
                     int batch_max, cudaStream_t stream) {
  dim3 grid(ceil_div(batch_max, kTileM), ceil_div(p.n_out, kLayerN));
  if (backward) DFM_CUDA_CHECK(launch_pdl(kPdlBwd, mlp_layer_kernel<true>, grid, dim3(kLayerThreads), mlp_layer_smem(), stream, tmA, tmB, p));
  else          DFM_CUDA_CHECK(launch_pdl(kPdlFwd, mlp_layer_kernel<false>, grid, dim3(kLayerThreads), mlp_layer_smem(), stream, tmA, tmB, p));
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_loss_init(const LossParams& p, int rows_pad, cudaStream_t stream) {
  const int lpe = p.n_last <= 64 ? 4 : p.n_last <= 128 ? 8 : p.n_last <= 256 ? 16 : 32;
  const dim3 grid(ceil_div(rows_pad, 8 * (32 / lpe)));
  if (lpe == 4)       DFM_CUDA_CHECK(launch_pdl(kPdlLoss, loss_init_kernel<4>, grid, dim3(256), 0, stream, p));
  else if (lpe == 8)  DFM_CUDA_CHECK(launch_pdl(kPdlLoss, loss_init_kernel<8>, grid, dim3(256), 0, stream, p));
  else if (lpe == 16) DFM_CUDA_CHECK(launch_pdl(kPdlLoss, loss_init_kernel<16>, grid, dim3(256), 0, stream, p));
  else                DFM_CUDA_CHECK(launch_pdl(kPdlLoss, loss_init_kernel<32>, grid, dim3(256), 0, stream, p));
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
