// The middle of the training step in ONE launch per 128-example tile:
//
//   forward layers 1..L-1   h_{l+1} = dropout_keep(relu(h_l W_l + b_l))                   (HVD.py:223-235)
//   output layer + loss     y_d = h_L . w_out + b_out ; sigmoid ; BCE ; dlogit              (HVD.py:246-268)
//   backward layers L-1..1  dz_{l-1} = (h_l > 0) * (dz_l W_l^T) / keep_{l-1}               (autodiff of the above)
//
// These are 2(L-1)+1 small dependent steps; as separate kernels each costs a launch-to-launch latency of 3-4 us for
// ~1.7 us of work per CTA (profiles/README.md), 18 us of a 58 us step.  Here the weights of layers 1..L-1 are
// TMA-loaded into shared memory once (the same swizzled tiles serve as MN-major B operand in the forward GEMM and
// as K-major B operand in the backward GEMM), every GEMM is tcgen05.mma with the accumulator in TMEM, and every
// epilogue writes its result both to global memory (the weight-gradient GEMM and embed_bwd need it) and, in the
// UMMA K-major 128-byte-swizzled layout, into shared memory as the A operand of the next GEMM.
//
// Warp roles: warp 0 = TMA + tcgen05.mma issuer (one elected thread), warp 1 = TMEM owner, warps 2..9 = epilogue:
// two warps per TMEM lane quarter, each taking half of the 16-column groups of a row.
// Used when L >= 2, every width of layers 0..L-1 is <= 256 and the tiles fit 227 KB; otherwise the engine falls
// back to the per-layer kernels (mlp_layer.cu).  The first layer (K_0 = F*Ke, wide) keeps its own kernel.
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kMidThreads = 320;

// per-stage timestamps of CTA 0: [4s] A operand ready, [4s+1] MMAs issued, [4s+2] accumulator ready, [4s+3] epilogue done
__device__ __forceinline__ void mid_stamp(const MidParams& p, int i) {
  if (p.dbg && blockIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.dbg[i] = t;
  }
}

__global__ void __launch_bounds__(kMidThreads, 1)
mlp_mid_kernel(const __grid_constant__ FwdMaps maps, const __grid_constant__ MidParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_buf0 = smem;
  uint8_t* a_buf1 = smem + p.a_bytes;
  uint8_t* w_reg = smem + 2 * size_t(p.a_bytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(w_reg + p.w_bytes);
  uint64_t* bar_w = bars;                               // weight tiles landed
  uint64_t* bar_h = bars + 1;                           // h_1 tile landed
  uint64_t* bar_a = bars + 2;                           // [stage] A operand written by the epilogue warps
  uint64_t* bar_acc = bars + 2 + kMidMaxStages;         // [stage] accumulator complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 + 2 * kMidMaxStages);
  float* yd_part = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [2][128]
  float* vec = yd_part + 256;                           // biases of layers 1..L-1, then w_out and b_out (<= 8 KB)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kTileM;
  const int count = p.st->count;                        // set before the step starts, never by a kernel of the step
  const int L = p.sh.L;
  const int S = 2 * (L - 1);
  pdl_launch_dependents();
  if (m0 >= count) return;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.act[1]);
    for (int l = 1; l < L; ++l) tma_prefetch_desc(&maps.w[l]);
    mbar_init(bar_w, 1);
    mbar_init(bar_h, 1);
    for (int s = 0; s < S; ++s) {
      mbar_init(&bar_a[s], 8);                          // one arrival per epilogue warp
      mbar_init(&bar_acc[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // The bf16 weights were written by the previous step's optimizer kernel: they may be fetched before this
      // kernel's predecessor (the first layer) has finished.
      mbar_arrive_expect_tx(bar_w, p.w_bytes);
      for (int l = 1; l < L; ++l) {
        const int nK = mid_tiles(p.sh.K[l]), nN = mid_tiles(p.sh.N[l]);
        for (int kc = 0; kc < nK; ++kc)
          for (int nc = 0; nc < nN; ++nc)
            tma_load_2d(w_reg + p.w_off[l] + size_t(nc * nK + kc) * 8192, &maps.w[l], bar_w, nc * 64, kc * 64);
      }
      pdl_wait();
      kstamp(p.st, kStampFwd0 + 1);
      const int nk1 = mid_tiles(p.sh.K[1]);
      mbar_arrive_expect_tx(bar_h, uint32_t(nk1) * 16384u);
      for (int j = 0; j < nk1; ++j) tma_load_2d(a_buf0 + size_t(j) * 16384, &maps.act[1], bar_h, j * 64, m0);
      mbar_wait(bar_w, 0, 400);
      for (int s = 0; s < S; ++s) {
        if (s == 0) mbar_wait(bar_h, 0, 401);
        else        mbar_wait(&bar_a[s], 0, 410 + s);
        tc_fence_after_sync();
        mid_stamp(p, 4 * s);
        const uint32_t a_addr = smem_u32((s & 1) ? a_buf1 : a_buf0);
        if (s < L - 1) {
          // forward layer l: D[128, N_l] = A[128, K_l] (K-major) x W_l[K_l, N_l] (MN-major; the 64-wide n-atoms of one
          // k chunk are nK tiles apart: weight tiles are stored [n chunk][k chunk])
          const int l = s + 1;
          const int nK = mid_tiles(p.sh.K[l]);
          const uint32_t w_addr = smem_u32(w_reg + p.w_off[l]);
          const uint32_t idesc = umma_idesc_bf16(kTileM, p.sh.N[l], 0, 1);
          const int k16n = p.sh.K[l] >> 4;
          for (int k16 = 0; k16 < k16n; ++k16) {
            const int kc = k16 >> 2, kk = k16 & 3;
            umma_bf16(tmem_base, umma_smem_desc(a_addr + kc * 16384 + kk * 32, 1024, 0),
                      umma_smem_desc(w_addr + uint32_t(kc) * 8192u + kk * 2048, 1024, uint32_t(nK) * 8192u), idesc, k16 != 0);
          }
        } else {
          // backward layer l: D[128, K_l] = dz_l[128, N_l] (K-major) x W_l^T: for one reduction chunk (n chunk of W_l)
          // the rows of W_l are consecutive 128-byte rows across the k-chunk tiles = one K-major B operand of K_l rows
          const int l = 2 * L - 2 - s;
          const int nK = mid_tiles(p.sh.K[l]);
          const uint32_t w_addr = smem_u32(w_reg + p.w_off[l]);
          const int k16n = p.sh.N[l] >> 4;
          const uint32_t idesc = umma_idesc_bf16(kTileM, p.sh.K[l], 0, 0);
          for (int k16 = 0; k16 < k16n; ++k16) {
            const int rc = k16 >> 2, kk = k16 & 3;
            umma_bf16(tmem_base, umma_smem_desc(a_addr + rc * 16384 + kk * 32, 1024, 0),
                      umma_smem_desc(w_addr + uint32_t(rc * nK) * 8192u + kk * 32, 1024, 0), idesc, k16 != 0);
          }
        }
        umma_commit(&bar_acc[s]);
        mid_stamp(p, 4 * s + 1);
      }
    }
  } else if (warp >= 2) {
    const int q = warp & 3;                       // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;             // which half of a row's 16-column groups
    const int row = q * 32 + lane;
    const int b = m0 + row;
    const bool valid = b < count;
    const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16);
    // biases and output weights (written by the previous step's optimizer kernel) go to shared memory once: a
    // global load per 16-column group inside the epilogues would put its latency on the critical path each time
    {
      const int te = threadIdx.x - 64;
      int off = 0;
      for (int l = 1; l < L; ++l) {
        for (int i = te; i < p.sh.N[l]; i += 256) vec[off + i] = __ldg(p.bias[l] + i);
        off += p.sh.N[l];
      }
      for (int i = te; i <= p.sh.N[L - 1]; i += 256) vec[off + i] = __ldg(p.w_out + i);
      asm volatile("bar.sync 5, 256;" ::: "memory");
    }
    int wout_off = 0;
    for (int l = 1; l < L; ++l) wout_off += p.sh.N[l];
    const float* w_out_s = vec + wout_off;
    // every thread that reads what earlier kernels of this step wrote (the step counter that keys the dropout
    // masks, h_1, the masks) waits for the programmatic dependency itself
    pdl_wait();
    const bool train = p.st->train != 0;
    const uint32_t step = uint32_t(p.st->step);
    const float fm_b = valid ? __ldg(p.fm_part + b) : 0.f;      // needed by the loss stage; fetched now
    const float label_b = valid ? __ldg(p.labels + b) : 0.f;
    for (int s = 0; s < S; ++s) {
      uint8_t* a_next = (s & 1) ? a_buf0 : a_buf1;
      const bool fwd = s < L - 1;
      const int l = fwd ? s + 1 : 2 * L - 2 - s;
      const int n_out = fwd ? p.sh.N[l] : p.sh.K[l];
      const int G = n_out >> 4;
      const int g_lo = half == 0 ? 0 : (G + 1) / 2, g_hi = half == 0 ? (G + 1) / 2 : G;
      if (fwd) {
        const float keep = p.sh.keep[l];
        const bool drop = train && keep < 1.f;
        const float inv_keep = drop ? 1.f / keep : 1.f;
        const uint32_t thr = uint32_t(keep * 65536.f);
        const bool last = l == L - 1;
        __nv_bfloat16* out_row = p.act[l + 1] + size_t(b) * p.sh.pitch[l + 1];
        int bias_off = 0;
        for (int j = 1; j < l; ++j) bias_off += p.sh.N[j];
        const float* bias = vec + bias_off;
        // the dropout decisions do not depend on the accumulator: 16 keep bits per group, computed before the wait
        uint32_t keep_bits[4] = {0u, 0u, 0u, 0u};
        if (drop) {
#pragma unroll
          for (int gi = 0; gi < 8; ++gi) {
            const int g = g_lo + gi;
            if (g >= g_hi) break;
            uint32_t bits = 0u;
#pragma unroll
            for (int h8 = 0; h8 < 2; ++h8) {
              uint32_t rnd[4];
              philox4x32(uint32_t(p.seed), uint32_t(p.seed >> 32), uint32_t(b), uint32_t(l), uint32_t(2 * g + h8), step,
                         rnd);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                bits |= ((rnd[i] & 0xffffu) < thr ? 1u : 0u) << (8 * h8 + 2 * i);
                bits |= ((rnd[i] >> 16) < thr ? 1u : 0u) << (8 * h8 + 2 * i + 1);
              }
            }
            keep_bits[gi >> 1] |= bits << (16 * (gi & 1));
          }
        }
        float yd = 0.f;
        mbar_wait(&bar_acc[s], 0, 420 + s);
        if (threadIdx.x == 64) mid_stamp(p, 4 * s + 2);
        tc_fence_after_sync();
#pragma unroll
        for (int gi = 0; gi < 8; ++gi) {               // <= 8 groups per half (widths <= 256); unrolled: register arrays
          const int g = g_lo + gi;
          if (g >= g_hi) break;
          const int c0 = g << 4;
          uint32_t r[16];
          tmem_ld_x16(t_row + c0, r);
          tmem_ld_wait_regs(r);
          float a[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 bv = *(reinterpret_cast<const float4*>(bias + c0) + i);
            a[4 * i] = fmaxf(__uint_as_float(r[4 * i]) + bv.x, 0.f);
            a[4 * i + 1] = fmaxf(__uint_as_float(r[4 * i + 1]) + bv.y, 0.f);
            a[4 * i + 2] = fmaxf(__uint_as_float(r[4 * i + 2]) + bv.z, 0.f);
            a[4 * i + 3] = fmaxf(__uint_as_float(r[4 * i + 3]) + bv.w, 0.f);
          }
          if (drop) {
            // same counters and bit assignment as mlp_layer_kernel (one Philox call = eight 16-bit uniforms)
            const uint32_t kb = keep_bits[gi >> 1] >> (16 * (gi & 1));
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = ((kb >> i) & 1u) ? a[i] * inv_keep : 0.f;
          }
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) pk[i] = valid ? pack_bf16x2(a[2 * i], a[2 * i + 1]) : 0u;
          if (valid) st_global_v8(out_row + c0, pk);
          if (last) {
            // the output layer reads h_L as the bf16 values that were stored
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&pk[i]);
              yd += __low2float(v) * w_out_s[c0 + 2 * i] + __high2float(v) * w_out_s[c0 + 2 * i + 1];
            }
          }
          uint8_t* tile = a_next + size_t(c0 >> 6) * 16384;
          const uint32_t gran = uint32_t(c0 & 63) >> 3;
          *reinterpret_cast<uint4*>(tile + sw128_offset(row, gran)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          *reinterpret_cast<uint4*>(tile + sw128_offset(row, gran + 1)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        }
        if (last) {
          // ---- output layer, loss, dlogit; then dz_{L-1} replaces h_L in the A buffer ----
          if (threadIdx.x == 64) kstamp_any(p.st, kStampLoss);
          yd_part[half * 128 + row] = yd;
          asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
          const int n_last = p.sh.N[l];
          float dl = 0.f, loss = 0.f;
          if (valid) {
            const float z = fm_b + yd_part[row] + yd_part[128 + row] + w_out_s[n_last];
            const float pr = 1.f / (1.f + __expf(-z));
            const float y = label_b;
            dl = (pr - y) / float(count);
            if (half == 0) {
              loss = fmaxf(z, 0.f) - z * y + log1pf(__expf(-fabsf(z)));
              p.logit[b] = z;
              p.prob[b] = pr;
              p.dlogit[b] = dl;
            }
          }
          if (half == 0) {
            const uint32_t dzl[8] = {pack_bf16x2(dl, 0.f), 0u, 0u, 0u, 0u, 0u, 0u, 0u};
            st_global_v8(p.dz[L] + size_t(b) * 16, dzl);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
            if (lane == 0 && loss != 0.f) atomicAdd(&p.st->loss_sum, loss);
          }
          __nv_bfloat16* dz_row = p.dz[l] + size_t(b) * n_last;
          for (int g = g_lo; g < g_hi; ++g) {
            const int c0 = g << 4;
            uint8_t* tile = a_next + size_t(c0 >> 6) * 16384;
            const uint32_t gran = uint32_t(c0 & 63) >> 3;
            uint4* s0 = reinterpret_cast<uint4*>(tile + sw128_offset(row, gran));
            uint4* s1 = reinterpret_cast<uint4*>(tile + sw128_offset(row, gran + 1));
            const uint4 h0 = *s0, h1 = *s1;
            const uint32_t hv[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
            uint32_t pk[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&hv[i]);
              const float d0 = __low2float(v) > 0.f ? dl * w_out_s[c0 + 2 * i] * inv_keep : 0.f;   // same rounding order as loss_init
              const float d1 = __high2float(v) > 0.f ? dl * w_out_s[c0 + 2 * i + 1] * inv_keep : 0.f;
              pk[i] = pack_bf16x2(d0, d1);
            }
            st_global_v8(dz_row + c0, pk);          // rows past the batch carry zeros (dl == 0)
            *s0 = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            *s1 = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          }
        }
      } else {
        // backward layer l: dz_{l-1} = (h_l > 0) * acc / keep_{l-1}
        const float keep = p.sh.keep[l - 1];
        const float inv_keep = (train && keep < 1.f) ? 1.f / keep : 1.f;
        const __nv_bfloat16* mask_row = p.act[l] + size_t(valid ? b : 0) * p.sh.pitch[l];
        __nv_bfloat16* dz_row = p.dz[l - 1] + size_t(b) * n_out;
        const bool feeds_next = s + 1 < S;
        // masks: h_{L-1} is still in the A buffer this stage overwrites (same thread, same granules); the masks of
        // the other layers are fetched from global memory into registers before the accumulator wait
        const bool mask_in_smem = l == L - 1;
        uint4 mk[16];
        if (!mask_in_smem) {
#pragma unroll
          for (int gi = 0; gi < 8; ++gi) {
            mk[2 * gi] = mk[2 * gi + 1] = make_uint4(0u, 0u, 0u, 0u);
            if (valid && g_lo + gi < g_hi) {
              mk[2 * gi] = __ldg(reinterpret_cast<const uint4*>(mask_row + 16 * (g_lo + gi)));
              mk[2 * gi + 1] = __ldg(reinterpret_cast<const uint4*>(mask_row + 16 * (g_lo + gi)) + 1);
            }
          }
        }
        if (threadIdx.x == 64) kstamp_any(p.st, kStampBwd0 + l);
        mbar_wait(&bar_acc[s], 0, 420 + s);
        if (threadIdx.x == 64) mid_stamp(p, 4 * s + 2);
        tc_fence_after_sync();
#pragma unroll
        for (int gi = 0; gi < 8; ++gi) {
          const int g = g_lo + gi;
          if (g >= g_hi) break;
          const int c0 = g << 4;
          uint8_t* mtile = a_next + size_t(c0 >> 6) * 16384;
          const uint32_t mgran = uint32_t(c0 & 63) >> 3;
          uint4 m0v, m1v;
          if (mask_in_smem) {
            m0v = *reinterpret_cast<const uint4*>(mtile + sw128_offset(row, mgran));
            m1v = *reinterpret_cast<const uint4*>(mtile + sw128_offset(row, mgran + 1));
          } else {
            m0v = mk[2 * gi];
            m1v = mk[2 * gi + 1];
          }
          uint32_t r[16];
          tmem_ld_x16(t_row + c0, r);
          tmem_ld_wait_regs(r);
          const uint32_t hv[8] = {m0v.x, m0v.y, m0v.z, m0v.w, m1v.x, m1v.y, m1v.z, m1v.w};
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&hv[i]);
            const float d0 = __low2float(v) > 0.f ? __uint_as_float(r[2 * i]) * inv_keep : 0.f;
            const float d1 = __high2float(v) > 0.f ? __uint_as_float(r[2 * i + 1]) * inv_keep : 0.f;
            pk[i] = valid ? pack_bf16x2(d0, d1) : 0u;
          }
          st_global_v8(dz_row + c0, pk);
          if (feeds_next) {
            uint8_t* tile = a_next + size_t(c0 >> 6) * 16384;
            const uint32_t gran = uint32_t(c0 & 63) >> 3;
            *reinterpret_cast<uint4*>(tile + sw128_offset(row, gran)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            *reinterpret_cast<uint4*>(tile + sw128_offset(row, gran + 1)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          }
        }
      }
      if (threadIdx.x == 64) mid_stamp(p, 4 * s + 3);
      if (s + 1 < S) {
        // generic-proxy writes of the next A operand -> visible to the tensor core; TMEM reads of this stage done
        fence_proxy_async_smem();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_a[s + 1]);
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// ------------------------------------- host side -------------------------------------------
// Fills the shared-memory plan; returns the dynamic shared memory the kernel needs, or -1 if this network does
// not fit the fused middle (the engine then uses the per-layer kernels).
int mlp_mid_plan(const MlpShape& sh, MidParams* mp) {
  if (sh.L < 2) return -1;
  int maxw = 0;
  for (int l = 0; l < sh.L; ++l) {
    if (sh.N[l] > kMidMaxW || sh.N[l] % 16 != 0) return -1;
    maxw = std::max(maxw, sh.N[l]);
  }
  uint32_t off = 0;
  for (int l = 1; l < sh.L; ++l) {
    mp->w_off[l] = off;
    off += uint32_t(mid_tiles(sh.K[l]) * mid_tiles(sh.N[l])) * 8192u;
  }
  mp->w_bytes = off;
  mp->a_bytes = uint32_t(kTileM) * uint32_t(mid_tiles(maxw) * 64) * 2u;
  const uint32_t bytes = mid_smem_bytes(mp->a_bytes, mp->w_bytes);
  return bytes <= 227u * 1024u ? int(bytes) : -1;
}

int mlp_mid_set_smem(int bytes) {
  DFM_CUDA_CHECK(cudaFuncSetAttribute(mlp_mid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return DFM_OK;
}

int launch_mlp_mid(const FwdMaps& maps, const MidParams& prm, int batch_max, int smem_bytes, cudaStream_t stream) {
  DFM_CUDA_CHECK(launch_pdl(kPdlFwd, mlp_mid_kernel, dim3(ceil_div(batch_max, kTileM)), dim3(kMidThreads), size_t(smem_bytes),
                            stream, maps, prm));
  return DFM_OK;
}

}  // namespace dfm
