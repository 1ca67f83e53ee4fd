// dfm_gemm_bf16: one 128 x n_tile output tile per CTA, warp-specialised:
//   warp 0      TMA producer (one elected lane)
//   warp 1      TMEM allocator + UMMA issuer (one elected lane)
//   warps 2..5  epilogue: tcgen05.ld accumulator rows -> bias/relu -> global
// K is streamed in 64-element chunks through a kStages-deep smem ring guarded by
// full/empty mbarriers; the accumulator (128 lanes x n_tile fp32 columns) lives in TMEM.
//
// This is the reference implementation of the pipeline the fused MLP kernels
// (mlp_fwd.cu / mlp_bwd.cu / wgrad.cu) specialise, and the op used when a layer
// has to run un-fused (batch-norm path).  Replaces cuBLAS SGEMM under
// tf.contrib.layers.fully_connected (HVD.py:224-229, 237-243).
#include "common.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kGemmStages = 4;
constexpr int kBlockM = 128;
constexpr int kBlockK = 64;
constexpr int kGemmThreads = 192;

struct GemmArgs {
  float* C;
  const float* bias;
  int M, N, K, ldc;
  int n_tile;   // columns per CTA (multiple of 16, <= 256; multiple of 64 if b_mn)
  int a_mn, b_mn;
  int relu;
};

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1 KiB alignment is required by SWIZZLE_128B tiles.
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const int n_tile = args.n_tile;
  const uint32_t a_bytes = kBlockM * kBlockK * 2;               // 16 KiB
  const uint32_t b_bytes = uint32_t(n_tile) * kBlockK * 2;      // n_tile * 128
  const uint32_t b_slot = (b_bytes + 1023u) & ~1023u;
  const uint32_t stage_bytes = a_bytes + b_slot;

  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kGemmStages * stage_bytes);
  uint64_t* empty_bar = full_bar + kGemmStages;
  uint64_t* acc_bar = empty_bar + kGemmStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kBlockM;
  const int n0 = blockIdx.y * n_tile;
  const int num_k = (args.K + kBlockK - 1) / kBlockK;

  // TMEM columns: power of two >= 32 covering n_tile.
  uint32_t tmem_cols = 32;
  while (tmem_cols < uint32_t(n_tile)) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kGemmStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(acc_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      for (int kc = 0; kc < num_k; ++kc) {
        const int s = kc % kGemmStages;
        const uint32_t ph = (kc / kGemmStages) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1, 100 + s);
        uint8_t* a_dst = smem + s * stage_bytes;
        uint8_t* b_dst = a_dst + a_bytes;
        mbar_arrive_expect_tx(&full_bar[s], a_bytes + b_bytes);
        const int k0 = kc * kBlockK;
        if (!args.a_mn) {
          tma_load_2d(a_dst, &tmA, &full_bar[s], k0, m0);
        } else {
          for (int a = 0; a < kBlockM / 64; ++a)
            tma_load_2d(a_dst + a * 8192, &tmA, &full_bar[s], m0 + a * 64, k0);
        }
        if (!args.b_mn) {
          tma_load_2d(b_dst, &tmB, &full_bar[s], k0, n0);
        } else {
          for (int a = 0; a < n_tile / 64; ++a)
            tma_load_2d(b_dst + a * 8192, &tmB, &full_bar[s], n0 + a * 64, k0);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------- UMMA issuer ------------------------------
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kBlockM, n_tile, args.a_mn, args.b_mn);
      for (int kc = 0; kc < num_k; ++kc) {
        const int s = kc % kGemmStages;
        const uint32_t ph = (kc / kGemmStages) & 1;
        mbar_wait(&full_bar[s], ph, 200 + s);
        tc_fence_after_sync();
        const uint32_t a_addr = smem_u32(smem + s * stage_bytes);
        const uint32_t b_addr = a_addr + a_bytes;
#pragma unroll
        for (int kk = 0; kk < kBlockK / 16; ++kk) {
          // K-major: +32 B per 16-element K step inside the 128 B swizzle row.
          // MN-major: +16 k-rows * 128 B = 2048 B; LBO = 8192 B between 64-wide MN atoms.
          const uint64_t adesc = args.a_mn ? umma_smem_desc(a_addr + kk * 2048, 1024, 8192)
                                           : umma_smem_desc(a_addr + kk * 32, 1024, 0);
          const uint64_t bdesc = args.b_mn ? umma_smem_desc(b_addr + kk * 2048, 1024, 8192)
                                           : umma_smem_desc(b_addr + kk * 32, 1024, 0);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (kc | kk) != 0);
        }
        umma_commit(&empty_bar[s]);   // smem slot reusable once these MMAs retire
      }
      umma_commit(acc_bar);           // accumulator complete
    }
  } else {
    // -------------------------------- epilogue --------------------------------
    const int q = warp & 3;           // TMEM lane quadrant this warp may read
    const int row = q * 32 + lane;
    const int gm = m0 + row;
    mbar_wait(acc_bar, 0, 300);
    tc_fence_after_sync();
    const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16);
    for (int c0 = 0; c0 < n_tile; c0 += 16) {
      uint32_t r[16];
      tmem_ld_x16(t_row + c0, r);
      tmem_ld_wait();
      if (gm < args.M) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int gn = n0 + c0 + i;
          if (gn < args.N) {
            float v = __uint_as_float(r[i]);
            if (args.bias) v += __ldg(args.bias + gn);
            if (args.relu) v = fmaxf(v, 0.f);
            args.C[size_t(gm) * args.ldc + gn] = v;
          }
        }
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace dfm

extern "C" int dfm_gemm_bf16(const void* A, int lda, int a_mn_major, const void* B, int ldb,
                             int b_mn_major, float* C, int ldc, int M, int N, int K,
                             const float* bias, int relu, dfm_stream_t stream_) {
  using namespace dfm;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  DFM_REQUIRE(M > 0 && N > 0 && K > 0, "dfm_gemm_bf16: empty problem %dx%dx%d", M, N, K);
  DFM_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "dfm_gemm_bf16: lda/ldb must be multiples of 8");
  // columns per CTA
  int n_tile = N >= 256 ? 256 : round_up(N, 16);
  if (b_mn_major) n_tile = N >= 256 ? 256 : round_up(N, 64);

  CUtensorMap tmA, tmB;
  int rc;
  if (!a_mn_major) rc = make_tmap_bf16_2d(&tmA, A, K, M, uint64_t(lda) * 2, 64, kBlockM);
  else             rc = make_tmap_bf16_2d(&tmA, A, M, K, uint64_t(lda) * 2, 64, 64);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tmB, B, K, N, uint64_t(ldb) * 2, 64, n_tile);
  else             rc = make_tmap_bf16_2d(&tmB, B, N, K, uint64_t(ldb) * 2, 64, 64);
  if (rc) return rc;

  GemmArgs args;
  args.C = C; args.bias = bias; args.M = M; args.N = N; args.K = K; args.ldc = ldc;
  args.n_tile = n_tile; args.a_mn = a_mn_major; args.b_mn = b_mn_major; args.relu = relu;

  const uint32_t b_slot = (uint32_t(n_tile) * 128u + 1023u) & ~1023u;
  const size_t smem = size_t(kGemmStages) * (16384 + b_slot) + 256 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    DFM_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        227 * 1024));
    attr_set = true;
  }
  dim3 grid(ceil_div(M, kBlockM), ceil_div(N, n_tile));
  gemm_bf16_kernel<<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, args);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}
