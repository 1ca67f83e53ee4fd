// Host-side plumbing shared by every translation unit of libdfm_cuda.so:
// error reporting through the C ABI, TMA tensor-map encoding, launch checks.
#pragma once

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/deepfm_b200.h"

namespace dfm {

// Last error text, per host thread (read with dfm_last_error()).
void set_error(const char* fmt, ...);

// Encodes a 2-D bf16 tensor map with 128-byte swizzle.
//   inner/outer      : tensor extent in elements (inner is contiguous)
//   pitch_bytes      : bytes between consecutive outer indices (multiple of 16)
//   box_inner/outer  : tile extent in elements (box_inner*2 must be 128)
// Returns 0 on success.
int make_tmap_bf16_2d(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer,
                      uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_outer);

#define DFM_CUDA_CHECK(expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ::dfm::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,               \
                       cudaGetErrorString(_e));                                           \
      return DFM_ERR_CUDA;                                                                \
    }                                                                                     \
  } while (0)

#define DFM_LAUNCH_CHECK()                                                                \
  do {                                                                                    \
    cudaError_t _e = cudaGetLastError();                                                  \
    if (_e != cudaSuccess) {                                                              \
      ::dfm::set_error("kernel launch failed at %s:%d: %s", __FILE__, __LINE__,           \
                       cudaGetErrorString(_e));                                           \
      return DFM_ERR_CUDA;                                                                \
    }                                                                                     \
  } while (0)

#define DFM_REQUIRE(cond, ...)                                                            \
  do {                                                                                    \
    if (!(cond)) {                                                                        \
      ::dfm::set_error(__VA_ARGS__);                                                      \
      return DFM_ERR_ARG;                                                                 \
    }                                                                                     \
  } while (0)

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

// ---------------------------------------------------------------------------
// Philox4x32-10 counter RNG (dropout masks).  Same stream on every replay of a
// (seed, step, layer, row, column-group) tuple, so forward and tests agree.
// ---------------------------------------------------------------------------
__host__ __device__ inline void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint64_t p0 = 0xD2511F53ull * c[0];
  const uint64_t p1 = 0xCD9E8D57ull * c[2];
  const uint32_t hi0 = static_cast<uint32_t>(p0 >> 32), lo0 = static_cast<uint32_t>(p0);
  const uint32_t hi1 = static_cast<uint32_t>(p1 >> 32), lo1 = static_cast<uint32_t>(p1);
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}

__host__ __device__ inline void philox4x32(uint32_t seed_lo, uint32_t seed_hi, uint32_t c0,
                                           uint32_t c1, uint32_t c2, uint32_t c3,
                                           uint32_t (&out)[4]) {
  uint32_t c[4] = {c0, c1, c2, c3};
  uint32_t k0 = seed_lo, k1 = seed_hi;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
}

// Dropout keep decisions of 8 consecutive activation columns [8*colgroup, 8*colgroup+8) of example `row` after
// hidden layer `layer` at optimizer step `step`: bit i set = column 8*colgroup+i is kept (probability `keep`).
// One Philox call yields eight 16-bit uniforms.  Every kernel that needs a mask recomputes it from here.
__host__ __device__ inline uint32_t dropout_keep_bits8(unsigned long long seed, uint32_t row, uint32_t layer,
                                                       uint32_t colgroup, uint32_t step, float keep) {
  const uint32_t thr = uint32_t(keep * 65536.f);
  uint32_t rnd[4];
  philox4x32(uint32_t(seed), uint32_t(seed >> 32), row, layer, colgroup, step, rnd);
  uint32_t bits = 0u;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    bits |= ((rnd[i] & 0xffffu) < thr ? 1u : 0u) << (2 * i);
    bits |= ((rnd[i] >> 16) < thr ? 1u : 0u) << (2 * i + 1);
  }
  return bits;
}

// uniform in [0,1)
__host__ __device__ inline float u32_to_unit(uint32_t x) { return (x >> 8) * (1.0f / 16777216.0f); }

// Launch with the programmatic-stream-serialization attribute (programmatic dependent launch): the kernel may be
// scheduled while the previous kernel of the stream is still running and must call pdl_wait() (ptx.cuh) before it
// touches anything that kernel wrote.  Inside a stream capture this becomes a programmatic graph edge.
// DFM_PDL=0 in the environment turns the attribute off (plain stream order).
// DFM_PDL is a bit mask (default: all): 1 forward layers, 2 loss, 4 backward layers, 8 dense_opt, 16 sparse_apply
enum : int { kPdlFwd = 1, kPdlLoss = 2, kPdlBwd = 4, kPdlDense = 8, kPdlSparse = 16 };
bool pdl_enabled(int which);
#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int which, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled(which) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

}  // namespace dfm
