// Batch normalisation of the deep tower (tf.contrib.layers.batch_norm, HVD.py:230-233 + 306-328).
//
// The reference normalises AFTER the ReLU (fully_connected has the ReLU built in): per hidden layer
//     a = relu(h W + b)                       written by the tcgen05 layer kernel (bf16)
//     y = gamma * (a - mean) * rstd + beta    mean / variance over the batch (training) or the moving averages
//     h' = dropout_keep(y)                    the next layer's input
// and in training also moves the averages in place (updates_collections=None): m <- d m + (1 - d) batch.
// Batch statistics need every row of a column, which no 128-row GEMM tile has, so BN is its own small launch between
// two layer kernels: one CTA per 8 columns (one 16-byte granule of every row), 256 threads striding the rows; the
// [B, 8] slab (32 KiB at B = 2048) stays in L1/L2 between the statistics pass and the apply pass.
//
// Backward (autodiff of the above), per column with sums over the batch of size B:
//     dbeta = sum dy,  dgamma = sum dy * xhat
//     da = gamma * rstd * (dy - dbeta / B - xhat * dgamma / B);   dz = (a > 0) ? da : 0
// dy arrives with the dropout of this layer already applied (layer kernel / loss kernel epilogues).
#include "engine.cuh"
#include "ptx.cuh"

namespace dfm {

constexpr int kBnThreads = 256;
constexpr int kBnCols = 8;

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
    f[2 * i] = __low2float(v);
    f[2 * i + 1] = __high2float(v);
  }
}

// sums `v[0..NV)` over the CTA; every thread returns with the totals
template <int NV>
__device__ __forceinline__ void block_sum(float (&v)[NV], float* smem /* [8][NV] */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
  __syncthreads();                       // smem may still be read from a previous call
  if (lane == 0)
#pragma unroll
    for (int i = 0; i < NV; ++i) smem[warp * NV + i] = v[i];
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kBnThreads / 32; ++w) t += smem[w * NV + i];
    v[i] = t;
  }
}

__global__ void __launch_bounds__(kBnThreads)
bn_fwd_kernel(const BnParams p) {
  __shared__ float red[(kBnThreads / 32) * 8];
  const int c0 = blockIdx.x * kBnCols;
  const int count = p.st->count;
  const bool train = p.st->train != 0;
  float mean[8], rstd[8];
  if (train) {
    float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = threadIdx.x; r < count; r += kBnThreads) {
      float a[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(p.pre + size_t(r) * p.n + c0)), a);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] += a[i];
    }
    block_sum<8>(s, red);
    const float inv = 1.f / float(count);
#pragma unroll
    for (int i = 0; i < 8; ++i) mean[i] = s[i] * inv;
    float q[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = threadIdx.x; r < count; r += kBnThreads) {
      float a[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(p.pre + size_t(r) * p.n + c0)), a);
#pragma unroll
      for (int i = 0; i < 8; ++i) q[i] += (a[i] - mean[i]) * (a[i] - mean[i]);
    }
    block_sum<8>(q, red);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float var = q[i] * inv;      // biased, like tf.nn.moments
      rstd[i] = rsqrtf(var + p.eps);
      if (threadIdx.x == 0) {
        p.save[c0 + i] = mean[i];
        p.save[p.n + c0 + i] = rstd[i];
        p.moving[c0 + i] = p.decay * p.moving[c0 + i] + (1.f - p.decay) * mean[i];
        p.moving[p.n + c0 + i] = p.decay * p.moving[p.n + c0 + i] + (1.f - p.decay) * var;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      mean[i] = __ldg(p.moving + c0 + i);
      rstd[i] = rsqrtf(__ldg(p.moving + p.n + c0 + i) + p.eps);
    }
  }
  float sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sc[i] = __ldg(p.gamma + c0 + i) * rstd[i];
    sh[i] = __ldg(p.beta + c0 + i) - mean[i] * sc[i];
  }
  const bool drop = train && p.keep < 1.f;
  const float inv_keep = drop ? 1.f / p.keep : 1.f;
  const uint32_t step = uint32_t(p.st->step);
  for (int r = threadIdx.x; r < count; r += kBnThreads) {
    float a[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(p.pre + size_t(r) * p.n + c0)), a);
    const uint32_t kb = drop ? dropout_keep_bits8(p.seed, uint32_t(r), uint32_t(p.layer), uint32_t(c0 >> 3), step, p.keep) : 0xffu;
    float y[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) y[i] = ((kb >> i) & 1u) ? (a[i] * sc[i] + sh[i]) * inv_keep : 0.f;
    *reinterpret_cast<uint4*>(p.out + size_t(r) * p.out_pitch + c0) =
        make_uint4(pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]), pack_bf16x2(y[4], y[5]), pack_bf16x2(y[6], y[7]));
  }
}

__global__ void __launch_bounds__(kBnThreads)
bn_bwd_kernel(const BnParams p, int rows_pad) {
  __shared__ float red[(kBnThreads / 32) * 16];
  const int c0 = blockIdx.x * kBnCols;
  const int count = p.st->count;
  float mean[8], rstd[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    mean[i] = p.save[c0 + i];
    rstd[i] = p.save[p.n + c0 + i];
  }
  float s[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) s[i] = 0.f;
  for (int r = threadIdx.x; r < count; r += kBnThreads) {
    float a[8], dy[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(p.pre + size_t(r) * p.n + c0)), a);
    unpack8(*reinterpret_cast<const uint4*>(p.dy + size_t(r) * p.n + c0), dy);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      s[i] += dy[i];
      s[8 + i] += dy[i] * (a[i] - mean[i]) * rstd[i];
    }
  }
  block_sum<16>(s, red);
  if (threadIdx.x < 8) {
    p.dgamma[c0 + threadIdx.x] = s[8 + threadIdx.x];          // dgamma
    p.dgamma[p.n + c0 + threadIdx.x] = s[threadIdx.x];        // dbeta
  }
  const float inv = 1.f / float(count);
  float g[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) g[i] = __ldg(p.gamma + c0 + i) * rstd[i];
  // rows past the batch lie inside the weight-gradient GEMM's batch range: they must read as zero
  for (int r = threadIdx.x; r < rows_pad; r += kBnThreads) {
    float dz[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (r < count) {
      float a[8], dy[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(p.pre + size_t(r) * p.n + c0)), a);
      unpack8(*reinterpret_cast<const uint4*>(p.dy + size_t(r) * p.n + c0), dy);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float xhat = (a[i] - mean[i]) * rstd[i];
        dz[i] = a[i] > 0.f ? g[i] * (dy[i] - s[i] * inv - xhat * s[8 + i] * inv) : 0.f;
      }
    }
    *reinterpret_cast<uint4*>(p.dy + size_t(r) * p.n + c0) =
        make_uint4(pack_bf16x2(dz[0], dz[1]), pack_bf16x2(dz[2], dz[3]), pack_bf16x2(dz[4], dz[5]), pack_bf16x2(dz[6], dz[7]));
  }
}

int launch_bn_fwd(const BnParams& p, cudaStream_t stream) {
  DFM_REQUIRE(p.n % kBnCols == 0 && p.out_pitch % 8 == 0, "batch norm: layer width %d must be a multiple of 8", p.n);
  bn_fwd_kernel<<<p.n / kBnCols, kBnThreads, 0, stream>>>(p);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

int launch_bn_bwd(const BnParams& p, int rows_pad, cudaStream_t stream) {
  bn_bwd_kernel<<<p.n / kBnCols, kBnThreads, 0, stream>>>(p, rows_pad);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm
