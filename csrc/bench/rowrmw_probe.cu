// Microbenchmark: what does a warp-per-row read-modify-write of scattered 128-byte rows cost on B200,
// and which part of it (loads, vector stores, scalar stores, dependent index load) is expensive?
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)

template <int MODE>
__global__ void __launch_bounds__(256) k(const int* __restrict__ idx, int n, float* p, float* m, float* v, float* g,
                                         float* w, float* wm, float* wv, int* slot) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < n; s += warps) {
    const int id = (MODE == 9) ? s : idx[s];
    const size_t pi = size_t(id) * 32 + lane, gi = size_t(s) * 32 + lane;
    if (MODE == 0 || MODE == 9) {          // full RMW of 3 row arrays + G, plus lane-0 scalars
      float a = p[pi], b = m[pi], c = v[pi], d = g[gi];
      g[gi] = 0.f;
      b = 0.9f * b + 0.1f * d; c = 0.999f * c + 0.001f * d * d; a -= 0.01f * b / (sqrtf(c) + 1e-8f);
      p[pi] = a; m[pi] = b; v[pi] = c;
      if (lane == 0) { float x = w[id], y = wm[id], z = wv[id]; y += x; z += y; w[id] = x - z; wm[id] = y; wv[id] = z; slot[id] = -1; }
    } else if (MODE == 1) {   // vector part only
      float a = p[pi], b = m[pi], c = v[pi], d = g[gi];
      g[gi] = 0.f;
      b = 0.9f * b + 0.1f * d; c = 0.999f * c + 0.001f * d * d; a -= 0.01f * b / (sqrtf(c) + 1e-8f);
      p[pi] = a; m[pi] = b; v[pi] = c;
    } else if (MODE == 2) {   // loads only (one store so the work is not optimised away)
      float a = p[pi] + m[pi] + v[pi] + g[gi];
      if (a == 123.456f) g[gi] = a;
    } else if (MODE == 3) {   // stores only
      p[pi] = 1.f; m[pi] = 2.f; v[pi] = 3.f; g[gi] = 0.f;
    } else if (MODE == 4) {   // single array RMW
      p[pi] = p[pi] * 0.99f;
    } else if (MODE == 5) {   // packed record: p,m,v contiguous (3*32 floats per id)
      float* r = p + size_t(id) * 96;
      float a = r[lane], b = r[32 + lane], c = r[64 + lane], d = g[gi];
      g[gi] = 0.f;
      b = 0.9f * b + 0.1f * d; c = 0.999f * c + 0.001f * d * d; a -= 0.01f * b / (sqrtf(c) + 1e-8f);
      r[lane] = a; r[32 + lane] = b; r[64 + lane] = c;
    } else if (MODE == 6) {   // scalar lane-0 part only
      if (lane == 0) { float x = w[id], y = wm[id], z = wv[id]; y += x; z += y; w[id] = x - z; wm[id] = y; wv[id] = z; slot[id] = -1; }
    }
  }
}

template <int MODE> float run(const int* idx, int n, int grid, float* p, float* m, float* v, float* g, float* w, float* wm, float* wv, int* slot) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 5; ++i) k<MODE><<<grid, 256>>>(idx, n, p, m, v, g, w, wm, wv, slot);
  cudaEventRecord(e0);
  for (int i = 0; i < 50; ++i) k<MODE><<<grid, 256>>>(idx, n, p, m, v, g, w, wm, wv, slot);
  cudaEventRecord(e1); CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms * 1000.f / 50;
}

int main() {
  const int V = 117581, n = 16850;
  float *p, *m, *v, *g, *w, *wm, *wv; int *slot, *idx;
  CK(cudaMalloc(&p, size_t(V) * 96 * 4)); CK(cudaMalloc(&m, size_t(V) * 32 * 4)); CK(cudaMalloc(&v, size_t(V) * 32 * 4));
  CK(cudaMalloc(&g, size_t(80000) * 32 * 4)); CK(cudaMalloc(&w, V * 4)); CK(cudaMalloc(&wm, V * 4)); CK(cudaMalloc(&wv, V * 4));
  CK(cudaMalloc(&slot, V * 4)); CK(cudaMalloc(&idx, n * 4));
  CK(cudaMemset(p, 0, size_t(V) * 96 * 4)); CK(cudaMemset(m, 0, size_t(V) * 128)); CK(cudaMemset(v, 0, size_t(V) * 128)); CK(cudaMemset(g, 0, 80000 * 128));
  std::vector<int> h(V); for (int i = 0; i < V; ++i) h[i] = i;
  std::mt19937 rng(1); std::shuffle(h.begin(), h.end(), rng);
  CK(cudaMemcpy(idx, h.data(), n * 4, cudaMemcpyHostToDevice));
  const char* names[] = {"full_rmw", "vector_only", "loads_only", "stores_only", "single_array_rmw", "packed_record", "scalar_only"};
  for (int grid : {148, 592, 1184, 2368}) {
    printf("grid %4d:", grid);
    printf(" %s=%.2f", names[0], run<0>(idx, n, grid, p, m, v, g, w, wm, wv, slot));
    printf(" %s=%.2f", names[1], run<1>(idx, n, grid, p, m, v, g, w, wm, wv, slot));
    printf(" %s=%.2f", names[2], run<2>(idx, n, grid, p, m, v, g, w, wm, wv, slot));
    printf(" %s=%.2f", names[3], run<3>(idx, n, grid, p, m, v, g, w, wm, wv, slot));
    printf(" %s=%.2f", names[4], run<4>(idx, n, grid, p, m, v, g, w, wm, wv, slot));
    printf(" %s=%.2f", names[5], run<5>(idx, n, grid, p, m, v, g, w, wm, wv, slot));
    printf(" %s=%.2f", names[6], run<6>(idx, n, grid, p, m, v, g, w, wm, wv, slot));
    printf(" sequential_ids=%.2f", run<9>(idx, n, grid, p, m, v, g, w, wm, wv, slot));
    printf("  (us)\n");
  }
  return 0;
}
