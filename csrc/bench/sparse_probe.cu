// Runs the real sparse_apply_kernel in isolation with synthetic state to bisect its cost.
#include "../cuda/optim.cu"
#include <vector>
#include <algorithm>
#include <random>
using namespace dfm;
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)
namespace dfm { void set_error(const char*, ...) {} }

int main() {
  const int V = 117581, n = 16850, Ke = 32;
  SparseArgs a{};
  CK(cudaMalloc(&a.fm_v, size_t(V) * 128)); CK(cudaMalloc(&a.v_s0, size_t(V) * 128)); CK(cudaMalloc(&a.v_s1, size_t(V) * 128));
  CK(cudaMalloc(&a.fm_w, V * 4)); CK(cudaMalloc(&a.w_s0, V * 4)); CK(cudaMalloc(&a.w_s1, V * 4));
  CK(cudaMalloc(&a.G, size_t(80000) * 128)); CK(cudaMalloc(&a.Gw, 80000 * 4)); CK(cudaMalloc(&a.slot_of, V * 4));
  int* uniq; CK(cudaMalloc(&uniq, 80000 * 4)); a.uniq_ids = uniq;
  CK(cudaMalloc(&a.st, sizeof(StepState)));
  CK(cudaMemset(a.fm_v, 0, size_t(V) * 128)); CK(cudaMemset(a.v_s0, 0, size_t(V) * 128)); CK(cudaMemset(a.v_s1, 0, size_t(V) * 128));
  CK(cudaMemset(a.fm_w, 0, V * 4)); CK(cudaMemset(a.w_s0, 0, V * 4)); CK(cudaMemset(a.w_s1, 0, V * 4));
  CK(cudaMemset(a.G, 0, size_t(80000) * 128)); CK(cudaMemset(a.Gw, 0, 80000 * 4));
  a.Ke = Ke; a.V = V; a.grad_scale = 1.f;
  std::vector<int> h(V); for (int i = 0; i < V; ++i) h[i] = i;
  std::mt19937 rng(1); std::shuffle(h.begin(), h.end(), rng);
  CK(cudaMemcpy(uniq, h.data(), n * 4, cudaMemcpyHostToDevice));
  OptHyper hy{}; hy.beta1 = 0.9f; hy.beta2 = 0.999f; hy.eps = 1e-8f; hy.momentum = 0.95f; hy.ftrl_lr_power = -0.5f; hy.l2_reg = 1e-4f;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int kind : {DFM_OPT_ADAM, DFM_OPT_GD, DFM_OPT_MOMENTUM}) {
    hy.kind = kind;
    for (int grid : {592, 1184}) {
      float tot = 0;
      for (int it = 0; it < 25; ++it) {
        StepState st{}; st.count = 2048; st.step = 5; st.lr = 5e-4f; st.n_unique = n; st.train = 1;
        CK(cudaMemcpy(a.st, &st, sizeof(st), cudaMemcpyHostToDevice));
        cudaEventRecord(e0);
        sparse_apply_kernel<<<grid, 256>>>(a, hy);
        cudaEventRecord(e1); CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (it >= 5) tot += ms * 1000.f / 20;
      }
      printf("kind %d grid %4d: %.2f us\n", kind, grid, tot);
    }
  }
  return 0;
}
