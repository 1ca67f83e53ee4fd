// Bisects the cost of sparse_apply_kernel: variants remove one feature at a time.
#include "../cuda/optim.cu"
#include <vector>
#include <algorithm>
#include <random>
using namespace dfm;
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)
namespace dfm { void set_error(const char*, ...) {} }

// VAR bits: 1 = n passed by value (no st reads for n), 2 = constants passed by value (no st reads at all),
//           4 = skip lane-0 scalar part, 8 = Adam inlined without switch, 16 = no l2 term, 32 = compile-time Ke
template <int VAR, int SC = 255>
__global__ void __launch_bounds__(256) variant(const SparseArgs a, const OptHyper hyper, int n_arg, OptStepConsts c_arg) {
  OptStepConsts c = (VAR & 2) ? c_arg : make_opt_consts(hyper, a.st->lr, a.st->step);
  const int n = (VAR & 1) ? n_arg : a.st->n_unique;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const int Ke = (VAR & 32) ? 32 : a.Ke;
  const float l2 = (VAR & 16) ? 0.f : hyper.l2_reg;
  for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < n; s += warps) {
    const int id = a.uniq_ids[s];
    for (int k = lane; k < Ke; k += 32) {
      const size_t gi = size_t(s) * Ke + k, pi = size_t(id) * Ke + k;
      float p = a.fm_v[pi], m = a.v_s0[pi], v = a.v_s1[pi];
      const float g = a.G[gi] * a.grad_scale + l2 * p;
      a.G[gi] = 0.f;
      if (VAR & 8) { m = c.b1 * m + (1.f - c.b1) * g; v = c.b2 * v + (1.f - c.b2) * g * g; p -= c.lr_t * m / (sqrtf(v) + c.eps); }
      else opt_apply(c, g, p, m, v);
      a.fm_v[pi] = p; a.v_s0[pi] = m; a.v_s1[pi] = v;
    }
    if (!(VAR & 4) && lane == 0) {
      // SC bits: 1 fm_w RMW, 2 w_s0 RMW, 4 w_s1 RMW, 8 Gw read+zero, 16 slot_of reset, 32 math
      float p = (SC & 1) ? a.fm_w[id] : 1.f, m = (SC & 2) ? a.w_s0[id] : 0.f, v = (SC & 4) ? a.w_s1[id] : 0.f;
      float g = l2 * p;
      if (SC & 8) { g += a.Gw[s] * a.grad_scale; a.Gw[s] = 0.f; }
      if (SC & 32) { m = c.b1 * m + (1.f - c.b1) * g; v = c.b2 * v + (1.f - c.b2) * g * g; p -= c.lr_t * m / (sqrtf(v) + c.eps); }
      else { m += g; v += m; p -= v; }
      if (SC & 1) a.fm_w[id] = p;
      if (SC & 2) a.w_s0[id] = m;
      if (SC & 4) a.w_s1[id] = v;
      if (SC & 16) a.slot_of[id] = -1;
      if (!(SC & 7) && p == 123.f) a.Gw[s] = p;
    }
  }
}

template <int VAR, int SC = 255> void run(const char* name, SparseArgs a, OptHyper hy, int n, int grid) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  OptStepConsts c{}; c.kind = hy.kind; c.lr = 5e-4f; c.lr_t = 4e-4f; c.b1 = 0.9f; c.b2 = 0.999f; c.eps = 1e-8f;
  float tot = 0;
  for (int it = 0; it < 25; ++it) {
    cudaEventRecord(e0);
    variant<VAR, SC><<<grid, 256>>>(a, hy, n, c);
    cudaEventRecord(e1); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (it >= 5) tot += ms * 1000.f / 20;
  }
  printf("  %-44s %.2f us\n", name, tot);
}

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
  OptHyper hy{}; hy.kind = DFM_OPT_ADAM; hy.beta1 = 0.9f; hy.beta2 = 0.999f; hy.eps = 1e-8f; hy.momentum = 0.95f; hy.ftrl_lr_power = -0.5f; hy.l2_reg = 1e-4f;
  StepState st{}; st.count = 2048; st.step = 5; st.lr = 5e-4f; st.n_unique = n; st.train = 1;
  CK(cudaMemcpy(a.st, &st, sizeof(st), cudaMemcpyHostToDevice));
  for (int grid : {592}) {
    printf("grid %d\n", grid);
    run<3 | 8, 255>("all scalar ops", a, hy, n, grid);
    run<3 | 8, 1 | 32>("only fm_w RMW (+math)", a, hy, n, grid);
    run<3 | 8, 1>("only fm_w RMW (trivial math)", a, hy, n, grid);
    run<3 | 8, 7>("fm_w,w_s0,w_s1 RMW (trivial math)", a, hy, n, grid);
    run<3 | 8, 7 | 32>("fm_w,w_s0,w_s1 RMW (+math)", a, hy, n, grid);
    run<3 | 8, 8>("only Gw read+zero", a, hy, n, grid);
    run<3 | 8, 16>("only slot_of reset", a, hy, n, grid);
    run<3 | 8, 32>("only math on lane 0", a, hy, n, grid);
    run<3 | 8, 0>("empty lane-0 block", a, hy, n, grid);
    run<3 | 4 | 8>("no scalar part", a, hy, n, grid);
  }
  return 0;
}
