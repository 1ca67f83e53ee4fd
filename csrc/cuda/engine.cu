// The step engine: owns every per-step buffer, the TMA descriptors and the launch order of
// the fused kernels, and exposes forward / train_step through the C ABI.  Parameters and
// optimizer slots are owned by the caller (torch tensors); the engine keeps the bf16
// tensor-core copy of the dense weights and the gradient buffers.
//
// One training step is seven launches on two streams (captured into a CUDA graph by the
// caller if it wishes; nothing here synchronises or allocates after create()):
//
//   embed_fm_fwd -> mlp_fwd -> mlp_bwd_chain -+-> embed_bwd ----------------+-> sparse_apply
//                                             +-> wgrad -> dense_opt -------+
//
// It stands in for one session.run(train_op) of the reference's Estimator loop
// (HVD.py:430-435: forward, autodiff, Adam apply; SURVEY 3.1 hot loop).
#include <vector>

#include "engine.cuh"

namespace dfm {

struct Engine {
  dfm_engine_config cfg;
  dfm_engine_tensors t;
  MlpShape sh;
  OptHyper hyper;
  int Bpad = 0;
  int n_dense = 0;
  size_t off[kMaxLayers + 2];           // flat offsets of P[l]; off[L+1] = fm_bias
  std::vector<void*> allocs;

  __nv_bfloat16* Wb = nullptr;
  float* dP = nullptr;
  __nv_bfloat16* act[kMaxLayers + 1] = {};
  __nv_bfloat16* dz[kMaxLayers + 1] = {};
  float *s_sum = nullptr, *fm_part = nullptr, *logit = nullptr, *prob = nullptr, *dlogit = nullptr;
  float *G = nullptr, *Gw = nullptr;
  int *slot_of = nullptr, *uniq_ids = nullptr;
  StepState* st = nullptr;
  WgradItem* items = nullptr;
  int n_items = 0;

  FwdMaps fmaps; BwdMaps bmaps; WgradMaps wmaps;
  int fwd_smem = 0, fwd_stages = 0, fwd_bslot = 0, fwd_chunks = 0;
  int bwd_smem = 0, bwd_stages = 0, bwd_bslot = 0, bwd_chunks = 0;

  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;

  // optional: the captured train step and the double-buffered host->device input pipeline
  static constexpr int kPipe = 2;
  cudaGraphExec_t graph[kPipe] = {nullptr, nullptr};
  uint8_t* in_slab[kPipe] = {nullptr, nullptr};     // [ids | vals | labels] for batch_max rows
  size_t slab_bytes = 0, off_vals = 0, off_labels = 0;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_h2d[kPipe] = {nullptr, nullptr}, ev_done[kPipe] = {nullptr, nullptr};
  long long pipe_step = 0;

  template <typename T>
  int alloc(T** p, size_t n, int fill_byte = 0) {
    void* q = nullptr;
    const size_t bytes = std::max<size_t>(n * sizeof(T), 256);
    DFM_CUDA_CHECK(cudaMalloc(&q, bytes));
    DFM_CUDA_CHECK(cudaMemset(q, fill_byte, bytes));
    allocs.push_back(q);
    *p = static_cast<T*>(q);
    return DFM_OK;
  }
};

static int engine_build(Engine& E) {
  const dfm_engine_config& c = E.cfg;
  DFM_REQUIRE(c.n_layers >= 1 && c.n_layers <= kMaxLayers, "n_layers %d outside [1,%d]", c.n_layers, kMaxLayers);
  DFM_REQUIRE(c.embedding_size % 16 == 0 && c.embedding_size <= 128,
              "the fused engine needs embedding_size %% 16 == 0 and <= 128 (got %d)", c.embedding_size);
  DFM_REQUIRE(c.batch_max >= 1 && c.field_size >= 1 && c.feature_size >= 1, "bad batch/field/feature size");
  MlpShape& sh = E.sh;
  sh.L = c.n_layers;
  sh.K[0] = c.field_size * c.embedding_size;
  for (int l = 0; l < sh.L; ++l) {
    DFM_REQUIRE(c.layers[l] % 16 == 0 && c.layers[l] >= 16 && c.layers[l] <= 512,
                "deep layer width %d must be a multiple of 16 in [16,512] (pad it on the host)", c.layers[l]);
    sh.N[l] = c.layers[l];
    sh.K[l + 1] = c.layers[l];
    sh.keep[l] = c.keep_prob[l] > 0.f ? c.keep_prob[l] : 1.f;
  }
  sh.N[sh.L] = 1;
  for (int l = 0; l <= sh.L; ++l) sh.pitch[l] = round_up(sh.K[l] + 1, 8);
  E.Bpad = round_up(c.batch_max, kTileM);

  size_t o = 0;
  for (int l = 0; l <= sh.L; ++l) {
    E.off[l] = o;
    o += size_t(sh.K[l] + 1) * sh.N[l];
  }
  E.off[sh.L + 1] = o;       // fm_bias
  E.n_dense = int(o + 1);

  E.hyper.kind = c.optimizer;
  E.hyper.beta1 = c.beta1; E.hyper.beta2 = c.beta2; E.hyper.eps = c.eps; E.hyper.momentum = c.momentum;
  E.hyper.ftrl_lr_power = c.ftrl_lr_power; E.hyper.ftrl_l1 = c.ftrl_l1; E.hyper.ftrl_l2 = c.ftrl_l2;
  E.hyper.l2_reg = c.l2_reg;

  int rc;
  const size_t rows = E.Bpad;
  if ((rc = E.alloc(&E.Wb, size_t(E.n_dense) + 8))) return rc;
  if ((rc = E.alloc(&E.dP, size_t(E.n_dense) + 8))) return rc;
  for (int l = 0; l <= sh.L; ++l) {
    if ((rc = E.alloc(&E.act[l], rows * sh.pitch[l]))) return rc;
    if ((rc = launch_fill_ones_column(E.act[l], int(rows), sh.pitch[l], sh.K[l], nullptr))) return rc;
    const int ncol = l < sh.L ? sh.N[l] : 16;
    if ((rc = E.alloc(&E.dz[l], rows * ncol))) return rc;
  }
  if ((rc = E.alloc(&E.s_sum, rows * c.embedding_size))) return rc;
  if ((rc = E.alloc(&E.fm_part, rows))) return rc;
  if ((rc = E.alloc(&E.logit, rows))) return rc;
  if ((rc = E.alloc(&E.prob, rows))) return rc;
  if ((rc = E.alloc(&E.dlogit, rows))) return rc;
  const size_t max_slots = size_t(c.batch_max) * c.field_size;
  if ((rc = E.alloc(&E.G, max_slots * c.embedding_size))) return rc;
  if ((rc = E.alloc(&E.Gw, max_slots))) return rc;
  if ((rc = E.alloc(&E.uniq_ids, max_slots))) return rc;
  if ((rc = E.alloc(&E.slot_of, size_t(c.feature_size), 0xFF))) return rc;   // all -1
  if ((rc = E.alloc(&E.st, 1))) return rc;

  // ---- shared-memory plans ----
  E.fwd_smem = mlp_fwd_plan(sh, &E.fwd_stages, &E.fwd_bslot, &E.fwd_chunks);
  DFM_REQUIRE(E.fwd_smem > 0, "deep_layers too wide for the fused forward kernel's shared-memory plan");
  E.bwd_smem = mlp_bwd_plan(sh, &E.bwd_stages, &E.bwd_bslot, &E.bwd_chunks);
  DFM_REQUIRE(E.bwd_smem > 0, "deep_layers too wide for the fused backward kernel's shared-memory plan");
  if ((rc = mlp_fwd_set_smem(E.fwd_smem))) return rc;
  if ((rc = mlp_bwd_set_smem(E.bwd_smem))) return rc;
  if ((rc = wgrad_set_smem())) return rc;

  // ---- TMA descriptors ----
  if ((rc = make_tmap_bf16_2d(&E.fmaps.e, E.act[0], sh.K[0], rows, size_t(sh.pitch[0]) * 2, 64, kTileM))) return rc;
  for (int l = 0; l < sh.L; ++l) {
    const __nv_bfloat16* w = E.Wb + E.off[l];
    if ((rc = make_tmap_bf16_2d(&E.fmaps.w[l], w, sh.N[l], sh.K[l], size_t(sh.N[l]) * 2, 64, 64))) return rc;
    if ((rc = make_tmap_bf16_2d(&E.bmaps.w[l], w, sh.N[l], sh.K[l], size_t(sh.N[l]) * 2, 64, std::min(256, sh.K[l]))))
      return rc;
  }
  if ((rc = make_tmap_bf16_2d(&E.bmaps.dz0, E.dz[0], sh.N[0], rows, size_t(sh.N[0]) * 2, 64, kTileM))) return rc;
  for (int l = 0; l <= sh.L; ++l) {
    if ((rc = make_tmap_bf16_2d(&E.wmaps.act[l], E.act[l], sh.K[l] + 1, rows, size_t(sh.pitch[l]) * 2, 64, 64))) return rc;
    const int ncol = l < sh.L ? sh.N[l] : 16;
    if ((rc = make_tmap_bf16_2d(&E.wmaps.dz[l], E.dz[l], ncol, rows, size_t(ncol) * 2, 64, 64))) return rc;
  }

  // ---- weight-gradient work list: (layer, 128-row tile, 256-col slab) x split-K over the batch ----
  std::vector<WgradItem> base;
  for (int l = 0; l <= sh.L; ++l) {
    const int ncol = l < sh.L ? sh.N[l] : 16;
    for (int m0 = 0; m0 < sh.K[l] + 1; m0 += kTileM)
      for (int n0 = 0; n0 < ncol; n0 += 256)
        base.push_back(WgradItem{l, m0, n0, std::min(256, round_up(ncol - n0, 16)), 0, 0});
  }
  const int k_chunks = ceil_div(int(rows), kChunkK);
  int ksplit = std::max(1, std::min(k_chunks, 148 / std::max<int>(1, int(base.size()))));
  const int per = ceil_div(k_chunks, ksplit);
  ksplit = ceil_div(k_chunks, per);
  std::vector<WgradItem> items;
  for (int s = 0; s < ksplit; ++s)
    for (auto it : base) {
      it.k_begin = s * per * kChunkK;
      it.k_end = std::min(k_chunks, (s + 1) * per) * kChunkK;
      items.push_back(it);
    }
  E.n_items = int(items.size());
  if ((rc = E.alloc(&E.items, items.size()))) return rc;
  DFM_CUDA_CHECK(cudaMemcpy(E.items, items.data(), items.size() * sizeof(WgradItem), cudaMemcpyHostToDevice));

  DFM_CUDA_CHECK(cudaStreamCreateWithFlags(&E.side, cudaStreamNonBlocking));
  DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E.ev_fork, cudaEventDisableTiming));
  DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E.ev_join, cudaEventDisableTiming));

  StepState init{};
  init.count = c.batch_max; init.step = 1; init.lr = c.lr; init.train = 1;
  DFM_CUDA_CHECK(cudaMemcpy(E.st, &init, sizeof(init), cudaMemcpyHostToDevice));
  if ((rc = launch_cast_weights(E.t.dense, E.Wb, E.n_dense, nullptr))) return rc;
  DFM_CUDA_CHECK(cudaDeviceSynchronize());
  return DFM_OK;
}

static int engine_forward(Engine& E, const int* ids, const float* vals, const float* labels, int dedupe,
                          cudaStream_t s) {
  const dfm_engine_config& c = E.cfg;
  const float* fm_bias = E.t.dense + E.off[E.sh.L + 1];
  int rc = launch_embed_fm_fwd(ids, vals, E.t.fm_w, E.t.fm_v, fm_bias, E.act[0], E.sh.pitch[0], E.s_sum, E.fm_part,
                               c.field_size, c.embedding_size, c.batch_max, E.st, E.slot_of, E.uniq_ids, dedupe, s);
  if (rc) return rc;
  FwdParams fp{};
  fp.sh = E.sh;
  for (int l = 0; l <= E.sh.L; ++l) {
    fp.P[l] = E.t.dense + E.off[l];
    fp.act[l] = E.act[l];
  }
  fp.fm_part = E.fm_part; fp.labels = labels; fp.logit = E.logit; fp.prob = E.prob;
  fp.dlogit = dedupe ? E.dlogit : nullptr;
  fp.st = E.st; fp.seed = c.seed; fp.stages = E.fwd_stages; fp.b_slot_bytes = E.fwd_bslot;
  fp.a_region_chunks = E.fwd_chunks;
  return launch_mlp_fwd(E.fmaps, fp, c.batch_max, E.fwd_smem, s);
}

static int engine_backward_update(Engine& E, const int* ids, const float* vals, cudaStream_t s) {
  const dfm_engine_config& c = E.cfg;
  BwdParams bp{};
  bp.sh = E.sh;
  for (int l = 0; l <= E.sh.L; ++l) {
    bp.act[l] = E.act[l];
    bp.dz[l] = E.dz[l];
  }
  bp.w_out = E.t.dense + E.off[E.sh.L];
  bp.dlogit = E.dlogit; bp.st = E.st; bp.stages = E.bwd_stages; bp.b_slot_bytes = E.bwd_bslot;
  bp.a_region_chunks = E.bwd_chunks;
  bp.ids = ids; bp.vals = vals; bp.s_sum = E.s_sum; bp.slot_of = E.slot_of; bp.G = E.G; bp.Gw = E.Gw;
  bp.F = c.field_size; bp.Ke = c.embedding_size;
  int rc = launch_mlp_bwd_chain(E.bmaps, bp, c.batch_max, E.bwd_smem, s);
  if (rc) return rc;

  // fork: embedding gradient on the side stream, weight gradients + dense update on the main one
  DFM_CUDA_CHECK(cudaEventRecord(E.ev_fork, s));
  DFM_CUDA_CHECK(cudaStreamWaitEvent(E.side, E.ev_fork, 0));
  if ((rc = launch_embed_bwd(E.bmaps, bp, c.batch_max, E.side))) return rc;
  DFM_CUDA_CHECK(cudaEventRecord(E.ev_join, E.side));

  WgradParams wp{};
  wp.sh = E.sh;
  for (int l = 0; l <= E.sh.L; ++l) wp.dP[l] = E.dP + E.off[l];
  wp.items = E.items; wp.st = E.st;
  if ((rc = launch_wgrad(E.wmaps, wp, E.n_items, s))) return rc;
  const float gscale = 1.f;
  if ((rc = launch_dense_opt(E.t.dense, E.t.dense_s0, E.t.dense_s1, E.dP, E.Wb, E.n_dense, E.st, E.hyper, gscale, s)))
    return rc;

  DFM_CUDA_CHECK(cudaStreamWaitEvent(s, E.ev_join, 0));
  SparseArgs sa{};
  sa.fm_v = E.t.fm_v; sa.v_s0 = E.t.fm_v_s0; sa.v_s1 = E.t.fm_v_s1;
  sa.fm_w = E.t.fm_w; sa.w_s0 = E.t.fm_w_s0; sa.w_s1 = E.t.fm_w_s1;
  sa.G = E.G; sa.Gw = E.Gw; sa.slot_of = E.slot_of; sa.uniq_ids = E.uniq_ids; sa.st = E.st;
  sa.Ke = c.embedding_size; sa.V = c.feature_size; sa.grad_scale = gscale;
  return launch_sparse_apply(sa, E.hyper, c.batch_max * c.field_size, c.table_update_mode == DFM_TABLE_DENSE_EXACT, s);
}

}  // namespace dfm

using namespace dfm;

extern "C" int dfm_engine_create(const dfm_engine_config* cfg, const dfm_engine_tensors* tensors, void** out) {
  DFM_REQUIRE(cfg && tensors && out, "dfm_engine_create: null argument");
  Engine* E = new Engine();
  E->cfg = *cfg;
  E->t = *tensors;
  int rc = engine_build(*E);
  if (rc) {
    for (void* p : E->allocs) cudaFree(p);
    delete E;
    return rc;
  }
  *out = E;
  return DFM_OK;
}

extern "C" int dfm_engine_destroy(void* h) {
  if (!h) return DFM_OK;
  Engine* E = static_cast<Engine*>(h);
  cudaDeviceSynchronize();
  for (void* p : E->allocs) cudaFree(p);
  if (E->side) cudaStreamDestroy(E->side);
  if (E->ev_fork) cudaEventDestroy(E->ev_fork);
  if (E->ev_join) cudaEventDestroy(E->ev_join);
  for (int i = 0; i < Engine::kPipe; ++i) {
    if (E->graph[i]) cudaGraphExecDestroy(E->graph[i]);
    if (E->ev_h2d[i]) cudaEventDestroy(E->ev_h2d[i]);
    if (E->ev_done[i]) cudaEventDestroy(E->ev_done[i]);
  }
  if (E->copy_stream) cudaStreamDestroy(E->copy_stream);
  delete E;
  return DFM_OK;
}

extern "C" int dfm_engine_set_state(void* h, int count, float lr, int train, int step, dfm_stream_t s) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(count >= 0 && count <= E->cfg.batch_max, "count %d outside [0,%d]", count, E->cfg.batch_max);
  return launch_set_state(E->st, count, lr, train, step, static_cast<cudaStream_t>(s));
}

extern "C" int dfm_engine_sync_weights(void* h, dfm_stream_t s) {
  Engine* E = static_cast<Engine*>(h);
  return launch_cast_weights(E->t.dense, E->Wb, E->n_dense, static_cast<cudaStream_t>(s));
}

extern "C" int dfm_engine_forward(void* h, const int* ids, const float* vals, const float* labels, dfm_stream_t s) {
  return engine_forward(*static_cast<Engine*>(h), ids, vals, labels, 0, static_cast<cudaStream_t>(s));
}

extern "C" int dfm_engine_train_step(void* h, const int* ids, const float* vals, const float* labels,
                                     dfm_stream_t s_) {
  Engine* E = static_cast<Engine*>(h);
  cudaStream_t s = static_cast<cudaStream_t>(s_);
  DFM_REQUIRE(labels != nullptr, "dfm_engine_train_step: labels are required");
  int rc = engine_forward(*E, ids, vals, labels, 1, s);
  if (rc) return rc;
  return engine_backward_update(*E, ids, vals, s);
}

extern "C" int dfm_engine_launches_per_step(void* h) { (void)h; return 7; }

// ---- CUDA-graph replay + double-buffered input pipeline -------------------------------------
// Input slab layout (host and device): ids int32 [batch_max*F] | vals f32 [batch_max*F] | labels f32 [batch_max]
extern "C" int dfm_engine_slab_layout(void* h, long long* bytes, long long* off_vals, long long* off_labels) {
  Engine* E = static_cast<Engine*>(h);
  const size_t bf = size_t(E->cfg.batch_max) * E->cfg.field_size;
  *off_vals = (long long)(bf * 4);
  *off_labels = (long long)(bf * 8);
  *bytes = (long long)(bf * 8 + size_t(E->cfg.batch_max) * 4);
  return DFM_OK;
}

// Captures the full train step once per input slab.  `stream` must not be the legacy default stream.
// Uses the step state as it is at replay time (set count/lr/train first; the step counter advances on
// the device).
extern "C" int dfm_engine_capture(void* h, dfm_stream_t s_) {
  Engine* E = static_cast<Engine*>(h);
  cudaStream_t s = static_cast<cudaStream_t>(s_);
  long long bytes, ov, ol;
  dfm_engine_slab_layout(h, &bytes, &ov, &ol);
  E->slab_bytes = size_t(bytes); E->off_vals = size_t(ov); E->off_labels = size_t(ol);
  for (int i = 0; i < Engine::kPipe; ++i) {
    if (!E->in_slab[i]) {
      int rc = E->alloc(&E->in_slab[i], E->slab_bytes);
      if (rc) return rc;
    }
    if (!E->ev_h2d[i]) {
      DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E->ev_h2d[i], cudaEventDisableTiming));
      DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E->ev_done[i], cudaEventDisableTiming));
    }
  }
  if (!E->copy_stream) DFM_CUDA_CHECK(cudaStreamCreateWithFlags(&E->copy_stream, cudaStreamNonBlocking));
  for (int i = 0; i < Engine::kPipe; ++i) {
    if (E->graph[i]) { cudaGraphExecDestroy(E->graph[i]); E->graph[i] = nullptr; }
    cudaGraph_t g = nullptr;
    DFM_CUDA_CHECK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    const int* ids = reinterpret_cast<const int*>(E->in_slab[i]);
    const float* vals = reinterpret_cast<const float*>(E->in_slab[i] + E->off_vals);
    const float* labels = reinterpret_cast<const float*>(E->in_slab[i] + E->off_labels);
    int rc = engine_forward(*E, ids, vals, labels, 1, s);
    if (!rc) rc = engine_backward_update(*E, ids, vals, s);
    cudaError_t ce = cudaStreamEndCapture(s, &g);
    if (rc) { if (g) cudaGraphDestroy(g); return rc; }
    if (ce != cudaSuccess) { set_error("cudaStreamEndCapture: %s", cudaGetErrorString(ce)); return DFM_ERR_CUDA; }
    ce = cudaGraphInstantiate(&E->graph[i], g, 0);
    cudaGraphDestroy(g);
    if (ce != cudaSuccess) { set_error("cudaGraphInstantiate: %s", cudaGetErrorString(ce)); return DFM_ERR_CUDA; }
  }
  E->pipe_step = 0;
  return DFM_OK;
}

// Device-resident replay (inputs already in slab `which`).
extern "C" int dfm_engine_replay(void* h, int which, dfm_stream_t s) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(which >= 0 && which < Engine::kPipe && E->graph[which], "dfm_engine_replay: capture first");
  DFM_CUDA_CHECK(cudaGraphLaunch(E->graph[which], static_cast<cudaStream_t>(s)));
  return DFM_OK;
}

extern "C" int dfm_engine_slab_ptr(void* h, int which, void** ptr) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(which >= 0 && which < Engine::kPipe && E->in_slab[which], "dfm_engine_slab_ptr: capture first");
  *ptr = E->in_slab[which];
  return DFM_OK;
}

// One end-to-end step from pinned host memory: H2D of the batch on the copy stream (overlapping the
// previous step's kernels), graph replay on `stream`, D2H of the loss sum into `loss_host` (pinned).
extern "C" int dfm_engine_step_from_host(void* h, const void* host_slab, float* loss_host, dfm_stream_t s_) {
  Engine* E = static_cast<Engine*>(h);
  cudaStream_t s = static_cast<cudaStream_t>(s_);
  const int w = int(E->pipe_step % Engine::kPipe);
  DFM_REQUIRE(E->graph[w] != nullptr, "dfm_engine_step_from_host: capture first");
  // slab w was last read by step (pipe_step - kPipe)
  if (E->pipe_step >= Engine::kPipe) DFM_CUDA_CHECK(cudaStreamWaitEvent(E->copy_stream, E->ev_done[w], 0));
  DFM_CUDA_CHECK(cudaMemcpyAsync(E->in_slab[w], host_slab, E->slab_bytes, cudaMemcpyHostToDevice, E->copy_stream));
  DFM_CUDA_CHECK(cudaEventRecord(E->ev_h2d[w], E->copy_stream));
  DFM_CUDA_CHECK(cudaStreamWaitEvent(s, E->ev_h2d[w], 0));
  DFM_CUDA_CHECK(cudaGraphLaunch(E->graph[w], s));
  DFM_CUDA_CHECK(cudaEventRecord(E->ev_done[w], s));
  if (loss_host)
    DFM_CUDA_CHECK(cudaMemcpyAsync(loss_host, &E->st->loss_sum, sizeof(float), cudaMemcpyDeviceToHost, s));
  E->pipe_step++;
  return DFM_OK;
}

extern "C" int dfm_engine_buffer(void* h, const char* name, int index, void** ptr, long long* n_elems) {
  Engine* E = static_cast<Engine*>(h);
  const MlpShape& sh = E->sh;
  const long long rows = E->Bpad;
  void* p = nullptr;
  long long n = 0;
  if (!strcmp(name, "logit")) { p = E->logit; n = rows; }
  else if (!strcmp(name, "prob")) { p = E->prob; n = rows; }
  else if (!strcmp(name, "dlogit")) { p = E->dlogit; n = rows; }
  else if (!strcmp(name, "fm_part")) { p = E->fm_part; n = rows; }
  else if (!strcmp(name, "s_sum")) { p = E->s_sum; n = rows * E->cfg.embedding_size; }
  else if (!strcmp(name, "state")) { p = E->st; n = sizeof(StepState) / 4; }
  else if (!strcmp(name, "dense_grad")) { p = E->dP; n = E->n_dense; }
  else if (!strcmp(name, "dense_bf16")) { p = E->Wb; n = E->n_dense; }
  else if (!strcmp(name, "G")) { p = E->G; n = (long long)E->cfg.batch_max * E->cfg.field_size * E->cfg.embedding_size; }
  else if (!strcmp(name, "Gw")) { p = E->Gw; n = (long long)E->cfg.batch_max * E->cfg.field_size; }
  else if (!strcmp(name, "uniq_ids")) { p = E->uniq_ids; n = (long long)E->cfg.batch_max * E->cfg.field_size; }
  else if (!strcmp(name, "slot_of")) { p = E->slot_of; n = E->cfg.feature_size; }
  else if (!strcmp(name, "act")) {
    DFM_REQUIRE(index >= 0 && index <= sh.L, "act index %d out of range", index);
    p = E->act[index]; n = rows * sh.pitch[index];
  } else if (!strcmp(name, "dz")) {
    DFM_REQUIRE(index >= 0 && index <= sh.L, "dz index %d out of range", index);
    p = E->dz[index]; n = rows * (index < sh.L ? sh.N[index] : 16);
  } else {
    set_error("dfm_engine_buffer: unknown buffer '%s'", name);
    return DFM_ERR_ARG;
  }
  *ptr = p;
  *n_elems = n;
  return DFM_OK;
}

extern "C" int dfm_engine_layout(void* h, long long* offsets, int* pitches, int* n_dense, int* bpad) {
  Engine* E = static_cast<Engine*>(h);
  for (int l = 0; l <= E->sh.L + 1; ++l) offsets[l] = (long long)E->off[l];
  for (int l = 0; l <= E->sh.L; ++l) pitches[l] = E->sh.pitch[l];
  *n_dense = E->n_dense;
  *bpad = E->Bpad;
  return DFM_OK;
}
