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
#include <algorithm>
#include <cstdlib>
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
  // batch norm (cfg.batch_norm): a_l = relu(.) before normalisation, per-batch {mean, rstd}, offsets of
  // {gamma, beta} inside the flat dense vector and of {moving mean, moving variance} inside t.bn_moving
  bool bn = false;
  __nv_bfloat16* pre[kMaxLayers] = {};
  float* bn_save[kMaxLayers] = {};
  size_t off_bn[kMaxLayers] = {};
  size_t off_mov[kMaxLayers] = {};
  float *s_sum = nullptr, *fm_part = nullptr, *logit = nullptr, *prob = nullptr, *dlogit = nullptr;
  float *G = nullptr, *Gw = nullptr;
  int *slot_of = nullptr, *uniq_ids = nullptr;
  StepState* st = nullptr;
  WgradItem* items = nullptr;
  int n_items = 0;
  int ksplit = 1;
  size_t part_stride = 0;      // dP holds [ksplit][part_stride] split-K slices of the dense gradient

  FwdMaps fmaps; WgradMaps wmaps;

  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_fork0 = nullptr, ev_dedupe = nullptr;

  // optional: the captured train step and the double-buffered host->device input pipeline
  static constexpr int kPipe = 2;
  cudaGraphExec_t graph[kPipe] = {nullptr, nullptr};
  uint8_t* in_slab[kPipe] = {nullptr, nullptr};     // [ids | vals | labels] for batch_max rows
  size_t slab_bytes = 0, off_vals = 0, off_labels = 0;
  cudaEvent_t ev_done[kPipe] = {nullptr, nullptr};
  long long pipe_step = 0;
  // pipelined flavours of the step graph: beside the kernels that train on slab w, a memcpy node prefetches the
  // NEXT batch into slab 1-w, and another copies the loss sum into a pinned ring -- no cross-stream events and no
  // stream-ordered copies between two graph launches (they cost 11-18 us per step, see scripts/gpu_replay_floor.py)
  enum { kSrcHost = 0, kSrcDevice = 1 };
  cudaGraph_t pf_graph[kPipe][2] = {{nullptr, nullptr}, {nullptr, nullptr}};   // kept alive: they own the node handles
  cudaGraphNode_t pf_node[kPipe][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
  cudaGraphExec_t pf_exec[kPipe][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
  cudaStream_t pf_stream = nullptr;
  cudaEvent_t ev_pf_fork = nullptr, ev_pf_join = nullptr, ev_loss = nullptr;
  uint8_t* dummy_host = nullptr;               // pinned placeholder source used at capture time
  float* loss_ring = nullptr;                  // pinned [kPipe]
  const void* staged_src[kPipe] = {nullptr, nullptr};   // what the prefetch node last copied into each slab

  // multi-GPU: views of every rank's symmetric block and tables (see comm.cu)
  XchgArgs xa{};
  bool comm_ready = false;
  bool emu = false;                           // single-GPU emulation of N ranks: the collectives are launched for all
                                              // virtual ranks at once by dfm_engine_emu_collectives
  XchgArgs* emu_args = nullptr;               // device array [emu_ranks] (owned by virtual rank 0)
  int emu_ranks = 0;
  unsigned long long* comm_dbg = nullptr;
  unsigned long long* kern_dbg = nullptr;
  int dedupe_after = 0;          // forward layer after which dedupe is forked (-1: right after embed_fwd); DFM_DEDUPE_AFTER

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
  DFM_REQUIRE(c.embedding_size % 16 == 0 && c.embedding_size <= 64,
              "the fused engine needs embedding_size %% 16 == 0 and <= 64 (got %d)", c.embedding_size);
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
  for (int l = 0; l <= sh.L; ++l) sh.pitch[l] = round_up(sh.K[l] + 1, 16);   // rows start on 32-byte boundaries (256-bit stores)
  E.Bpad = round_up(c.batch_max, kTileM);

  E.bn = c.batch_norm != 0;
  DFM_REQUIRE(!E.bn || E.t.bn_moving, "batch_norm needs the bn_moving tensor");
  DFM_REQUIRE(c.loss_type == DFM_LOSS_LOG || c.loss_type == DFM_LOSS_SQUARE, "unknown loss_type %d", c.loss_type);
  size_t o = 0, om = 0;
  for (int l = 0; l < sh.L; ++l) {
    E.off[l] = o;
    o += size_t(sh.K[l] + 1) * sh.N[l];
  }
  if (E.bn)
    for (int l = 0; l < sh.L; ++l) {     // [gamma | beta] per layer, between the hidden layers and the out layer
      E.off_bn[l] = o;   o += 2 * size_t(sh.N[l]);
      E.off_mov[l] = om; om += 2 * size_t(sh.N[l]);
    }
  E.off[sh.L] = o;
  o += size_t(sh.K[sh.L] + 1) * sh.N[sh.L];
  E.off[sh.L + 1] = o;       // fm_bias
  E.n_dense = int(o + 1);

  E.hyper.kind = c.optimizer;
  E.hyper.beta1 = c.beta1; E.hyper.beta2 = c.beta2; E.hyper.eps = c.eps; E.hyper.momentum = c.momentum;
  E.hyper.ftrl_lr_power = c.ftrl_lr_power; E.hyper.ftrl_l1 = c.ftrl_l1; E.hyper.ftrl_l2 = c.ftrl_l2;
  E.hyper.l2_reg = c.l2_reg;

  int rc;
  const size_t rows = E.Bpad;
  if ((rc = E.alloc(&E.Wb, size_t(E.n_dense) + 8))) return rc;
  for (int l = 0; l <= sh.L; ++l) {
    if ((rc = E.alloc(&E.act[l], rows * sh.pitch[l]))) return rc;
    if ((rc = launch_fill_ones_column(E.act[l], int(rows), sh.pitch[l], sh.K[l], nullptr))) return rc;
    const int ncol = l < sh.L ? sh.N[l] : 16;
    if ((rc = E.alloc(&E.dz[l], rows * ncol))) return rc;
  }
  if (E.bn)
    for (int l = 0; l < sh.L; ++l) {
      if ((rc = E.alloc(&E.pre[l], rows * sh.N[l]))) return rc;
      if ((rc = E.alloc(&E.bn_save[l], 2 * size_t(sh.N[l])))) return rc;
    }
  if ((rc = E.alloc(&E.s_sum, rows * c.embedding_size))) return rc;
  if ((rc = E.alloc(&E.fm_part, rows))) return rc;
  if ((rc = E.alloc(&E.logit, rows))) return rc;
  if ((rc = E.alloc(&E.prob, rows))) return rc;
  if ((rc = E.alloc(&E.dlogit, rows))) return rc;
  const size_t max_slots = size_t(c.batch_max) * c.field_size;
  if ((rc = E.alloc(&E.G, max_slots * g_pitch(c.embedding_size)))) return rc;
  if ((rc = E.alloc(&E.Gw, max_slots))) return rc;
  if ((rc = E.alloc(&E.uniq_ids, max_slots))) return rc;
  if ((rc = E.alloc(&E.slot_of, size_t(c.feature_size), 0xFF))) return rc;   // all -1
  if ((rc = E.alloc(&E.st, 1))) return rc;
  if ((rc = E.alloc(&E.kern_dbg, 512))) return rc;

  // ---- kernel attributes ----
  if ((rc = mlp_layer_set_smem())) return rc;
  if ((rc = embed_bwd_set_smem(c.embedding_size))) return rc;
  if ((rc = wgrad_set_smem())) return rc;
  // ---- TMA descriptors ----
  for (int l = 0; l < sh.L; ++l) {
    const __nv_bfloat16* w = E.Wb + E.off[l];
    if ((rc = make_tmap_bf16_2d(&E.fmaps.act[l], E.act[l], sh.K[l], rows, size_t(sh.pitch[l]) * 2, 64, kTileM))) return rc;
    if ((rc = make_tmap_bf16_2d(&E.fmaps.w[l], w, sh.N[l], sh.K[l], size_t(sh.N[l]) * 2, 64, 64))) return rc;
    if ((rc = make_tmap_bf16_2d(&E.fmaps.dz[l], E.dz[l], sh.N[l], rows, size_t(sh.N[l]) * 2, 64, kTileM))) return rc;
  }
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
        base.push_back(WgradItem{l, m0, n0, std::min(256, round_up(ncol - n0, 16)), 0, 0, 0});
  }
  const int k_chunks = ceil_div(int(rows), kChunkK);
  // wgrad runs beside embed_bwd (side stream); both kernels need an SM to themselves per CTA (shared
  // memory), so together they must fit the chip or they take turns and the overlap is lost.
  const int eb_ctas = ceil_div(c.batch_max, kTileM) * ceil_div(sh.K[0], 256);
  const int wg_budget = std::max(int(base.size()), 148 - std::min(eb_ctas, 100));
  int ksplit = std::max(1, std::min(k_chunks, wg_budget / std::max<int>(1, int(base.size()))));
  const int per = ceil_div(k_chunks, ksplit);
  ksplit = ceil_div(k_chunks, per);
  std::vector<WgradItem> items;
  for (int s = 0; s < ksplit; ++s)
    for (auto it : base) {
      it.k_begin = s * per * kChunkK;
      it.k_end = std::min(k_chunks, (s + 1) * per) * kChunkK;
      it.ks = s;
      items.push_back(it);
    }
  E.ksplit = ksplit;
  E.part_stride = size_t(round_up(E.n_dense, 4)) + 8;
  if ((rc = E.alloc(&E.dP, E.part_stride * ksplit))) return rc;
  E.n_items = int(items.size());
  if ((rc = E.alloc(&E.items, items.size()))) return rc;
  DFM_CUDA_CHECK(cudaMemcpy(E.items, items.data(), items.size() * sizeof(WgradItem), cudaMemcpyHostToDevice));

  DFM_CUDA_CHECK(cudaStreamCreateWithFlags(&E.side, cudaStreamNonBlocking));
  DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E.ev_fork, cudaEventDisableTiming));
  DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E.ev_join, cudaEventDisableTiming));
  DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E.ev_fork0, cudaEventDisableTiming));
  DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E.ev_dedupe, cudaEventDisableTiming));

  StepState init{};
  init.count = c.batch_max; init.step = 0; init.lr = c.lr; init.train = 1;
  if (E.hyper.kind == DFM_OPT_ADAM) { init.adam_b1 = E.hyper.beta1; init.adam_b2 = E.hyper.beta2; init.lr_t = c.lr; }
  DFM_CUDA_CHECK(cudaMemcpy(E.st, &init, sizeof(init), cudaMemcpyHostToDevice));
  if ((rc = launch_cast_weights(E.t.dense, E.Wb, E.n_dense, nullptr))) return rc;
  DFM_CUDA_CHECK(cudaDeviceSynchronize());
  return DFM_OK;
}

static BnParams bn_params(const Engine& E, int l) {
  BnParams b{};
  b.st = E.st; b.pre = E.pre[l]; b.out = E.act[l + 1]; b.dy = E.dz[l];
  b.gamma = E.t.dense + E.off_bn[l]; b.beta = b.gamma + E.sh.N[l];
  b.moving = E.t.bn_moving + E.off_mov[l]; b.save = E.bn_save[l];
  b.dgamma = E.dP + E.off_bn[l];                 // slice 0 of the split-K gradient buffer; the other slices stay zero
  b.n = E.sh.N[l]; b.out_pitch = E.sh.pitch[l + 1]; b.layer = l;
  b.keep = E.sh.keep[l]; b.decay = E.cfg.bn_decay; b.eps = E.cfg.bn_eps; b.seed = E.cfg.seed;
  return b;
}

static int engine_forward(Engine& E, const int* ids, const float* vals, const float* labels, int train,
                          cudaStream_t s) {
  const dfm_engine_config& c = E.cfg;
  const MlpShape& sh = E.sh;
  int rc;
  const float* fm_bias = E.t.dense + E.off[sh.L + 1];
  // embed_fm_fwd opens the step: it also clears the per-step counters (n_unique, n_seg, loss_sum)
  if ((rc = launch_embed_fm_fwd(ids, vals, E.t.fm_w, E.t.fm_v, fm_bias, E.act[0], sh.pitch[0], E.s_sum, E.fm_part,
                                c.field_size, c.embedding_size, c.batch_max, E.st, train,
                                W0Cast{E.t.dense + E.off[0], E.Wb + E.off[0], sh.K[0] * sh.N[0]}, c.feature_size, s)))
    return rc;
  // Sparse-gradient bookkeeping (dedupe) runs on the side stream beside the MLP; its results are first needed by
  // embed_bwd, which is queued behind it there.  It is forked after layer `dedupe_after` (default 0): next to the
  // first, widest layer its 80k atomics-heavy threads slowed that layer's TMA ring from ~8 to ~24 us, while the
  // later layers only occupy 16-32 SMs.
  auto fork_dedupe = [&]() -> int {
    DFM_CUDA_CHECK(cudaEventRecord(E.ev_fork0, s));
    DFM_CUDA_CHECK(cudaStreamWaitEvent(E.side, E.ev_fork0, 0));
    const int r = launch_dedupe(ids, c.field_size, E.st, E.slot_of, E.uniq_ids, c.batch_max,
                                E.comm_ready ? E.xa.cv.world : 1, c.batch_max * c.field_size, c.feature_size, E.side);
    if (r) return r;
    DFM_CUDA_CHECK(cudaEventRecord(E.ev_dedupe, E.side));
    return DFM_OK;
  };
  const int dedupe_after = std::min(E.dedupe_after, sh.L - 1);
  if (train && dedupe_after < 0 && (rc = fork_dedupe())) return rc;
  for (int l = 0; l < sh.L; ++l) {
    LayerParams lp{};
    lp.st = E.st; lp.bias = E.t.dense + E.off[l] + size_t(sh.K[l]) * sh.N[l];
    lp.out = E.act[l + 1]; lp.out_pitch = sh.pitch[l + 1];
    lp.n_out = sh.N[l]; lp.k_red = sh.K[l]; lp.b_box_rows = 64; lp.layer = l; lp.keep = sh.keep[l]; lp.seed = c.seed;
    if (E.bn) { lp.out = E.pre[l]; lp.out_pitch = sh.N[l]; lp.keep = 1.f; }   // BN kernel normalises and drops out
    lp.dbg = E.kern_dbg + 384 + 8 * l;            // kern_dbg: [0,328) embed_bwd, [384,480) 8 stamps per layer kernel
    if ((rc = launch_mlp_layer(false, E.fmaps.act[l], E.fmaps.w[l], lp, c.batch_max, s))) return rc;
    if (E.bn && (rc = launch_bn_fwd(bn_params(E, l), s))) return rc;
    if (train && l == dedupe_after && (rc = fork_dedupe())) return rc;
  }
  LossParams q{};
  q.st = E.st; q.h_last = E.act[sh.L]; q.h_pitch = sh.pitch[sh.L]; q.n_last = sh.K[sh.L];
  q.w_out = E.t.dense + E.off[sh.L]; q.fm_part = E.fm_part; q.labels = labels;
  q.logit = E.logit; q.prob = E.prob; q.dlogit = train ? E.dlogit : nullptr;
  q.dz_last = train ? E.dz[sh.L - 1] : nullptr; q.dz_out = E.dz[sh.L];
  q.keep_last = sh.keep[sh.L - 1]; q.rows_pad = E.Bpad;
  q.loss_type = c.loss_type; q.bn = E.bn; q.seed = c.seed; q.last_layer = sh.L - 1;
  return launch_loss_init(q, train ? E.Bpad : c.batch_max, s);
}

static SparseArgs sparse_args(const Engine& E) {
  SparseArgs sa{};
  sa.fm_v = E.t.fm_v; sa.v_s0 = E.t.fm_v_s0; sa.v_s1 = E.t.fm_v_s1;
  sa.fm_w = E.t.fm_w; sa.w_s0 = E.t.fm_w_s0; sa.w_s1 = E.t.fm_w_s1;
  sa.G = E.G; sa.Gw = E.Gw; sa.slot_of = E.slot_of; sa.uniq_ids = E.uniq_ids; sa.st = E.st;
  sa.Ke = E.cfg.embedding_size; sa.V = E.cfg.feature_size; sa.grad_scale = 1.f;
  return sa;
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
  bp.dlogit = E.dlogit; bp.st = E.st;
  bp.ids = ids; bp.vals = vals; bp.s_sum = E.s_sum; bp.slot_of = E.slot_of; bp.G = E.G; bp.Gw = E.Gw;
  bp.F = c.field_size; bp.Ke = c.embedding_size; bp.V = int(c.feature_size); bp.dbg = E.kern_dbg;
  int rc;
  // batch norm: the loss kernel / each backward layer kernel leaves the gradient w.r.t. the BN OUTPUT in dz[l];
  // bn_bwd turns it into the gradient w.r.t. the layer's pre-activation, in place, and writes dgamma / dbeta
  if (E.bn && (rc = launch_bn_bwd(bn_params(E, E.sh.L - 1), E.Bpad, s))) return rc;
  for (int l = E.sh.L - 1; l >= 1; --l) {       // dz_{l-1} = (h_l > 0) * (dz_l W_l^T) / keep_{l-1}
    LayerParams lp{};
    lp.st = E.st; lp.mask = E.act[l]; lp.mask_pitch = E.sh.pitch[l];
    lp.out = E.dz[l - 1]; lp.out_pitch = E.sh.N[l - 1];
    lp.n_out = E.sh.K[l]; lp.k_red = E.sh.N[l]; lp.b_box_rows = 64; lp.layer = l; lp.keep = E.sh.keep[l - 1];
    lp.seed = c.seed; lp.bn = E.bn;
    lp.dbg = E.kern_dbg + 384 + 8 * (kMaxLayers + l);
    if ((rc = launch_mlp_layer(true, E.fmaps.dz[l], E.fmaps.w[l], lp, c.batch_max, s))) return rc;
    if (E.bn && (rc = launch_bn_bwd(bn_params(E, l - 1), E.Bpad, s))) return rc;
  }

  // fork: the embedding gradient and the sparse update stay on the main stream (they are the critical path, and a
  // same-stream kernel->kernel edge is ~2 us cheaper than a cross-stream one); the weight gradients and the dense
  // update run beside them on the side stream, behind dedupe
  DFM_CUDA_CHECK(cudaEventRecord(E.ev_fork, s));
  DFM_CUDA_CHECK(cudaStreamWaitEvent(E.side, E.ev_fork, 0));
  // wgrad is queued first: its grid is sized to what embed_bwd leaves free (or to a third of the GPU when
  // embed_bwd alone would fill it), and it heads the longer dependent chain of the two on wide networks
  WgradParams wp{};
  wp.sh = E.sh;
  for (int l = 0; l <= E.sh.L; ++l) wp.dP[l] = E.dP + E.off[l];
  wp.items = E.items; wp.st = E.st; wp.part_stride = E.part_stride;
  if ((rc = launch_wgrad(E.wmaps, wp, E.n_items, E.side))) return rc;
  DFM_CUDA_CHECK(cudaStreamWaitEvent(s, E.ev_dedupe, 0));
  if ((rc = launch_embed_bwd(E.fmaps, bp, c.batch_max, s))) return rc;
  const float gscale = 1.f;
  if (E.comm_ready) {
    // Multi-GPU.  The dense all-reduce + optimizer follows the weight-gradient GEMM on the side stream and so runs
    // under embed_bwd; the sparse exchange + owner-sharded update follows embed_bwd on the main stream.  In the
    // single-GPU emulation both are launched once for all virtual ranks by dfm_engine_emu_collectives.
    if (!E.emu && (rc = launch_dense_sync(E.xa, nullptr, 0, E.side))) return rc;
    DFM_CUDA_CHECK(cudaEventRecord(E.ev_join, E.side));
    if (!E.emu && (rc = launch_exchange_apply(E.xa, nullptr, 0, s))) return rc;
    if (!E.emu && E.xa.cv.exact &&
        (rc = launch_table_dense_from(sparse_args(E), E.hyper, E.xa.cv.tab_g[E.xa.cv.rank], 1.f / float(E.xa.cv.world), s)))
      return rc;
    DFM_CUDA_CHECK(cudaStreamWaitEvent(s, E.ev_join, 0));
    return DFM_OK;
  }
  // (the dense update leaves the bf16 copy of W_0 alone -- embed_bwd's GEMM may still be reading it; see W0Cast)
  if ((rc = launch_dense_opt(E.t.dense, E.t.dense_s0, E.t.dense_s1, E.dP, E.ksplit, E.part_stride, E.Wb, E.n_dense, E.st,
                             E.hyper, gscale, int(E.off[0]), int(E.off[0]) + E.sh.K[0] * E.sh.N[0], E.side)))
    return rc;
  DFM_CUDA_CHECK(cudaEventRecord(E.ev_join, E.side));

  SparseArgs sa = sparse_args(E);
  sa.grad_scale = gscale;
  rc = launch_sparse_apply(sa, E.hyper, c.batch_max * c.field_size, c.table_update_mode == DFM_TABLE_DENSE_EXACT, s);
  DFM_CUDA_CHECK(cudaStreamWaitEvent(s, E.ev_join, 0));
  return rc;
}

}  // namespace dfm

using namespace dfm;

extern "C" int dfm_engine_create(const dfm_engine_config* cfg, const dfm_engine_tensors* tensors, void** out) {
  DFM_REQUIRE(cfg && tensors && out, "dfm_engine_create: null argument");
  Engine* E = new Engine();
  E->cfg = *cfg;
  E->t = *tensors;
  if (const char* v = getenv("DFM_DEDUPE_AFTER")) E->dedupe_after = atoi(v);
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
  if (E->ev_fork0) cudaEventDestroy(E->ev_fork0);
  if (E->ev_dedupe) cudaEventDestroy(E->ev_dedupe);
  for (int i = 0; i < Engine::kPipe; ++i) {
    if (E->graph[i]) cudaGraphExecDestroy(E->graph[i]);
    if (E->ev_done[i]) cudaEventDestroy(E->ev_done[i]);
    for (int k = 0; k < 2; ++k) {
      if (E->pf_exec[i][k]) cudaGraphExecDestroy(E->pf_exec[i][k]);
      if (E->pf_graph[i][k]) cudaGraphDestroy(E->pf_graph[i][k]);
    }
  }
  if (E->pf_stream) cudaStreamDestroy(E->pf_stream);
  if (E->ev_pf_fork) cudaEventDestroy(E->ev_pf_fork);
  if (E->ev_pf_join) cudaEventDestroy(E->ev_pf_join);
  if (E->ev_loss) cudaEventDestroy(E->ev_loss);
  if (E->dummy_host) cudaFreeHost(E->dummy_host);
  if (E->loss_ring) cudaFreeHost(E->loss_ring);
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

extern "C" int dfm_engine_launches_per_step(void* h) {
  Engine* E = static_cast<Engine*>(h);
  // dedupe, embed_fwd, L forward layers, loss, L-1 backward layers, embed_bwd, wgrad, dense opt, sparse (1 or 2)
  // single GPU: + dense opt + sparse apply; multi GPU: + one exchange kernel
  const int mlp = E->sh.L + 1 + (E->sh.L - 1);
  return 2 + mlp + 2 + 2 + (E->bn ? 2 * E->sh.L : 0);   // + dense optimizer + sparse apply (multi-GPU: dense all-reduce + sparse exchange)
}

// ---- phase timing (warm, CUDA events) ----------------------------------------------------------
// ms_out[0] = forward (dedupe beside embed+FM, MLP layers, loss), ms_out[1] = backward + update,
// ms_out[2] = whole step.  Means over `iters` real training steps.  Single-GPU only.
extern "C" int dfm_engine_profile(void* h, const int* ids, const float* vals, const float* labels, int iters,
                                  float* ms_out) {
  Engine& E = *static_cast<Engine*>(h);
  DFM_REQUIRE(!E.comm_ready, "dfm_engine_profile is single-GPU only");
  cudaStream_t s = nullptr;
  DFM_CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  cudaEvent_t ev[3];
  for (auto& e : ev) DFM_CUDA_CHECK(cudaEventCreate(&e));
  for (int i = 0; i < 3; ++i) ms_out[i] = 0.f;
  int rc = DFM_OK;
  for (int it = -3; it < iters && !rc; ++it) {
    cudaEventRecord(ev[0], s);
    rc = engine_forward(E, ids, vals, labels, 1, s);
    cudaEventRecord(ev[1], s);
    if (!rc) rc = engine_backward_update(E, ids, vals, s);
    cudaEventRecord(ev[2], s);
    DFM_CUDA_CHECK(cudaStreamSynchronize(s));
    if (it >= 0) {
      float a = 0.f, b = 0.f;
      cudaEventElapsedTime(&a, ev[0], ev[1]);
      cudaEventElapsedTime(&b, ev[1], ev[2]);
      ms_out[0] += a / iters; ms_out[1] += b / iters; ms_out[2] += (a + b) / iters;
    }
  }
  for (auto& e : ev) cudaEventDestroy(e);
  cudaStreamDestroy(s);
  return rc;
}

// ---- multi-GPU attachment ---------------------------------------------------------------------
// The caller (parallel/symmetric.py) owns the symmetric memory: it allocates, on every rank, one block of
// dfm_engine_comm_bytes() zeroed bytes, makes every rank's block and embedding tables addressable from every rank
// (torch symmetric memory with an NVLS multicast address, or CUDA IPC), and hands the pointers to comm_attach.
static size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

static size_t comm_layout(const Engine& E, int world, CommView* cv) {
  const size_t cap = size_t(E.cfg.batch_max) * E.cfg.field_size;
  const size_t n4 = size_t(round_up(E.n_dense, 4)) + 8;
  size_t o = 0;
  CommView tmp{};
  CommView& v = cv ? *cv : tmp;
  v.off_dP = o;        o = align256(o + n4 * 4);
  v.flags_bytes = align256(size_t(world) * kCommMaxCtas * 4);
  v.off_flags = o;     o += 4 * v.flags_bytes;
  v.off_in_count = o;  o = align256(o + size_t(world) * kCommMaxCtas * 4);
  v.off_in_tails = o;  o = align256(o + size_t(world) * cap * 8);
  v.off_in_grads = o;  o = align256(o + size_t(world) * cap * E.cfg.embedding_size * 4);
  v.off_entries = o;   o = align256(o + (size_t(E.cfg.feature_size) + world) * 8);
  v.off_counter = o;   o += 256;
  return o;
}

extern "C" int dfm_engine_comm_bytes(void* h, int world, long long* bytes) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(world >= 2 && world <= kCommMaxRanks, "comm: world %d unsupported (2..%d)", world, kCommMaxRanks);
  *bytes = (long long)comm_layout(*E, world, nullptr);
  return DFM_OK;
}

// Buffers whose size depends on the world size, and the argument block of the collective kernels.
static int comm_prepare(Engine& E, int world, int rank, int ctas, int dense_ctas) {
  DFM_REQUIRE(world >= 2 && world <= kCommMaxRanks && rank >= 0 && rank < world, "comm: world %d rank %d unsupported",
              world, rank);
  DFM_REQUIRE(!E.comm_ready, "comm already attached");
  // the merge phase gives one lane per sender out of the embedding_size/4 lanes that serve a record
  DFM_REQUIRE(world <= E.cfg.embedding_size / 4, "multi-GPU fused path: world %d needs embedding_size >= %d", world,
              4 * world);
  DFM_REQUIRE(ctas >= world && ctas <= kCommMaxCtas && dense_ctas >= 1 && dense_ctas <= kCommMaxCtas, "comm: bad grid");
  const size_t cap = size_t(E.cfg.batch_max) * E.cfg.field_size;
  int rc;
  // one slot segment per owner rank (dedupe_kernel): worst case every id of a batch has the same owner
  if ((rc = E.alloc(&E.G, size_t(world) * cap * g_pitch(E.cfg.embedding_size)))) return rc;
  if ((rc = E.alloc(&E.uniq_ids, size_t(world) * cap))) return rc;
  XchgArgs& x = E.xa;
  x = XchgArgs{};
  if ((rc = E.alloc(&E.comm_dbg, 32))) return rc;
  if ((rc = E.alloc(&x.bar_local, 64))) return rc;
  CommView& cv = x.cv;
  comm_layout(E, world, &cv);
  cv.world = world; cv.rank = rank; cv.cap = int(cap);
  cv.ctas = ctas; cv.dense_ctas = dense_ctas;
  cv.dbg = E.comm_dbg;
  cv.exact = E.cfg.table_update_mode == DFM_TABLE_DENSE_EXACT;
  {
    StepState now{};
    DFM_CUDA_CHECK(cudaDeviceSynchronize());
    DFM_CUDA_CHECK(cudaMemcpy(&now, E.st, sizeof(now), cudaMemcpyDeviceToHost));
    cv.epoch_base = unsigned(now.comm_epoch);
  }
  double secs = 120.0;     // a rank may legitimately stall this long on the host (checkpoint, input) while its peers wait
  if (const char* v = getenv("DFM_COMM_TIMEOUT_S")) secs = atof(v);
  cv.watchdog = (long long)(std::max(secs, 0.001) * 1.9e9);
  x.a = sparse_args(E);
  x.d.P = E.t.dense; x.d.s0 = E.t.dense_s0; x.d.s1 = E.t.dense_s1; x.d.part = E.dP; x.d.ksplit = E.ksplit;
  x.d.part_stride = E.part_stride; x.d.Wb = E.Wb; x.d.n = E.n_dense;
  x.d.wb_skip_lo = int(E.off[0]); x.d.wb_skip_hi = int(E.off[0]) + E.sh.K[0] * E.sh.N[0];
  x.hyper = E.hyper;
  return DFM_OK;
}

// block_ptrs / fm_v_ptrs / fm_w_ptrs: `world` device pointers each (entry `rank` = this rank's own buffer); *_mc: the
// NVLS multicast address of the same buffers or NULL.  grad_ptrs / grad_mc: only for the dense_exact table update
// (per-rank [V, embedding_size + 8] fp32 zeroed tables that receive the summed sparse gradients), else NULL.
extern "C" int dfm_engine_comm_attach(void* h, int world, int rank, void* const* block_ptrs, void* block_mc,
                                      void* const* fm_v_ptrs, void* fm_v_mc, void* const* fm_w_ptrs, void* fm_w_mc,
                                      void* const* grad_ptrs, void* grad_mc) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(block_ptrs && fm_v_ptrs && fm_w_ptrs, "dfm_engine_comm_attach: null pointer table");
  // the dense kernel shares the GPU with the sparse exchange (one 1024-thread CTA per SM each): enough CTAs that its
  // optimizer pass (26 bytes per parameter) ends before the exchange does, no more
  int dense_ctas = std::min(48, std::max(24, E->n_dense / 11000));
  if (const char* v = getenv("DFM_DENSE_CTAS")) dense_ctas = atoi(v);
  // one 1024-thread CTA per SM, all resident at once, beside the dense kernel's CTAs
  int xchg_ctas = 148 - dense_ctas;
  if (const char* v = getenv("DFM_XCHG_CTAS")) xchg_ctas = std::max(world, std::min(atoi(v), 148 - dense_ctas));
  int rc = comm_prepare(*E, world, rank, xchg_ctas, dense_ctas);
  if (rc) return rc;
  CommView& cv = E->xa.cv;
  DFM_REQUIRE(fm_v_ptrs[rank] == E->t.fm_v && fm_w_ptrs[rank] == E->t.fm_w,
              "dfm_engine_comm_attach: entry %d of the table pointers must be this engine's own fm_v / fm_w", rank);
  DFM_REQUIRE(!cv.exact || grad_ptrs, "dense_exact on several GPUs needs the summed-gradient tables");
  for (int r = 0; r < world; ++r) {
    cv.base[r] = static_cast<uint8_t*>(block_ptrs[r]);
    cv.tab_v[r] = static_cast<float*>(fm_v_ptrs[r]);
    cv.tab_w[r] = static_cast<float*>(fm_w_ptrs[r]);
    cv.tab_g[r] = grad_ptrs ? static_cast<float*>(grad_ptrs[r]) : nullptr;
  }
  const char* no_mc = getenv("DFM_NO_MULTICAST");
  const bool use_mc = !(no_mc && no_mc[0] == '1');
  cv.mc_base = use_mc ? static_cast<uint8_t*>(block_mc) : nullptr;
  // the tables take multicast stores only if both have an alias
  const bool tab_mc = use_mc && fm_v_mc && fm_w_mc && (!cv.exact || grad_mc);
  cv.mc_v = tab_mc ? static_cast<float*>(fm_v_mc) : nullptr;
  cv.mc_w = tab_mc ? static_cast<float*>(fm_w_mc) : nullptr;
  cv.mc_g = tab_mc && cv.exact ? static_cast<float*>(grad_mc) : nullptr;
  E->comm_ready = true;
  return DFM_OK;
}

// 1 if the attached views use NVLS multicast for the dense all-reduce / for the table rows (bit 0 / bit 1)
extern "C" int dfm_engine_comm_multicast(void* h) {
  Engine* E = static_cast<Engine*>(h);
  return (E->xa.cv.mc_base ? 1 : 0) | (E->xa.cv.mc_v ? 2 : 0);
}

// ---- single-GPU emulation of N ranks ---------------------------------------------------------
// N engines on ONE device stand for N ranks: each runs its own forward / backward, and the two collective kernels
// are launched once, cooperatively, over all virtual ranks' data (blockIdx.x / ctas = rank).  This is how the
// multi-rank protocol is tested (and profiled under ncu) on a box with a single GPU; separate launches per rank that
// wait on one another on one GPU would not be guaranteed to run at the same time.
extern "C" int dfm_engine_emu_attach(void* const* engines, int n, void* const* grad_tables) {
  DFM_REQUIRE(engines && n >= 2 && n <= kCommMaxRanks, "dfm_engine_emu_attach: 2..%d engines", kCommMaxRanks);
  int dev = 0, sms = 0;
  DFM_CUDA_CHECK(cudaGetDevice(&dev));
  DFM_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int ctas = std::max(n, std::min(kCommMaxCtas, sms / n));     // all n * ctas CTAs must be co-resident
  const int dense_ctas = std::max(1, std::min(16, sms / n));
  std::vector<uint8_t*> blocks(n);
  for (int r = 0; r < n; ++r) {
    Engine* E = static_cast<Engine*>(engines[r]);
    int rc = comm_prepare(*E, n, r, ctas, dense_ctas);
    if (rc) return rc;
    if ((rc = E->alloc(&blocks[r], comm_layout(*E, n, nullptr)))) return rc;
  }
  std::vector<XchgArgs> all(n);
  for (int r = 0; r < n; ++r) {
    Engine* E = static_cast<Engine*>(engines[r]);
    CommView& cv = E->xa.cv;
    DFM_REQUIRE(!cv.exact || grad_tables, "dense_exact emulation needs the summed-gradient tables");
    for (int q = 0; q < n; ++q) {
      Engine* Q = static_cast<Engine*>(engines[q]);
      cv.base[q] = blocks[q];
      cv.tab_v[q] = Q->t.fm_v;
      cv.tab_w[q] = Q->t.fm_w;
      cv.tab_g[q] = grad_tables ? static_cast<float*>(grad_tables[q]) : nullptr;
    }
    E->emu = true;
    E->emu_ranks = n;
    E->comm_ready = true;
    all[r] = E->xa;
  }
  Engine* E0 = static_cast<Engine*>(engines[0]);
  int rc = E0->alloc(&E0->emu_args, size_t(n));
  if (rc) return rc;
  DFM_CUDA_CHECK(cudaMemcpy(E0->emu_args, all.data(), sizeof(XchgArgs) * n, cudaMemcpyHostToDevice));
  return DFM_OK;
}

// Call after dfm_engine_train_step of every virtual rank (same stream): the collectives of that step, all ranks.
extern "C" int dfm_engine_emu_collectives(void* const* engines, int n, dfm_stream_t s_) {
  Engine* E0 = static_cast<Engine*>(engines[0]);
  DFM_REQUIRE(E0->emu && E0->emu_args && E0->emu_ranks == n, "dfm_engine_emu_collectives: call dfm_engine_emu_attach first");
  cudaStream_t s = static_cast<cudaStream_t>(s_);
  int rc = launch_dense_sync(E0->xa, E0->emu_args, n, s);
  if (rc) return rc;
  if ((rc = launch_exchange_apply(E0->xa, E0->emu_args, n, s))) return rc;
  if (E0->cfg.table_update_mode == DFM_TABLE_DENSE_EXACT) {
    for (int r = 0; r < n; ++r) {
      Engine* E = static_cast<Engine*>(engines[r]);
      if ((rc = launch_table_dense_from(sparse_args(*E), E->hyper, E->xa.cv.tab_g[r], 1.f / float(n), s))) return rc;
    }
  }
  return DFM_OK;
}

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
    if (!E->ev_done[i]) DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E->ev_done[i], cudaEventDisableTiming));
  }
  if (!E->pf_stream) {
    DFM_CUDA_CHECK(cudaStreamCreateWithFlags(&E->pf_stream, cudaStreamNonBlocking));
    DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E->ev_pf_fork, cudaEventDisableTiming));
    DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E->ev_pf_join, cudaEventDisableTiming));
    DFM_CUDA_CHECK(cudaEventCreateWithFlags(&E->ev_loss, cudaEventDisableTiming));
    DFM_CUDA_CHECK(cudaMallocHost(reinterpret_cast<void**>(&E->dummy_host), E->slab_bytes));
    DFM_CUDA_CHECK(cudaMallocHost(reinterpret_cast<void**>(&E->loss_ring), Engine::kPipe * sizeof(float)));
    memset(E->dummy_host, 0, E->slab_bytes);
    memset(E->loss_ring, 0, Engine::kPipe * sizeof(float));
  }
  for (int i = 0; i < Engine::kPipe; ++i) {
    if (E->graph[i]) { cudaGraphExecDestroy(E->graph[i]); E->graph[i] = nullptr; }
    for (int k = 0; k < 2; ++k) {
      if (E->pf_exec[i][k]) { cudaGraphExecDestroy(E->pf_exec[i][k]); E->pf_exec[i][k] = nullptr; }
      if (E->pf_graph[i][k]) { cudaGraphDestroy(E->pf_graph[i][k]); E->pf_graph[i][k] = nullptr; }
    }
    E->staged_src[i] = nullptr;
    cudaGraph_t g = nullptr;
    uint8_t* other = E->in_slab[1 - i];
    DFM_CUDA_CHECK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    // branch 1: prefetch of the next batch into the other slab (source patched before every launch)
    DFM_CUDA_CHECK(cudaEventRecord(E->ev_pf_fork, s));
    DFM_CUDA_CHECK(cudaStreamWaitEvent(E->pf_stream, E->ev_pf_fork, 0));
    DFM_CUDA_CHECK(cudaMemcpyAsync(other, E->dummy_host, E->slab_bytes, cudaMemcpyHostToDevice, E->pf_stream));
    const int* ids = reinterpret_cast<const int*>(E->in_slab[i]);
    const float* vals = reinterpret_cast<const float*>(E->in_slab[i] + E->off_vals);
    const float* labels = reinterpret_cast<const float*>(E->in_slab[i] + E->off_labels);
    int rc = engine_forward(*E, ids, vals, labels, 1, s);
    // branch 2: the loss sum is final once the loss kernel has run; it leaves for the pinned ring from there
    DFM_CUDA_CHECK(cudaEventRecord(E->ev_loss, s));
    DFM_CUDA_CHECK(cudaStreamWaitEvent(E->pf_stream, E->ev_loss, 0));
    DFM_CUDA_CHECK(cudaMemcpyAsync(E->loss_ring + i, &E->st->loss_sum, sizeof(float), cudaMemcpyDeviceToHost, E->pf_stream));
    DFM_CUDA_CHECK(cudaEventRecord(E->ev_pf_join, E->pf_stream));
    if (!rc) rc = engine_backward_update(*E, ids, vals, s);
    DFM_CUDA_CHECK(cudaStreamWaitEvent(s, E->ev_pf_join, 0));
    cudaError_t ce = cudaStreamEndCapture(s, &g);
    if (rc) { if (g) cudaGraphDestroy(g); return rc; }
    if (ce != cudaSuccess) { set_error("cudaStreamEndCapture: %s", cudaGetErrorString(ce)); return DFM_ERR_CUDA; }

    // find the prefetch node (the memcpy whose destination is the other slab)
    size_t n_nodes = 0;
    DFM_CUDA_CHECK(cudaGraphGetNodes(g, nullptr, &n_nodes));
    std::vector<cudaGraphNode_t> nodes(n_nodes);
    DFM_CUDA_CHECK(cudaGraphGetNodes(g, nodes.data(), &n_nodes));
    cudaGraphNode_t pf = nullptr;
    for (cudaGraphNode_t nd : nodes) {
      cudaGraphNodeType ty;
      DFM_CUDA_CHECK(cudaGraphNodeGetType(nd, &ty));
      if (ty != cudaGraphNodeTypeMemcpy) continue;
      cudaMemcpy3DParms mp{};
      DFM_CUDA_CHECK(cudaGraphMemcpyNodeGetParams(nd, &mp));
      if (mp.dstPtr.ptr == other) pf = nd;
    }
    DFM_REQUIRE(pf != nullptr, "dfm_engine_capture: prefetch node not found in the captured graph");

    // flavour 0: pinned-host source (as captured)
    E->pf_graph[i][Engine::kSrcHost] = g;
    E->pf_node[i][Engine::kSrcHost] = pf;
    DFM_CUDA_CHECK(cudaGraphInstantiate(&E->pf_exec[i][Engine::kSrcHost], g, 0));
    // flavour 1: device source
    cudaGraph_t gd = nullptr;
    cudaGraphNode_t pfd = nullptr;
    DFM_CUDA_CHECK(cudaGraphClone(&gd, g));
    DFM_CUDA_CHECK(cudaGraphNodeFindInClone(&pfd, pf, gd));
    DFM_CUDA_CHECK(cudaGraphMemcpyNodeSetParams1D(pfd, other, E->in_slab[i], E->slab_bytes, cudaMemcpyDeviceToDevice));
    E->pf_graph[i][Engine::kSrcDevice] = gd;
    E->pf_node[i][Engine::kSrcDevice] = pfd;
    DFM_CUDA_CHECK(cudaGraphInstantiate(&E->pf_exec[i][Engine::kSrcDevice], gd, 0));
    // plain flavour: no prefetch
    cudaGraph_t gp = nullptr;
    cudaGraphNode_t pfp = nullptr;
    DFM_CUDA_CHECK(cudaGraphClone(&gp, g));
    DFM_CUDA_CHECK(cudaGraphNodeFindInClone(&pfp, pf, gp));
    DFM_CUDA_CHECK(cudaGraphDestroyNode(pfp));
    ce = cudaGraphInstantiate(&E->graph[i], gp, 0);
    cudaGraphDestroy(gp);
    if (ce != cudaSuccess) { set_error("cudaGraphInstantiate: %s", cudaGetErrorString(ce)); return DFM_ERR_CUDA; }
  }
  E->pipe_step = 0;
  return DFM_OK;
}

// One pipelined step: trains on `cur` and, beside it, copies `next` (pinned host or device memory, or null)
// into the other input slab for the following call.  If `cur` is not what the previous call prefetched it is
// copied first, in stream order (correct, but the copy is then on the critical path).  `*slot` receives the
// index of the pinned loss-ring entry this step writes (read it after the work queued here has finished).
extern "C" int dfm_engine_step_pipelined(void* h, const void* cur, const void* next, int* slot, dfm_stream_t s_) {
  Engine* E = static_cast<Engine*>(h);
  cudaStream_t s = static_cast<cudaStream_t>(s_);
  const int w = int(E->pipe_step % Engine::kPipe);
  DFM_REQUIRE(E->graph[w] != nullptr, "dfm_engine_step_pipelined: capture first");
  DFM_REQUIRE(cur != nullptr, "dfm_engine_step_pipelined: null batch");
  if (E->staged_src[w] != cur)
    DFM_CUDA_CHECK(cudaMemcpyAsync(E->in_slab[w], cur, E->slab_bytes, cudaMemcpyDefault, s));
  E->staged_src[w] = nullptr;
  int kind = -1;
  if (next) {
    cudaPointerAttributes at{};
    DFM_CUDA_CHECK(cudaPointerGetAttributes(&at, next));
    if (at.type == cudaMemoryTypeHost) kind = Engine::kSrcHost;
    else if (at.type == cudaMemoryTypeDevice) kind = Engine::kSrcDevice;
  }
  if (kind < 0) {
    DFM_CUDA_CHECK(cudaGraphLaunch(E->graph[w], s));
  } else {
    DFM_CUDA_CHECK(cudaGraphExecMemcpyNodeSetParams1D(
        E->pf_exec[w][kind], E->pf_node[w][kind], E->in_slab[1 - w], next, E->slab_bytes,
        kind == Engine::kSrcHost ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice));
    DFM_CUDA_CHECK(cudaGraphLaunch(E->pf_exec[w][kind], s));
    E->staged_src[1 - w] = next;
  }
  DFM_CUDA_CHECK(cudaEventRecord(E->ev_done[w], s));
  if (slot) *slot = w;
  E->pipe_step++;
  return DFM_OK;
}

extern "C" int dfm_engine_reset_pipeline(void* h) {
  Engine* E = static_cast<Engine*>(h);
  E->staged_src[0] = E->staged_src[1] = nullptr;
  return DFM_OK;
}

extern "C" int dfm_engine_loss_ring(void* h, float** ring, int* n) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(E->loss_ring != nullptr, "dfm_engine_loss_ring: capture first");
  *ring = E->loss_ring;
  *n = Engine::kPipe;
  return DFM_OK;
}

// Device-resident replay (inputs already in slab `which`).
extern "C" int dfm_engine_replay(void* h, int which, dfm_stream_t s) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(which >= 0 && which < Engine::kPipe && E->graph[which], "dfm_engine_replay: capture first");
  E->staged_src[0] = E->staged_src[1] = nullptr;
  DFM_CUDA_CHECK(cudaGraphLaunch(E->graph[which], static_cast<cudaStream_t>(s)));
  return DFM_OK;
}

extern "C" int dfm_engine_slab_ptr(void* h, int which, void** ptr) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(which >= 0 && which < Engine::kPipe && E->in_slab[which], "dfm_engine_slab_ptr: capture first");
  *ptr = E->in_slab[which];
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
  else if (!strcmp(name, "dense_grad")) { p = E->dP; n = (long long)E->part_stride * E->ksplit; }
  else if (!strcmp(name, "dense_bf16")) { p = E->Wb; n = E->n_dense; }
  else if (!strcmp(name, "G")) { p = E->G; n = (long long)(E->comm_ready ? E->xa.cv.world : 1) * E->cfg.batch_max * E->cfg.field_size * g_pitch(E->cfg.embedding_size); }
  else if (!strcmp(name, "Gw")) { p = E->Gw; n = (long long)E->cfg.batch_max * E->cfg.field_size; }
  else if (!strcmp(name, "uniq_ids")) { p = E->uniq_ids; n = (long long)(E->comm_ready ? E->xa.cv.world : 1) * E->cfg.batch_max * E->cfg.field_size; }
  else if (!strcmp(name, "slot_of")) { p = E->slot_of; n = E->cfg.feature_size; }
  else if (!strcmp(name, "kern_dbg")) { p = E->kern_dbg; n = 1024; }
  else if (!strcmp(name, "comm_dbg")) { p = E->comm_dbg; n = 64; }   /* 32 x u64 phase timestamps, as int32 pairs */
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

// offsets of [gamma | beta] of layer l inside the dense vector and of [moving_mean | moving_variance] inside bn_moving
extern "C" int dfm_engine_bn_layout(void* h, long long* dense_offsets, long long* moving_offsets) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(E->bn, "dfm_engine_bn_layout: the engine was created without batch_norm");
  for (int l = 0; l < E->sh.L; ++l) {
    dense_offsets[l] = (long long)E->off_bn[l];
    moving_offsets[l] = (long long)E->off_mov[l];
  }
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

// ---- plain device allocations + CUDA IPC (fallback symmetric-memory provider) ----------------
extern "C" int dfm_dev_alloc(long long bytes, void** ptr) {
  DFM_REQUIRE(bytes > 0 && ptr, "dfm_dev_alloc: bad arguments");
  void* p = nullptr;
  DFM_CUDA_CHECK(cudaMalloc(&p, size_t(bytes)));
  DFM_CUDA_CHECK(cudaMemset(p, 0, size_t(bytes)));
  *ptr = p;
  return DFM_OK;
}

extern "C" int dfm_dev_free(void* ptr) {
  if (ptr) DFM_CUDA_CHECK(cudaFree(ptr));
  return DFM_OK;
}

extern "C" int dfm_ipc_export(const void* ptr, void* handle_out) {
  DFM_REQUIRE(ptr && handle_out, "dfm_ipc_export: null argument");
  cudaIpcMemHandle_t hd;
  static_assert(sizeof(hd) == 64, "IPC handle size");
  DFM_CUDA_CHECK(cudaIpcGetMemHandle(&hd, const_cast<void*>(ptr)));
  memcpy(handle_out, &hd, 64);
  return DFM_OK;
}

// The mapping stays open for the life of the process (the engines that use it do not own it).
extern "C" int dfm_ipc_open(const void* handle, void** ptr) {
  DFM_REQUIRE(handle && ptr, "dfm_ipc_open: null argument");
  cudaIpcMemHandle_t hd;
  memcpy(&hd, handle, 64);
  void* p = nullptr;
  DFM_CUDA_CHECK(cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
  *ptr = p;
  return DFM_OK;
}

// ---- the sparse exchange on its own (collective sweep / profiling) ---------------------------
namespace dfm {
__global__ void open_exchange_kernel(StepState* st, int n) {
  st->count = n;                 // dedupe walks count * F positions (F = 1 here)
  st->step += 1;
  st->comm_epoch += 1;
  open_step_consts(st);
  for (int o = 0; o < 8; ++o) st->n_seg[o] = 0;
}
__global__ void fill_segments_kernel(const StepState* st, float* G, int cap, int gp, int Ke, int world) {
  for (int o = 0; o < world; ++o)
    for (long long i = blockIdx.x * blockDim.x + threadIdx.x; i < (long long)st->n_seg[o] * (Ke + 1); i += (long long)gridDim.x * blockDim.x) {
      const long long s = i / (Ke + 1);
      G[(size_t(o) * cap + s) * gp + (i - s * (Ke + 1))] = 1.f;
    }
}
}  // namespace dfm

extern "C" int dfm_engine_exchange_prepare(void* h, const int* ids, int n, dfm_stream_t s_) {
  Engine* E = static_cast<Engine*>(h);
  cudaStream_t s = static_cast<cudaStream_t>(s_);
  DFM_REQUIRE(E->comm_ready && !E->emu, "dfm_engine_exchange_prepare: attach the multi-GPU views first");
  const int cap = E->cfg.batch_max * E->cfg.field_size;
  DFM_REQUIRE(ids && n > 0 && n <= cap, "dfm_engine_exchange_prepare: n %d outside (0, %d]", n, cap);
  open_exchange_kernel<<<1, 1, 0, s>>>(E->st, n);
  DFM_LAUNCH_CHECK();
  int rc = launch_dedupe(ids, 1, E->st, E->slot_of, E->uniq_ids, n, E->xa.cv.world, cap, E->cfg.feature_size, s);
  if (rc) return rc;
  fill_segments_kernel<<<296, 256, 0, s>>>(E->st, E->G, cap, g_pitch(E->cfg.embedding_size), E->cfg.embedding_size, E->xa.cv.world);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

extern "C" int dfm_engine_exchange_run(void* h, dfm_stream_t s_) {
  Engine* E = static_cast<Engine*>(h);
  DFM_REQUIRE(E->comm_ready && !E->emu, "dfm_engine_exchange_run: attach the multi-GPU views first");
  return launch_exchange_apply(E->xa, nullptr, 0, static_cast<cudaStream_t>(s_));
}
