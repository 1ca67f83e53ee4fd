/*
 * deepfm_b200.h -- C ABI of libdfm_cuda.so, the sm_100a kernel library of the
 * DeepFM framework.  Every entry point
 *   - takes raw device pointers (the Python layer passes tensor.data_ptr()),
 *   - enqueues work on the given stream and returns without synchronising,
 *   - returns DFM_OK or a negative DFM_ERR_* code; dfm_last_error() gives the
 *     text for the calling host thread.
 *
 * Citations of the form HVD.py:NNN point at
 * /root/reference/2-hvd-gpu/DeepFM-hvd-tfrecord-vectorized-map.py, the TF1
 * script whose graph these kernels replace.
 */
#ifndef DEEPFM_B200_H_
#define DEEPFM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DFM_OK 0
#define DFM_ERR_ARG (-1)
#define DFM_ERR_CUDA (-2)
#define DFM_ERR_UNSUPPORTED (-3)

typedef void* dfm_stream_t; /* cudaStream_t */

const char* dfm_last_error(void);
int dfm_version(void);

/* ------------------------------------------------------------------------
 * Optimizers (tf.train.*Optimizer of HVD.py:281-294) and table update modes
 * ---------------------------------------------------------------------- */
#define DFM_OPT_ADAM 0
#define DFM_OPT_ADAGRAD 1
#define DFM_OPT_MOMENTUM 2
#define DFM_OPT_FTRL 3
#define DFM_OPT_GD 4

#define DFM_TABLE_LAZY 0        /* touched rows only (sparse-gradient path)            */
#define DFM_TABLE_DENSE_EXACT 1 /* every row every step, g = sparse + l2*row (HVD.py:266-267) */

#define DFM_LOSS_LOG 0
#define DFM_LOSS_SQUARE 1

#define DFM_MAX_LAYERS 6

typedef struct {
  int32_t batch_max;       /* per-GPU batch capacity                                       */
  int32_t field_size;      /* F   (--field_size)                                            */
  int32_t embedding_size;  /* K   (--embedding_size), multiple of 16, <= 128                */
  int32_t n_layers;        /* hidden layers of the deep tower (--deep_layers)               */
  int32_t layers[DFM_MAX_LAYERS];   /* widths, multiples of 16, <= 512                      */
  float keep_prob[DFM_MAX_LAYERS];  /* --dropout values are KEEP probabilities (HVD.py:235) */
  int64_t feature_size;    /* V   (--feature_size), rows of fm_w / fm_v                     */
  int32_t optimizer;       /* DFM_OPT_*                                                     */
  int32_t table_update_mode; /* DFM_TABLE_*                                                 */
  float lr;                /* already scaled by world size (HVD.py:171)                     */
  float l2_reg;            /* applies to fm_w and fm_v only (HVD.py:266-267)                */
  float beta1, beta2, eps; /* Adam                                                          */
  float momentum;          /* Momentum                                                      */
  float ftrl_lr_power, ftrl_l1, ftrl_l2;
  uint64_t seed;           /* dropout stream                                                */
  int32_t batch_norm;      /* --batch_norm: FC -> ReLU -> BN -> dropout (HVD.py:224-235,306-328) */
  int32_t loss_type;       /* DFM_LOSS_LOG (HVD.py:264) or DFM_LOSS_SQUARE: mean (sigmoid(y) - label)^2 */
  float bn_decay;          /* --batch_norm_decay, moving-average decay                      */
  float bn_eps;            /* 1e-3, tf.contrib.layers.batch_norm default                    */
} dfm_engine_config;

/* Caller-owned fp32 device tensors (parameters and their optimizer slots).
 * dense* are flat vectors laid out as
 *   [P_0 | P_1 | ... | P_{L-1} | P_out | fm_bias],  P_l = [K_l + 1, N_l] row-major,
 * rows 0..K_l-1 = Deep-part/mlp{l}/weights, row K_l = .../biases (P_out: deep_out).
 * With batch_norm the vector is [P_0 | ... | P_{L-1} | BN_0 | ... | BN_{L-1} | P_out | fm_bias] with
 * BN_l = [gamma (N_l) | beta (N_l)] (Deep-part/bn_{l}/gamma, beta), and bn_moving holds, per layer,
 * [moving_mean (N_l) | moving_variance (N_l)]; dfm_engine_layout reports the offsets. */
typedef struct {
  float* fm_w;    float* fm_w_s0;  float* fm_w_s1;    /* [V]        */
  float* fm_v;    float* fm_v_s0;  float* fm_v_s1;    /* [V, K]     */
  float* dense;   float* dense_s0; float* dense_s1;   /* [n_dense]  */
  float* bn_moving;                                   /* [2 * sum N_l] or NULL without batch_norm */
} dfm_engine_tensors;

int dfm_engine_create(const dfm_engine_config* cfg, const dfm_engine_tensors* tensors, void** engine);
int dfm_engine_destroy(void* engine);
/* offsets[L+2] (P_0..P_out, fm_bias) in elements, pitches[L+1] of the activation buffers. */
int dfm_engine_layout(void* engine, long long* offsets, int* pitches, int* n_dense, int* batch_pad);
/* batch_norm: per hidden layer, offset of [gamma | beta] in `dense` and of [moving_mean | moving_variance] in bn_moving. */
int dfm_engine_bn_layout(void* engine, long long* dense_offsets, long long* moving_offsets);
/* Step scalars live on the device (graph-replay safe): valid rows, lr, train flag;
 * step >= 0 also resets the 1-based optimizer step counter. */
int dfm_engine_set_state(void* engine, int count, float lr, int train, int step, dfm_stream_t stream);
/* Refresh the bf16 tensor-core copy after the caller changed `dense` (init, restore, broadcast). */
int dfm_engine_sync_weights(void* engine, dfm_stream_t stream);
/* Forward only: embed+FM kernel, fused MLP kernel.  ids int32 [B,F], vals f32 [B,F],
 * labels f32 [B] or NULL.  Results in the "prob"/"logit" buffers, loss sum in "state". */
int dfm_engine_forward(void* engine, const int* ids, const float* vals, const float* labels,
                       dfm_stream_t stream);
/* Forward + backward + optimizer update of every parameter (one reference train_op). */
int dfm_engine_train_step(void* engine, const int* ids, const float* vals, const float* labels,
                          dfm_stream_t stream);
int dfm_engine_launches_per_step(void* engine);

/* CUDA-graph replay and the double-buffered input pipeline.
 * Input slab (host and device): ids int32 [batch_max*F] | vals f32 [batch_max*F] | labels f32 [batch_max]. */
int dfm_engine_slab_layout(void* engine, long long* bytes, long long* off_vals, long long* off_labels);
/* Records one train step per device input slab (2 slabs) on `stream` (not the legacy default stream). */
int dfm_engine_capture(void* engine, dfm_stream_t stream);
int dfm_engine_slab_ptr(void* engine, int which, void** ptr);
int dfm_engine_replay(void* engine, int which, dfm_stream_t stream);
/* Pipelined step: trains on `cur` (a full batch slab, pinned host or device memory) and copies `next` (same
 * kinds, or NULL) into the other input slab beside the kernels, as a node of the same graph; the next call
 * should then pass that pointer as `cur`.  A `cur` that was not prefetched is copied in stream order first.
 * `next` must stay unchanged until the work queued by this call has finished.  *slot = entry of the pinned
 * loss ring (dfm_engine_loss_ring) that receives this step's loss sum. */
int dfm_engine_step_pipelined(void* engine, const void* cur, const void* next, int* slot, dfm_stream_t stream);
int dfm_engine_loss_ring(void* engine, float** ring, int* n_slots);
/* Forget which source pointers are staged (call when slab addresses may be reused with new contents). */
int dfm_engine_reset_pipeline(void* engine);
/* Device pointer + element count of an internal buffer:
 * "prob","logit","dlogit","fm_part","s_sum","state","dense_grad","dense_bf16","G","Gw",
 * "uniq_ids","slot_of", and "act"/"dz" with a layer index. */
int dfm_engine_buffer(void* engine, const char* name, int index, void** ptr, long long* n_elems);

/* Warm CUDA-event timings (ms, means over `iters` real training steps): ms_out[0] forward,
 * ms_out[1] backward + update, ms_out[2] whole step.  Single-GPU only.  Synchronises. */
int dfm_engine_profile(void* engine, const int* ids, const float* vals, const float* labels, int iters,
                       float* ms_out);

/* Multi-GPU (synchronous data parallel, hvd.DistributedOptimizer, HVD.py:296) without NCCL on the step path.
 * The caller owns the symmetric memory: on every rank it allocates one zero-filled block of
 * dfm_engine_comm_bytes() bytes, makes every rank's block and every rank's fm_v / fm_w (the tensors given
 * to dfm_engine_create) addressable from every rank -- CUDA IPC, or cuMem / torch symmetric memory, which
 * also yields one NVLS multicast address per buffer -- and passes the pointers to comm_attach:
 *   *_ptrs  `world` device pointers, entry `rank` being this rank's own buffer;
 *   *_mc    multicast address of the same buffers, or NULL (the kernels then loop over the peers);
 *   grad_*  only for table_update_mode DENSE_EXACT: zero-filled fp32 tables [feature_size, embedding_size + 8].
 * From then on train_step runs the dense gradient all-reduce + optimizer as one kernel beside the embedding
 * backward (multimem.ld_reduce / multimem.st where a multicast address is given) and exchanges the sparse
 * embedding gradients owner-sharded: row id belongs to rank id % world, which sums the ranks' gradients,
 * applies the lazy update to its copy of the optimizer slots and stores the new row into every replica
 * (csrc/cuda/comm.cu).  Optimizer slots of a table row are therefore current on its owner only.
 * All ranks must call train_step the same number of times; a rank that waits longer than
 * DFM_COMM_TIMEOUT_S (default 120) seconds for a peer traps.
 * dfm_engine_comm_multicast: bit 0 = dense all-reduce uses NVLS, bit 1 = table rows are stored through NVLS. */
int dfm_engine_comm_bytes(void* engine, int world, long long* bytes);
int dfm_engine_comm_attach(void* engine, int world, int rank, void* const* block_ptrs, void* block_mc,
                           void* const* fm_v_ptrs, void* fm_v_mc, void* const* fm_w_ptrs, void* fm_w_mc,
                           void* const* grad_ptrs, void* grad_mc);
int dfm_engine_comm_multicast(void* engine);
/* Single-GPU emulation of `n` ranks, for tests and profiling on a one-GPU box: `n` engines created on the
 * same device act as the ranks.  After emu_attach each engine's train_step stops before the collectives;
 * once every engine has stepped, emu_collectives launches the two collective kernels ONCE, cooperatively, over
 * all virtual ranks (separate launches that wait on one another on one GPU are not guaranteed to co-run). */
int dfm_engine_emu_attach(void* const* engines, int n, void* const* grad_tables);
int dfm_engine_emu_collectives(void* const* engines, int n, dfm_stream_t stream);
/* Stand-alone sum all-reduce of n fp32 values (n % 4 == 0) held in symmetric memory, in place in every rank's copy:
 * the dense half of the step without the optimizer (ncclAllReduce's role in Horovod's fusion buffer, HVD.py:296).
 * data_ptrs / flag_ptrs: `world` peer-mapped pointers each; data_mc: NVLS multicast alias of the copies or NULL;
 * flags: 2 * world * 160 zero-initialised uint32 per rank; epoch = 1, 2, 3, ... per call, the same on every rank. */
int dfm_allreduce_f32(int world, int rank, void* const* data_ptrs, void* data_mc, void* const* flag_ptrs,
                      long long n, unsigned int epoch, dfm_stream_t stream);
/* The sparse exchange + owner-sharded lazy update of a multi-GPU engine on its own (collective sweep, profiling):
 * prepare de-duplicates `n` ids (n <= batch_max * field_size) and gives every distinct id a gradient row of ones;
 * run launches the exchange kernel for them.  Both are collective: every rank calls them the same number of times. */
int dfm_engine_exchange_prepare(void* engine, const int* ids, int n, dfm_stream_t stream);
int dfm_engine_exchange_run(void* engine, dfm_stream_t stream);
/* Plain device allocations with CUDA IPC export / import (the fallback symmetric-memory provider). */
int dfm_dev_alloc(long long bytes, void** ptr);            /* cudaMalloc + zero fill */
int dfm_dev_free(void* ptr);
int dfm_ipc_export(const void* ptr, void* handle_out /* 64 bytes */);
int dfm_ipc_open(const void* handle /* 64 bytes */, void** ptr);

/* ------------------------------------------------------------------------
 * De-duplication of a batch's feature ids (the step that turns B*F row gradients into n_unique
 * rows before they are communicated or applied; TF's IndexedSlices path has no such step,
 * HVD.py:197,201 + SURVEY K13).  ids int32 [n] in [0, V).  Outputs: uniq_out [<= n] distinct ids,
 * pos_slot_out [n] with uniq_out[pos_slot_out[i]] == ids[i], count_out [1].
 * `workspace` is caller-owned scratch of dfm_unique_slots_workspace_bytes(V) bytes, initialised once
 * with dfm_unique_slots_init; every call leaves it ready for the next one.
 * ---------------------------------------------------------------------- */
long long dfm_unique_slots_workspace_bytes(long long V);
int dfm_unique_slots_init(void* workspace, long long V, dfm_stream_t stream);
int dfm_unique_slots(const int* ids, int n, long long V, void* workspace, int* uniq_out, int* pos_slot_out,
                     int* count_out, dfm_stream_t stream);

/* dfm_unique_slots without a workspace argument and without any device allocation, for callers that
 * may not allocate GPU memory: plug-ins living in a process whose GPU memory is already owned by
 * another framework's allocator, or loops that cannot take the device-wide synchronisation of
 * cudaMalloc/cudaFree.  (No counterpart in the reference: TensorFlow allocates for its ops.)
 * The scratch is a 4 MiB array compiled into the library and resident on every GPU that loads it,
 * which caps the id space at DFM_UNIQUE_STATIC_MAX_V; a larger V is rejected with an error.
 * The first call on each device looks the array up and fills it with -1 on `stream`: make that call
 * before any stream capture.  That set-up is guarded by a host mutex, and later calls on other streams or
 * host threads are ordered behind the fill; a call that fails midway makes the next call fill the map again.
 * All calls on a device share the array: calls issued back to back on one stream are fine, calls whose
 * kernels overlap on the GPU (different streams with no ordering between them) race on it.
 * Same inputs/outputs as dfm_unique_slots. */
#define DFM_UNIQUE_STATIC_MAX_V (1 << 20)
long long dfm_unique_slots_static_max_v(void);
int dfm_unique_slots_static(const int* ids, int n, long long V, int* uniq_out, int* pos_slot_out,
                            int* count_out, dfm_stream_t stream);

/* tf.metrics.auc accumulators (HVD.py:271): hist[2][num_thresholds+1] uint64, row 0 positives. */
int dfm_auc_hist(const float* prob, const float* label, int n, unsigned long long* hist,
                 int num_thresholds, dfm_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DEEPFM_B200_H_ */
