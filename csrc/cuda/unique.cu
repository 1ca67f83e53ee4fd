// dfm_unique_slots: the de-duplication step of the sparse-gradient path as a standalone op.
//
// Given n ids in [0, V) it returns the distinct ids (in claim order), the compact slot of every
// input position (uniq[pos_slot[i]] == ids[i]) and the number of distinct ids.  It is what lets the
// embedding backward reduce row gradients into a dense [n_unique, K] buffer before anything is
// communicated or applied -- the dedupe TensorFlow's IndexedSlices path never did (SURVEY 2.7 K13,
// HVD.py:197,201).  The engine runs the same claim kernel (dedupe_kernel, embed_fm.cu) on its side
// stream; this entry point exposes it for tests and for callers that build their own step.
//
// Scratch: a workspace of (kUniqueHeaderInts + V) ints owned by the caller: a small header (the
// StepState the claim kernel counts in) followed by the id -> slot map, which must read -1 everywhere
// on entry and is restored to -1 on exit (only touched entries are rewritten).
#include "unique.cuh"

namespace dfm {

// pos_slot[i] = slot of ids[i]
__global__ void __launch_bounds__(256)
unique_lookup_kernel(const int* __restrict__ ids, int n, int V, const int* __restrict__ slot_of,
                     int* __restrict__ pos_slot) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pos_slot[i] = slot_index(slot_of[min(max(__ldg(ids + i), 0), V - 1)]);   // same clamp as the claim kernel
}

// restores the map and publishes the count; one thread per distinct id
__global__ void __launch_bounds__(256)
unique_reset_kernel(const StepState* __restrict__ st, const int* __restrict__ uniq, int* __restrict__ slot_of,
                    int* __restrict__ count_out) {
  const int n = st->n_unique;
  for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < n; s += gridDim.x * blockDim.x) slot_of[uniq[s]] = -1;
  if (blockIdx.x == 0 && threadIdx.x == 0) count_out[0] = n;
}

int unique_workspace_init(void* workspace, long long V, cudaStream_t stream) {
  int* ws = static_cast<int*>(workspace);
  DFM_CUDA_CHECK(cudaMemsetAsync(ws, 0, kUniqueHeaderInts * sizeof(int), stream));
  DFM_CUDA_CHECK(cudaMemsetAsync(ws + kUniqueHeaderInts, 0xFF, size_t(V) * sizeof(int), stream));
  return DFM_OK;
}

// The three launches shared by the caller-workspace and the static-workspace entry points.
int unique_slots_run(const int* ids, int n, long long V, void* workspace, int* uniq_out, int* pos_slot_out,
                     int* count_out, cudaStream_t stream) {
  int* ws = static_cast<int*>(workspace);
  StepState* st = reinterpret_cast<StepState*>(ws);
  int* slot_of = ws + kUniqueHeaderInts;
  int rc;
  // count = n (one "field" per "example"), n_unique = 0
  if ((rc = launch_set_state(st, n, 0.f, 0, -1, stream))) return rc;
  DFM_CUDA_CHECK(cudaMemsetAsync(&st->n_unique, 0, sizeof(int), stream));
  if ((rc = launch_dedupe(ids, 1, st, slot_of, uniq_out, n, 1, n, V, stream))) return rc;
  unique_lookup_kernel<<<ceil_div(n, 256), 256, 0, stream>>>(ids, n, int(V), slot_of, pos_slot_out);
  DFM_LAUNCH_CHECK();
  unique_reset_kernel<<<std::max(1, std::min(592, ceil_div(n, 256))), 256, 0, stream>>>(st, uniq_out, slot_of, count_out);
  DFM_LAUNCH_CHECK();
  return DFM_OK;
}

}  // namespace dfm

using namespace dfm;

extern "C" long long dfm_unique_slots_workspace_bytes(long long V) {
  return (kUniqueHeaderInts + V) * (long long)sizeof(int);
}

extern "C" int dfm_unique_slots_init(void* workspace, long long V, dfm_stream_t stream) {
  DFM_REQUIRE(workspace && V > 0, "dfm_unique_slots_init: bad arguments");
  return unique_workspace_init(workspace, V, static_cast<cudaStream_t>(stream));
}

extern "C" int dfm_unique_slots(const int* ids, int n, long long V, void* workspace, int* uniq_out,
                                int* pos_slot_out, int* count_out, dfm_stream_t stream) {
  DFM_REQUIRE(ids && workspace && uniq_out && pos_slot_out && count_out, "dfm_unique_slots: null argument");
  DFM_REQUIRE(n > 0 && V > 0, "dfm_unique_slots: n and V must be positive (n=%d, V=%lld)", n, V);
  return unique_slots_run(ids, n, V, workspace, uniq_out, pos_slot_out, count_out, static_cast<cudaStream_t>(stream));
}
