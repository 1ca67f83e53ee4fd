// dfm_unique_slots_static: dfm_unique_slots for callers that must not allocate device memory.
//
// Who that is: code that runs inside a process whose GPU memory already belongs to someone else's
// allocator -- a preprocessing or serving plug-in next to a framework that has reserved the whole
// device for its caching pool, where a cudaMalloc from the outside can fail or carve the pool up --
// and host loops that cannot afford the device-wide synchronisation cudaMalloc / cudaFree imply.
// Such a caller can still hand in its input and output buffers (it got those from its host
// framework), but it has nowhere to put the id -> slot map the op needs as scratch.
//
// So this variant brings its own scratch: a file-scope __device__ array, 4 MiB + 512 bytes, which
// every GPU that loads libdfm_cuda.so holds from load until the process exits.  That fixes the
// largest id space it can serve: feature_size <= DFM_UNIQUE_STATIC_MAX_V (1,048,576).  Larger
// tables need dfm_unique_slots with a caller-owned workspace.
//
// Set-up happens inside the first call on each device: the array's device address is looked up with
// cudaGetSymbolAddress (in host code the array's name is only a stand-in, not an address a kernel
// can use) and the map is filled with -1 on the caller's stream.  Make that first call outside any
// stream capture.  The set-up is serialised by a host mutex per device, and an event recorded behind
// the fill is waited for by every later call, so a call on another stream (or from another host
// thread) can never run ahead of the fill.  If a call fails after it started using the map, the next
// call fills the map again.  No call here -- first or later -- uses cudaMalloc, cudaFree or any
// allocating library routine (cudaEventCreate allocates no device memory).
//
// All calls on a device share the one array.  A call leaves the map at -1 again (it rewrites only
// the entries it touched) and clears the counter it uses before counting, so back-to-back calls on
// one stream are fine; two calls whose kernels overlap on the GPU (different streams, no ordering
// between them) race on the map and the counter and give wrong answers.  Ids outside [0, V) are
// treated as the nearest valid id, so a bad id never reaches memory outside the array.
#include <mutex>

#include "unique.cuh"

namespace dfm {

constexpr long long kUniqueStaticMaxV = DFM_UNIQUE_STATIC_MAX_V;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) int g_unique_scratch[kUniqueHeaderInts + kUniqueStaticMaxV];

// Per-device state, guarded by `mu`: device address of g_unique_scratch, whether the map holds -1 everywhere,
// and the event recorded behind the fill that made it so.
struct StaticScratch {
  std::mutex mu;
  void* ptr = nullptr;
  bool clean = false;
  bool fill_seen_done = false;   // the fill's event has been observed complete: later calls need not wait for it
  cudaEvent_t filled = nullptr;
};
static StaticScratch g_scratch[kMaxDevices];

}  // namespace dfm

using namespace dfm;

extern "C" long long dfm_unique_slots_static_max_v(void) { return kUniqueStaticMaxV; }

extern "C" int dfm_unique_slots_static(const int* ids, int n, long long V, int* uniq_out, int* pos_slot_out,
                                       int* count_out, dfm_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  DFM_REQUIRE(ids && uniq_out && pos_slot_out && count_out, "dfm_unique_slots_static: null argument");
  DFM_REQUIRE(n > 0 && V > 0, "dfm_unique_slots_static: n and V must be positive (n=%d, V=%lld)", n, V);
  DFM_REQUIRE(V <= kUniqueStaticMaxV,
              "dfm_unique_slots_static: feature_size %lld exceeds the built-in scratch (max %lld ids); "
              "use dfm_unique_slots with a caller-owned workspace",
              V, kUniqueStaticMaxV);
  int dev = -1;
  DFM_CUDA_CHECK(cudaGetDevice(&dev));
  DFM_REQUIRE(dev >= 0 && dev < kMaxDevices, "dfm_unique_slots_static: device ordinal %d not supported", dev);
  StaticScratch& sc = g_scratch[dev];
  std::lock_guard<std::mutex> lock(sc.mu);      // host-side only: enqueueing is serialised, the GPU work is not
  if (!sc.ptr) {
    void* p = nullptr;
    DFM_CUDA_CHECK(cudaGetSymbolAddress(&p, g_unique_scratch));
    DFM_CUDA_CHECK(cudaEventCreateWithFlags(&sc.filled, cudaEventDisableTiming));
    sc.ptr = p;
  }
  if (!sc.clean) {
    int rc = unique_workspace_init(sc.ptr, kUniqueStaticMaxV, stream);
    if (rc) return rc;
    DFM_CUDA_CHECK(cudaEventRecord(sc.filled, stream));
    sc.clean = true;
    sc.fill_seen_done = false;
  }
  if (!sc.fill_seen_done) {                      // the fill may have been queued on another stream: order behind it
    if (cudaEventQuery(sc.filled) == cudaSuccess) sc.fill_seen_done = true;
    else DFM_CUDA_CHECK(cudaStreamWaitEvent(stream, sc.filled, 0));
  }
  sc.clean = false;                              // until the three launches are queued the map cannot be trusted
  int rc = unique_slots_run(ids, n, V, sc.ptr, uniq_out, pos_slot_out, count_out, stream);
  if (rc) return rc;                             // next call fills the map again
  sc.clean = true;
  return DFM_OK;
}
