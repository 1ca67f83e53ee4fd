// Error text, version, and the TMA tensor-map encoder (driver entry point is
// resolved at run time so the library links against cudart only and can be
// built on a machine without a GPU driver).
#include "common.cuh"

#include <cudaTypedefs.h>

namespace dfm {

bool pdl_enabled(int which) {
  static const int mask = [] {
    const char* v = getenv("DFM_PDL");
    return v ? atoi(v) : 0xff;
  }();
  return (mask & which) != 0;
}


static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}

int make_tmap_bf16_2d(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer,
                      uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_outer) {
  auto fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return DFM_ERR_CUDA;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (pitch_bytes & 15u) || box_inner * 2 != 128 ||
      box_outer == 0 || box_outer > 256) {
    set_error("make_tmap_bf16_2d: bad alignment/box (base %p pitch %llu box %ux%u)", base,
              (unsigned long long)pitch_bytes, box_inner, box_outer);
    return DFM_ERR_ARG;
  }
  cuuint64_t gdim[2] = {inner, outer};
  cuuint64_t gstride[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (inner %llu outer %llu pitch %llu box %ux%u)",
              int(r), (unsigned long long)inner, (unsigned long long)outer,
              (unsigned long long)pitch_bytes, box_inner, box_outer);
    return DFM_ERR_CUDA;
  }
  return DFM_OK;
}

}  // namespace dfm

extern "C" const char* dfm_last_error(void) { return dfm::g_err; }
extern "C" int dfm_version(void) { return 100; }
