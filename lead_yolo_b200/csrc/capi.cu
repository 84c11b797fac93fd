// Library-wide plumbing of libleadyolo.so: error text, device check, TMA descriptor encoding.
#include <cuda.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "ly_common.cuh"
#include "ly_tmap.h"

namespace ly {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

bool pdl_enabled() {
  static const bool on = getenv("LY_NO_PDL") == nullptr;   // dev knob: plain stream-ordered launches
  return on;
}

int device_sm_count() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}

// cuTensorMapEncodeTiled is a driver entry point; resolve it through the runtime so the library
// links against libcudart only (libcuda.so is absent on build machines without a GPU).
typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn g_encode = nullptr;
static std::once_flag g_encode_once;

static void resolve_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
      q == cudaDriverEntryPointSuccess)
    g_encode = (encode_tiled_fn)fn;
}

int tmap_encode_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box) {
  std::call_once(g_encode_once, resolve_encode);
  if (!g_encode) return fail(LY_EUNSUPPORTED, "cuTensorMapEncodeTiled not available from the CUDA driver");
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0)
    return fail(LY_EINVAL, "TMA source %p is not 16-byte aligned", base);
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bdim[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
    if (dims[i] == 0) return fail(LY_EINVAL, "TMA dim %d is zero", i);
    if (box[i] == 0 || box[i] > 256) return fail(LY_EINVAL, "TMA box %d = %u out of range", i, box[i]);
  }
  for (int i = 0; i + 1 < rank; ++i) {
    gstr[i] = strides_bytes[i];
    if (strides_bytes[i] % 16 != 0)
      return fail(LY_EINVAL, "TMA stride %d = %llu bytes is not a multiple of 16", i,
                  (unsigned long long)strides_bytes[i]);
  }
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim,
                        gstr, bdim, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(LY_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return LY_OK;
}

}  // namespace ly

extern "C" {

const char* ly_last_error(void) { return ly::g_err; }

int ly_version(void) { return 100; }

int ly_device_check(void) {
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  int major = 0;
  LY_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10)
    return ly::fail(LY_EUNSUPPORTED, "device %d has compute capability %d.x; libleadyolo needs sm_100 (B200)", dev,
                    major);
  return LY_OK;
}

}  // extern "C"
