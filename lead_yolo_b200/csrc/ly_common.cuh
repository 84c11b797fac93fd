// Shared host/device helpers for libleadyolo.so.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/leadyolo.h"

namespace ly {

// thread-local error text (set by ly_fail, read by ly_last_error)
void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);

#define LY_REQUIRE(cond, ...)                                   \
  do {                                                          \
    if (!(cond)) return ::ly::fail(LY_EINVAL, __VA_ARGS__);     \
  } while (0)

#define LY_CUDA(...)                                                                                     \
  do {                                                                                                   \
    cudaError_t e__ = (__VA_ARGS__);                                                                     \
    if (e__ != cudaSuccess)                                                                              \
      return ::ly::fail(LY_ECUDA, "%s failed: %s (%s:%d)", #__VA_ARGS__, cudaGetErrorString(e__), __FILE__, \
                        __LINE__);                                                                       \
  } while (0)

#define LY_LAUNCH_CHECK() LY_CUDA(cudaGetLastError())

int device_sm_count();
// grid size (blocks of `block` threads) whose total thread count is a multiple of V (see elementwise.cu)
int flat_grid(long M, int V, int block);

static inline int ceil_div(long a, long b) { return (int)((a + b - 1) / b); }

// Programmatic dependent launch.  A kernel started through ly::launch may become resident while the kernel before it
// in the stream is still draining its last wave: its launch latency and its prologue (barrier init, TMEM allocation,
// descriptor prefetch) then overlap that tail.  Contract for every kernel launched this way:
//   * pdl_wait() before the first access to memory another kernel may have written (or may still read) — it returns
//     when the preceding grid has completed and its writes are visible;
//   * pdl_trigger() once the kernel holds its TMEM (a dependant that allocated first would starve it), else first thing.
// Both are no-ops for a launch without the attribute (LY_NO_PDL=1, or <<<>>>).
bool pdl_enabled();

#if defined(__CUDACC__)
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() {
  pdl_trigger();
  pdl_wait();
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                          Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<Args&&>(args)...);
}
// sigmoid through the hardware tanh: one MUFU op instead of ex2 + rcp (relative error ~2^-11, below bf16)
__device__ __forceinline__ float fast_sigmoid(float v) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * v));
  return fmaf(0.5f, t, 0.5f);
}
template <int ACT>
__device__ __forceinline__ float act_apply_t(float v) {
  if (ACT == LY_ACT_RELU) return fmaxf(v, 0.f);
  if (ACT == LY_ACT_SILU) return v * fast_sigmoid(v);
  if (ACT == LY_ACT_HSWISH) return v * fminf(fmaxf(v + 3.f, 0.f), 6.f) * (1.f / 6.f);
  if (ACT == LY_ACT_SIGMOID) return fast_sigmoid(v);
  return v;
}
__device__ __forceinline__ float act_apply(float v, int act) {
  switch (act) {
    case LY_ACT_RELU: return act_apply_t<LY_ACT_RELU>(v);
    case LY_ACT_SILU: return act_apply_t<LY_ACT_SILU>(v);
    case LY_ACT_HSWISH: return act_apply_t<LY_ACT_HSWISH>(v);
    case LY_ACT_SIGMOID: return act_apply_t<LY_ACT_SIGMOID>(v);
    default: return v;
  }
}
// d act(v) / dv
__device__ __forceinline__ float act_grad(float v, int act) {
  switch (act) {
    case LY_ACT_RELU: return v > 0.f ? 1.f : 0.f;
    case LY_ACT_SILU: {
      float s = fast_sigmoid(v);
      return s * (1.f + v * (1.f - s));
    }
    case LY_ACT_HSWISH: {
      float r = fminf(fmaxf(v + 3.f, 0.f), 6.f) * (1.f / 6.f);
      return r + ((v > -3.f && v < 3.f) ? v * (1.f / 6.f) : 0.f);
    }
    case LY_ACT_SIGMOID: {
      float s = fast_sigmoid(v);
      return s * (1.f - s);
    }
    default: return 1.f;
  }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}
// 8 bf16 <-> 8 floats through one 16-byte vector
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
  return u;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
#endif

}  // namespace ly
