// Optimiser step of the training loop on flat fp32 buffers: gradient-norm clipping, SGD with Nesterov momentum and
// the exponential moving average of the weights in ONE pass over the parameters — reference train.py:330-341
// (unscale, clip_grad_norm_(max_norm=10), optimizer.step, ema.update), utils/torch_utils.py:318-346 (smart_optimizer:
// torch.optim.SGD(momentum, nesterov=True) with three parameter groups: weights with decay | BatchNorm weights |
// biases) and 404-432 (ModelEMA.update).  The reference runs this as ~1000 small foreach / elementwise launches.
//
// Layout: all parameters live in one flat fp32 buffer ordered by group: [0, n0) decayed weights, [n0, n1) BatchNorm
// weights, [n1, n) biases.  Hyper-parameters are read from DEVICE memory so that a captured CUDA graph picks up a
// new learning rate / EMA decay without re-capture:
//   hp = { lr0, lr1, lr2, momentum, weight_decay (group 0), max_norm (<= 0: no clipping), ema_decay, grad_scale }
// grad_scale multiplies every gradient first (1/loss-scale, or 1/world for a summed all-reduce).
#include "ly_common.cuh"

namespace ly {

__global__ void __launch_bounds__(256) sq_sum_kernel(const float* __restrict__ g, long n, float* __restrict__ out) {
  pdl_enter();
  __shared__ float sm[8];
  float a = 0.f;
  const long stride = (long)gridDim.x * blockDim.x * 4;
  long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  for (; i + 3 < n; i += stride) {
    const float4 v = *reinterpret_cast<const float4*>(g + i);
    a += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (i < n)
    for (long j = i; j < n && j < i + 4; ++j) a += g[j] * g[j];
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x < 32) {
    a = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.f;
    a = warp_sum(a);
    if (threadIdx.x == 0) atomicAdd(out, a);
  }
}

__global__ void __launch_bounds__(256) sgd_ema_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                      float* __restrict__ mom, float* __restrict__ ema, long n, long n0,
                                                      long n1, const float* __restrict__ hp,
                                                      const float* __restrict__ grad_sq) {
  pdl_enter();
  const float momentum = hp[3], wd = hp[4], max_norm = hp[5], d = hp[6], gs = hp[7];
  float clip = 1.f;
  if (max_norm > 0.f && grad_sq != nullptr) {
    // torch.nn.utils.clip_grad_norm_: coef = max_norm / (total_norm + 1e-6), clamped to 1
    const float total = sqrtf(grad_sq[0]) * fabsf(gs);
    clip = fminf(max_norm / (total + 1e-6f), 1.f);
  }
  const float k = gs * clip;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const int grp = i < n0 ? 0 : (i < n1 ? 1 : 2);
    const float lr = hp[grp];
    float w = p[i];
    float gr = g[i] * k;
    if (grp == 0) gr = fmaf(wd, w, gr);
    const float b = fmaf(momentum, mom[i], gr);
    mom[i] = b;
    gr = fmaf(momentum, b, gr);
    w = fmaf(-lr, gr, w);
    p[i] = w;
    if (ema != nullptr) ema[i] = fmaf(d, ema[i] - w, w);   // d*ema + (1-d)*w
  }
}

__global__ void __launch_bounds__(256) ema_kernel(float* __restrict__ ema, const float* __restrict__ src, long n,
                                                  const float* __restrict__ hp) {
  pdl_enter();
  const float d = hp[6];
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    ema[i] = fmaf(d, ema[i] - src[i], src[i]);
}

}  // namespace ly

using namespace ly;

extern "C" int ly_sq_sum(const float* g, long n, float* out, ly_stream_t st) {
  LY_REQUIRE(g && out && n > 0, "ly_sq_sum: bad arguments");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(g) % 16 == 0, "ly_sq_sum: pointer not 16-byte aligned");
  const int grid = (int)((n / 4 + 255) / 256 < device_sm_count() * 4 ? (n / 4 + 255) / 256 + 1 : device_sm_count() * 4);
  LY_CUDA(launch(sq_sum_kernel, grid, 256, 0, (cudaStream_t)st, g, n, out));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_sgd_ema_step(float* params, const float* grads, float* momentum_buf, float* ema, long n, long n0,
                               long n1, const float* hp, const float* grad_sq, ly_stream_t st) {
  LY_REQUIRE(params && grads && momentum_buf && hp && n > 0, "ly_sgd_ema_step: null pointer");
  LY_REQUIRE(0 <= n0 && n0 <= n1 && n1 <= n, "ly_sgd_ema_step: group boundaries %ld, %ld outside [0, %ld]", n0, n1, n);
  const int grid = (int)((n + 255) / 256 < device_sm_count() * 8 ? (n + 255) / 256 : device_sm_count() * 8);
  LY_CUDA(launch(sgd_ema_kernel, grid, 256, 0, (cudaStream_t)st, params, grads, momentum_buf, ema, n, n0, n1, hp, grad_sq));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_ema_update(float* ema, const float* src, long n, const float* hp, ly_stream_t st) {
  LY_REQUIRE(ema && src && hp && n > 0, "ly_ema_update: bad arguments");
  const int grid = (int)((n + 255) / 256 < device_sm_count() * 8 ? (n + 255) / 256 : device_sm_count() * 8);
  LY_CUDA(launch(ema_kernel, grid, 256, 0, (cudaStream_t)st, ema, src, n, hp));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
