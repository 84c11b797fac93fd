// Per-channel streaming kernels over NHWC bf16 matrices [M pixels][C channels] (pixel stride ld):
// BatchNorm statistics / finalize / apply (+activation) and their backward passes.
//
// Access pattern shared by all kernels here: the matrix is a flat stream of 16-byte vectors (8 channels).
// The launch makes (gridDim.x * blockDim.x) a multiple of V = C/8, so every thread keeps ONE fixed channel
// vector (its scale/shift/mean live in registers) while consecutive threads touch consecutive vectors
// (fully coalesced).  HBM-bound: algorithmic bytes = each operand read or written once.
#include <stdlib.h>

#include "ly_common.cuh"

namespace ly {

static int gcd_i(int a, int b) { return b == 0 ? a : gcd_i(b, a % b); }

// grid size whose total thread count is a multiple of V
int flat_grid(long M, int V, int block) {
  const int m = V / gcd_i(V, block);
  long want = (M * V + (long)block * 4 - 1) / ((long)block * 4);
  const long cap = (long)device_sm_count() * 8;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  long g = (want + m - 1) / m * m;
  return (int)g;
}

// Grid for the REDUCING kernels (statistics, BatchNorm-backward sums): every block ends with 2*C global atomics, so the
// block count is kept low — at least 16 rows per thread, at most four blocks per SM (1024 threads x 4-8 loads of 16 B
// in flight per SM: the kernels are HBM-latency bound below that).
int flat_grid_reduce(long M, int V, int block) {
  const int m = V / gcd_i(V, block);
  long want = (M * V + (long)block * 16 - 1) / ((long)block * 16);
  const long cap = (long)device_sm_count() * 4;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  long g = (want + m - 1) / m * m;
  return (int)g;
}

#define LY_FLAT_SETUP(V)                                                         \
  const long T_ = (long)gridDim.x * blockDim.x;                                  \
  const long t_ = (long)blockIdx.x * blockDim.x + threadIdx.x;                   \
  const int v = (int)(t_ % (V));                                                 \
  const long row0 = t_ / (V);                                                    \
  const long rstep = T_ / (V);

// Block-level reduction of per-thread channel-vector partials a[8], b[8] (thread t owns channel vector (t_ % V)) followed
// by ONE global atomic per (block, channel, statistic).  The partials go through shared memory and each of the 2*C
// result slots is summed by one thread over the block's threads that own its vector — no shared-memory atomics, which
// serialise 256 ways when V is small (CoordAtt's 8-channel matrices: 60 us for a 160 KB tensor).
__device__ __forceinline__ void block_channel_reduce(const float (&a)[8], const float (&b)[8], int V, long t0,
                                                     float* __restrict__ out_a, float* __restrict__ out_b,
                                                     float* __restrict__ parts = nullptr) {
  __shared__ float part[256][17];
  const int tid = threadIdx.x;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    part[tid][j] = a[j];
    part[tid][8 + j] = b[j];
  }
  __syncthreads();
  const int C = V * 8;
  const int first = (int)((V - (t0 % V)) % V);   // block thread index that owns vector 0
  for (int i = tid; i < 2 * C; i += blockDim.x) {
    const int which = i >= C, c = which ? i - C : i;
    const int vc = c >> 3, j = c & 7;
    float sum = 0.f;
    for (int tt = (first + vc) % V; tt < (int)blockDim.x; tt += V) sum += part[tt][which * 8 + j];
    if (parts != nullptr) parts[(long)blockIdx.x * 2 * C + i] = sum;   // ordered finalize follows (deterministic)
    else if (which == 0) atomicAdd(&out_a[c], sum);
    else if (out_b != nullptr) atomicAdd(&out_b[c], sum);
  }
}

// s1 | s2 = Σ_b parts[b][0..2C): one warp per output, lane l adds blocks l, l+32, ... in order and the 32 lane sums are
// folded by a fixed shuffle tree — the same order on every run.
__global__ void __launch_bounds__(256) ordered_finalize_kernel(const float* __restrict__ parts, int B, int C,
                                                               float* __restrict__ s1, float* __restrict__ s2) {
  pdl_enter();
  const int i = (int)(((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= 2 * C) return;
  float a = 0.f;
  for (int b = lane; b < B; b += 32) a += parts[(long)b * 2 * C + i];
  a = warp_sum(a);
  if (lane == 0) {
    if (i < C) s1[i] = a;
    else s2[i - C] = a;
  }
}

// Same, but the block sums are added as 64-bit FIXED-POINT integers (value * 2^fa / 2^fb): integer addition is
// associative, so the total does not depend on the order in which the blocks arrive — training-mode BatchNorm
// statistics, and with them the whole forward pass, are bit-reproducible from run to run.  Per-thread and per-block
// partial sums are fp32 in a fixed order; only the cross-block step is fixed point.
__device__ __forceinline__ void block_channel_reduce_fx(const float (&a)[8], const float (&b)[8], int V, long t0,
                                                        long long* __restrict__ out_a, long long* __restrict__ out_b,
                                                        float scale_a, float scale_b) {
  __shared__ float part[256][17];
  const int tid = threadIdx.x;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    part[tid][j] = a[j];
    part[tid][8 + j] = b[j];
  }
  __syncthreads();
  const int C = V * 8;
  const int first = (int)((V - (t0 % V)) % V);
  for (int i = tid; i < 2 * C; i += blockDim.x) {
    const int which = i >= C, c = which ? i - C : i;
    const int vc = c >> 3, j = c & 7;
    float sum = 0.f;
    for (int tt = (first + vc) % V; tt < (int)blockDim.x; tt += V) sum += part[tt][which * 8 + j];
    const long long q = __double2ll_rn((double)sum * (double)(which ? scale_b : scale_a));
    atomicAdd(reinterpret_cast<unsigned long long*>(which ? &out_b[c] : &out_a[c]), (unsigned long long)q);
  }
}

__device__ __forceinline__ uint4 ld16(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void st16(__nv_bfloat16* p, const uint4& u) { *reinterpret_cast<uint4*>(p) = u; }

// ------------------------------------------------------------------------------------------------
// sum[c] += Σ x, sq[c] += Σ x²   (sq may be null)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) col_stats_kernel(const __nv_bfloat16* __restrict__ x, long ld, long M, int V,
                                                        float* __restrict__ sum, float* __restrict__ sq) {
  pdl_enter();
  LY_FLAT_SETUP(V)
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  long row = row0;
  for (; row + 3 * rstep < M; row += 4 * rstep) {
    uint4 u0 = ld16(x + row * ld + v * 8);
    uint4 u1 = ld16(x + (row + rstep) * ld + v * 8);
    uint4 u2 = ld16(x + (row + 2 * rstep) * ld + v * 8);
    uint4 u3 = ld16(x + (row + 3 * rstep) * ld + v * 8);
    float f[8];
    unpack8(u0, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
    unpack8(u1, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
    unpack8(u2, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
    unpack8(u3, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
  }
  for (; row < M; row += rstep) {
    float f[8];
    unpack8(ld16(x + row * ld + v * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
  }
  block_channel_reduce(s, q, V, (long)blockIdx.x * blockDim.x, sum, sq);
}

// ------------------------------------------------------------------------------------------------
// BatchNorm finalize: batch mean / biased var -> (mean, rstd, scale, shift) and running-stat update.
// nn.BatchNorm2d semantics: running_var gets the unbiased variance.
// ------------------------------------------------------------------------------------------------
__global__ void bn_finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sq, int C, float n,
                                   float eps, float momentum, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* running_mean, float* running_var,
                                   float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                   float* __restrict__ scale_out, float* __restrict__ shift_out) {
  pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = sum[c] / n;
  const float var = fmaxf(sq[c] / n - mean * mean, 0.f);
  const float rstd = rsqrtf(var + eps);
  const float g = gamma != nullptr ? gamma[c] : 1.f;
  const float b = beta != nullptr ? beta[c] : 0.f;
  mean_out[c] = mean;
  rstd_out[c] = rstd;
  scale_out[c] = g * rstd;
  shift_out[c] = b - mean * g * rstd;
  if (running_mean != nullptr) {
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    const float unbiased = n > 1.f ? var * n / (n - 1.f) : var;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

// Inference fold: scale = gamma / sqrt(running_var + eps); shift = beta - running_mean*scale + bias*scale
__global__ void bn_fold_kernel(int C, float eps, const float* __restrict__ gamma, const float* __restrict__ beta,
                               const float* __restrict__ rmean, const float* __restrict__ rvar,
                               const float* __restrict__ conv_bias, float* __restrict__ mean_out,
                               float* __restrict__ rstd_out, float* __restrict__ scale_out,
                               float* __restrict__ shift_out) {
  pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float rstd = rsqrtf(rvar[c] + eps);
  const float sc = gamma[c] * rstd;
  float sh = beta[c] - rmean[c] * sc;
  if (conv_bias != nullptr) sh += conv_bias[c] * sc;
  if (mean_out != nullptr) mean_out[c] = rmean[c];
  if (rstd_out != nullptr) rstd_out[c] = rstd;
  scale_out[c] = sc;
  shift_out[c] = sh;
}

// ------------------------------------------------------------------------------------------------
// y = act(x * scale[c] + shift[c])
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) scale_shift_act_kernel(const __nv_bfloat16* __restrict__ x, long ldx, long M,
                                                              int V, const float* __restrict__ scale,
                                                              const float* __restrict__ shift, int act,
                                                              __nv_bfloat16* __restrict__ y, long ldy) {
  pdl_enter();
  LY_FLAT_SETUP(V)
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = scale != nullptr ? scale[v * 8 + j] : 1.f;
    sh[j] = shift != nullptr ? shift[v * 8 + j] : 0.f;
  }
  for (long row = row0; row < M; row += rstep) {
    float f[8];
    unpack8(ld16(x + row * ldx + v * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = act_apply(fmaf(f[j], sc[j], sh[j]), act);
    st16(y + row * ldy + v * 8, pack8(f));
  }
}

// ------------------------------------------------------------------------------------------------
// Backward of y = act(scale*z + shift) with scale = gamma*rstd, shift = beta - mean*scale.
// pass 1: s1[c] += Σ dact, s2[c] += Σ dact * xhat     (dact = dy * act'(scale*z+shift), xhat = (z-mean)*rstd)
// pass 2: dz = scale * (dact - s1/n - xhat*s2/n)      (training) |  dz = scale * dact  (eval / plain affine)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bn_act_bwd_reduce_kernel(
    const __nv_bfloat16* __restrict__ dy, long lddy, const __nv_bfloat16* __restrict__ z, long ldz, long M, int V,
    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean,
    const float* __restrict__ rstd, int act, float* __restrict__ s1, float* __restrict__ s2,
    float* __restrict__ parts) {
  pdl_enter();
  LY_FLAT_SETUP(V)
  float sc[8], sh[8], mu[8], rs[8], a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = scale[v * 8 + j]; sh[j] = shift[v * 8 + j];
    mu[j] = mean[v * 8 + j]; rs[j] = rstd[v * 8 + j];
    a1[j] = a2[j] = 0.f;
  }
  auto accum = [&](const uint4& ug, const uint4& uf) {
    float g[8], f[8];
    unpack8(ug, g);
    unpack8(uf, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = g[j] * act_grad(fmaf(f[j], sc[j], sh[j]), act);
      a1[j] += d;
      a2[j] = fmaf(d, (f[j] - mu[j]) * rs[j], a2[j]);
    }
  };
  long row = row0;
  for (; row + 3 * rstep < M; row += 4 * rstep) {       // eight 16-byte loads in flight per thread
    uint4 ug[4], uf[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      ug[i] = ld16(dy + (row + i * rstep) * lddy + v * 8);
      uf[i] = ld16(z + (row + i * rstep) * ldz + v * 8);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) accum(ug[i], uf[i]);
  }
  for (; row < M; row += rstep) accum(ld16(dy + row * lddy + v * 8), ld16(z + row * ldz + v * 8));
  block_channel_reduce(a1, a2, V, (long)blockIdx.x * blockDim.x, s1, s2, parts);
}

__global__ void __launch_bounds__(256) bn_act_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ dy, long lddy, const __nv_bfloat16* __restrict__ z, long ldz, long M, int V,
    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean,
    const float* __restrict__ rstd, int act, const float* __restrict__ s1, const float* __restrict__ s2,
    float inv_n, int training, __nv_bfloat16* __restrict__ dz, long lddz) {
  pdl_enter();
  LY_FLAT_SETUP(V)
  float sc[8], sh[8], mu[8], rs[8], m1[8], m2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = scale[v * 8 + j]; sh[j] = shift[v * 8 + j];
    mu[j] = mean != nullptr ? mean[v * 8 + j] : 0.f;
    rs[j] = rstd != nullptr ? rstd[v * 8 + j] : 1.f;
    m1[j] = training ? s1[v * 8 + j] * inv_n : 0.f;
    m2[j] = training ? s2[v * 8 + j] * inv_n : 0.f;
  }
  auto apply = [&](long row, const uint4& ug, const uint4& uf) {
    float g[8], f[8];
    unpack8(ug, g);
    unpack8(uf, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = g[j] * act_grad(fmaf(f[j], sc[j], sh[j]), act);
      const float xh = (f[j] - mu[j]) * rs[j];
      f[j] = sc[j] * (d - m1[j] - xh * m2[j]);
    }
    st16(dz + row * lddz + v * 8, pack8(f));
  };
  long row = row0;
  for (; row + rstep < M; row += 2 * rstep) {           // four loads in flight per thread
    const uint4 g0 = ld16(dy + row * lddy + v * 8), f0 = ld16(z + row * ldz + v * 8);
    const uint4 g1 = ld16(dy + (row + rstep) * lddy + v * 8), f1 = ld16(z + (row + rstep) * ldz + v * 8);
    apply(row, g0, f0);
    apply(row + rstep, g1, f1);
  }
  if (row < M) apply(row, ld16(dy + row * lddy + v * 8), ld16(z + row * ldz + v * 8));
}

// dz = dy * act'(.) where the derivative is taken from the OUTPUT y (relu, sigmoid, none).
__global__ void __launch_bounds__(256) act_bwd_from_out_kernel(const __nv_bfloat16* __restrict__ dy, long lddy,
                                                               const __nv_bfloat16* __restrict__ y, long ldy, long M,
                                                               int V, int act, __nv_bfloat16* __restrict__ dz,
                                                               long lddz) {
  pdl_enter();
  LY_FLAT_SETUP(V)
  for (long row = row0; row < M; row += rstep) {
    float g[8], f[8];
    unpack8(ld16(dy + row * lddy + v * 8), g);
    unpack8(ld16(y + row * ldy + v * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float d = 1.f;
      if (act == LY_ACT_RELU) d = f[j] > 0.f ? 1.f : 0.f;
      else if (act == LY_ACT_SIGMOID) d = f[j] * (1.f - f[j]);
      g[j] *= d;
    }
    st16(dz + row * lddz + v * 8, pack8(g));
  }
}

static int check_mat(const void* p, long ld, long M, int C, const char* what) {
  LY_REQUIRE(p != nullptr, "%s: null pointer", what);
  LY_REQUIRE(M > 0 && C > 0 && C % 8 == 0, "%s: bad shape M=%ld C=%d (C must be a multiple of 8)", what, M, C);
  LY_REQUIRE(ld >= C && ld % 8 == 0, "%s: ld=%ld invalid", what, ld);
  LY_REQUIRE(reinterpret_cast<uintptr_t>(p) % 16 == 0, "%s: pointer not 16-byte aligned", what);
  LY_REQUIRE(C <= 4096, "%s: C=%d too large", what, C);
  return LY_OK;
}

}  // namespace ly

using namespace ly;

extern "C" int ly_col_stats(const void* x, long ld, long M, int C, float* sum, float* sq, ly_stream_t st) {
  int rc = check_mat(x, ld, M, C, "ly_col_stats");
  if (rc) return rc;
  LY_REQUIRE(sum != nullptr, "ly_col_stats: null sum");
  const int V = C / 8;
  LY_CUDA(launch(col_stats_kernel, flat_grid_reduce(M, V, 256), 256, 0, (cudaStream_t)st, 
      (const __nv_bfloat16*)x, ld, M, V, sum, sq));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_bn_finalize(const float* sum, const float* sq, int C, long n, float eps, float momentum,
                              const float* gamma, const float* beta, float* running_mean, float* running_var,
                              float* mean_out, float* rstd_out, float* scale_out, float* shift_out,
                              ly_stream_t st) {
  LY_REQUIRE(sum && sq && mean_out && rstd_out && scale_out && shift_out && C > 0 && n > 0, "ly_bn_finalize: bad args");
  LY_CUDA(launch(bn_finalize_kernel, ceil_div(C, 128), 128, 0, (cudaStream_t)st, sum, sq, C, (float)n, eps, momentum, gamma,
                                                                      beta, running_mean, running_var, mean_out,
                                                                      rstd_out, scale_out, shift_out));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_bn_fold(int C, float eps, const float* gamma, const float* beta, const float* running_mean,
                          const float* running_var, const float* conv_bias, float* mean_out, float* rstd_out,
                          float* scale_out, float* shift_out, ly_stream_t st) {
  LY_REQUIRE(C > 0 && gamma && beta && running_mean && running_var && scale_out && shift_out, "ly_bn_fold: bad args");
  LY_CUDA(launch(bn_fold_kernel, ceil_div(C, 128), 128, 0, (cudaStream_t)st, C, eps, gamma, beta, running_mean, running_var,
                                                                  conv_bias, mean_out, rstd_out, scale_out, shift_out));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_scale_shift_act(const void* x, long ldx, long M, int C, const float* scale, const float* shift,
                                  int act, void* y, long ldy, ly_stream_t st) {
  int rc = check_mat(x, ldx, M, C, "ly_scale_shift_act(x)");
  if (rc) return rc;
  rc = check_mat(y, ldy, M, C, "ly_scale_shift_act(y)");
  if (rc) return rc;
  const int V = C / 8;
  LY_CUDA(launch(scale_shift_act_kernel, flat_grid(M, V, 256), 256, 0, (cudaStream_t)st, 
      (const __nv_bfloat16*)x, ldx, M, V, scale, shift, act, (__nv_bfloat16*)y, ldy));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_reduce_blocks(long M, int C) {
  if (M <= 0 || C <= 0 || C % 8 != 0) return -1;
  return flat_grid_reduce(M, C / 8, 256);
}

extern "C" int ly_act_bwd_from_out(const void* dy, long lddy, const void* y, long ldy, long M, int C, int act,
                                   void* dz, long lddz, ly_stream_t st) {
  int rc = check_mat(dy, lddy, M, C, "ly_act_bwd_from_out(dy)");
  if (rc) return rc;
  rc = check_mat(y, ldy, M, C, "ly_act_bwd_from_out(y)");
  if (rc) return rc;
  rc = check_mat(dz, lddz, M, C, "ly_act_bwd_from_out(dz)");
  if (rc) return rc;
  LY_REQUIRE(act == LY_ACT_NONE || act == LY_ACT_RELU || act == LY_ACT_SIGMOID,
             "ly_act_bwd_from_out: activation %d has no output-only derivative", act);
  const int V = C / 8;
  LY_CUDA(launch(act_bwd_from_out_kernel, flat_grid(M, V, 256), 256, 0, (cudaStream_t)st, 
      (const __nv_bfloat16*)dy, lddy, (const __nv_bfloat16*)y, ldy, M, V, act, (__nv_bfloat16*)dz, lddz));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// ------------------------------------------------------------------------------------------------
// Training-mode BatchNorm in two launches (statistics, then normalise + activation) with the finalize step folded
// into the second one.  The sums are taken about a per-channel pivot K = x[0][c] (shifted-data algorithm):
// var = Σ(x-K)²/n − (Σ(x-K)/n)² does not cancel when |mean| >> std, which the plain E[x²] − mean² form does.
// acc: [C] fp32 pivot K | [C] int64 Σ(x-K)·2^24 | [C] int64 Σ(x-K)²·2^20  (5*C floats of space, zeroed by the caller).
// ------------------------------------------------------------------------------------------------
namespace ly {

// fixed-point scales of the two sums: |Σ(x-K)| < 5e11 and Σ(x-K)² < 8e12 are representable (activations up to the
// thousands on 1.6 M pixels); resolutions 6e-8 and 1e-6
constexpr float kBnScale1 = 16777216.f;   // 2^24
constexpr float kBnScale2 = 1048576.f;    // 2^20

__global__ void __launch_bounds__(256) bn_stats_kernel(const __nv_bfloat16* __restrict__ x, long ld, long M, int V,
                                                       float* __restrict__ acc) {
  pdl_enter();
  const int C = V * 8;
  LY_FLAT_SETUP(V)
  float s[8], q[8], K[8];
  unpack8(ld16(x + v * 8), K);
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  if (t_ < V) {  // the V threads of row 0 publish the pivot (V may exceed one block)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[v * 8 + j] = K[j];
  }
  long row = row0;
  for (; row + 3 * rstep < M; row += 4 * rstep) {
    uint4 u[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) u[i] = ld16(x + (row + i * rstep) * ld + v * 8);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float f[8];
      unpack8(u[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = f[j] - K[j];
        s[j] += d;
        q[j] = fmaf(d, d, q[j]);
      }
    }
  }
  for (; row < M; row += rstep) {
    float f[8];
    unpack8(ld16(x + row * ld + v * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = f[j] - K[j];
      s[j] += d;
      q[j] = fmaf(d, d, q[j]);
    }
  }
  long long* fx = reinterpret_cast<long long*>(acc + C);
  block_channel_reduce_fx(s, q, V, (long)blockIdx.x * blockDim.x, fx, fx + C, kBnScale1, kBnScale2);
}

__device__ __forceinline__ void bn_moments(const float* __restrict__ acc, int C, int c, float inv_n, float& mean,
                                           float& var) {
  const long long* fx = reinterpret_cast<const long long*>(acc + C);
  const double s1 = (double)fx[c] * (1.0 / (double)kBnScale1), s2 = (double)fx[C + c] * (1.0 / (double)kBnScale2);
  const float d = (float)(s1 * (double)inv_n);
  mean = acc[c] + d;
  var = fmaxf((float)(s2 * (double)inv_n) - d * d, 0.f);
}

__global__ void __launch_bounds__(256) bn_apply_train_kernel(
    const __nv_bfloat16* __restrict__ z, long ldz, long M, int V, const float* __restrict__ acc,
    const float* __restrict__ gamma, const float* __restrict__ beta, float eps, float momentum, float* running_mean,
    float* running_var, float* __restrict__ stats, int act, __nv_bfloat16* __restrict__ y, long ldy) {
  pdl_enter();
  LY_FLAT_SETUP(V)
  const int C = V * 8;
  const float n = (float)M, inv_n = 1.f / n;
  if (blockIdx.x == 0) {
    // saved statistics (mean, rstd, scale, shift) for the backward pass + running-statistics update
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float mean, var;
      bn_moments(acc, C, c, inv_n, mean, var);
      const float rstd = rsqrtf(var + eps);
      const float g = gamma != nullptr ? gamma[c] : 1.f, b = beta != nullptr ? beta[c] : 0.f;
      stats[c] = mean;
      stats[C + c] = rstd;
      stats[2 * C + c] = g * rstd;
      stats[3 * C + c] = b - mean * g * rstd;
      if (running_mean != nullptr) {
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
        const float unbiased = n > 1.f ? var * n / (n - 1.f) : var;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
      }
    }
  }
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = v * 8 + j;
    float mean, var;
    bn_moments(acc, C, c, inv_n, mean, var);
    const float rstd = rsqrtf(var + eps);
    const float g = gamma != nullptr ? gamma[c] : 1.f, b = beta != nullptr ? beta[c] : 0.f;
    sc[j] = g * rstd;
    sh[j] = b - mean * sc[j];
  }
  auto apply = [&](long row, const uint4& u) {
    float f[8];
    unpack8(u, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = act_apply(fmaf(f[j], sc[j], sh[j]), act);
    st16(y + row * ldy + v * 8, pack8(f));
  };
  long row = row0;
  for (; row + 3 * rstep < M; row += 4 * rstep) {
    uint4 u[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) u[i] = ld16(z + (row + i * rstep) * ldz + v * 8);
#pragma unroll
    for (int i = 0; i < 4; ++i) apply(row + i * rstep, u[i]);
  }
  for (; row < M; row += rstep) apply(row, ld16(z + row * ldz + v * 8));
}


// ------------------------------------------------------------------------------------------------------------------
// Training-mode BatchNorm BACKWARD of a large tensor as ONE kernel: reduce | grid barrier | ordered finalize | grid
// barrier | apply.  The grid is launched cooperatively (all blocks co-resident); the barrier is an arrival counter in
// zeroed memory, bounded like every wait in this library.  What a later phase reads that an earlier phase of the SAME
// kernel wrote goes through __ldcg (L2), never through the read-only path.
// Measured (tests/dev/bn_time.py, in a replayed graph): the fused kernel wins 8-15 % from 50 M elements up (349 -> 299 us
// at 118 M) and loses ~5 us per layer below (two barriers, no programmatic overlap with its neighbours, a smaller grid);
// the same fusion of the FORWARD pair (statistics | barrier | normalise) lost at every size and was removed.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void grid_barrier(unsigned* counter) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    const long long t0 = clock64();
    unsigned seen;
    do {
      asm volatile("ld.acquire.gpu.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
      if (seen >= gridDim.x) break;
      __nanosleep(32);
      if (clock64() - t0 > 4000000000LL) __trap();   // a grid that is not co-resident must not hang the GPU
    } while (true);
  }
  __syncthreads();
}

// sync: two zeroed barrier counters.  parts: [gridDim.x][2*C] block partials (ordered sum -> s1, s2).
__global__ void __launch_bounds__(256, 2) bn_bwd_fused_kernel(
    const __nv_bfloat16* __restrict__ dy, long lddy, const __nv_bfloat16* __restrict__ z, long ldz, long M, int V,
    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean,
    const float* __restrict__ rstd, int act, int training, float* s1, float* s2, float* parts, unsigned* sync,
    __nv_bfloat16* __restrict__ dz, long lddz) {
  pdl_enter();
  const int C = V * 8;
  LY_FLAT_SETUP(V)
  float sc[8], sh[8], mu[8], rs[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = scale[v * 8 + j]; sh[j] = shift[v * 8 + j];
    mu[j] = mean[v * 8 + j]; rs[j] = rstd[v * 8 + j];
  }
  {   // ---- phase 1: Σ d, Σ d·x̂ per block (bn_act_bwd_reduce_kernel)
    float a1[8], a2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a1[j] = a2[j] = 0.f;
    auto accum = [&](const uint4& ug, const uint4& uf) {
      float g[8], f[8];
      unpack8(ug, g);
      unpack8(uf, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = g[j] * act_grad(fmaf(f[j], sc[j], sh[j]), act);
        a1[j] += d;
        a2[j] = fmaf(d, (f[j] - mu[j]) * rs[j], a2[j]);
      }
    };
    long row = row0;
    for (; row + 3 * rstep < M; row += 4 * rstep) {
      uint4 ug[4], uf[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ug[i] = ld16(dy + (row + i * rstep) * lddy + v * 8);
        uf[i] = ld16(z + (row + i * rstep) * ldz + v * 8);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) accum(ug[i], uf[i]);
    }
    for (; row < M; row += rstep) accum(ld16(dy + row * lddy + v * 8), ld16(z + row * ldz + v * 8));
    block_channel_reduce(a1, a2, V, (long)blockIdx.x * blockDim.x, nullptr, nullptr, parts);
  }
  grid_barrier(sync);
  {   // ---- phase 2: ordered sum of the block partials, one warp per output (ordered_finalize_kernel)
    const int lane = threadIdx.x & 31;
    const int B = (int)gridDim.x;
    for (int i = (int)blockIdx.x * 8 + (threadIdx.x >> 5); i < 2 * C; i += B * 8) {
      float a = 0.f;
      for (int b = lane; b < B; b += 32) a += __ldcg(parts + (long)b * 2 * C + i);
      a = warp_sum(a);
      if (lane == 0) {
        if (i < C) s1[i] = a;
        else s2[i - C] = a;
      }
    }
  }
  grid_barrier(sync + 1);
  // ---- phase 3: dz (bn_act_bwd_apply_kernel)
  const float inv_n = 1.f / (float)M;
  float m1[8], m2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    m1[j] = training ? __ldcg(s1 + v * 8 + j) * inv_n : 0.f;
    m2[j] = training ? __ldcg(s2 + v * 8 + j) * inv_n : 0.f;
  }
  auto apply = [&](long row, const uint4& ug, const uint4& uf) {
    float g[8], f[8];
    unpack8(ug, g);
    unpack8(uf, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = g[j] * act_grad(fmaf(f[j], sc[j], sh[j]), act);
      const float xh = (f[j] - mu[j]) * rs[j];
      f[j] = sc[j] * (d - m1[j] - xh * m2[j]);
    }
    st16(dz + row * lddz + v * 8, pack8(f));
  };
  long row = row0;
  for (; row + rstep < M; row += 2 * rstep) {
    const uint4 g0 = ld16(dy + row * lddy + v * 8), f0 = ld16(z + row * ldz + v * 8);
    const uint4 g1 = ld16(dy + (row + rstep) * lddy + v * 8), f1 = ld16(z + (row + rstep) * ldz + v * 8);
    apply(row, g0, f0);
    apply(row + rstep, g1, f1);
  }
  if (row < M) apply(row, ld16(dy + row * lddy + v * 8), ld16(z + row * ldz + v * 8));
}

// Cooperative launch: every block of the grid is resident at once (the launch fails otherwise), which is what the
// grid barrier needs.  `blocks_per_sm` of the kernel is queried once per device.
template <typename... KArgs, typename... Args>
static cudaError_t launch_coop(void (*kernel)(KArgs...), int grid, int block, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<Args&&>(args)...);
}

template <typename K>
static int coop_capacity(K kernel, int block) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, 0) != cudaSuccess || per_sm < 1) return 0;
  return per_sm * device_sm_count();
}

// largest grid <= cap that keeps the (row, channel-vector) mapping of LY_FLAT_SETUP consistent (multiple of V / gcd)
static int fused_grid(long M, int V, int cap) {
  const int m = V / gcd_i(V, 256);
  long want = flat_grid_reduce(M, V, 256);
  if (want > cap) want = cap / m * m;
  return (int)want;
}

static const bool g_bn_unfused = getenv("LY_BN_UNFUSED") != nullptr;   // dev knob: the separate kernels
constexpr long kBnFuseMin = 48L << 20;                                  // elements from which the fused backward wins

}  // namespace ly

extern "C" int ly_bn_stats(const void* x, long ld, long M, int C, float* acc, ly_stream_t st) {
  int rc = check_mat(x, ld, M, C, "ly_bn_stats");
  if (rc) return rc;
  LY_REQUIRE(acc != nullptr, "ly_bn_stats: null accumulator");
  const int V = C / 8;
  LY_CUDA(launch(bn_stats_kernel, flat_grid_reduce(M, V, 256), 256, 0, (cudaStream_t)st, (const __nv_bfloat16*)x, ld,
                                                                                          M, V, acc));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_bn_apply_train(const void* z, long ldz, long M, int C, const float* acc, const float* gamma,
                                 const float* beta, float eps, float momentum, float* running_mean,
                                 float* running_var, float* stats, int act, void* y, long ldy, ly_stream_t st) {
  int rc = check_mat(z, ldz, M, C, "ly_bn_apply_train(z)");
  if (rc) return rc;
  rc = check_mat(y, ldy, M, C, "ly_bn_apply_train(y)");
  if (rc) return rc;
  LY_REQUIRE(acc && stats, "ly_bn_apply_train: null statistics");
  LY_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "ly_bn_apply_train: running statistics come in pairs");
  const int V = C / 8;
  LY_CUDA(launch(bn_apply_train_kernel, flat_grid(M, V, 256), 256, 0, (cudaStream_t)st, 
      (const __nv_bfloat16*)z, ldz, M, V, acc, gamma, beta, eps, momentum, running_mean, running_var, stats, act,
      (__nv_bfloat16*)y, ldy));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_bn_act_bwd(const void* dy, long lddy, const void* z, long ldz, long M, int C, const float* scale,
                             const float* shift, const float* mean, const float* rstd, int act, int training,
                             float* s1, float* s2, float* parts, unsigned* sync, void* dz, long lddz, ly_stream_t st) {
  int rc = check_mat(dy, lddy, M, C, "ly_bn_act_bwd(dy)");
  if (rc) return rc;
  rc = check_mat(z, ldz, M, C, "ly_bn_act_bwd(z)");
  if (rc) return rc;
  rc = check_mat(dz, lddz, M, C, "ly_bn_act_bwd(dz)");
  if (rc) return rc;
  LY_REQUIRE(scale && shift && mean && rstd && s1 && s2, "ly_bn_act_bwd: null statistics");
  const int V = C / 8;
  if (parts != nullptr && sync != nullptr && !g_bn_unfused && M * (long)C >= kBnFuseMin) {
    // one cooperative kernel: reduce | grid barrier | ordered finalize | grid barrier | apply
    static const int cap = coop_capacity(bn_bwd_fused_kernel, 256);
    const int fg = fused_grid(M, V, cap);
    if (fg >= 1) {
      LY_CUDA(launch_coop(bn_bwd_fused_kernel, fg, 256, (cudaStream_t)st, (const __nv_bfloat16*)dy, lddy,
                          (const __nv_bfloat16*)z, ldz, M, V, scale, shift, mean, rstd, act, training, s1, s2, parts, sync,
                          (__nv_bfloat16*)dz, lddz));
      LY_LAUNCH_CHECK();
      return LY_OK;
    }
  }
  const int grid = flat_grid(M, V, 256);
  const int rgrid = flat_grid_reduce(M, V, 256);
  LY_CUDA(launch(bn_act_bwd_reduce_kernel, rgrid, 256, 0, (cudaStream_t)st, 
      (const __nv_bfloat16*)dy, lddy, (const __nv_bfloat16*)z, ldz, M, V, scale, shift, mean, rstd, act, s1, s2, parts));
  LY_LAUNCH_CHECK();
  if (parts != nullptr) {
    LY_CUDA(launch(ordered_finalize_kernel, ceil_div(2L * C * 32, 256), 256, 0, (cudaStream_t)st, parts, rgrid, C, s1, s2));
    LY_LAUNCH_CHECK();
  }
  LY_CUDA(launch(bn_act_bwd_apply_kernel, grid, 256, 0, (cudaStream_t)st, 
      (const __nv_bfloat16*)dy, lddy, (const __nv_bfloat16*)z, ldz, M, V, scale, shift, mean, rstd, act, s1, s2,
      1.f / (float)M, training, (__nv_bfloat16*)dz, lddz));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
