// Inference pipeline of RFCBAMConv (reference models/rfa.py:113-128) without materialising the receptive-field
// matrix G in HBM:
//   pass 1  rf_stats_kernel   recomputes g = relu(bn(dwconv(x))) on the fly and emits only max_c / mean_c per
//                             (pixel, tap) plus the per-(image, channel) sums of x for the SE gate
//   tiny    rf_se_att_kernel  SE gate s = sigmoid(W2 relu(W1 mean(x)))  and  att = sigmoid(conv3x3(maps))
//   pass 2  rf_apply_kernel   recomputes g, multiplies by s*att and writes the GEMM operand A (bf16) once
// HBM traffic: x twice + A once (the GEMM reads A once more) instead of five G-sized passes.
//
// Thread mapping of both passes: a warp = 32 consecutive output pixels x ONE 8-channel vector, so the k*k x k*k
// depthwise weights, BN scale/shift of that vector are warp-uniform loads (one sector per warp instead of one per
// lane); the kernels are then bound by the FMA pipe (k^4 MACs per output channel), not by weight traffic.
#include <atomic>

#include "ly_common.cuh"

namespace ly {

__device__ __forceinline__ uint4 ldx16(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// Asynchronous input staging: each warp copies the k*k input vectors of its 32 pixels (16 B per lane and tap) into a
// warp-private shared-memory buffer with cp.async (zero-filled outside the image), one pixel group ahead of the
// arithmetic, so the global-load latency is off the critical path even at the low occupancy these register-heavy
// kernels run at.
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src, bool valid) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst_smem);
  const int bytes = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int K>
__device__ __forceinline__ void rf_prefetch(uint8_t* buf, const __nv_bfloat16* __restrict__ x, long ldx, long M, int H,
                                            int W, int s, int ho, int wo, long m, int v, int lane) {
  const bool pact = m < M;
  const int w = pact ? (int)(m % wo) : 0, h = pact ? (int)((m / wo) % ho) : 0;
  const long n = pact ? m / ((long)wo * ho) : 0;
#pragma unroll
  for (int i = 0; i < K; ++i)
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int yy = h * s + i - K / 2, xx = w * s + j - K / 2;
      const bool ok = pact && yy >= 0 && yy < H && xx >= 0 && xx < W;
      const __nv_bfloat16* src = ok ? x + ((n * H + yy) * W + xx) * ldx + v * 8 : x;
      cp_async16(buf + ((i * K + j) * 32 + lane) * 16, src, ok);
    }
}

template <int K>
__device__ __forceinline__ void rf_fetch(const uint8_t* buf, int lane, float (&xin)[K * K][8]) {
#pragma unroll
  for (int ij = 0; ij < K * K; ++ij) unpack8(*reinterpret_cast<const uint4*>(buf + (ij * 32 + lane) * 16), xin[ij]);
}

template <int K>
__device__ __forceinline__ void rf_load_inputs(const __nv_bfloat16* __restrict__ x, long ldx, int H, int W, int s,
                                               int n, int h, int w, int v, bool act, float (&xin)[K * K][8]) {
#pragma unroll
  for (int i = 0; i < K; ++i)
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int yy = h * s + i - K / 2, xx = w * s + j - K / 2;
      if (act && yy >= 0 && yy < H && xx >= 0 && xx < W) {
        unpack8(ldx16(x + (((long)n * H + yy) * W + xx) * ldx + v * 8), xin[i * K + j]);
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) xin[i * K + j][e] = 0.f;
      }
    }
}

// Shared-memory copy of the depthwise weights / folded BN of the block's VB channel vectors:
//   wsm[(t*KK + ij)*VB*8 + vl*8 + e],  ssm[(t*VB + vl)*8 + e] (scale), hsm[...] (shift)      (vl = local vector index)
// Reads are warp-uniform (broadcast): one shared-memory wavefront feeds 4 FMA instructions.
template <int K>
__device__ __forceinline__ void rf_stage_weights(const float* __restrict__ wgk, const float* __restrict__ scale,
                                                 const float* __restrict__ shift, int C, int v0, int VB, int V,
                                                 float* wsm, float* ssm, float* hsm) {
  constexpr int KK = K * K;
  const int n8 = VB * 8;
  for (int i = threadIdx.x; i < KK * KK * n8; i += blockDim.x) {
    const int r = i / n8, c = i % n8;
    wsm[i] = (v0 * 8 + c < C) ? wgk[(long)r * C + v0 * 8 + c] : 0.f;
  }
  for (int i = threadIdx.x; i < KK * n8; i += blockDim.x) {
    const int t = i / n8, c = i % n8;
    const bool ok = v0 * 8 + c < C;
    ssm[i] = ok ? scale[(long)t * C + v0 * 8 + c] : 0.f;
    hsm[i] = ok ? shift[(long)t * C + v0 * 8 + c] : 0.f;
  }
}

// g[e] for tap t of this thread's 8 channels (folded BN + ReLU, rounded to bf16 like the stored operand)
template <int K>
__device__ __forceinline__ void rf_tap(const float (&xin)[K * K][8], const float* wsm, const float* ssm,
                                       const float* hsm, int VB, int vl, int t, float (&g)[8]) {
  constexpr int KK = K * K;
  const int n8 = VB * 8;
#pragma unroll
  for (int e = 0; e < 8; ++e) g[e] = 0.f;
#pragma unroll
  for (int ij = 0; ij < KK; ++ij) {
    const float4 a0 = *reinterpret_cast<const float4*>(wsm + (t * KK + ij) * n8 + vl * 8);
    const float4 a1 = *reinterpret_cast<const float4*>(wsm + (t * KK + ij) * n8 + vl * 8 + 4);
    g[0] = fmaf(xin[ij][0], a0.x, g[0]); g[1] = fmaf(xin[ij][1], a0.y, g[1]);
    g[2] = fmaf(xin[ij][2], a0.z, g[2]); g[3] = fmaf(xin[ij][3], a0.w, g[3]);
    g[4] = fmaf(xin[ij][4], a1.x, g[4]); g[5] = fmaf(xin[ij][5], a1.y, g[5]);
    g[6] = fmaf(xin[ij][6], a1.z, g[6]); g[7] = fmaf(xin[ij][7], a1.w, g[7]);
  }
  const float4 s0 = *reinterpret_cast<const float4*>(ssm + t * n8 + vl * 8);
  const float4 s1 = *reinterpret_cast<const float4*>(ssm + t * n8 + vl * 8 + 4);
  const float4 h0 = *reinterpret_cast<const float4*>(hsm + t * n8 + vl * 8);
  const float4 h1 = *reinterpret_cast<const float4*>(hsm + t * n8 + vl * 8 + 4);
  const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
  const float sh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
  for (int e = 0; e < 8; ++e)
    g[e] = __bfloat162float(__float2bfloat16(fmaxf(fmaf(g[e], sc[e], sh[e]), 0.f)));
}

// grid = (persistent blocks, channel-vector groups); block = WPB warps, warp w owns vector blockIdx.y*WPB + w and
// the block strides over groups of 32 pixels.  Dynamic smem: weights + scale/shift of the block's WPB vectors.
template <int K>
__global__ void __launch_bounds__(512, 1) rf_stats_kernel(const __nv_bfloat16* __restrict__ x, long ldx, long M, int H,
                                                          int W, int V, int WPB, int s, int ho, int wo,
                                                          const float* __restrict__ wgk,
                                                          const float* __restrict__ scale,
                                                          const float* __restrict__ shift, float* __restrict__ maps,
                                                          float* __restrict__ gap_sum, int use_atomics,
                                                          __nv_bfloat16* __restrict__ G) {
  pdl_enter();
  constexpr int KK = K * K;
  extern __shared__ float dsm[];
  __shared__ int smax[KK][32];
  __shared__ float ssum[KK][32];
  const int C = V * 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* wsm = dsm;
  float* ssm = wsm + KK * KK * WPB * 8;
  float* hsm = ssm + KK * WPB * 8;
  uint8_t* stage = reinterpret_cast<uint8_t*>(hsm + KK * WPB * 8) + (size_t)warp * 2 * KK * 512;
  rf_stage_weights<K>(wgk, scale, shift, C, blockIdx.y * WPB, WPB, V, wsm, ssm, hsm);
  const int v = blockIdx.y * WPB + warp;
  const bool vact = v < V;
  const long groups = (M + 31) / 32;
  int pbuf = 0;
  if (vact && (long)blockIdx.x < groups) rf_prefetch<K>(stage, x, ldx, M, H, W, s, ho, wo, (long)blockIdx.x * 32 + lane, v, lane);
  cp_async_commit();
  for (long pg = blockIdx.x; pg < groups; pg += gridDim.x, pbuf ^= 1) {
    if (vact && pg + gridDim.x < groups)
      rf_prefetch<K>(stage + (pbuf ^ 1) * KK * 512, x, ldx, M, H, W, s, ho, wo, (pg + gridDim.x) * 32 + lane, v, lane);
    cp_async_commit();
    __syncthreads();  // staged weights visible / previous iteration's smax, ssum consumed
    for (int i = threadIdx.x; i < KK * 32; i += blockDim.x) {
      smax[i / 32][i % 32] = 0;  // features are >= 0 after ReLU: 0 is the identity of the max
      ssum[i / 32][i % 32] = 0.f;
    }
    __syncthreads();
    const long m = pg * 32 + lane;
    const bool pact = m < M;
    const int w = pact ? (int)(m % wo) : 0, h = pact ? (int)((m / wo) % ho) : 0;
    const int n = pact ? (int)(m / ((long)wo * ho)) : 0;
    if (vact) {
      float xin[KK][8];
      cp_async_wait<1>();
      __syncwarp();
      rf_fetch<K>(stage + pbuf * KK * 512, lane, xin);
      // SE pooling: every input pixel is summed exactly once, by the output pixel whose s x s block contains it
      float gsum[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) gsum[e] = 0.f;
#pragma unroll
      for (int i = 0; i < K; ++i)
#pragma unroll
        for (int j = 0; j < K; ++j) {
          const bool own = (K == 1) || (i >= K / 2 && i < K / 2 + s && j >= K / 2 && j < K / 2 + s);
          if (own) {
#pragma unroll
            for (int e = 0; e < 8; ++e) gsum[e] += xin[i * K + j][e];
          }
        }
      const int n_first = __shfl_sync(0xffffffffu, n, 0);
      const int n_last = __shfl_sync(0xffffffffu, pact ? n : -1, 31);
      if (n_first == n_last) {  // whole warp inside one image: reduce over the 32 pixels first
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float r = warp_sum(gsum[e]);
          if (lane == 0) atomicAdd(&gap_sum[(long)n * C + v * 8 + e], r);
        }
      } else if (pact) {
#pragma unroll
        for (int e = 0; e < 8; ++e) atomicAdd(&gap_sum[(long)n * C + v * 8 + e], gsum[e]);
      }
#pragma unroll 1
      for (int t = 0; t < KK; ++t) {
        float g[8];
        rf_tap<K>(xin, wsm, ssm, hsm, WPB, warp, t, g);
        float mx = g[0], sm = g[0];
#pragma unroll
        for (int e = 1; e < 8; ++e) { mx = fmaxf(mx, g[e]); sm += g[e]; }
        atomicMax(&smax[t][lane], __float_as_int(mx));
        atomicAdd(&ssum[t][lane], sm);
        if (G != nullptr && pact) *reinterpret_cast<uint4*>(G + (m * KK + t) * (long)C + v * 8) = pack8(g);
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < KK * 32; i += blockDim.x) {
      const int t = i / 32, l = i % 32;
      const long mm = pg * 32 + l;
      if (mm >= M) continue;
      const int ww = (int)(mm % wo), hh = (int)((mm / wo) % ho);
      const long nn = mm / ((long)wo * ho);
      const int n1 = t / K, n2 = t % K;
      const long pos = (nn * (ho * K) + (hh * K + n1)) * (long)(wo * K) + (ww * K + n2);
      const float mean_part = ssum[t][l] / (float)C;
      if (use_atomics) {
        atomicMax(reinterpret_cast<int*>(maps + pos * 2), smax[t][l]);
        atomicAdd(maps + pos * 2 + 1, mean_part);
      } else {
        *reinterpret_cast<float2*>(maps + pos * 2) = make_float2(__int_as_float(smax[t][l]), mean_part);
      }
    }
  }
}

// blocks [0, N): SE gate of image n;  blocks [N, ...): spatial attention map
__global__ void __launch_bounds__(256) rf_se_att_kernel(const float* __restrict__ gap_sum, int N, int C, float inv_hw,
                                                        const float* __restrict__ w1 /*[R][C]*/,
                                                        const float* __restrict__ w2 /*[C][R]*/, int R,
                                                        float* __restrict__ se, const float* __restrict__ maps,
                                                        int HK, int WK, const float* __restrict__ wa,
                                                        float* __restrict__ att) {
  pdl_enter();
  if ((int)blockIdx.x < N) {
    __shared__ float mean[512];
    __shared__ float hid[64];
    const int n = blockIdx.x;
    for (int c = threadIdx.x; c < C; c += blockDim.x) mean[c] = gap_sum[(long)n * C + c] * inv_hw;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < R; r += (blockDim.x >> 5)) {
      float a = 0.f;
      for (int c = lane; c < C; c += 32) a = fmaf(w1[(long)r * C + c], mean[c], a);
      a = warp_sum(a);
      if (lane == 0) hid[r] = fmaxf(a, 0.f);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float a = 0.f;
      for (int r = 0; r < R; ++r) a = fmaf(w2[(long)c * R + r], hid[r], a);
      se[(long)n * C + c] = 1.f / (1.f + __expf(-a));
    }
    return;
  }
  const long total = (long)N * HK * WK;
  const long t = (long)(blockIdx.x - N) * blockDim.x + threadIdx.x;
  if (t >= total) return;
  int xq, y;
  long n;
  if (total < (1L << 31)) {   // 32-bit divisions: the 64-bit ones were most of this thread's instructions
    const unsigned tu = (unsigned)t, row = tu / (unsigned)WK;
    xq = (int)(tu - row * (unsigned)WK);
    n = (long)(row / (unsigned)HK);
    y = (int)(row - (unsigned)n * (unsigned)HK);
  } else {
    xq = (int)(t % WK);
    y = (int)((t / WK) % HK);
    n = t / ((long)WK * HK);
  }
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int yy = y + i - 1, xx = xq + j - 1;
      if (yy < 0 || yy >= HK || xx < 0 || xx >= WK) continue;
      const float2 mm = __ldg(reinterpret_cast<const float2*>(maps + ((n * HK + yy) * (long)WK + xx) * 2));
      acc = fmaf(mm.x, __ldg(wa + i * 3 + j), acc);
      acc = fmaf(mm.y, __ldg(wa + 9 + i * 3 + j), acc);
    }
  att[t] = 1.f / (1.f + __expf(-acc));
}

// pass 2: A[m][t*C+c] = g * se[n,c] * att[n, h*k+n1, w*k+n2].  grid = (persistent blocks, groups of 8 channel
// vectors); block = 8 warps, warp w owns vector blockIdx.y*8 + w (8 x 16 B = one full 128 B line per (pixel, tap)).
template <int K>
__global__ void __launch_bounds__(256) rf_apply_kernel(const __nv_bfloat16* __restrict__ x, long ldx, long M, int H,
                                                       int W, int V, int s, int ho, int wo,
                                                       const float* __restrict__ wgk,
                                                       const float* __restrict__ scale,
                                                       const float* __restrict__ shift,
                                                       const float* __restrict__ se, const float* __restrict__ att,
                                                       __nv_bfloat16* __restrict__ A) {
  pdl_enter();
  constexpr int KK = K * K;
  constexpr int VB = 8;
  extern __shared__ float dsm[];
  const int C = V * 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* wsm = dsm;
  float* ssm = wsm + KK * KK * VB * 8;
  float* hsm = ssm + KK * VB * 8;
  uint8_t* stage = reinterpret_cast<uint8_t*>(hsm + KK * VB * 8) + (size_t)warp * 2 * KK * 512;
  rf_stage_weights<K>(wgk, scale, shift, C, blockIdx.y * VB, VB, V, wsm, ssm, hsm);
  __syncthreads();
  const int v = blockIdx.y * VB + warp;
  if (v >= V) return;
  const long groups = (M + 31) / 32;
  int pbuf = 0;
  if ((long)blockIdx.x < groups) rf_prefetch<K>(stage, x, ldx, M, H, W, s, ho, wo, (long)blockIdx.x * 32 + lane, v, lane);
  cp_async_commit();
  for (long pg = blockIdx.x; pg < groups; pg += gridDim.x, pbuf ^= 1) {
    if (pg + gridDim.x < groups)
      rf_prefetch<K>(stage + (pbuf ^ 1) * KK * 512, x, ldx, M, H, W, s, ho, wo, (pg + gridDim.x) * 32 + lane, v, lane);
    cp_async_commit();
    const long m = pg * 32 + lane;
    const bool pact = m < M;
    const int w = pact ? (int)(m % wo) : 0, h = pact ? (int)((m / wo) % ho) : 0;
    const int n = pact ? (int)(m / ((long)wo * ho)) : 0;
    float xin[KK][8];
    cp_async_wait<1>();
    __syncwarp();
    rf_fetch<K>(stage + pbuf * KK * 512, lane, xin);
    __syncwarp();
    if (!pact) continue;
    float s8[8];
    {
      const float4 a0 = __ldg(reinterpret_cast<const float4*>(se + (long)n * C + v * 8));
      const float4 a1 = __ldg(reinterpret_cast<const float4*>(se + (long)n * C + v * 8 + 4));
      s8[0] = a0.x; s8[1] = a0.y; s8[2] = a0.z; s8[3] = a0.w; s8[4] = a1.x; s8[5] = a1.y; s8[6] = a1.z; s8[7] = a1.w;
    }
    __nv_bfloat16* arow = A + m * (long)KK * C + v * 8;
#pragma unroll 1
    for (int t = 0; t < KK; ++t) {
      float g[8];
      rf_tap<K>(xin, wsm, ssm, hsm, VB, warp, t, g);
      const int n1 = t / K, n2 = t % K;
      const float a = __ldg(att + ((long)n * (ho * K) + (h * K + n1)) * (long)(wo * K) + (w * K + n2));
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] *= s8[e] * a;
      *reinterpret_cast<uint4*>(arow + (long)t * C) = pack8(g);
    }
  }
}

// pass 2 when pass 1 stored the features: A = G * se[n,c] * att (pure streaming, in place allowed)
__global__ void __launch_bounds__(256) rf_gate2_kernel(const __nv_bfloat16* __restrict__ G, long M, int KK, int K,
                                                       int V, int ho, int wo, const float* __restrict__ se,
                                                       const float* __restrict__ att, __nv_bfloat16* __restrict__ A) {
  pdl_enter();
  const long total = M * KK * V;
  const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int v = (int)(tid % V);
  const int t = (int)((tid / V) % KK);
  const long m = tid / ((long)V * KK);
  const int w = (int)(m % wo), h = (int)((m / wo) % ho);
  const long n = m / ((long)wo * ho);
  const float a = __ldg(att + (n * (ho * K) + (h * K + t / K)) * (long)(wo * K) + (w * K + t % K));
  float g[8];
  unpack8(__ldg(reinterpret_cast<const uint4*>(G + tid * 8)), g);
  const float4 s0 = __ldg(reinterpret_cast<const float4*>(se + n * (V * 8) + v * 8));
  const float4 s1 = __ldg(reinterpret_cast<const float4*>(se + n * (V * 8) + v * 8 + 4));
  g[0] *= s0.x * a; g[1] *= s0.y * a; g[2] *= s0.z * a; g[3] *= s0.w * a;
  g[4] *= s1.x * a; g[5] *= s1.y * a; g[6] *= s1.z * a; g[7] *= s1.w * a;
  *reinterpret_cast<uint4*>(A + tid * 8) = pack8(g);
}

// ---- k = 1 (generate is a per-channel scale): lean streaming kernels --------------------------------------------
// pass 1: g = relu(w'[c]*x + shift[c]);  max_c / mean_c via warp shuffles;  SE sums accumulated per warp.
// One warp per pixel, four pixels in flight per warp (four independent 16-byte loads per lane before any arithmetic).
template <bool TWO>   // TWO: a second channel vector per lane (C > 256)
__global__ void __launch_bounds__(256) rf1_stats_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int HW, int V,
                                                        const float* __restrict__ wgk, const float* __restrict__ scale,
                                                        const float* __restrict__ shift, float* __restrict__ maps,
                                                        float* __restrict__ gap_sum) {
  pdl_enter();
  const int C = V * 8;
  const int n = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const long warps = (long)gridDim.x * wpb;
  const long w0 = (long)blockIdx.x * wpb + (threadIdx.x >> 5);
  constexpr int NQ = TWO ? 2 : 1;
  constexpr bool two = TWO;
  float wv[NQ][8], hv[NQ][8], gs[NQ][8];
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    const int v = lane + q * 32;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const bool ok = v < V;
      wv[q][e] = ok ? wgk[v * 8 + e] * (scale != nullptr ? scale[v * 8 + e] : 1.f) : 0.f;
      hv[q][e] = ok ? shift[v * 8 + e] : 0.f;
      gs[q][e] = 0.f;
    }
  }
  const float invC = 1.f / (float)C;
  const __nv_bfloat16* img = x + (long)n * HW * ldx;
  for (long p0 = w0; p0 < HW; p0 += 4 * warps) {
    uint4 u[4][NQ];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long p = p0 + j * warps;
      u[j][0] = (p < HW && lane < V) ? ldx16(img + p * ldx + lane * 8) : make_uint4(0, 0, 0, 0);
      if (two) u[j][NQ - 1] = (p < HW && lane + 32 < V) ? ldx16(img + p * ldx + (lane + 32) * 8) : make_uint4(0, 0, 0, 0);
    }
    float mx[4], sm[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      mx[j] = 0.f; sm[j] = 0.f;
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        float f[8];
        unpack8(u[j][q], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          gs[q][e] += f[e];                       // zero vectors (inactive lanes / pixels) add nothing
          const float g = fmaxf(fmaf(f[e], wv[q][e], hv[q][e]), 0.f);
          mx[j] = fmaxf(mx[j], g);
          sm[j] += g;
        }
      }
      if (lane >= V) { mx[j] = 0.f; sm[j] = 0.f; }   // inactive lanes would contribute relu(shift)
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], o));
        sm[j] += __shfl_xor_sync(0xffffffffu, sm[j], o);
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long p = p0 + j * warps;
      if (lane == j && p < HW) *reinterpret_cast<float2*>(maps + ((long)n * HW + p) * 2) = make_float2(mx[j], sm[j] * invC);
    }
  }
  // SE sums: the block's warps are reduced in shared memory first (8x fewer same-address atomics)
  __shared__ float red[8][NQ * 256];
  const int wib = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < NQ; ++q)
#pragma unroll
    for (int e = 0; e < 8; ++e) red[wib][(q * 32 + lane) * 8 + e] = gs[q][e];
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f;
    for (int w = 0; w < wpb; ++w) a += red[w][c];
    atomicAdd(&gap_sum[(long)n * C + c], a);
  }
}

// pass 2: A = relu(w'*x + shift) * se[n,c] * att[n,pixel] = relu(x * (w'*se) + shift*se) * att   (se > 0).
// A thread keeps ONE 8-channel vector for its whole life, so the per-channel coefficients are registers and the loop
// body is one 16-byte load, one broadcast attention load, 16 FMAs/MAXs and one 16-byte store; block = rows x V threads.
__global__ void __launch_bounds__(256) rf1_apply_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int HW, int V,
                                                        const float* __restrict__ wgk, const float* __restrict__ scale,
                                                        const float* __restrict__ shift, const float* __restrict__ se,
                                                        const float* __restrict__ att, __nv_bfloat16* __restrict__ A) {
  pdl_enter();
  const int n = blockIdx.y;
  const int rows = 256 / V;
  const int v = threadIdx.x % V, row = threadIdx.x / V;
  if (row >= rows) return;
  const int C = V * 8;
  float wse[8], hse[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int c = v * 8 + e;
    const float g = __ldg(se + (long)n * C + c);
    wse[e] = __ldg(wgk + c) * (scale != nullptr ? __ldg(scale + c) : 1.f) * g;
    hse[e] = __ldg(shift + c) * g;
  }
  const __nv_bfloat16* xin = x + (long)n * HW * ldx + v * 8;
  __nv_bfloat16* out = A + (long)n * HW * C + v * 8;
  const float* at = att + (long)n * HW;
  const long stride = (long)gridDim.x * rows;
  long p = (long)blockIdx.x * rows + row;
  for (; p + 3 * stride < HW; p += 4 * stride) {
    uint4 u[4];
    float a[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { u[j] = ldx16(xin + (p + j * stride) * ldx); a[j] = __ldg(at + p + j * stride); }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float f[8];
      unpack8(u[j], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = fmaxf(fmaf(f[e], wse[e], hse[e]), 0.f) * a[j];
      *reinterpret_cast<uint4*>(out + (p + j * stride) * C) = pack8(f);
    }
  }
  for (; p < HW; p += stride) {
    float f[8];
    unpack8(ldx16(xin + p * ldx), f);
    const float a = __ldg(at + p);
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = fmaxf(fmaf(f[e], wse[e], hse[e]), 0.f) * a;
    *reinterpret_cast<uint4*>(out + p * C) = pack8(f);
  }
}

}  // namespace ly

using namespace ly;

extern "C" int ly_rfcbam_front_infer(const void* x, long ldx, int N, int H, int W, int C, int k, int s, int ho,
                                     int wo, const float* wgk, const float* scale, const float* shift,
                                     const float* w1, const float* w2, int R, const float* wa, float* maps,
                                     float* gap_sum, float* se, float* att, void* A, ly_stream_t st) {
  LY_REQUIRE(x && wgk && scale && shift && w1 && w2 && wa && maps && gap_sum && se && att && A,
             "ly_rfcbam_front_infer: null pointer");
  LY_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0 && C <= 512, "ly_rfcbam_front_infer: bad sizes (C=%d)", C);
  LY_REQUIRE((k == 1 && s == 1) || (k == 3 && (s == 1 || s == 2)), "ly_rfcbam_front_infer: k=%d s=%d unsupported", k, s);
  LY_REQUIRE(ho == (H + 2 * (k / 2) - k) / s + 1 && wo == (W + 2 * (k / 2) - k) / s + 1,
             "ly_rfcbam_front_infer: output size mismatch");
  // the SE sums visit every input pixel exactly once only if the s x s blocks tile the input
  LY_REQUIRE(s == 1 || (H % 2 == 0 && W % 2 == 0), "ly_rfcbam_front_infer: stride 2 needs even H, W");
  const auto al16 = [](const void* p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  LY_REQUIRE(al16(x) && al16(A) && al16(wgk) && al16(se) && al16(scale) && al16(shift) && al16(maps) &&
                 ldx % 8 == 0 && ldx >= C,
             "ly_rfcbam_front_infer: alignment / ld");
  LY_REQUIRE(R > 0 && R <= 64, "ly_rfcbam_front_infer: R=%d out of range", R);
  cudaStream_t stream = (cudaStream_t)st;
  const int V = C / 8;
  const long M = (long)N * ho * wo;
  const int vgroups = ceil_div(V, 16);
  const int WPB = ceil_div(V, vgroups);
  LY_CUDA(cudaMemsetAsync(gap_sum, 0, sizeof(float) * (size_t)N * C, stream));
  if (vgroups > 1 && k != 1) LY_CUDA(cudaMemsetAsync(maps, 0, sizeof(float) * (size_t)N * ho * k * wo * k * 2, stream));
  const int KK = k * k;
  const long groups = (M + 31) / 32;
  const int sms = device_sm_count();
  {
    long bx = groups < (long)sms ? groups : (long)sms;      // one block per SM (512 threads, 128 registers)
    dim3 grid1((unsigned)bx, (unsigned)vgroups);
    const size_t sm1 = sizeof(float) * (size_t)(KK * KK + 2 * KK) * WPB * 8 + (size_t)WPB * 2 * KK * 512;
    if (k == 1) {
      int bx1 = ceil_div((long)sms * 8, N);
      if (bx1 > ceil_div(H * W, 32)) bx1 = ceil_div(H * W, 32);
      if (bx1 < 1) bx1 = 1;
      if (V > 32)
        LY_CUDA(launch(rf1_stats_kernel<true>, dim3((unsigned)bx1, (unsigned)N), 256, 0, stream, (const __nv_bfloat16*)x, ldx, H * W,
                                                                                    V, wgk, scale, shift, maps, gap_sum));
      else
        LY_CUDA(launch(rf1_stats_kernel<false>, dim3((unsigned)bx1, (unsigned)N), 256, 0, stream, (const __nv_bfloat16*)x, ldx, H * W,
                                                                                     V, wgk, scale, shift, maps, gap_sum));
    } else {
      static std::atomic<bool> attr_done[64];   // zero-initialised; set only after the attribute call succeeded
      int dev = 0;
      LY_CUDA(cudaGetDevice(&dev));
      if (!attr_done[dev & 63].load()) {
        LY_CUDA(cudaFuncSetAttribute(rf_stats_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024));
        LY_CUDA(cudaFuncSetAttribute(rf_apply_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
        attr_done[dev & 63].store(true);
      }
      // k = 3: the features cost k^4 = 81 MACs per channel; store them once (in A) instead of recomputing in pass 2
      LY_CUDA(launch(rf_stats_kernel<3>, grid1, WPB * 32, sm1, stream, (const __nv_bfloat16*)x, ldx, M, H, W, V, WPB, s, ho, wo,
                                                          wgk, scale, shift, maps, gap_sum, vgroups > 1,
                                                          (__nv_bfloat16*)A));
    }
    LY_LAUNCH_CHECK();
  }
  const long tot_att = (long)N * ho * k * wo * k;
  LY_CUDA(launch(rf_se_att_kernel, N + ceil_div(tot_att, 256), 256, 0, stream, gap_sum, N, C, 1.f / (float)(H * W), w1, w2, R, se,
                                                                  maps, ho * k, wo * k, wa, att));
  LY_LAUNCH_CHECK();
  {
    const int vg8 = ceil_div(V, 8);
    long bx = ceil_div((long)sms * 4, vg8);                 // ~4 resident blocks of 256 threads per SM
    if (bx > groups) bx = groups;
    if (bx < 1) bx = 1;
    dim3 grid2((unsigned)bx, (unsigned)vg8);
    const size_t sm2 = sizeof(float) * (size_t)(KK * KK + 2 * KK) * 8 * 8 + (size_t)8 * 2 * KK * 512;
    if (k == 1)
    {
      const int rows = 256 / V;
      long bxa = ceil_div((long)H * W, (long)rows * 8);     // ~8 pixels per thread
      const long cap = ceil_div((long)sms * 16, (long)N);
      if (bxa > cap) bxa = cap;
      if (bxa < 1) bxa = 1;
      LY_CUDA(launch(rf1_apply_kernel, dim3((unsigned)bxa, (unsigned)N), 256, 0, stream, (const __nv_bfloat16*)x, ldx, H * W, V, wgk,
                                                                            scale, shift, se, att, (__nv_bfloat16*)A));
    }
    else
      LY_CUDA(launch(rf_gate2_kernel, ceil_div(M * KK * V, 256), 256, 0, stream, (const __nv_bfloat16*)A, M, KK, k, V, ho, wo, se,
                                                                    att, (__nv_bfloat16*)A));
  }
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// ---- building blocks of the k = 3 inference pipeline whose feature generation runs on the tensor cores
// (csrc/rf_generate_tc.cu): SE channel sums, SE gate + spatial attention, gating of the stored features.
namespace ly {
// block = (Vp lanes over 16-byte channel vectors) x (256 / Vp pixel rows); 4 independent loads in flight per thread
template <bool PARTS>
__global__ void __launch_bounds__(256) rf_gap_sum_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int HW, int V,
                                                         int Vp, float* __restrict__ gap_sum) {
  pdl_enter();
  __shared__ float red[256 * 8];
  const int n = blockIdx.y;
  const int v = threadIdx.x % Vp, row = threadIdx.x / Vp, rows = 256 / Vp;
  const long stride = (long)gridDim.x * rows;
  float gs[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) gs[e] = 0.f;
  if (v < V) {
    const __nv_bfloat16* base = x + (long)n * HW * ldx + v * 8;
    long p = (long)blockIdx.x * rows + row;
    for (; p + 3 * stride < HW; p += 4 * stride) {
      uint4 u[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) u[j] = ldx16(base + (p + j * stride) * ldx);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float f[8];
        unpack8(u[j], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) gs[e] += f[e];
      }
    }
    for (; p < HW; p += stride) {
      float f[8];
      unpack8(ldx16(base + p * ldx), f);
#pragma unroll
      for (int e = 0; e < 8; ++e) gs[e] += f[e];
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[e * 256 + threadIdx.x] = gs[e];
  __syncthreads();
  // thread (e, v) sums the rows of its column
  for (int i = threadIdx.x; i < V * 8; i += 256) {
    const int e = i % 8, vv = i / 8;
    float a = 0.f;
    for (int r = 0; r < rows; ++r) a += red[e * 256 + r * Vp + vv];
    if (PARTS) gap_sum[((long)blockIdx.x * gridDim.y + n) * V * 8 + vv * 8 + e] = a;   // this block's partial
    else atomicAdd(&gap_sum[(long)n * V * 8 + vv * 8 + e], a);
  }
}

// out[n][c] = bf16( (Σ_b parts[b][n][c]) / HW ), the partials added in index order
__global__ void gap_finish_kernel(const float* __restrict__ parts, int B, long NC, float inv, __nv_bfloat16* __restrict__ out) {
  pdl_enter();
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= NC) return;
  float a = 0.f;
  for (int b = 0; b < B; ++b) a += parts[(long)b * NC + i];
  out[i] = __float2bfloat16(a * inv);
}
}  // namespace ly

// gap_sum[n][c] = sum over the H*W pixels of x (fp32; zeroed here first)
extern "C" int ly_rf_gap_sum(const void* x, long ldx, int N, int HW, int C, float* gap_sum, ly_stream_t st) {
  LY_REQUIRE(x && gap_sum && C % 8 == 0 && C <= 512 && ldx % 8 == 0 && ldx >= C, "ly_rf_gap_sum: bad args");
  cudaStream_t stream = (cudaStream_t)st;
  LY_CUDA(cudaMemsetAsync(gap_sum, 0, sizeof(float) * (size_t)N * C, stream));
  const int V = C / 8;
  int Vp = 1;
  while (Vp < V) Vp *= 2;
  const int rows = 256 / Vp;
  long bx = ceil_div((long)HW, (long)rows * 8);
  const long cap = ceil_div(8L * device_sm_count(), (long)N);
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  LY_CUDA(launch(rf_gap_sum_kernel<false>, dim3((unsigned)bx, (unsigned)N), 256, 0, stream, (const __nv_bfloat16*)x, ldx, HW, V, Vp,
                                                                               gap_sum));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// Deterministic global average pool for the training path (rfa.SE, models/rfa.py:80,90): LY_GAP_PARTS blocks per image
// store their partial sums, a second launch adds them in order and writes the bf16 mean [N][C].
extern "C" int ly_gap_mean(const void* x, long ldx, int N, int HW, int C, float* parts, void* out, ly_stream_t st) {
  LY_REQUIRE(x && parts && out && C % 8 == 0 && C <= 512 && ldx % 8 == 0 && ldx >= C && N > 0 && HW > 0,
             "ly_gap_mean: bad args");
  cudaStream_t stream = (cudaStream_t)st;
  const int V = C / 8;
  int Vp = 1;
  while (Vp < V) Vp *= 2;
  LY_CUDA(launch(rf_gap_sum_kernel<true>, dim3(LY_GAP_PARTS, (unsigned)N), 256, 0, stream, (const __nv_bfloat16*)x, ldx, HW, V, Vp,
                                                                              parts));
  LY_LAUNCH_CHECK();
  const long NC = (long)N * C;
  LY_CUDA(launch(gap_finish_kernel, ceil_div(NC, 256), 256, 0, stream, parts, LY_GAP_PARTS, NC, 1.f / (float)HW,
                                                           (__nv_bfloat16*)out));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// se[n][c] = sigmoid(W2 relu(W1 gap_sum/HW));  att = sigmoid(conv3x3(maps))   (rfa.py:88-92, 106, 127)
extern "C" int ly_rf_se_att(const float* gap_sum, int N, int C, int HW, const float* w1, const float* w2, int R,
                            const float* maps, int HK, int WK, const float* wa, float* se, float* att,
                            ly_stream_t st) {
  LY_REQUIRE(gap_sum && w1 && w2 && maps && wa && se && att && R > 0 && R <= 64 && C <= 512, "ly_rf_se_att: bad args");
  const long tot = (long)N * HK * WK;
  LY_CUDA(launch(rf_se_att_kernel, N + ceil_div(tot, 256), 256, 0, (cudaStream_t)st, gap_sum, N, C, 1.f / (float)HW, w1, w2, R, se,
                                                                        maps, HK, WK, wa, att));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
