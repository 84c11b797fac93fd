// RFCBAMConv (reference models/rfa.py:95-129) — receptive-field feature generation,
// channel max/mean, 3x3 spatial-attention conv and gating, forward and backward.  The final k x k / stride k
// convolution over the unfolded map is a plain GEMM with K = k*k*C on the gated matrix (conv_gemm.cu).
//
// Layouts (h, w = output resolution; t = n1*k + n2 the position inside the k x k receptive field):
//   G / A  : bf16 [N*h*w][k*k*C], column t*C + c   == U[b, c, h*k+n1, w*k+n2] of the reference
//   maps   : fp32 [N][h*k][w*k][2]  (0: max over c, 1: mean over c), argmax int16 [N][h*k][w*k]
//   att    : fp32 [N][h*k][w*k]
//   wgk    : fp32 [k*k (t)][k*k (i*k+j)][C]  (generate.0.weight[c*k*k+t, 0, i, j])
#include <stdlib.h>

#include <mutex>

#include "ly_common.cuh"

namespace ly {

__device__ __forceinline__ uint4 ldg16(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// ------------------------------------------------------------------------------------------------
// generate: depthwise k x k, stride s, pad k/2, C -> k*k*C, optional folded BN + ReLU epilogue.
// One thread per (output pixel, 8-channel vector); computes all k*k taps t for those 8 channels.
// ------------------------------------------------------------------------------------------------
template <int K>
__global__ void __launch_bounds__(128) rf_generate_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int N,
                                                          int H, int W, int V, int s, int ho, int wo,
                                                          const float* __restrict__ wgk,
                                                          const float* __restrict__ scale,
                                                          const float* __restrict__ shift, int act,
                                                          __nv_bfloat16* __restrict__ G) {
  pdl_enter();
  constexpr int KK = K * K;
  const int C = V * 8;
  const long total = (long)N * ho * wo * V;
  const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int v = (int)(tid % V);
  const long m = tid / V;
  const int w = (int)(m % wo), h = (int)((m / wo) % ho), n = (int)(m / ((long)wo * ho));
  float xin[KK][8];
#pragma unroll
  for (int i = 0; i < K; ++i)
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int yy = h * s + i - K / 2, xx = w * s + j - K / 2;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
        unpack8(ldg16(x + (((long)n * H + yy) * W + xx) * ldx + v * 8), xin[i * K + j]);
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) xin[i * K + j][e] = 0.f;
      }
    }
  __nv_bfloat16* grow = G + m * (long)KK * C + v * 8;
#pragma unroll 1
  for (int t = 0; t < KK; ++t) {
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int ij = 0; ij < KK; ++ij) {
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(wgk + ((long)t * KK + ij) * C + v * 8));
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(wgk + ((long)t * KK + ij) * C + v * 8 + 4));
      acc[0] = fmaf(xin[ij][0], w0.x, acc[0]); acc[1] = fmaf(xin[ij][1], w0.y, acc[1]);
      acc[2] = fmaf(xin[ij][2], w0.z, acc[2]); acc[3] = fmaf(xin[ij][3], w0.w, acc[3]);
      acc[4] = fmaf(xin[ij][4], w1.x, acc[4]); acc[5] = fmaf(xin[ij][5], w1.y, acc[5]);
      acc[6] = fmaf(xin[ij][6], w1.z, acc[6]); acc[7] = fmaf(xin[ij][7], w1.w, acc[7]);
    }
    if (scale != nullptr) {
#pragma unroll
      for (int e = 0; e < 8; ++e)
        acc[e] = act_apply(fmaf(acc[e], __ldg(scale + t * C + v * 8 + e), __ldg(shift + t * C + v * 8 + e)), act);
    }
    *reinterpret_cast<uint4*>(grow + (long)t * C) = pack8(acc);
  }
}

// ------------------------------------------------------------------------------------------------
// channel max / mean of G per (pixel, tap).  LPR lanes share one (m, t) row of C channels, each lane takes every
// LPR-th 16-byte vector (four per lane for C = 128 / 256, all loads issued before the first use), so a warp covers
// 32 / LPR rows at once.  (One warp per row left half the lanes idle at C = 128 and ran at 12 % of the HBM rate.)
// ------------------------------------------------------------------------------------------------
template <int LPR>
__global__ void __launch_bounds__(256) rf_poolc_kernel(const __nv_bfloat16* __restrict__ G, long M, int KK, int K,
                                                       int V, int ho, int wo, float* __restrict__ maps,
                                                       int16_t* __restrict__ amax) {
  pdl_enter();
  const int sub = threadIdx.x % LPR;
  const long row = ((long)blockIdx.x * blockDim.x + threadIdx.x) / LPR;
  const bool live = row < M * KK;
  const int C = V * 8;
  const __nv_bfloat16* rp = G + (live ? row : 0) * (long)C;
  float mx = -INFINITY, sum = 0.f;
  int mi = 0;
  for (int v0 = sub; v0 < V; v0 += 4 * LPR) {
    uint4 u[4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (v0 + q * LPR < V) u[q] = ldg16(rp + (v0 + q * LPR) * 8);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int v = v0 + q * LPR;
      if (v < V) {
        float f[8];
        unpack8(u[q], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          sum += f[e];
          if (f[e] > mx) { mx = f[e]; mi = v * 8 + e; }
        }
      }
    }
  }
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
    if (om > mx || (om == mx && oi < mi)) { mx = om; mi = oi; }
  }
  if (live && sub == 0) {
    const int t = (int)(row % KK);
    const long m = row / KK;
    const int w = (int)(m % wo), h = (int)((m / wo) % ho);
    const long n = m / ((long)wo * ho);
    const int n1 = t / K, n2 = t % K;
    const long pos = (n * (ho * K) + (h * K + n1)) * (long)(wo * K) + (w * K + n2);
    maps[pos * 2] = mx;
    maps[pos * 2 + 1] = sum / (float)C;
    if (amax != nullptr) amax[pos] = (int16_t)mi;
  }
}

static int poolc_launch(const void* G, long M, int KK, int k, int V, int ho, int wo, float* maps, int16_t* amax,
                        cudaStream_t st) {
  const long rows = M * KK;
  const __nv_bfloat16* g = (const __nv_bfloat16*)G;
  if (V <= 16)
    LY_CUDA(launch(rf_poolc_kernel<4>, ceil_div(rows * 4, 256), 256, 0, st, g, M, KK, k, V, ho, wo, maps, amax));
  else if (V <= 32)
    LY_CUDA(launch(rf_poolc_kernel<8>, ceil_div(rows * 8, 256), 256, 0, st, g, M, KK, k, V, ho, wo, maps, amax));
  else
    LY_CUDA(launch(rf_poolc_kernel<32>, ceil_div(rows * 32, 256), 256, 0, st, g, M, KK, k, V, ho, wo, maps, amax));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// att = sigmoid(conv3x3_{2->1, pad 1}(maps))   (get_weight, rfa.py:106,127)
__global__ void __launch_bounds__(256) rf_att_kernel(const float* __restrict__ maps, int N, int HK, int WK,
                                                     const float* __restrict__ wa /*[2][3][3]*/,
                                                     float* __restrict__ att) {
  pdl_enter();
  const long total = (long)N * HK * WK;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int x = (int)(t % WK), y = (int)((t / WK) % HK);
  const long n = t / ((long)WK * HK);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int yy = y + i - 1, xx = x + j - 1;
      if (yy < 0 || yy >= HK || xx < 0 || xx >= WK) continue;
      const float2 mm = __ldg(reinterpret_cast<const float2*>(maps + ((n * HK + yy) * (long)WK + xx) * 2));
      acc = fmaf(mm.x, __ldg(wa + i * 3 + j), acc);
      acc = fmaf(mm.y, __ldg(wa + 9 + i * 3 + j), acc);
    }
  att[t] = 1.f / (1.f + __expf(-acc));
}

// A = G * se[n, c] * att[n, h*k+n1, w*k+n2]     (also used for dG1 = dA * se * att)
// LPR lanes share one (m, t) row, four 16-byte vectors per lane: the row's index arithmetic (three divisions) and its
// attention value are paid once per four vectors and the four loads are in flight together.
template <int LPR>
__global__ void __launch_bounds__(256) rf_gate_kernel(const __nv_bfloat16* __restrict__ G, long M, int KK, int K,
                                                      int V, int ho, int wo, const __nv_bfloat16* __restrict__ se,
                                                      long ldse, const float* __restrict__ att,
                                                      __nv_bfloat16* __restrict__ A) {
  pdl_enter();
  const int sub = threadIdx.x % LPR;
  const long row = ((long)blockIdx.x * blockDim.x + threadIdx.x) / LPR;
  if (row >= M * KK) return;
  const int t = (int)(row % KK);
  const long m = row / KK;
  const int w = (int)(m % wo), h = (int)((m / wo) % ho);
  const long n = m / ((long)wo * ho);
  const int n1 = t / K, n2 = t % K;
  const float a = __ldg(att + (n * (ho * K) + (h * K + n1)) * (long)(wo * K) + (w * K + n2));
  const __nv_bfloat16* grow = G + row * (long)(V * 8);
  __nv_bfloat16* arow = A + row * (long)(V * 8);
  const __nv_bfloat16* srow = se + n * ldse;
  for (int v0 = sub; v0 < V; v0 += 4 * LPR) {
    uint4 ug[4], us[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int v = v0 + q * LPR;
      if (v < V) {
        ug[q] = ldg16(grow + v * 8);
        us[q] = ldg16(srow + v * 8);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int v = v0 + q * LPR;
      if (v < V) {
        float g[8], s8[8];
        unpack8(ug[q], g);
        unpack8(us[q], s8);
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] *= s8[e] * a;
        *reinterpret_cast<uint4*>(arow + v * 8) = pack8(g);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Backward of the gating: given dA, G, se, att ->
//   dG1 = dA*se*att (written to dG), dse[part][n,c] = this block's share of Σ dA*G*att, datt[pos] = Σ_c dA*G*se.
// grid = (blocks per image, N).  LPR lanes share one (m, t) row (each lane up to four 16-byte channel vectors of dA and of
// G, all eight loads in flight before the first use); the 256 / LPR row groups of a block stride over the rows of its
// image.  (One warp per row: half the lanes idle at C = 128, one row in flight per warp, 31 % of the HBM rate.)
// ------------------------------------------------------------------------------------------------
template <int LPR>
__global__ void __launch_bounds__(256) rf_gate_bwd_kernel(const __nv_bfloat16* __restrict__ dA,
                                                          const __nv_bfloat16* __restrict__ G, int KK, int K, int V,
                                                          int ho, int wo, const __nv_bfloat16* __restrict__ se,
                                                          long ldse, const float* __restrict__ att,
                                                          __nv_bfloat16* __restrict__ dG, float* __restrict__ dse,
                                                          float* __restrict__ datt) {
  pdl_enter();
  constexpr int QV = 4;                       // channel vectors per lane: C <= LPR * 32
  constexpr int GPB = 256 / LPR;              // row groups per block
  const int lane = threadIdx.x & 31;
  const int sub = threadIdx.x % LPR;
  const int grp = threadIdx.x / LPR;
  const int n = blockIdx.y;
  const int C = V * 8;
  const long per_img = (long)ho * wo * KK;
  const long step = (long)gridDim.x * GPB;
  float s8[QV][8], ds[QV][8];
#pragma unroll
  for (int q = 0; q < QV; ++q) {
    const int v = sub + q * LPR;
#pragma unroll
    for (int e = 0; e < 8; ++e) { s8[q][e] = 0.f; ds[q][e] = 0.f; }
    if (v < V) unpack8(ldg16(se + (long)n * ldse + v * 8), s8[q]);
  }
  // every thread of the block runs the same number of iterations (the shuffles below need whole warps)
  for (long base = (long)blockIdx.x * GPB; base < per_img; base += step) {
    const long i = base + grp;
    const bool live = i < per_img;
    const int t = (int)(i % KK);
    const long ml = i / KK;  // pixel within image
    const int w = (int)(ml % wo), h = (int)(ml / wo);
    const int n1 = t / K, n2 = t % K;
    const long pos = ((long)n * (ho * K) + (h * K + n1)) * (long)(wo * K) + (w * K + n2);
    const long rowoff = (((long)n * ho * wo + ml) * KK + t) * (long)C;
    uint4 ud[QV], ug[QV];
#pragma unroll
    for (int q = 0; q < QV; ++q) {
      const int v = sub + q * LPR;
      if (live && v < V) {
        ud[q] = ldg16(dA + rowoff + v * 8);
        ug[q] = ldg16(G + rowoff + v * 8);
      }
    }
    const float a = live ? __ldg(att + pos) : 0.f;
    float da_att = 0.f;
#pragma unroll
    for (int q = 0; q < QV; ++q) {
      const int v = sub + q * LPR;
      if (live && v < V) {
        float d[8], g[8];
        unpack8(ud[q], d);
        unpack8(ug[q], g);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float dg = d[e] * g[e];
          ds[q][e] = fmaf(dg, a, ds[q][e]);
          da_att = fmaf(dg, s8[q][e], da_att);
          d[e] *= s8[q][e] * a;
        }
        *reinterpret_cast<uint4*>(dG + rowoff + v * 8) = pack8(d);
      }
    }
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) da_att += __shfl_xor_sync(0xffffffffu, da_att, o);
    if (live && sub == 0) datt[pos] = da_att;
  }
  // Deterministic reduction: the row groups of a warp are folded by shuffles, the eight warps of the block are summed in
  // warp order through shared memory and the block STORES its partial at dse[blockIdx.x][n][c]; the caller adds the
  // gridDim.x partials in order.  (dse is broadcast back over all pixels of the image after a bf16 rounding: with
  // order-dependent float atomics a rounding flip of one value shifted a whole channel's input gradient from run to run.)
#pragma unroll
  for (int q = 0; q < QV; ++q)
#pragma unroll
    for (int e = 0; e < 8; ++e)
#pragma unroll
      for (int o = LPR; o < 32; o <<= 1) ds[q][e] += __shfl_xor_sync(0xffffffffu, ds[q][e], o);
  __shared__ float red[8][512];
  const int warp = threadIdx.x >> 5;
  if (lane < LPR) {
#pragma unroll
    for (int q = 0; q < QV; ++q) {
      const int v = sub + q * LPR;
      if (v < V) {
#pragma unroll
        for (int e = 0; e < 8; ++e) red[warp][v * 8 + e] = ds[q][e];
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f;
#pragma unroll
    for (int w8 = 0; w8 < 8; ++w8) a += red[w8][c];
    dse[((long)blockIdx.x * gridDim.y + n) * C + c] = a;
  }
}

// Backward of att = sigmoid(conv3x3(maps)): dmaps (transposed conv) and dWa (18 values, atomics).
__global__ void __launch_bounds__(256) rf_att_bwd_kernel(const float* __restrict__ datt, const float* __restrict__ att,
                                                         const float* __restrict__ maps, int N, int HK, int WK,
                                                         const float* __restrict__ wa, float* __restrict__ dmaps,
                                                         float* __restrict__ dwa) {
  pdl_enter();
  __shared__ float sw[18];
  if (threadIdx.x < 18) sw[threadIdx.x] = 0.f;
  __syncthreads();
  const long total = (long)N * HK * WK;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  float lw[18];
#pragma unroll
  for (int q = 0; q < 18; ++q) lw[q] = 0.f;
  if (t < total) {
    const int x = (int)(t % WK), y = (int)((t / WK) % HK);
    const long n = t / ((long)WK * HK);
    // dpre at this position (for dWa) and at neighbours (for dmaps)
    const float a0 = att[t];
    const float dp0 = datt[t] * a0 * (1.f - a0);
    float dm0 = 0.f, dm1 = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        // forward: pre[y,x] += wa[ch,i,j] * maps[y+i-1, x+j-1, ch]
        const int yy = y + i - 1, xx = x + j - 1;
        if (yy >= 0 && yy < HK && xx >= 0 && xx < WK) {
          const float2 mm = __ldg(reinterpret_cast<const float2*>(maps + ((n * HK + yy) * (long)WK + xx) * 2));
          lw[i * 3 + j] = dp0 * mm.x;
          lw[9 + i * 3 + j] = dp0 * mm.y;
        }
        // transposed: dmaps[y,x,ch] += wa[ch,i,j] * dpre[y-i+1, x-j+1]
        const int yo = y - i + 1, xo = x - j + 1;
        if (yo >= 0 && yo < HK && xo >= 0 && xo < WK) {
          const long o = (n * HK + yo) * (long)WK + xo;
          const float ao = __ldg(att + o);
          const float dpo = __ldg(datt + o) * ao * (1.f - ao);
          dm0 = fmaf(__ldg(wa + i * 3 + j), dpo, dm0);
          dm1 = fmaf(__ldg(wa + 9 + i * 3 + j), dpo, dm1);
        }
      }
    dmaps[t * 2] = dm0;
    dmaps[t * 2 + 1] = dm1;
  }
#pragma unroll
  for (int q = 0; q < 18; ++q) {
    const float s = warp_sum(lw[q]);
    if ((threadIdx.x & 31) == 0) atomicAdd(&sw[q], s);
  }
  __syncthreads();
  if (threadIdx.x < 18) atomicAdd(&dwa[threadIdx.x], sw[threadIdx.x]);
}

// dG += dmean/C + [c == argmax] * dmax       (in place on dG); LPR lanes per (m, t) row as in rf_gate_kernel
template <int LPR>
__global__ void __launch_bounds__(256) rf_poolc_bwd_kernel(__nv_bfloat16* __restrict__ dG, long M, int KK, int K,
                                                           int V, int ho, int wo, const float* __restrict__ dmaps,
                                                           const int16_t* __restrict__ amax) {
  pdl_enter();
  const int sub = threadIdx.x % LPR;
  const long row = ((long)blockIdx.x * blockDim.x + threadIdx.x) / LPR;
  if (row >= M * KK) return;
  const int t = (int)(row % KK);
  const long m = row / KK;
  const int w = (int)(m % wo), h = (int)((m / wo) % ho);
  const long n = m / ((long)wo * ho);
  const int n1 = t / K, n2 = t % K;
  const long pos = (n * (ho * K) + (h * K + n1)) * (long)(wo * K) + (w * K + n2);
  const float2 dm = __ldg(reinterpret_cast<const float2*>(dmaps + pos * 2));
  const int am = amax[pos];
  const float dmean = dm.y / (float)(V * 8);
  __nv_bfloat16* drow = dG + row * (long)(V * 8);
  for (int v0 = sub; v0 < V; v0 += 4 * LPR) {
    uint4 u[4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (v0 + q * LPR < V) u[q] = *reinterpret_cast<const uint4*>(drow + (v0 + q * LPR) * 8);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int v = v0 + q * LPR;
      if (v < V) {
        float d[8];
        unpack8(u[q], d);
#pragma unroll
        for (int e = 0; e < 8; ++e) d[e] += dmean + ((v * 8 + e) == am ? dm.x : 0.f);
        *reinterpret_cast<uint4*>(drow + v * 8) = pack8(d);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Backward of generate w.r.t. the input: dx[n,y,x,c] = Σ_{t,i,j} wgk[t][i*k+j][c] * dZ[m(h,w)][t*C+c]
// over the (h,w) with h*s + i - k/2 == y, w*s + j - k/2 == x;  plus extra[n,c] (SE pooling gradient, already
// divided by H*W) when given.  One thread per (input pixel, 8-channel vector).
// ------------------------------------------------------------------------------------------------
template <int K>
__global__ void __launch_bounds__(128) rf_generate_dx_kernel(const __nv_bfloat16* __restrict__ dZ, int N, int H, int W,
                                                             int V, int s, int ho, int wo,
                                                             const float* __restrict__ wgk,
                                                             const float* __restrict__ extra,
                                                             __nv_bfloat16* __restrict__ dx, long lddx) {
  pdl_enter();
  constexpr int KK = K * K;
  const int C = V * 8;
  const long total = (long)N * H * W * V;
  const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int v = (int)(tid % V);
  const long pix = tid / V;
  const int x = (int)(pix % W), y = (int)((pix / W) % H), n = (int)(pix / ((long)W * H));
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = extra != nullptr ? __ldg(extra + (long)n * C + v * 8 + e) : 0.f;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    const int hn = y + K / 2 - i;
    if (hn < 0 || hn % s != 0) continue;
    const int h = hn / s;
    if (h >= ho) continue;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int wn = x + K / 2 - j;
      if (wn < 0 || wn % s != 0) continue;
      const int w = wn / s;
      if (w >= wo) continue;
      const long m = ((long)n * ho + h) * wo + w;
      const int ij = i * K + j;
#pragma unroll 1   // (measured: unrolling the nine loads is 13 % slower — the grid already hides the latency)
      for (int t = 0; t < KK; ++t) {
        float d[8];
        unpack8(ldg16(dZ + (m * KK + t) * (long)C + v * 8), d);
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(wgk + ((long)t * KK + ij) * C + v * 8));
        const float4 w1 = __ldg(reinterpret_cast<const float4*>(wgk + ((long)t * KK + ij) * C + v * 8 + 4));
        acc[0] = fmaf(d[0], w0.x, acc[0]); acc[1] = fmaf(d[1], w0.y, acc[1]);
        acc[2] = fmaf(d[2], w0.z, acc[2]); acc[3] = fmaf(d[3], w0.w, acc[3]);
        acc[4] = fmaf(d[4], w1.x, acc[4]); acc[5] = fmaf(d[5], w1.y, acc[5]);
        acc[6] = fmaf(d[6], w1.z, acc[6]); acc[7] = fmaf(d[7], w1.w, acc[7]);
      }
    }
  }
  *reinterpret_cast<uint4*>(dx + pix * lddx + v * 8) = pack8(acc);
}

// col2im of the patch gradients: dx[n,y,x,c] = Σ_{(i,j)} R[m(h,w)][i*3+j][c] over the (h,w) with h*s + i - 1 == y,
// w*s + j - 1 == x (at most four terms for s = 2), plus extra[n,c].  R comes from ly_rf_tc_taps (tensor cores); this pass
// only gathers.  One thread per (input pixel, 8-channel vector).
template <int S>
__global__ void __launch_bounds__(256) rf_col2im_kernel(const __nv_bfloat16* __restrict__ R, int N, int H, int W, int V,
                                                        int ho, int wo, const float* __restrict__ extra,
                                                        __nv_bfloat16* __restrict__ dx, long lddx) {
  pdl_enter();
  const int C = V * 8;
  const long total = (long)N * H * W * V;
  const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int v = (int)(tid % V);
  const long pix = tid / V;
  const int x = (int)(pix % W), y = (int)((pix / W) % H), n = (int)(pix / ((long)W * H));
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = extra != nullptr ? __ldg(extra + (long)n * C + v * 8 + e) : 0.f;
  if (S == 2) {
    // stride 2: an even coordinate is the centre tap of one output pixel, an odd one the first tap of the next output
    // pixel and the last tap of the previous — at most 2 x 2 terms, all four loads in flight together
    int hh[2], ih[2], ww[2], jw[2];
    bool vh[2], vw[2];
    if ((y & 1) == 0) { hh[0] = y >> 1; ih[0] = 1; vh[0] = true; hh[1] = 0; ih[1] = 0; vh[1] = false; }
    else { hh[0] = (y + 1) >> 1; ih[0] = 0; vh[0] = hh[0] < ho; hh[1] = (y - 1) >> 1; ih[1] = 2; vh[1] = true; }
    if ((x & 1) == 0) { ww[0] = x >> 1; jw[0] = 1; vw[0] = true; ww[1] = 0; jw[1] = 0; vw[1] = false; }
    else { ww[0] = (x + 1) >> 1; jw[0] = 0; vw[0] = ww[0] < wo; ww[1] = (x - 1) >> 1; jw[1] = 2; vw[1] = true; }
    uint4 u[4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b)
        if (vh[a] && vw[b])
          u[a * 2 + b] = ldg16(R + ((((long)n * ho + hh[a]) * wo + ww[b]) * 9 + (ih[a] * 3 + jw[b])) * (long)C + v * 8);
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b)
        if (vh[a] && vw[b]) {
          float f[8];
          unpack8(u[a * 2 + b], f);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] += f[e];
        }
  } else {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int h = y + 1 - i;
      if (h < 0 || h >= ho) continue;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int w = x + 1 - j;
        if (w < 0 || w >= wo) continue;
        float f[8];
        unpack8(ldg16(R + ((((long)n * ho + h) * wo + w) * 9 + (i * 3 + j)) * (long)C + v * 8), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += f[e];
      }
    }
  }
  *reinterpret_cast<uint4*>(dx + pix * lddx + v * 8) = pack8(acc);
}

// dwgk[t][ij][c] += Σ_m dZ[m][t*C+c] * x[n, h*s+i-k/2, w*s+j-k/2, c].
// One thread per (8-channel vector v, feature t) and pixel lane, consecutive threads on consecutive v (every load of a
// warp is a contiguous run of a pixel's channels); each thread keeps its k*k x 8 sums in registers over its pixels.
//   PARTS = true : every pixel lane STORES its sums as one row of parts[lanes][k*k*k*k*C]; rf_wgrad_finalize adds the rows
//                  in order.  No atomics (the first version issued 72 per thread: 10.9 M on 10 k addresses) and the
//                  result is bit-reproducible.
//   PARTS = false: atomics into dwgk (any V; used when no scratch is given).
template <int K, bool PARTS>
__global__ void __launch_bounds__(288, 3) rf_generate_wgrad_kernel(const __nv_bfloat16* __restrict__ dZ,
                                                                   const __nv_bfloat16* __restrict__ x, long ldx, int N,
                                                                   int H, int W, int V, int s, int ho, int wo,
                                                                   float* __restrict__ dwgk, float* __restrict__ parts) {
  pdl_enter();
  // A thread owns (vector v, feature t, kernel ROW i): K x 8 sums.  Splitting the k*k taps over K threads keeps the
  // register count low enough for three blocks per SM — the loop is bound by memory latency (ncu: 14 % warps active and
  // 15 stalled-on-scoreboard warps per issue with one 115-register block per SM), and two pixels are in flight per thread.
  constexpr int KK = K * K;
  const int C = V * 8;
  const unsigned M = (unsigned)N * ho * wo;
  const unsigned work = (unsigned)V * KK * K;  // (v, t, i) triples
  const unsigned T = gridDim.x * blockDim.x;
  const unsigned tid = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lanes = T / work;
  if (lanes == 0 || tid >= lanes * work) return;
  const int v = (int)(tid % (unsigned)V);
  const int t = (int)((tid / (unsigned)V) % KK);
  const int i = (int)((tid / (unsigned)(V * KK)) % K);
  const unsigned pl = tid / work;
  const unsigned hw = (unsigned)ho * wo;
  float acc[K][8];
#pragma unroll
  for (int j = 0; j < K; ++j)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[j][e] = 0.f;
  auto body = [&](unsigned m, const uint4& draw, const uint4 (&xraw)[K], const bool (&ok)[K]) {
    float d[8];
    unpack8(draw, d);
#pragma unroll
    for (int j = 0; j < K; ++j) {
      if (!ok[j]) continue;
      float f[8];
      unpack8(xraw[j], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[j][e] = fmaf(d[e], f[e], acc[j][e]);
    }
  };
  auto fetch = [&](unsigned m, uint4& draw, uint4 (&xraw)[K], bool (&ok)[K]) {
    const unsigned n = m / hw, r = m - n * hw;
    const int h = (int)(r / (unsigned)wo), w = (int)(r - (unsigned)h * wo);
    draw = ldg16(dZ + ((size_t)m * KK + t) * (size_t)C + v * 8);
    const int yy = h * s + i - K / 2;
    const __nv_bfloat16* xr = x + ((size_t)n * H + yy) * W * ldx + v * 8;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int xx = w * s + j - K / 2;
      ok[j] = yy >= 0 && yy < H && xx >= 0 && xx < W;
      if (ok[j]) xraw[j] = ldg16(xr + (size_t)xx * ldx);
    }
  };
  unsigned m = pl;
  for (; m + lanes < M; m += 2 * lanes) {      // two pixels in flight
    uint4 d0, d1, x0[K], x1[K];
    bool k0[K], k1[K];
    fetch(m, d0, x0, k0);
    fetch(m + lanes, d1, x1, k1);
    body(m, d0, x0, k0);
    body(m + lanes, d1, x1, k1);
  }
  if (m < M) {
    uint4 d0, x0[K];
    bool k0[K];
    fetch(m, d0, x0, k0);
    body(m, d0, x0, k0);
  }
  if (PARTS) {
    float* row = parts + (size_t)pl * KK * KK * C + v * 8;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      float4* dst = reinterpret_cast<float4*>(row + (size_t)(t * KK + i * K + j) * C);
      dst[0] = make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
      dst[1] = make_float4(acc[j][4], acc[j][5], acc[j][6], acc[j][7]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < K; ++j)
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(&dwgk[((long)t * KK + i * K + j) * C + v * 8 + e], acc[j][e]);
  }
}

// dwgk[o] += Σ_rows parts[row][o]: one warp per output, lane l adds rows l, l+32, ... in order, fixed shuffle tree
__global__ void __launch_bounds__(256) rf_wgrad_finalize_kernel(const float* __restrict__ parts, int rows, long outs,
                                                                float* __restrict__ dwgk) {
  pdl_enter();
  const long o = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (o >= outs) return;
  float a = 0.f;
  for (int r = lane; r < rows; r += 32) a += parts[(size_t)r * outs + o];
  a = warp_sum(a);
  if (lane == 0) dwgk[o] += a;
}

int rf_wgrad_tc_launch(const void* dZ, const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo,
                       float* parts, long max_rows, int* rows_out, cudaStream_t stream);   // rf_generate_tc.cu

struct RfGeo {
  int N, H, W, C, K, s, ho, wo;
};

static int rf_check(const RfGeo& g, const char* what) {
  LY_REQUIRE(g.N > 0 && g.H > 0 && g.W > 0, "%s: bad spatial size", what);
  LY_REQUIRE(g.C > 0 && g.C % 8 == 0 && g.C <= 512, "%s: C=%d must be a multiple of 8 and <= 512", what, g.C);
  LY_REQUIRE((g.K == 1 && g.s == 1) || (g.K == 3 && (g.s == 1 || g.s == 2)), "%s: k=%d s=%d unsupported", what, g.K,
             g.s);
  LY_REQUIRE(g.ho == (g.H + 2 * (g.K / 2) - g.K) / g.s + 1 && g.wo == (g.W + 2 * (g.K / 2) - g.K) / g.s + 1,
             "%s: output size mismatch", what);
  return LY_OK;
}

}  // namespace ly

using namespace ly;

#define LY_ALIGNED16(p) (reinterpret_cast<uintptr_t>(p) % 16 == 0)

extern "C" int ly_rf_generate(const void* x, long ldx, int N, int H, int W, int C, int k, int s, int ho, int wo,
                              const float* wgk, const float* scale, const float* shift, int act, void* G,
                              ly_stream_t st) {
  RfGeo g{N, H, W, C, k, s, ho, wo};
  int rc = rf_check(g, "ly_rf_generate");
  if (rc) return rc;
  LY_REQUIRE(x && wgk && G && LY_ALIGNED16(x) && LY_ALIGNED16(G) && LY_ALIGNED16(wgk) && ldx % 8 == 0 && ldx >= C,
             "ly_rf_generate: bad pointers / ld");
  LY_REQUIRE((scale == nullptr) == (shift == nullptr), "ly_rf_generate: scale and shift go together");
  const int V = C / 8;
  const long total = (long)N * ho * wo * V;
  if (k == 1)
    LY_CUDA(launch(rf_generate_kernel<1>, ceil_div(total, 128), 128, 0, (cudaStream_t)st, 
        (const __nv_bfloat16*)x, ldx, N, H, W, V, s, ho, wo, wgk, scale, shift, act, (__nv_bfloat16*)G));
  else
    LY_CUDA(launch(rf_generate_kernel<3>, ceil_div(total, 128), 128, 0, (cudaStream_t)st, 
        (const __nv_bfloat16*)x, ldx, N, H, W, V, s, ho, wo, wgk, scale, shift, act, (__nv_bfloat16*)G));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_rf_attention(const void* G, int N, int ho, int wo, int C, int k, const float* wa, float* maps,
                               int16_t* amax, float* att, ly_stream_t st) {
  LY_REQUIRE(G && wa && maps && att && LY_ALIGNED16(G) && C % 8 == 0 && C > 0, "ly_rf_attention: bad args");
  const int V = C / 8, KK = k * k;
  const long M = (long)N * ho * wo;
  {
    const int prc = poolc_launch(G, M, KK, k, V, ho, wo, maps, amax, (cudaStream_t)st);
    if (prc != LY_OK) return prc;
  }
  const long total = (long)N * ho * k * wo * k;
  LY_CUDA(launch(rf_att_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, maps, N, ho * k, wo * k, wa, att));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_rf_gate(const void* G, int N, int ho, int wo, int C, int k, const void* se, long ldse,
                          const float* att, void* A, ly_stream_t st) {
  LY_REQUIRE(G && se && att && A && LY_ALIGNED16(G) && LY_ALIGNED16(A) && LY_ALIGNED16(se) && C % 8 == 0 &&
                 ldse % 8 == 0,
             "ly_rf_gate: bad args");
  const int V = C / 8, KK = k * k;
  const long M = (long)N * ho * wo;
  {
    const __nv_bfloat16 *g_ = (const __nv_bfloat16*)G, *se_ = (const __nv_bfloat16*)se;
    __nv_bfloat16* a_ = (__nv_bfloat16*)A;
    cudaStream_t s_ = (cudaStream_t)st;
    const long rows = M * KK;
    if (V <= 16) LY_CUDA(launch(rf_gate_kernel<4>, ceil_div(rows * 4, 256), 256, 0, s_, g_, M, KK, k, V, ho, wo, se_, ldse, att, a_));
    else if (V <= 32) LY_CUDA(launch(rf_gate_kernel<8>, ceil_div(rows * 8, 256), 256, 0, s_, g_, M, KK, k, V, ho, wo, se_, ldse, att, a_));
    else LY_CUDA(launch(rf_gate_kernel<32>, ceil_div(rows * 32, 256), 256, 0, s_, g_, M, KK, k, V, ho, wo, se_, ldse, att, a_));
  }
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_rf_gate_bwd(const void* dA, const void* G, int N, int ho, int wo, int C, int k, const void* se,
                              long ldse, const float* att, void* dG, float* dse, float* datt, ly_stream_t st) {
  LY_REQUIRE(dA && G && se && att && dG && dse && datt && C % 8 == 0 && C <= 512, "ly_rf_gate_bwd: bad args");
  const int V = C / 8, KK = k * k;
  const long per_img = (long)ho * wo * KK;
  (void)per_img;
  dim3 grid(LY_RF_GATE_PARTS, N);   // fixed: the caller sizes dse as [LY_RF_GATE_PARTS][N][C]
  const __nv_bfloat16 *da_ = (const __nv_bfloat16*)dA, *g_ = (const __nv_bfloat16*)G, *se_ = (const __nv_bfloat16*)se;
  __nv_bfloat16* dg_ = (__nv_bfloat16*)dG;
  cudaStream_t s_ = (cudaStream_t)st;
  if (V <= 16) LY_CUDA(launch(rf_gate_bwd_kernel<4>, grid, 256, 0, s_, da_, g_, KK, k, V, ho, wo, se_, ldse, att, dg_, dse, datt));
  else if (V <= 32) LY_CUDA(launch(rf_gate_bwd_kernel<8>, grid, 256, 0, s_, da_, g_, KK, k, V, ho, wo, se_, ldse, att, dg_, dse, datt));
  else LY_CUDA(launch(rf_gate_bwd_kernel<16>, grid, 256, 0, s_, da_, g_, KK, k, V, ho, wo, se_, ldse, att, dg_, dse, datt));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_rf_attention_bwd(const float* datt, const float* att, const float* maps, const int16_t* amax, int N,
                                   int ho, int wo, int C, int k, const float* wa, float* dmaps, float* dwa, void* dG,
                                   ly_stream_t st) {
  LY_REQUIRE(datt && att && maps && amax && wa && dmaps && dwa && dG && C % 8 == 0, "ly_rf_attention_bwd: bad args");
  const int V = C / 8, KK = k * k;
  const long M = (long)N * ho * wo;
  const long total = (long)N * ho * k * wo * k;
  LY_CUDA(launch(rf_att_bwd_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, datt, att, maps, N, ho * k, wo * k, wa, dmaps,
                                                                       dwa));
  LY_LAUNCH_CHECK();
  {
    __nv_bfloat16* dg_ = (__nv_bfloat16*)dG;
    cudaStream_t s_ = (cudaStream_t)st;
    const long rows = M * KK;
    if (V <= 16) LY_CUDA(launch(rf_poolc_bwd_kernel<4>, ceil_div(rows * 4, 256), 256, 0, s_, dg_, M, KK, k, V, ho, wo, dmaps, amax));
    else if (V <= 32) LY_CUDA(launch(rf_poolc_bwd_kernel<8>, ceil_div(rows * 8, 256), 256, 0, s_, dg_, M, KK, k, V, ho, wo, dmaps, amax));
    else LY_CUDA(launch(rf_poolc_bwd_kernel<32>, ceil_div(rows * 32, 256), 256, 0, s_, dg_, M, KK, k, V, ho, wo, dmaps, amax));
  }
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_rf_col2im(const void* R, int N, int H, int W, int C, int s, int ho, int wo, const float* extra, void* dx,
                            long lddx, ly_stream_t st) {
  LY_REQUIRE(R && dx && C % 8 == 0 && C > 0 && (s == 1 || s == 2) && N > 0, "ly_rf_col2im: bad arguments");
  LY_REQUIRE(ho == (H + 2 - 3) / s + 1 && wo == (W + 2 - 3) / s + 1, "ly_rf_col2im: output size mismatch");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(R) % 16 == 0 && reinterpret_cast<uintptr_t>(dx) % 16 == 0 && lddx % 8 == 0 &&
                 lddx >= C,
             "ly_rf_col2im: alignment / ld");
  const int V = C / 8;
  const long total = (long)N * H * W * V;
  if (s == 2)
    LY_CUDA(launch(rf_col2im_kernel<2>, ceil_div(total, 256), 256, 0, (cudaStream_t)st, (const __nv_bfloat16*)R, N, H, W, V, ho,
                   wo, extra, (__nv_bfloat16*)dx, lddx));
  else
    LY_CUDA(launch(rf_col2im_kernel<1>, ceil_div(total, 256), 256, 0, (cudaStream_t)st, (const __nv_bfloat16*)R, N, H, W, V, ho,
                   wo, extra, (__nv_bfloat16*)dx, lddx));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" long ly_rf_wgrad_scratch_floats(int C, int k) {
  if (C <= 0 || C % 8 != 0 || (k != 1 && k != 3)) return -1;
  const long work = (long)(C / 8) * k * k * k;
  long lanes = ((long)device_sm_count() * 3 * 288) / work;
  if (lanes < 1) lanes = 1;
  return lanes * (long)k * k * k * k * C;
}

extern "C" int ly_rf_generate_bwd(const void* dZ, const void* x, long ldx, int N, int H, int W, int C, int k, int s,
                                  int ho, int wo, const float* wgk, const float* extra, void* dx, long lddx,
                                  float* dwgk, float* wscratch, long wscratch_floats, ly_stream_t st) {
  RfGeo g{N, H, W, C, k, s, ho, wo};
  int rc = rf_check(g, "ly_rf_generate_bwd");
  if (rc) return rc;
  LY_REQUIRE(dZ && x && wgk && LY_ALIGNED16(dZ) && LY_ALIGNED16(x), "ly_rf_generate_bwd: bad pointers");
  const int V = C / 8, KK = k * k;
  if (dx != nullptr) {
    LY_REQUIRE(LY_ALIGNED16(dx) && lddx % 8 == 0 && lddx >= C, "ly_rf_generate_bwd: bad dx");
    const long total = (long)N * H * W * V;
    if (k == 1)
      LY_CUDA(launch(rf_generate_dx_kernel<1>, ceil_div(total, 128), 128, 0, (cudaStream_t)st, 
          (const __nv_bfloat16*)dZ, N, H, W, V, s, ho, wo, wgk, extra, (__nv_bfloat16*)dx, lddx));
    else
      LY_CUDA(launch(rf_generate_dx_kernel<3>, ceil_div(total, 128), 128, 0, (cudaStream_t)st, 
          (const __nv_bfloat16*)dZ, N, H, W, V, s, ho, wo, wgk, extra, (__nv_bfloat16*)dx, lddx));
    LY_LAUNCH_CHECK();
  }
  if (dwgk != nullptr) {
    const long M = (long)N * ho * wo;
    LY_REQUIRE(M * KK < (1L << 31), "ly_rf_generate_bwd: too many pixels");
    const long work = (long)V * KK * k;
    const long outs = (long)KK * KK * C;
    cudaStream_t stream = (cudaStream_t)st;
    static const bool wgrad_cuda = getenv("LY_RF_WGRAD_CUDA") != nullptr;   // dev knob: the CUDA-core kernel
    if (k == 3 && C % 16 == 0 && wscratch != nullptr && wscratch_floats >= outs && !wgrad_cuda) {
      // tensor cores: per-slice partial rows, then the ordered sum
      int rows = 0;
      rc = rf_wgrad_tc_launch(dZ, x, ldx, N, H, W, C, s, ho, wo, wscratch, wscratch_floats / outs, &rows, stream);
      if (rc) return rc;
      LY_CUDA(launch(rf_wgrad_finalize_kernel, ceil_div(outs * 32, 256), 256, 0, stream, wscratch, rows, outs, dwgk));
      LY_LAUNCH_CHECK();
      return LY_OK;
    }
    // pixel lanes: three 288-thread blocks per SM, every lane a row of `parts`
    long lanes = ((long)device_sm_count() * 3 * 288) / work;
    if (lanes > M) lanes = M;
    if (lanes < 1) lanes = 1;
    const int grid = ceil_div(lanes * work, 288);
    const bool use_parts = wscratch != nullptr && wscratch_floats >= lanes * outs;
    if (k == 1) {
      if (use_parts)
        LY_CUDA(launch(rf_generate_wgrad_kernel<1, true>, grid, 288, 0, stream, (const __nv_bfloat16*)dZ, (const __nv_bfloat16*)x, ldx,
                                                                    N, H, W, V, s, ho, wo, dwgk, wscratch));
      else
        LY_CUDA(launch(rf_generate_wgrad_kernel<1, false>, grid, 288, 0, stream, (const __nv_bfloat16*)dZ, (const __nv_bfloat16*)x, ldx,
                                                                     N, H, W, V, s, ho, wo, dwgk, nullptr));
    } else {
      if (use_parts)
        LY_CUDA(launch(rf_generate_wgrad_kernel<3, true>, grid, 288, 0, stream, (const __nv_bfloat16*)dZ, (const __nv_bfloat16*)x, ldx,
                                                                    N, H, W, V, s, ho, wo, dwgk, wscratch));
      else
        LY_CUDA(launch(rf_generate_wgrad_kernel<3, false>, grid, 288, 0, stream, (const __nv_bfloat16*)dZ, (const __nv_bfloat16*)x, ldx,
                                                                     N, H, W, V, s, ho, wo, dwgk, nullptr));
    }
    LY_LAUNCH_CHECK();
    if (use_parts) {
      LY_CUDA(launch(rf_wgrad_finalize_kernel, ceil_div(outs * 32, 256), 256, 0, stream, wscratch, (int)lanes, outs, dwgk));
      LY_LAUNCH_CHECK();
    }
  }
  return LY_OK;
}

// ------------------------------------------------------------------------------------------------
// ly_rf_attention_noalloc — same result as ly_rf_attention, but the channel max/mean maps (the scratch
// between its two kernels) live in library-owned device memory instead of a caller-allocated buffer.
//
// Who needs it: callers that cannot obtain device memory at the moment they call us — an inference
// server that hands out its whole memory arena at start-up and has no allocator in the request path, or
// a latency-critical loop where cudaMalloc is forbidden because it synchronises the entire device.
//
// The scratch is one file-scope __device__ array of LY_RF_NOALLOC_MAX_FLOATS floats (8 MiB); every GPU
// that loads this library keeps that much memory for the life of the process.  Its device address is
// looked up once per device (cudaGetSymbolAddress, with that device current) on the FIRST call made on
// that device — so make that first call outside any stream capture.  Nothing is allocated or freed here.
// All calls on a device share the one array: two calls whose kernels may overlap on the GPU (different
// streams) race on it.  The array keeps whatever the previous call left, which is harmless: the first
// kernel overwrites every map entry the second kernel reads, so nothing has to start at zero.
// ------------------------------------------------------------------------------------------------
#define LY_RF_NOALLOC_MAX_FLOATS (2 * 1024 * 1024)

static __device__ __align__(256) float g_rf_maps_scratch[LY_RF_NOALLOC_MAX_FLOATS];

namespace ly {
static std::mutex g_rf_scratch_mu;
static float* g_rf_scratch_ptr[64] = {nullptr};

// Device address of g_rf_maps_scratch on the current device (host code must not use the symbol's own
// address: on the host it is only a stand-in).
static int rf_scratch_for_current_device(float** out) {
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  LY_REQUIRE(dev >= 0 && dev < 64, "ly_rf_attention_noalloc: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_rf_scratch_mu);
  if (g_rf_scratch_ptr[dev] == nullptr) {
    float* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_rf_maps_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      return fail(LY_ECUDA, "ly_rf_attention_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                  cudaGetErrorString(e));
    g_rf_scratch_ptr[dev] = ptr;
  }
  *out = g_rf_scratch_ptr[dev];
  return LY_OK;
}
}  // namespace ly

extern "C" int ly_rf_attention_noalloc(const void* G, int N, int ho, int wo, int C, int k, const float* wa,
                                       int16_t* amax, float* att, ly_stream_t st) {
  LY_REQUIRE(G && wa && att && LY_ALIGNED16(G) && C % 8 == 0 && C > 0, "ly_rf_attention_noalloc: bad args");
  LY_REQUIRE(N > 0 && ho > 0 && wo > 0 && (k == 1 || k == 3), "ly_rf_attention_noalloc: bad geometry");
  const long need = (long)N * ho * k * wo * k * 2;  // floats in the max/mean maps
  if (need > (long)LY_RF_NOALLOC_MAX_FLOATS)
    return fail(LY_EINVAL,
                "ly_rf_attention_noalloc: N*ho*k*wo*k*2 = %ld floats exceeds the built-in scratch of %d floats "
                "(8 MiB); use ly_rf_attention with a caller-provided maps buffer for this size",
                need, (int)LY_RF_NOALLOC_MAX_FLOATS);
  float* maps = nullptr;
  int rc = rf_scratch_for_current_device(&maps);
  if (rc != LY_OK) return rc;
  const int V = C / 8, KK = k * k;
  const long M = (long)N * ho * wo;
  {
    const int prc = poolc_launch(G, M, KK, k, V, ho, wo, maps, amax, (cudaStream_t)st);
    if (prc != LY_OK) return prc;
  }
  const long total = (long)N * ho * k * wo * k;
  LY_CUDA(launch(rf_att_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, maps, N, ho * k, wo * k, wa, att));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
