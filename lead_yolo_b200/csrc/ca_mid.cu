// CoordAtt, inference: the tiny middle part between pooling and gating as ONE kernel (reference
// models/common.py:1601-1608): per pooled position r (row of the "HW rows" matrix)
//   y[m]   = h_swish(scale[m] * (sum_c W1[m][c] * P[r][c]) + shift[m])          m < mip (folded conv1 bias + bn1)
//   a[r,c] = sigmoid(sum_m Wx[c][m] * y[m] + bx[c]),   Wx/bx = conv_h for the first nh rows, conv_w afterwards
// One block per row: warp m reduces output m over the channels, then all threads produce the gates.
#include "ly_common.cuh"

namespace ly {

__global__ void __launch_bounds__(256) ca_mid_kernel(const __nv_bfloat16* __restrict__ P, long ldp, long rows, long nh,
                                                     int C, int mip, const float* __restrict__ w1 /*[mip][C]*/,
                                                     const float* __restrict__ scale, const float* __restrict__ shift,
                                                     const float* __restrict__ wh /*[C][mip]*/,
                                                     const float* __restrict__ bh, const float* __restrict__ ww,
                                                     const float* __restrict__ bw, __nv_bfloat16* __restrict__ gates,
                                                     long ldg) {
  pdl_enter();
  __shared__ float y[32];
  const long r = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const __nv_bfloat16* prow = P + r * ldp;
  for (int m = warp; m < mip; m += (blockDim.x >> 5)) {
    float a = 0.f;
    for (int c = lane * 8; c < C; c += 256) {
      float f[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(prow + c)), f);
      const float4 u0 = __ldg(reinterpret_cast<const float4*>(w1 + (long)m * C + c));
      const float4 u1 = __ldg(reinterpret_cast<const float4*>(w1 + (long)m * C + c + 4));
      a = fmaf(f[0], u0.x, a); a = fmaf(f[1], u0.y, a); a = fmaf(f[2], u0.z, a); a = fmaf(f[3], u0.w, a);
      a = fmaf(f[4], u1.x, a); a = fmaf(f[5], u1.y, a); a = fmaf(f[6], u1.z, a); a = fmaf(f[7], u1.w, a);
    }
    a = warp_sum(a);
    if (lane == 0) y[m] = act_apply_t<LY_ACT_HSWISH>(fmaf(a, scale[m], shift[m]));
  }
  __syncthreads();
  const bool is_h = r < nh;
  const float* wx = is_h ? wh : ww;
  const float* bx = is_h ? bh : bw;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = bx[c];
    for (int m = 0; m < mip; ++m) a = fmaf(wx[(long)c * mip + m], y[m], a);
    gates[r * ldg + c] = __float2bfloat16(1.f / (1.f + __expf(-a)));
  }
}

// Fast path for mip == 8 and C == 32 * CPL: one WARP per pooled row, no block barrier inside the row loop.  The block
// stages W1 [8][C] and Wx transposed to [8][C] in shared memory with coalesced loads (per-lane scalar loads of the
// weights made the first version L1-request-bound); lane l owns channels [l*CPL, (l+1)*CPL) and reads its slices as
// contiguous vectors.  The 8 hidden values come out of xor-shuffle reductions in every lane.
// Blocks [0, gh) take the "h" rows (conv_h), the rest the "w" rows.
template <int CPL>
__global__ void __launch_bounds__(256) ca_mid_warp_kernel(const __nv_bfloat16* __restrict__ P, long ldp, long rows, long nh,
                                                          int gh, const float* __restrict__ w1,
                                                          const float* __restrict__ scale, const float* __restrict__ shift,
                                                          const float* __restrict__ wh, const float* __restrict__ bh,
                                                          const float* __restrict__ ww, const float* __restrict__ bw,
                                                          __nv_bfloat16* __restrict__ gates, long ldg) {
  pdl_enter();
  constexpr int C = 32 * CPL;
  __shared__ __align__(16) float s_w1[8 * C], s_wx[8 * C], s_b[C], s_sc[8], s_sh[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool is_h = (int)blockIdx.x < gh;
  const float* wx = is_h ? wh : ww;
  const float* bx = is_h ? bh : bw;
  for (int i = threadIdx.x; i < 8 * C; i += 256) {
    s_w1[i] = __ldg(w1 + i);
    const int c = i >> 3, m = i & 7;        // wx is [C][8]: coalesced read, transposed write
    s_wx[m * C + c] = __ldg(wx + i);
  }
  for (int i = threadIdx.x; i < C; i += 256) s_b[i] = __ldg(bx + i);
  if (threadIdx.x < 8) { s_sc[threadIdx.x] = __ldg(scale + threadIdx.x); s_sh[threadIdx.x] = __ldg(shift + threadIdx.x); }
  __syncthreads();
  const long first = is_h ? 0 : nh, count = is_h ? nh : rows - nh;
  const long wid = (long)(is_h ? blockIdx.x : blockIdx.x - gh) * 8 + warp;
  const long nw = (long)(is_h ? gh : (int)gridDim.x - gh) * 8;
  const int c0 = lane * CPL;
  for (long i = wid; i < count; i += nw) {
    const long r = first + i;
    float f[CPL];
    const __nv_bfloat16* src = P + r * ldp + c0;
    if constexpr (CPL == 8) {
      unpack8(__ldg(reinterpret_cast<const uint4*>(src)), f);
    } else {
#pragma unroll
      for (int e = 0; e < CPL; e += 2) {
        const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(src + e);
        f[e] = __low2float(v); f[e + 1] = __high2float(v);
      }
    }
    float y[8];
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      float a = 0.f;
#pragma unroll
      for (int e = 0; e < CPL; ++e) a = fmaf(f[e], s_w1[m * C + c0 + e], a);
      y[m] = a;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int m = 0; m < 8; ++m) y[m] += __shfl_xor_sync(0xffffffffu, y[m], o);
#pragma unroll
    for (int m = 0; m < 8; ++m) y[m] = act_apply_t<LY_ACT_HSWISH>(fmaf(y[m], s_sc[m], s_sh[m]));
    float g[CPL];
#pragma unroll
    for (int e = 0; e < CPL; ++e) {
      float a = s_b[c0 + e];
#pragma unroll
      for (int m = 0; m < 8; ++m) a = fmaf(s_wx[m * C + c0 + e], y[m], a);
      g[e] = 1.f / (1.f + __expf(-a));
    }
    __nv_bfloat16* dst = gates + r * ldg + c0;
    if constexpr (CPL == 8) {
      *reinterpret_cast<uint4*>(dst) = pack8(g);
    } else {
#pragma unroll
      for (int e = 0; e < CPL; e += 2) *reinterpret_cast<__nv_bfloat162*>(dst + e) = __floats2bfloat162_rn(g[e], g[e + 1]);
    }
  }
}

}  // namespace ly

using namespace ly;

extern "C" int ly_ca_mid_infer(const void* pooled, long ldp, long rows, long nh, int C, int mip, const float* w1,
                               const float* scale, const float* shift, const float* wh, const float* bh,
                               const float* ww, const float* bw, void* gates, long ldg, ly_stream_t st) {
  LY_REQUIRE(pooled && w1 && scale && shift && wh && bh && ww && bw && gates, "ly_ca_mid_infer: null pointer");
  LY_REQUIRE(rows > 0 && nh >= 0 && nh <= rows && C % 8 == 0 && C > 0 && mip > 0 && mip <= 32,
             "ly_ca_mid_infer: bad sizes (C=%d mip=%d)", C, mip);
  LY_REQUIRE(ldp % 8 == 0 && ldp >= C && ldg >= C && reinterpret_cast<uintptr_t>(pooled) % 16 == 0 &&
                 reinterpret_cast<uintptr_t>(w1) % 16 == 0,
             "ly_ca_mid_infer: alignment / ld");
  if (mip == 8 && (C == 64 || C == 128 || C == 256) && ldg % 8 == 0 && reinterpret_cast<uintptr_t>(gates) % 16 == 0) {
    // rows per block: up to four per warp so the staged weights are reused, fewer when that would leave SMs idle
    // (the 20x20 level has 2560 rows: 80 blocks of 32 rows ran on 80 of 148 SMs)
    const long per = rows >= 32 * 296 ? 32 : (rows >= 16 * 296 ? 16 : 8);
    const int gh = (int)ceil_div(nh, per), gwb = (int)ceil_div(rows - nh, per);
    cudaStream_t s_ = (cudaStream_t)st;
#define LY_CAMID(CPL)                                                                                                   \
  (void)launch(ca_mid_warp_kernel<CPL>, gh + gwb, 256, 0, s_, (const __nv_bfloat16*)pooled, ldp, rows, nh, gh, w1, scale, shift, wh, \
                                                    bh, ww, bw, (__nv_bfloat16*)gates, ldg)
    if (C == 64) LY_CAMID(2);
    else if (C == 128) LY_CAMID(4);
    else LY_CAMID(8);
#undef LY_CAMID
    LY_LAUNCH_CHECK();
    return LY_OK;
  }
  LY_CUDA(launch(ca_mid_kernel, (unsigned)rows, 256, 0, (cudaStream_t)st, (const __nv_bfloat16*)pooled, ldp, rows, nh, C, mip, w1,
                                                             scale, shift, wh, bh, ww, bw, (__nv_bfloat16*)gates, ldg));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
