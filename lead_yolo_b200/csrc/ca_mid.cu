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
  ca_mid_kernel<<<(unsigned)rows, 256, 0, (cudaStream_t)st>>>((const __nv_bfloat16*)pooled, ldp, rows, nh, C, mip, w1,
                                                             scale, shift, wh, bh, ww, bw, (__nv_bfloat16*)gates, ldg);
  LY_LAUNCH_CHECK();
  return LY_OK;
}
