// FasterNet partial 3x3 convolution (reference models/common.py:1416,1432-1437) on CUDA cores, inference path.
// Only the first dim//4 channels (6/10/20/40 in LEAD-YOLO) are convolved; as an implicit GEMM that is K = 9 x 64
// zero-padded columns for <= 40 useful ones, so a direct kernel wins: a warp = 32 consecutive pixels x ONE group of
// 8 output channels, which makes every weight load warp-uniform (one sector per warp) and leaves the kernel bound by
// reading the 8-aligned channel slice of x once and writing t1 once.
//   t1[p][co] = sum_{tap, ci} wk[tap][ci][co] * x[p + tap][ci]      (zero padding, channels >= dim//4 have zero weights)
#include "ly_common.cuh"

namespace ly {

template <int CPP>
__global__ void __launch_bounds__(256) pconv3x3_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int N, int H,
                                                       int W, const float* __restrict__ wk /*[9][CPP][CPP]*/,
                                                       __nv_bfloat16* __restrict__ out, long ldo) {
  constexpr int VO = CPP / 8;
  extern __shared__ float wsm[];  // [9][CPP][CPP]: broadcast reads, one wavefront per 4 FMA instructions
  for (int i = threadIdx.x; i < 9 * CPP * CPP; i += blockDim.x) wsm[i] = wk[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long M = (long)N * H * W;
  const long items = ((M + 31) / 32) * VO;  // (pixel group, output vector) pairs, one per warp
  const long warps = ((long)gridDim.x * blockDim.x) >> 5;
  for (long gw = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; gw < items; gw += warps) {
    const int vo = (int)(gw % VO);
    const long m = (gw / VO) * 32 + lane;
    if (m >= M) continue;
    const int w = (int)(m % W), h = (int)((m / W) % H);
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int yy = h + tap / 3 - 1, xx = w + tap % 3 - 1;
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      const __nv_bfloat16* px = x + (m + (long)(tap / 3 - 1) * W + (tap % 3 - 1)) * ldx;
#pragma unroll
      for (int vi = 0; vi < VO; ++vi) {
        float f[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(px + vi * 8)), f);
#pragma unroll
        for (int ci = 0; ci < 8; ++ci) {
          const float* wr = wsm + (tap * CPP + vi * 8 + ci) * CPP + vo * 8;
          const float4 a0 = *reinterpret_cast<const float4*>(wr);
          const float4 a1 = *reinterpret_cast<const float4*>(wr + 4);
          acc[0] = fmaf(f[ci], a0.x, acc[0]); acc[1] = fmaf(f[ci], a0.y, acc[1]);
          acc[2] = fmaf(f[ci], a0.z, acc[2]); acc[3] = fmaf(f[ci], a0.w, acc[3]);
          acc[4] = fmaf(f[ci], a1.x, acc[4]); acc[5] = fmaf(f[ci], a1.y, acc[5]);
          acc[6] = fmaf(f[ci], a1.z, acc[6]); acc[7] = fmaf(f[ci], a1.w, acc[7]);
        }
      }
    }
    *reinterpret_cast<uint4*>(out + m * ldo + vo * 8) = pack8(acc);
  }
}

}  // namespace ly

using namespace ly;

// x: NHWC bf16 (pixel stride ldx), only channels [0, cpp) are read; wk: fp32 [9][cpp][cpp] (tap, ci, co), rows/cols
// beyond the real dim//4 zero; out: bf16 [pixels][ldo], channels [0, cpp).  cpp in {8, 16, 24, 32, 40}.
extern "C" int ly_pconv3x3_fwd(const void* x, long ldx, int N, int H, int W, int cpp, const float* wk, void* out,
                               long ldo, ly_stream_t st) {
  LY_REQUIRE(x && wk && out && N > 0 && H > 0 && W > 0, "ly_pconv3x3_fwd: bad args");
  LY_REQUIRE(ldx % 8 == 0 && ldx >= cpp && ldo % 8 == 0 && ldo >= cpp, "ly_pconv3x3_fwd: ld");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(out) % 16 == 0 &&
                 reinterpret_cast<uintptr_t>(wk) % 16 == 0,
             "ly_pconv3x3_fwd: alignment");
  const long M = (long)N * H * W;
  const int VO = cpp / 8;
  const long items = (long)ceil_div(M, 32) * VO;
  const size_t smem = sizeof(float) * 9 * (size_t)cpp * cpp;
  const int occ = smem > 48 * 1024 ? 3 : 6;                      // resident blocks per SM
  long g = ceil_div(items, 8);
  if (g > (long)device_sm_count() * occ) g = (long)device_sm_count() * occ;
  const int grid = (int)(g < 1 ? 1 : g);
  cudaStream_t stream = (cudaStream_t)st;
#define LY_PC(CP)                                                                                              \
  case CP:                                                                                                     \
    if (smem > 48 * 1024)                                                                                      \
      LY_CUDA(cudaFuncSetAttribute(pconv3x3_kernel<CP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    pconv3x3_kernel<CP><<<grid, 256, smem, stream>>>((const __nv_bfloat16*)x, ldx, N, H, W, wk,                \
                                                     (__nv_bfloat16*)out, ldo);                                \
    break;
  switch (cpp) {
    LY_PC(8) LY_PC(16) LY_PC(24) LY_PC(32) LY_PC(40)
    default: return fail(LY_EUNSUPPORTED, "ly_pconv3x3_fwd: cpp=%d not supported", cpp);
  }
#undef LY_PC
  LY_LAUNCH_CHECK();
  return LY_OK;
}
