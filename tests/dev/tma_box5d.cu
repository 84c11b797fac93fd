// Micro-benchmark: TMA service time of 5-D tap boxes (16 channels x Wt x Ht pixels) of an NHWC tensor (dev tool).
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ly_tc.cuh"
using namespace ly;

template <int NB>
__global__ void __launch_bounds__(32, 1) k(const __grid_constant__ CUtensorMap m, int batches, uint32_t bytes, int N, int tiles_h,
                                          int Ht, int groups, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[2];
  if (threadIdx.x == 0) {
    mbar_init(&bar[0], 1); mbar_init(&bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    int item = blockIdx.x * 48 * 3;
    for (int i = 0; i < batches; ++i, ++item) {
      const int s = i & 1;
      const int ti = item % 3, cg = (item / 3) % groups, mt = item / 3 / groups;
      const int th = mt % tiles_h, n = (mt / tiles_h) % N;
      if (i >= 2) mbar_wait(&bar[s], ((i >> 1) - 1) & 1);
      mbar_arrive_expect_tx(&bar[s], bytes * NB);
#pragma unroll
      for (int b = 0; b < NB; ++b) tma_load_5d(smem + (s * NB + b) * 4096, &m, &bar[s], cg * 16, b - 1, 0, th * Ht + ti - 1, n);
    }
    for (int i = batches; i < batches + 2; ++i) mbar_wait(&bar[i & 1], ((i >> 1) - 1) & 1);
    out[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  const int N = 64, H = 40, W = 40, C = 128;   // one parity plane of the 80x80x128 layer: pixel stride 2*C, row stride 2*80*C
  void* x;
  cudaMalloc(&x, (size_t)N * 80 * 80 * C * 2);
  cudaMemset(x, 1, (size_t)N * 80 * 80 * C * 2);
  long long* out;
  cudaMalloc(&out, 148 * 8);
  const int geo[3][2] = {{40, 3}, {16, 8}, {8, 16}};
  for (int g = 0; g < 3; ++g) {
    const int Wt = geo[g][0], Ht = geo[g][1];
    CUtensorMap m;
    cuuint64_t gd[5] = {(cuuint64_t)C, (cuuint64_t)W, 1, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t gs[4] = {(cuuint64_t)2 * C * 2, (cuuint64_t)80 * C * 2, (cuuint64_t)2 * 80 * C * 2, (cuuint64_t)80 * 80 * C * 2};
    cuuint32_t bd[5] = {16, (cuuint32_t)Wt, 1, (cuuint32_t)Ht, 1};
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, x, gd, gs, bd, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r) { printf("encode failed %d\n", (int)r); return 1; }
    const int batches = 48 * 3;
    cudaFuncSetAttribute(k<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 3 * 4096 + 1024);
    for (int rep = 0; rep < 3; ++rep)
      k<3><<<148, 32, 2 * 3 * 4096 + 1024>>>(m, batches, (uint32_t)(32 * Wt * Ht), N, (H + Ht - 1) / Ht, Ht, C / 16, out);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed\n"); return 1; }
    long long h[148];
    cudaMemcpy(h, out, 148 * 8, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += (double)h[i];
    avg /= 148;
    printf("box 16ch x %2d x %2d (%3d rows): %7.1f clk/box  %5.2f clk/row\n", Wt, Ht, Wt * Ht, avg / (batches * 3), avg / (batches * 3) / (Wt * Ht));
  }
  return 0;
}
