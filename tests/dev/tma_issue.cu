// Micro-benchmark: cost of issuing TMA boxes from one thread, nb boxes per mbarrier phase (dev tool, not shipped).
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ly_tc.cuh"
using namespace ly;

template <int NB>
__global__ void __launch_bounds__(32, 1) k(const __grid_constant__ CUtensorMap m, int batches, uint32_t bytes, int nrows_total,
                                          int rows, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[2];
  if (threadIdx.x == 0) {
    mbar_init(&bar[0], 1); mbar_init(&bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    int row = (blockIdx.x * 977) % (nrows_total - rows * NB);
    for (int i = 0; i < batches; ++i) {
      const int s = i & 1;
      if (i >= 2) mbar_wait(&bar[s], ((i >> 1) - 1) & 1);
      mbar_arrive_expect_tx(&bar[s], bytes * NB);
#pragma unroll
      for (int b = 0; b < NB; ++b) tma_load_2d(smem + (s * NB + b) * 4096, &m, &bar[s], 0, row + b * rows);
      row += rows * NB;
      if (row + rows * NB > nrows_total) row = 0;
    }
    for (int i = batches; i < batches + 2; ++i) mbar_wait(&bar[i & 1], ((i >> 1) - 1) & 1);
    out[blockIdx.x] = clock64() - t0;
  }
}

template <int NB>
void run(const CUtensorMap& m, int R, long long* out) {
  const int batches = 1800 / NB;
  cudaFuncSetAttribute(k<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * NB * 4096 + 1024);
  for (int rep = 0; rep < 2; ++rep) k<NB><<<148, 32, 2 * NB * 4096 + 1024>>>(m, batches, 4096u, R, 128, out);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed\n"); return; }
  long long h[148];
  cudaMemcpy(h, out, 148 * 8, cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += (double)h[i];
  avg /= 148;
  printf("boxes per barrier %2d: %7.1f clk/box  %6.2f B/clk/SM\n", NB, avg / (batches * NB), 4096.0 * batches * NB / avg);
}

int main() {
  const int C = 128, R = 1 << 18;
  void* x;
  cudaMalloc(&x, (size_t)R * C * 2);
  cudaMemset(x, 1, (size_t)R * C * 2);
  long long* out;
  cudaMalloc(&out, 148 * 8);
  CUtensorMap m;
  cuuint64_t gd[2] = {(cuuint64_t)C, (cuuint64_t)R};
  cuuint64_t gs[1] = {(cuuint64_t)C * 2};
  cuuint32_t bd[2] = {16, 128};
  cuuint32_t es[2] = {1, 1};
  CUresult r = cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, x, gd, gs, bd, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                      CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r) { printf("encode failed %d\n", (int)r); return 1; }
  run<1>(m, R, out);
  run<3>(m, R, out);
  run<9>(m, R, out);
  return 0;
}
