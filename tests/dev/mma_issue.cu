// Micro-benchmark: tcgen05.mma issue rate from one thread (SS mode, SW128 K-major operands, garbage data), alone and with
// 16 busy ALU warps on the same SM (dev tool, not shipped).
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ly_tc.cuh"
using namespace ly;

__global__ void __launch_bounds__(576, 1) k(int N, int nmma, int per_commit, int busy, long long* out, float* sink, int two) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar, bar2;
  __shared__ uint32_t tmem_s;
  __shared__ volatile int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&bar2, 1); fence_mbar_init(); stop = 0; }
  if (warp == 2) tmem_alloc(&tmem_s, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  if (warp == 0 && two) {
    // second issuing lane (another SM sub-partition): does the issue rate double?
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16_f32(128, N, 0, 0);
      const uint64_t a0 = umma_smem_desc_sw128(smem_u32(smem), 16, 1024);
      const uint64_t b0 = umma_smem_desc_sw128(smem_u32(smem + 65536), 16, 1024);
      for (int i = 0; i < nmma; i += per_commit) {
        for (int j = 0; j < per_commit; ++j)
          umma_bf16(tmem_s + 256u, a0 + (uint64_t)((j & 3) * 2 + ((i + j) & 12) * 256), b0 + (uint64_t)((j & 3) * 2), idesc, 1u);
        umma_commit(&bar2);
      }
      mbar_wait(&bar2, (nmma / per_commit - 1) & 1);
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16_f32(128, N, 0, 0);
      const uint64_t a0 = umma_smem_desc_sw128(smem_u32(smem), 16, 1024);
      const uint64_t b0 = umma_smem_desc_sw128(smem_u32(smem + 65536), 16, 1024);
      const long long t0 = clock64();
      uint32_t ph = 0;
      for (int i = 0; i < nmma; i += per_commit) {
        for (int j = 0; j < per_commit; ++j)
          umma_bf16(tmem_s, a0 + (uint64_t)((j & 3) * 2 + ((i + j) & 12) * 256), b0 + (uint64_t)((j & 3) * 2), idesc, 1u);
        umma_commit(&bar);
        if (per_commit >= 64) { mbar_wait(&bar, ph); ph ^= 1; }
        (void)ph;
      }
      if (per_commit < 64) {
        // barrier phases completed = nmma / per_commit; wait for the last one
        const int phases = nmma / per_commit;
        mbar_wait(&bar, (phases - 1) & 1);
      }
      out[blockIdx.x] = clock64() - t0;
      stop = 1;
    }
  } else if (warp >= 2 && busy) {
    float a = threadIdx.x, b = 1.0001f;
    while (!stop) {
#pragma unroll
      for (int j = 0; j < 64; ++j) a = fmaf(a, b, 0.5f);
    }
    if (a == 12345.f) sink[0] = a;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_s, 512);
}

int main() {
  long long* out; cudaMalloc(&out, 148 * 8);
  float* sink; cudaMalloc(&sink, 4);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
  for (int two = 0; two < 2; ++two)
  for (int busy = 0; busy < 2; ++busy)
    for (int N : {64, 128, 256})
      for (int pc : {4, 32}) {
        const int nmma = 4096;
        for (int rep = 0; rep < 2; ++rep) k<<<148, 576, 160 * 1024>>>(N, nmma, pc, busy, out, sink, two);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
        long long h[148];
        cudaMemcpy(h, out, 148 * 8, cudaMemcpyDeviceToHost);
        double avg = 0;
        for (int i = 0; i < 148; ++i) avg += (double)h[i];
        avg /= 148;
        printf("lanes=%d busy=%d N=%3d MMAs/commit=%2d: %6.1f clk per MMA of lane 1 (floor %d)\n", two + 1, busy, N, pc, avg / nmma, 128 * N / 256);
      }
  return 0;
}
