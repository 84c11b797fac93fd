// Micro-benchmark: TMA tiled-load service time per box as a function of the box row width (dev tool, not shipped).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I lead_yolo_b200/csrc tests/dev/tma_rowrate.cu -o /tmp/tma_rowrate -lcuda
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ly_tc.cuh"
using namespace ly;

__global__ void __launch_bounds__(256, 1) k(const __grid_constant__ CUtensorMap m, int boxes, uint32_t bytes, int nrows_total,
                                          int rows, int cstep, int ncol, long long* out, int depth) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[64];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 64; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  __syncthreads();
  const int wid = threadIdx.x >> 5;
  uint64_t* bar = bars + wid * 8;
  uint8_t* sm = smem + wid * depth * 4096;
  if ((threadIdx.x & 31) == 0) {
    long long t0 = clock64();
    int row = ((blockIdx.x * 8 + wid) * 977) % (nrows_total - rows);
    for (int i = 0; i < boxes; ++i) {
      const int s = i % depth;
      if (i >= depth) mbar_wait(&bar[s], ((i / depth) - 1) & 1);
      mbar_arrive_expect_tx(&bar[s], bytes);
      tma_load_2d(sm + s * 4096, &m, &bar[s], (i % ncol) * cstep, row);
      row += rows;
      if (row + rows > nrows_total) row = 0;
    }
    for (int i = boxes; i < boxes + depth; ++i) mbar_wait(&bar[i % depth], ((i / depth) - 1) & 1);
    if (wid == 0) out[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  const int C = 128, R = 1 << 18;   // 64 MB bf16 matrix [R][C]: L2 resident after the first pass
  void* x;
  cudaMalloc(&x, (size_t)R * C * 2);
  cudaMemset(x, 1, (size_t)R * C * 2);
  long long* out;
  cudaMalloc(&out, 148 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 16384 + 1024);
  for (int cw = 16; cw <= 16; cw *= 2) {
    for (int rows = 128; rows <= 128; rows *= 2) {
      CUtensorMap m;
      cuuint64_t gd[2] = {(cuuint64_t)C, (cuuint64_t)R};
      cuuint64_t gs[1] = {(cuuint64_t)C * 2};
      cuuint32_t bd[2] = {(cuuint32_t)cw, (cuuint32_t)rows};
      cuuint32_t es[2] = {1, 1};
      CUtensorMapSwizzle sw = cw == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : cw == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
      CUresult r = cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, x, gd, gs, bd, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                          sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r) { printf("encode failed %d\n", (int)r); return 1; }
      for (int warps : {1, 2, 4, 8})
      for (int depth : {2, 4}) {
        const int grid = 148;
        const int boxes = 2000;
        for (int rep = 0; rep < 2; ++rep)
          k<<<grid, 32 * warps, 8 * 16384 + 1024>>>(m, boxes, (uint32_t)(cw * 2 * rows), R, rows, cw, C / cw, out, depth);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed\n"); return 1; }
        long long h[148];
        cudaMemcpy(h, out, grid * 8, cudaMemcpyDeviceToHost);
        double avg = 0;
        for (int i = 0; i < grid; ++i) avg += (double)h[i];
        avg /= grid;
        printf("row %3d B x %3d rows, warps %d depth %3d: %7.1f clk/box  %6.2f clk/row  %6.2f B/clk/SM\n", cw * 2, rows, warps, depth,
               avg / boxes, avg / boxes / rows, cw * 2.0 * rows * boxes * warps / avg);
      }
    }
  }
  return 0;
}
