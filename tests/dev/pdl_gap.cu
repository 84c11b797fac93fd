// Micro-benchmark: cost of a kernel boundary inside a CUDA graph, with and without programmatic dependent launch.
// A chain of 64 persistent-style kernels (148 CTAs x 576 threads, 100 KB dynamic smem, ~W us of spinning each).
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(576, 1) k(int spin_clk, int pdl, float* sink) {
  extern __shared__ unsigned char smem[];
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // "prologue": touch smem
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) smem[i] = (unsigned char)i;
  __syncthreads();
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  const long long t0 = clock64();
  while (clock64() - t0 < spin_clk) {}
  if (smem[threadIdx.x & 1023] == 255 && spin_clk < 0) sink[0] = 1.f;
}

static float run(int pdl, int spin_clk, int chain) {
  cudaStream_t st; cudaStreamCreate(&st);
  float* sink; cudaMalloc(&sink, 4);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaGraph_t g; cudaGraphExec_t ge;
  cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
  for (int i = 0; i < chain; ++i) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(148); cfg.blockDim = dim3(576); cfg.dynamicSmemBytes = 100 * 1024; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, k, spin_clk, pdl, sink);
  }
  cudaStreamEndCapture(st, &g);
  if (cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) { printf("instantiate failed\n"); return -1; }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int w = 0; w < 3; ++w) cudaGraphLaunch(ge, st);
  cudaEventRecord(e0, st);
  for (int w = 0; w < 10; ++w) cudaGraphLaunch(ge, st);
  cudaEventRecord(e1, st);
  cudaStreamSynchronize(st);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms / 10 * 1000 / chain;
}

int main() {
  for (int spin : {2000, 20000, 60000}) {
    float a = run(0, spin, 64), b = run(1, spin, 64);
    printf("spin %6d clk (%.1f us at 1.9 GHz): %.2f us per kernel plain, %.2f us with PDL (err %s)\n", spin, spin / 1900.0, a, b,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
