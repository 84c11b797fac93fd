// Dev probe: where does each warp role of rf_tc_kernel wait?  Compiles the kernel with RF_TC_DEBUG and prints, per role,
// cycles spent in each mbarrier wait and in total (average over CTAs).  Not shipped, not part of the library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I lead_yolo_b200/csrc -I include tests/dev/rf_tc_probe.cu \
//        lead_yolo_b200/csrc/capi.cu -o tests/dev/rf_tc_probe.bin -lcuda
#define RF_TC_DEBUG 1
#include "../../lead_yolo_b200/csrc/rf_generate_tc.cu"
#include <cstdio>
#include <vector>

int main() {
  const int N = 64, H = 80, W = 80, C = 128, s = 2, ho = 40, wo = 40;
  void *x, *wc, *A;
  float *shift, *se, *att, *maps;
  cudaMalloc(&x, (size_t)N * H * W * C * 2);
  cudaMemset(x, 0x3c, (size_t)N * H * W * C * 2);
  cudaMalloc(&wc, (size_t)C / 16 * 1296 * 2);
  cudaMemset(wc, 0x3c, (size_t)C / 16 * 1296 * 2);
  cudaMalloc(&shift, C * 9 * 4);
  cudaMemset(shift, 0, C * 9 * 4);
  cudaMalloc(&se, N * C * 4);
  cudaMemset(se, 0, N * C * 4);
  cudaMalloc(&att, (size_t)N * ho * 3 * wo * 3 * 4);
  cudaMemset(att, 0, (size_t)N * ho * 3 * wo * 3 * 4);
  cudaMalloc(&maps, (size_t)N * ho * 3 * wo * 3 * 8);
  cudaMalloc(&A, (size_t)N * ho * wo * 9 * C * 2);
  const char* names[12] = {"producer wait empty", "producer total", "mma wait tempty", "mma wait b1_full", "mma wait full",
                           "mma total", "writer wait b1_empty", "writer total", "epi(w2) wait tfull", "epi(w2) total",
                           "epi(w14) wait tfull", "epi(w14) total"};
  for (int mode = 0; mode < 2; ++mode) {
    for (int rep = 0; rep < 3; ++rep) {
      int rc = mode == 0 ? ly_rf_tc_stats(x, C, N, H, W, C, s, ho, wo, wc, shift, maps, nullptr)
                         : ly_rf_tc_apply(x, C, N, H, W, C, s, ho, wo, wc, shift, se, att, A, nullptr);
      if (rc) { printf("error: %s\n", ly_last_error()); return 1; }
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed\n"); return 1; }
    static long long h[148][16];
    cudaMemcpyFromSymbol(h, ly::rf_tc_dbg, sizeof(h));
    printf("mode %d (%s), ~49 items per CTA\n", mode, mode ? "apply" : "stats");
    for (int z = 0; z < 12; ++z) {
      double a = 0;
      for (int b = 0; b < 147; ++b) a += (double)h[b][z];
      printf("  %-22s %9.0f clk  (%6.0f / item)\n", names[z], a / 147, a / 147 / 49);
    }
  }
  return 0;
}
