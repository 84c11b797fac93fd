// Dev probe: where does each warp role of rf_fused_kernel wait?  (see rf_tc_probe.cu)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I lead_yolo_b200/csrc -I include \
//        tests/dev/rf_fused_probe.cu lead_yolo_b200/csrc/capi.cu -o tests/dev/rf_fused_probe.bin
#define RF_TC_DEBUG 1
#include "../../lead_yolo_b200/csrc/rf_generate_tc.cu"
#include <cstdio>

int main() {
  for (int cfg = 0; cfg < 2; ++cfg) {
    const int N = 64, H = cfg ? 40 : 80, W = H, C = cfg ? 256 : 128, s = 2, ho = H / 2, wo = W / 2, Cout = C;
    void *x, *wc, *w2, *out;
    float *shift, *se, *att, *sc, *sf;
    cudaMalloc(&x, (size_t)N * H * W * C * 2); cudaMemset(x, 0x3c, (size_t)N * H * W * C * 2);
    cudaMalloc(&wc, (size_t)C / 16 * 1296 * 2); cudaMemset(wc, 0x3c, (size_t)C / 16 * 1296 * 2);
    cudaMalloc(&shift, C * 9 * 4); cudaMemset(shift, 0, C * 9 * 4);
    cudaMalloc(&se, N * C * 4); cudaMemset(se, 0, N * C * 4);
    cudaMalloc(&att, (size_t)N * ho * 3 * wo * 3 * 4); cudaMemset(att, 0, (size_t)N * ho * 3 * wo * 3 * 4);
    cudaMalloc(&w2, (size_t)9 * C * Cout * 2); cudaMemset(w2, 0x3c, (size_t)9 * C * Cout * 2);
    cudaMalloc(&sc, Cout * 4); cudaMemset(sc, 0, Cout * 4);
    cudaMalloc(&sf, Cout * 4); cudaMemset(sf, 0, Cout * 4);
    cudaMalloc(&out, (size_t)N * ho * wo * Cout * 2);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms = 0;
    for (int rep = 0; rep < 4; ++rep) {
      if (rep == 3) cudaEventRecord(e0);
      int rc = ly_rfcbam_fused_infer(x, C, N, H, W, C, s, ho, wo, wc, shift, se, att, w2, sc, sf, Cout, out, Cout, nullptr);
      if (rc) { printf("error: %s\n", ly_last_error()); return 1; }
      if (rep == 3) cudaEventRecord(e1);
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed\n"); return 1; }
    cudaEventElapsedTime(&ms, e0, e1);
    static long long h[148][16];
    cudaMemcpyFromSymbol(h, ly::rf_tc_dbg, sizeof(h));
    const int items = cfg ? 32 : 48;
    printf("C=%d Cout=%d %dx%d: %.1f us, ~%d items per CTA\n", C, Cout, H, W, ms * 1e3, items);
    const char* names[12] = {"", "", "mma wait gen_ready", "mma wait a_full", "mma wait g2_ready", "mma total",
                             "writer wait tfull", "writer total", "epi(w2) wait tfull", "epi(w2) wait w_empty",
                             "epi(w2) wait out_full", "epi(w2) total"};
    for (int z = 2; z < 12; ++z) {
      double a = 0;
      int cnt = 0;
      for (int b = 0; b < 148; ++b) if (h[b][5] > 0) { a += (double)h[b][z]; ++cnt; }
      printf("  %-24s %9.0f clk  (%6.0f / item)\n", names[z], a / cnt, a / cnt / items);
    }
  }
  return 0;
}
