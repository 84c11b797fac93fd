// Dev probe: where does each role of fasternet_block_kernel wait?  Not shipped.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I lead_yolo_b200/csrc -I include tests/dev/fnb_probe.cu \
//        lead_yolo_b200/csrc/capi.cu -o tests/dev/fnb_probe.bin -lcuda
#define FNB_DEBUG 1
#include "../../lead_yolo_b200/csrc/fasternet_block.cu"
#include <cstdio>
#include <cstdlib>

int main(int argc, char** argv) {
  const int C = argc > 1 ? atoi(argv[1]) : 24, HW = argc > 2 ? atoi(argv[2]) : 160, N = 64;
  const int C2 = 2 * C, N2 = (C + 15) / 16 * 16, chunks = (C2 + 63) / 64;
  int off_w2 = C2 * 128, off_wc = off_w2 + chunks * N2 * 128, w_bytes = off_wc + 9 * C2 * 32;
  if (C == 80) { off_w2 = ly::f80::kOffW2; off_wc = ly::f80::kOffWc; w_bytes = ly::f80::kWBytes; }
  void *x, *y, *w; float* sc;
  cudaMalloc(&x, (size_t)N * HW * HW * C * 2); cudaMemset(x, 0x3c, (size_t)N * HW * HW * C * 2);
  cudaMalloc(&y, (size_t)N * HW * HW * C * 2);
  cudaMalloc(&w, w_bytes); cudaMemset(w, 0x3c, w_bytes);
  cudaMalloc(&sc, C2 * 4); cudaMemset(sc, 0, C2 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 3; ++rep) {
    if (rep == 2) cudaEventRecord(e0);
    if (ly_fasternet_block_infer(x, y, N, HW, HW, C, w, w_bytes, off_w2, off_wc, sc, sc, nullptr)) { printf("error: %s\n", ly_last_error()); return 1; }
  }
  cudaEventRecord(e1);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed\n"); return 1; }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  static long long h[296][16];
  cudaMemcpyFromSymbol(h, ly::fnb_dbg, sizeof(h));
  const char* names[14] = {"producer wait x_free", "producer wait u_full(halo)", "producer total", "mma wait x_full", "mma wait taps_full",
                           "mma wait v_ready", "mma wait y_free", "mma total", "hidden wait u_full", "hidden wait y_full(prev)", "hidden total",
                           "output wait u_full", "output wait y_full", "output total"};
  const int idx[14] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13};
  int nb = 0; for (int b = 0; b < 296; ++b) if (h[b][7]) ++nb;
  const double tiles = (double)N * ((HW + 13) / 14) * ((HW + 7) / 8) / nb;   // pitch-16 tiles (14 x 8)
  printf("C=%d %dx%d: %.1f us, %d CTAs, ~%.1f tiles each\n", C, HW, HW, ms * 1e3, nb, tiles);
  for (int z = 0; z < 14; ++z) {
    double a = 0;
    for (int b = 0; b < 296; ++b) a += (double)h[b][idx[z]];
    printf("  %-28s %9.0f clk (%6.0f / tile)\n", names[z], a / nb, a / nb / tiles);
  }
  if (C == 80) {   // %globaltimer stamps at kernel entry / exit of every CTA
    long long t0 = h[0][14], t1 = h[0][15]; double in = 0;
    for (int b = 0; b < nb; ++b) { if (h[b][14] < t0) t0 = h[b][14]; if (h[b][15] > t1) t1 = h[b][15]; in += (double)(h[b][15] - h[b][14]); }
    long long e_last = 0; for (int b = 0; b < nb; ++b) if (h[b][14] > e_last) e_last = h[b][14];
    printf("  first entry -> last exit %.1f us; last entry at +%.1f us; mean CTA lifetime %.1f us\n", (t1 - t0) * 1e-3, (e_last - t0) * 1e-3, in / nb * 1e-3);
  }
  return 0;
}
