// Dev probe: where does each warp role of conv_gemm_kernel wait?  Not shipped, not part of the library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I lead_yolo_b200/csrc -I include tests/dev/conv_probe.cu \
//        lead_yolo_b200/csrc/capi.cu -o tests/dev/conv_probe.bin -lcuda
//   conv_probe.bin N H W k cout act res f32 cin0 [cin1 ...]
#define CONV_GEMM_DEBUG 1
#include "../../lead_yolo_b200/csrc/conv_gemm.cu"
#include <cstdio>
#include <cstdlib>

int main(int argc, char** argv) {
  if (argc < 10) { printf("usage: N H W k cout act res f32 cin...\n"); return 1; }
  ly_conv_desc d;
  memset(&d, 0, sizeof(d));
  d.N = atoi(argv[1]); d.H = atoi(argv[2]); d.W = atoi(argv[3]);
  const int k = atoi(argv[4]);
  d.kh = d.kw = k; d.stride = (k == 3 || k == 1) ? 1 : k; d.pad = k == 3 ? 1 : 0;
  d.Cout = atoi(argv[5]); d.act = atoi(argv[6]);
  const int res = atoi(argv[7]);
  d.out_fp32 = atoi(argv[8]);
  d.n_src = argc - 9;
  const long pix = (long)d.N * d.H * d.W;
  for (int s = 0; s < d.n_src; ++s) {
    d.src_c[s] = atoi(argv[9 + s]); d.src_ld[s] = d.src_c[s];
    void* x; cudaMalloc(&x, pix * d.src_c[s] * 2); cudaMemset(x, 0x3c, pix * d.src_c[s] * 2);
    d.src[s] = x;
  }
  const int nk = ly_conv_num_chunks(&d);
  void* w; cudaMalloc(&w, (size_t)d.Cout * nk * 64 * 2); cudaMemset(w, 0x3c, (size_t)d.Cout * nk * 64 * 2);
  d.w_packed = w;
  const long opix = d.stride > 1 ? pix / (d.stride * d.stride) : pix;
  void* out; cudaMalloc(&out, opix * d.Cout * 4);
  d.out = out; d.out_ld = d.Cout;
  if (res) { void* r; cudaMalloc(&r, opix * d.Cout * 2); cudaMemset(r, 0, opix * d.Cout * 2); d.residual = r; d.res_ld = d.Cout; }
  float* sc; cudaMalloc(&sc, d.Cout * 4); cudaMemset(sc, 0, d.Cout * 4);
  d.scale = sc; d.shift = sc;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  // a 256 MB scrub between repetitions evicts the operands from L2; the median of 5 timed launches is reported
  void* scrub; cudaMalloc(&scrub, 256 << 20);
  float times[5];
  for (int rep = 0; rep < 6; ++rep) {
    cudaMemsetAsync(scrub, rep, 256 << 20);
    cudaEventRecord(e0);
    int rc_;
    if (getenv("PROBE_DETECT")) {   // Detect head: na = 3, no = Cout / 3, outputs raw + z
      static float *raw_ = nullptr, *z_ = nullptr, *anc_ = nullptr;
      if (!raw_) { cudaMalloc(&raw_, opix * d.Cout * 4); cudaMalloc(&z_, opix * d.Cout * 4); cudaMalloc(&anc_, 64); cudaMemset(anc_, 0, 64); }
      rc_ = ly_detect_head_infer(d.src[0], d.src_ld[0], d.N, d.H, d.W, d.src_c[0], d.w_packed, sc, 3, d.Cout / 3, 8.f, anc_, getenv("PROBE_NORAW") ? nullptr : raw_, z_,
                                 (long)3 * d.H * d.W, 0, nullptr);
    } else {
      rc_ = ly_conv_fwd(&d, nullptr);
    }
    if (rc_) { printf("error: %s\n", ly_last_error()); return 1; }
    cudaEventRecord(e1);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed\n"); return 1; }
    if (rep > 0) cudaEventElapsedTime(&times[rep - 1], e0, e1);
  }
  for (int i = 0; i < 5; ++i) for (int j = i + 1; j < 5; ++j) if (times[j] < times[i]) { float t = times[i]; times[i] = times[j]; times[j] = t; }
  float ms = times[2];
  static long long h[148][16];
  cudaMemcpyFromSymbol(h, ly::conv_dbg, sizeof(h));
  const char* names[9] = {"producer wait empty", "producer total", "mma wait tmem_empty", "mma wait full", "mma total",
                          "epi(w2) wait tmem_full", "epi(w2) total", "epi(w6) wait tmem_full", "epi(w6) total"};
  printf("N=%d %dx%d k=%d cout=%d act=%d res=%d f32=%d nsrc=%d chunks=%d : %.1f us\n", d.N, d.H, d.W, k, d.Cout, d.act, res,
         d.out_fp32, d.n_src, nk, ms * 1e3);
  for (int z = 0; z < 9; ++z) {
    double a = 0;
    for (int b = 0; b < 148; ++b) a += (double)h[b][z];
    printf("  %-24s %9.0f clk\n", names[z], a / 148);
  }
  return 0;
}
