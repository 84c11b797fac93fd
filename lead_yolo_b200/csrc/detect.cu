// Detect head post-processing (reference models/yolo.py:95-125): from the 1x1-conv output of one level,
//   raw [N][na][ny][nx][no]   = view(bs, na, no, ny, nx).permute(0, 1, 3, 4, 2)   (what training / the loss consume)
//   z   [N][rows_total][no]   rows of this level at row_off + a*ny*nx + y*nx + x :
//        xy = (2*sigmoid(t) - 0.5 + grid) * stride,  wh = (2*sigmoid(t))^2 * anchor * stride,  rest = sigmoid(t)
// fp32 throughout (bf16 here costs box pixels).  One thread per (n, y, x, a).
#include "ly_common.cuh"

namespace ly {

__global__ void __launch_bounds__(256) detect_decode_kernel(const float* __restrict__ p /*[N][ny][nx][na*no]*/, int N,
                                                            int ny, int nx, int na, int no, float stride,
                                                            const float* __restrict__ anchors /*[na][2], stride units*/,
                                                            float* __restrict__ raw, float* __restrict__ z,
                                                            long rows_total, long row_off) {
  pdl_enter();
  const long total = (long)N * ny * nx * na;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int a = (int)(t % na);
  const int x = (int)((t / na) % nx);
  const int y = (int)((t / ((long)na * nx)) % ny);
  const long n = t / ((long)na * nx * ny);
  const float* src = p + ((n * ny + y) * nx + x) * (long)(na * no) + (long)a * no;
  float* r = raw != nullptr ? raw + (((n * na + a) * ny + y) * nx + x) * (long)no : nullptr;
  float* zz = z != nullptr ? z + (n * rows_total + row_off + ((long)a * ny + y) * nx + x) * (long)no : nullptr;
  const float aw = anchors[a * 2] * stride, ah = anchors[a * 2 + 1] * stride;
  for (int j = 0; j < no; ++j) {
    const float v = src[j];
    if (r != nullptr) r[j] = v;
    if (zz != nullptr) {
      const float s = 1.f / (1.f + expf(-v));
      float o;
      if (j == 0) o = (s * 2.f + ((float)x - 0.5f)) * stride;
      else if (j == 1) o = (s * 2.f + ((float)y - 0.5f)) * stride;
      else if (j == 2) o = (s * 2.f) * (s * 2.f) * aw;
      else if (j == 3) o = (s * 2.f) * (s * 2.f) * ah;
      else o = s;
      zz[j] = o;
    }
  }
}

}  // namespace ly

using namespace ly;

// p: fp32 NHWC conv output of one Detect level ([N][ny][nx][na*no], dense).  raw and/or z may be NULL.
extern "C" int ly_detect_decode(const float* p, int N, int ny, int nx, int na, int no, float stride,
                                const float* anchors, float* raw, float* z, long rows_total, long row_off,
                                ly_stream_t st) {
  LY_REQUIRE(p && anchors && (raw || z) && N > 0 && ny > 0 && nx > 0 && na > 0 && no >= 5, "ly_detect_decode: bad args");
  LY_REQUIRE(z == nullptr || (row_off >= 0 && row_off + (long)na * ny * nx <= rows_total),
             "ly_detect_decode: rows out of range");
  const long total = (long)N * ny * nx * na;
  LY_CUDA(launch(detect_decode_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, p, N, ny, nx, na, no, stride, anchors, raw,
                                                                          z, rows_total, row_off));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
