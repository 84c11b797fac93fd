// Spatial helper kernels (CUDA cores, HBM-bound): stem patch embedding from the raw image, 5x5 max pooling
// chain of SPPF, nearest 2x upsample / concat slice copies, global average pooling.
#include <atomic>

#include "ly_common.cuh"

namespace ly {

__device__ __forceinline__ uint4 ldv(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// ------------------------------------------------------------------------------------------------
// Stem: PatchEmbed_FasterNet.proj on the input image (reference models/common.py:1531,1537-1539):
// 4x4 stride-4 convolution, 3 -> COUT channels, straight from the NCHW image (fp32, or uint8 with the
// reference's /255 folded into in_scale — train.py:295) to NHWC bf16.  out = scale*conv + shift.
// One thread per output pixel; weights [48][COUT] in shared memory (broadcast reads).
// ------------------------------------------------------------------------------------------------
template <int COUT, typename Tin>
__global__ void __launch_bounds__(128) stem_kernel(const Tin* __restrict__ img, int N, int H, int W,
                                                   const float* __restrict__ wk /*[48][COUT]*/,
                                                   const float* __restrict__ scale, const float* __restrict__ shift,
                                                   float in_scale, __nv_bfloat16* __restrict__ out, long ld_out) {
  pdl_enter();
  __shared__ float sw[48 * COUT];
  for (int i = threadIdx.x; i < 48 * COUT; i += blockDim.x) sw[i] = wk[i] * in_scale;
  __syncthreads();
  const int Ho = H / 4, Wo = W / 4;
  const long total = (long)N * Ho * Wo;
  const long pix = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= total) return;
  const int wo = (int)(pix % Wo);
  const int ho = (int)((pix / Wo) % Ho);
  const int n = (int)(pix / ((long)Wo * Ho));
  float acc[COUT];
#pragma unroll
  for (int j = 0; j < COUT; ++j) acc[j] = 0.f;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int ky = 0; ky < 4; ++ky) {
      const Tin* src = img + (((long)n * 3 + c) * H + (ho * 4 + ky)) * W + wo * 4;
      float xv[4];
      if constexpr (sizeof(Tin) == 4) {
        const float4 f = __ldg(reinterpret_cast<const float4*>(src));
        xv[0] = f.x; xv[1] = f.y; xv[2] = f.z; xv[3] = f.w;
      } else {
        const uchar4 u = __ldg(reinterpret_cast<const uchar4*>(src));
        xv[0] = u.x; xv[1] = u.y; xv[2] = u.z; xv[3] = u.w;
      }
#pragma unroll
      for (int kx = 0; kx < 4; ++kx) {
        const float* wrow = &sw[((c * 4 + ky) * 4 + kx) * COUT];
#pragma unroll
        for (int j = 0; j < COUT; j += 4) {
          const float4 w4 = *reinterpret_cast<const float4*>(wrow + j);
          acc[j] = fmaf(xv[kx], w4.x, acc[j]);
          acc[j + 1] = fmaf(xv[kx], w4.y, acc[j + 1]);
          acc[j + 2] = fmaf(xv[kx], w4.z, acc[j + 2]);
          acc[j + 3] = fmaf(xv[kx], w4.w, acc[j + 3]);
        }
      }
    }
  }
  __nv_bfloat16* o = out + pix * ld_out;
#pragma unroll
  for (int j = 0; j < COUT; j += 8) {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float sc = scale != nullptr ? __ldg(scale + j + i) : 1.f;
      const float sh = shift != nullptr ? __ldg(shift + j + i) : 0.f;
      v[i] = fmaf(acc[j + i], sc, sh);
    }
    *reinterpret_cast<uint4*>(o + j) = pack8(v);
  }
}

// Stem weight gradient: dW[k][co] += Σ_pix x[pix][k] * dz[pix][co]   (k = (c,ky,kx), 48 values)
// block = 192 threads = 48 k  x  4 co-groups; pixels staged through shared memory in chunks of 64.
template <int COUT, typename Tin>
__global__ void __launch_bounds__(192) stem_wgrad_kernel(const Tin* __restrict__ img, int N, int H, int W,
                                                         const __nv_bfloat16* __restrict__ dz, long ld_dz,
                                                         float in_scale, float* __restrict__ dwk /*[48][COUT]*/) {
  pdl_enter();
  constexpr int P = 64;
  constexpr int CG = COUT / 4;
  __shared__ float sx[P][49];
  __shared__ float sd[P][COUT + 1];
  const int Ho = H / 4, Wo = W / 4;
  const long total = (long)N * Ho * Wo;
  const int k = threadIdx.x % 48, g = threadIdx.x / 48;
  float acc[CG];
#pragma unroll
  for (int j = 0; j < CG; ++j) acc[j] = 0.f;
  for (long base = (long)blockIdx.x * P; base < total; base += (long)gridDim.x * P) {
    __syncthreads();
    // one thread per (pixel, channel, kernel row): the four kx values are one aligned 4-element run of the image
    for (int i = threadIdx.x; i < P * 12; i += blockDim.x) {
      const int p = i / 12, ck = i % 12;
      const long pix = base + p;
      float v4[4] = {0.f, 0.f, 0.f, 0.f};
      if (pix < total) {
        const int wo = (int)(pix % Wo), ho = (int)((pix / Wo) % Ho), n = (int)(pix / ((long)Wo * Ho));
        const int c = ck >> 2, ky = ck & 3;
        const Tin* src = img + (((long)n * 3 + c) * H + (ho * 4 + ky)) * W + wo * 4;
        if (sizeof(Tin) == 1) {
          const uchar4 u = *reinterpret_cast<const uchar4*>(src);
          v4[0] = (float)u.x * in_scale; v4[1] = (float)u.y * in_scale;
          v4[2] = (float)u.z * in_scale; v4[3] = (float)u.w * in_scale;
        } else {
          const float4 f = *reinterpret_cast<const float4*>(src);
          v4[0] = f.x * in_scale; v4[1] = f.y * in_scale; v4[2] = f.z * in_scale; v4[3] = f.w * in_scale;
        }
      }
#pragma unroll
      for (int kx = 0; kx < 4; ++kx) sx[p][ck * 4 + kx] = v4[kx];
    }
    for (int i = threadIdx.x; i < P * COUT; i += blockDim.x) {
      const int p = i / COUT, co = i % COUT;
      const long pix = base + p;
      sd[p][co] = pix < total ? __bfloat162float(dz[pix * ld_dz + co]) : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int p = 0; p < P; ++p) {
      const float xv = sx[p][k];
#pragma unroll
      for (int j = 0; j < CG; ++j) acc[j] = fmaf(xv, sd[p][g * CG + j], acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < CG; ++j) atomicAdd(&dwk[k * COUT + g * CG + j], acc[j]);
}

// ------------------------------------------------------------------------------------------------
// 5x5 / stride 1 / pad 2 max pooling on an NHWC channel slice (SPPF, models/common.py:356,364-367).
// One thread per (pixel, 8-channel vector).  idx (optional) records the arg-max tap 0..24 (first maximum
// in row-major window order, as ATen) for the backward pass.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) maxpool5_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int N, int H,
                                                       int W, int V, __nv_bfloat16* __restrict__ y, long ldy,
                                                       uint8_t* __restrict__ idx) {
  pdl_enter();
  const long total = (long)N * H * W * V;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int v = (int)(t % V);
  const long pix = t / V;
  const int w = (int)(pix % W), h = (int)((pix / W) % H), n = (int)(pix / ((long)W * H));
  float best[8];
  int bi[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; }
  for (int dy = -2; dy <= 2; ++dy) {
    const int hh = h + dy;
    if (hh < 0 || hh >= H) continue;
    for (int dx = -2; dx <= 2; ++dx) {
      const int ww = w + dx;
      if (ww < 0 || ww >= W) continue;
      float f[8];
      unpack8(ldv(x + (((long)n * H + hh) * W + ww) * ldx + v * 8), f);
      const int tap = (dy + 2) * 5 + (dx + 2);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (f[j] > best[j]) { best[j] = f[j]; bi[j] = tap; }
    }
  }
  *reinterpret_cast<uint4*>(y + pix * ldy + v * 8) = pack8(best);
  if (idx != nullptr) {
    uint2 pk;
    pk.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
    pk.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
    *reinterpret_cast<uint2*>(idx + pix * (long)V * 8 + v * 8) = pk;
  }
}

// SPPF's three cascaded 5x5 pools in one launch (inference): y1 = mp5(x), y2 = mp5(y1) = mp9(x), y3 = mp5(y2) = mp13(x).
// One block per (image, 8-channel vector): the H x W plane of that vector sits in shared memory; a separable pass
// (rows, then columns) produces the 5 / 9 / 13 windows with packed bf16 maxima.  Out-of-image taps are skipped, which
// equals MaxPool2d's -inf padding.
__device__ __forceinline__ uint4 max8(const uint4& a, const uint4& b) {
  uint4 r;
  __nv_bfloat162 t;
  t = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a.x), *reinterpret_cast<const __nv_bfloat162*>(&b.x));
  r.x = *reinterpret_cast<uint32_t*>(&t);
  t = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a.y), *reinterpret_cast<const __nv_bfloat162*>(&b.y));
  r.y = *reinterpret_cast<uint32_t*>(&t);
  t = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a.z), *reinterpret_cast<const __nv_bfloat162*>(&b.z));
  r.z = *reinterpret_cast<uint32_t*>(&t);
  t = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a.w), *reinterpret_cast<const __nv_bfloat162*>(&b.w));
  r.w = *reinterpret_cast<uint32_t*>(&t);
  return r;
}

__global__ void __launch_bounds__(256) sppf_pool3_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int H, int W,
                                                         int V, __nv_bfloat16* __restrict__ y, long ldy, long ystep) {
  pdl_enter();
  extern __shared__ uint4 sp[];                 // [4][H*W]: input, row-max 5, 9, 13
  const int HW = H * W;
  uint4* sIn = sp;
  uint4* s5 = sp + HW;
  uint4* s9 = sp + 2 * HW;
  uint4* s13 = sp + 3 * HW;
  const int n = blockIdx.x / V, v = blockIdx.x - n * V;
  const __nv_bfloat16* src = x + (long)n * HW * ldx + v * 8;
  for (int p = threadIdx.x; p < HW; p += blockDim.x) sIn[p] = ldv(src + (long)p * ldx);
  __syncthreads();
  for (int p = threadIdx.x; p < HW; p += blockDim.x) {
    const int h = p / W, w = p - h * W;
    const uint4* row = sIn + h * W;
    uint4 m = row[w];
#pragma unroll
    for (int d = 1; d <= 6; ++d) {
      if (w - d >= 0) m = max8(m, row[w - d]);
      if (w + d < W) m = max8(m, row[w + d]);
      if (d == 2) s5[p] = m;
      if (d == 4) s9[p] = m;
    }
    s13[p] = m;
  }
  __syncthreads();
  __nv_bfloat16* dst = y + (long)n * HW * ldy + v * 8;
  for (int p = threadIdx.x; p < HW; p += blockDim.x) {
    const int h = p / W, w = p - h * W;
    uint4 m5 = s5[p], m9 = s9[p], m13 = s13[p];
#pragma unroll
    for (int d = 1; d <= 6; ++d) {
#pragma unroll
      for (int sgn = -1; sgn <= 1; sgn += 2) {
        const int hh = h + sgn * d;
        if (hh >= 0 && hh < H) {
          const int q = hh * W + w;
          if (d <= 2) m5 = max8(m5, s5[q]);
          if (d <= 4) m9 = max8(m9, s9[q]);
          m13 = max8(m13, s13[q]);
        }
      }
    }
    __nv_bfloat16* o = dst + (long)p * ldy;
    *reinterpret_cast<uint4*>(o) = m5;
    *reinterpret_cast<uint4*>(o + ystep) = m9;
    *reinterpret_cast<uint4*>(o + 2 * ystep) = m13;
  }
}

// dx[i] (+)= Σ_{outputs o whose recorded arg-max is i} dy[o]   (gather form: no atomics)
__global__ void __launch_bounds__(256) maxpool5_bwd_kernel(const __nv_bfloat16* __restrict__ dy, long lddy,
                                                           const uint8_t* __restrict__ idx, int N, int H, int W,
                                                           int V, __nv_bfloat16* __restrict__ dx, long lddx,
                                                           int accumulate) {
  pdl_enter();
  const long total = (long)N * H * W * V;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int v = (int)(t % V);
  const long pix = t / V;
  const int w = (int)(pix % W), h = (int)((pix / W) % H), n = (int)(pix / ((long)W * H));
  float acc[8];
  if (accumulate) {
    unpack8(*reinterpret_cast<const uint4*>(dx + pix * lddx + v * 8), acc);
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  }
  for (int dy_ = -2; dy_ <= 2; ++dy_) {
    const int ho = h + dy_;  // output row whose window contains (h,w)
    if (ho < 0 || ho >= H) continue;
    for (int dx_ = -2; dx_ <= 2; ++dx_) {
      const int wo = w + dx_;
      if (wo < 0 || wo >= W) continue;
      const long op = ((long)n * H + ho) * W + wo;
      // tap index of (h,w) inside the window of output (ho,wo)
      const int tap = (h - ho + 2) * 5 + (w - wo + 2);
      const uint2 pk = __ldg(reinterpret_cast<const uint2*>(idx + op * (long)V * 8 + v * 8));
      float g[8];
      unpack8(ldv(dy + op * lddy + v * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int b = (int)(((j < 4 ? pk.x : pk.y) >> ((j & 3) * 8)) & 0xffu);
        if (b == tap) acc[j] += g[j];
      }
    }
  }
  *reinterpret_cast<uint4*>(dx + pix * lddx + v * 8) = pack8(acc);
}

// ------------------------------------------------------------------------------------------------
// Channel-slice copy with optional nearest 2x upsample (nn.Upsample(None,2,'nearest') + Concat,
// LEAD-YOLO.yaml:29-34, models/common.py:531-538): dst[n, h, w, :C] = src[n, h/up, w/up, :C].
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) copy_up_kernel(const __nv_bfloat16* __restrict__ src, long lds, int N, int Hs,
                                                      int Ws, int V, int up, __nv_bfloat16* __restrict__ dst,
                                                      long ldd) {
  pdl_enter();
  const int H = Hs * up, W = Ws * up;
  const long total = (long)N * H * W * V;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int v = (int)(t % V);
  const long pix = t / V;
  const int w = (int)(pix % W), h = (int)((pix / W) % H), n = (int)(pix / ((long)W * H));
  const long sp = ((long)n * Hs + h / up) * Ws + w / up;
  *reinterpret_cast<uint4*>(dst + pix * ldd + v * 8) = ldv(src + sp * lds + v * 8);
}

// up == 2: one thread per SOURCE vector, two of them in flight: one load feeds four stores
__global__ void __launch_bounds__(256) copy_up2_kernel(const __nv_bfloat16* __restrict__ src, long lds, int N, int Hs,
                                                       int Ws, int V, __nv_bfloat16* __restrict__ dst, long ldd) {
  pdl_enter();
  const long total = (long)N * Hs * Ws * V;
  const long t0 = ((long)blockIdx.x * blockDim.x + threadIdx.x);
  const long stride = (long)gridDim.x * blockDim.x;
  const int W = Ws * 2;
  uint4 u[2];
  long sp[2];
  int vv[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const long t = t0 + k * stride;
    vv[k] = (int)(t % V);
    sp[k] = t / V;
    if (t < total) u[k] = ldv(src + sp[k] * lds + vv[k] * 8);
  }
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    if (t0 + k * stride < total) {
      const int w = (int)(sp[k] % Ws);
      const long nh = sp[k] / Ws;               // n * Hs + h
      __nv_bfloat16* o = dst + ((nh * 2) * W + 2 * w) * ldd + vv[k] * 8;
      *reinterpret_cast<uint4*>(o) = u[k];
      *reinterpret_cast<uint4*>(o + ldd) = u[k];
      *reinterpret_cast<uint4*>(o + (long)W * ldd) = u[k];
      *reinterpret_cast<uint4*>(o + (long)W * ldd + ldd) = u[k];
    }
  }
}

// Backward of the 2x nearest upsample: dst[n,h,w] = Σ_{a,b<2} src[n,2h+a,2w+b]
__global__ void __launch_bounds__(256) downsum2_kernel(const __nv_bfloat16* __restrict__ src, long lds, int N, int H,
                                                       int W, int V, __nv_bfloat16* __restrict__ dst, long ldd) {
  pdl_enter();
  const long total = (long)N * H * W * V;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int v = (int)(t % V);
  const long pix = t / V;
  const int w = (int)(pix % W), h = (int)((pix / W) % H), n = (int)(pix / ((long)W * H));
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      float f[8];
      unpack8(ldv(src + (((long)n * 2 * H + 2 * h + a) * (2 * W) + 2 * w + b) * lds + v * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  *reinterpret_cast<uint4*>(dst + pix * ldd + v * 8) = pack8(acc);
}

// ------------------------------------------------------------------------------------------------
// Global average pool over H*W per (image, channel) -> bf16 [N][C]  (rfa.SE.gap, models/rfa.py:80,90).
// block = (32 channel vectors) x (16 pixel lanes); grid = (ceil(V/32), N).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) gap_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int HW, int V,
                                                  __nv_bfloat16* __restrict__ out, long ldo) {
  pdl_enter();
  __shared__ float red[16][32][9];
  const int vl = threadIdx.x & 31, pl = threadIdx.x >> 5;
  const int v = blockIdx.x * 32 + vl;
  const int n = blockIdx.y;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (v < V) {
    const __nv_bfloat16* base = x + (long)n * HW * ldx + v * 8;
    for (int p = pl; p < HW; p += 16) {
      float f[8];
      unpack8(ldv(base + (long)p * ldx), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[pl][vl][j] = acc[j];
  __syncthreads();
  if (pl == 0 && v < V) {
    float s[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = 0.f;
    for (int q = 0; q < 16; ++q)
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += red[q][vl][j];
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] *= inv;
    *reinterpret_cast<uint4*>(out + (long)n * ldo + v * 8) = pack8(s);
  }
}

static int chk(const void* p, long ld, int C, const char* what) {
  LY_REQUIRE(p != nullptr, "%s: null pointer", what);
  LY_REQUIRE(C > 0 && C % 8 == 0 && ld >= C && ld % 8 == 0, "%s: C=%d ld=%ld invalid", what, C, ld);
  LY_REQUIRE(reinterpret_cast<uintptr_t>(p) % 16 == 0, "%s: pointer not 16-byte aligned", what);
  return LY_OK;
}

template <typename Tin>
static int stem_dispatch(const Tin* img, int N, int H, int W, int Cout, const float* wk, const float* scale,
                         const float* shift, float in_scale, __nv_bfloat16* out, long ld_out, cudaStream_t st) {
  const long total = (long)N * (H / 4) * (W / 4);
  const int grid = ceil_div(total, 128);
#define LY_STEM(CO)                                                                                          \
  case CO:                                                                                                   \
    LY_CUDA(launch(stem_kernel<CO, Tin>, grid, 128, 0, st, img, N, H, W, wk, scale, shift, in_scale, out, ld_out));       \
    break;
  switch (Cout) {
    LY_STEM(8) LY_STEM(16) LY_STEM(24) LY_STEM(32) LY_STEM(40) LY_STEM(48) LY_STEM(64)
    default: return fail(LY_EUNSUPPORTED, "ly_stem_fwd: Cout=%d not supported", Cout);
  }
#undef LY_STEM
  LY_LAUNCH_CHECK();
  return LY_OK;
}

template <typename Tin>
static int stem_wgrad_dispatch(const Tin* img, int N, int H, int W, int Cout, const __nv_bfloat16* dz, long ld_dz,
                               float in_scale, float* dwk, cudaStream_t st) {
  const int grid = device_sm_count() * 4;
#define LY_STEMW(CO)                                                                               \
  case CO:                                                                                         \
    LY_CUDA(launch(stem_wgrad_kernel<CO, Tin>, grid, 192, 0, st, img, N, H, W, dz, ld_dz, in_scale, dwk));      \
    break;
  switch (Cout) {
    LY_STEMW(8) LY_STEMW(16) LY_STEMW(24) LY_STEMW(32) LY_STEMW(40) LY_STEMW(48) LY_STEMW(64)
    default: return fail(LY_EUNSUPPORTED, "ly_stem_wgrad: Cout=%d not supported", Cout);
  }
#undef LY_STEMW
  LY_LAUNCH_CHECK();
  return LY_OK;
}


// Patch matrix of the stem for its weight gradient: P bf16 [N*(H/4)*(W/4)][64], column c*16 + ky*4 + kx = the raw pixel
// img[n][c][4y+ky][4x+kx] (uint8 values are exact in bf16), columns 48..63 zero.  The stem's weight gradient is then a
// 1x1 convolution's: dW[o][k] = Σ_pixels dz[pixel][o] * P[pixel][k] on the tensor cores (ly_conv_wgrad) instead of
// 48*Cout fp32 FMAs per pixel on the CUDA cores.  Four threads per patch, one per 32-byte quarter of its row.
template <typename Tin>
__global__ void __launch_bounds__(256) stem_im2col_kernel(const Tin* __restrict__ img, int N, int H, int W,
                                                          __nv_bfloat16* __restrict__ P) {
  pdl_enter();
  const int Ho = H / 4, Wo = W / 4;
  const long total = (long)N * Ho * Wo * 4;
  const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int c = (int)(tid & 3);
  const long pix = tid >> 2;
  const int x = (int)(pix % Wo), y = (int)((pix / Wo) % Ho);
  const long n = pix / ((long)Wo * Ho);
  uint32_t w[8];
  if (c == 3) {
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = 0u;
  } else {
    const Tin* base = img + ((n * 3 + c) * H + (long)y * 4) * W + (long)x * 4;
#pragma unroll
    for (int ky = 0; ky < 4; ++ky) {
      float f[4];
      if (sizeof(Tin) == 1) {
        const uchar4 u = *reinterpret_cast<const uchar4*>(base + (long)ky * W);
        f[0] = (float)u.x; f[1] = (float)u.y; f[2] = (float)u.z; f[3] = (float)u.w;
      } else {
        const float4 u = *reinterpret_cast<const float4*>(base + (long)ky * W);
        f[0] = u.x; f[1] = u.y; f[2] = u.z; f[3] = u.w;
      }
      w[ky * 2] = pack_bf16x2(f[0], f[1]);
      w[ky * 2 + 1] = pack_bf16x2(f[2], f[3]);
    }
  }
  uint4* dst = reinterpret_cast<uint4*>(P + pix * 64 + c * 16);
  dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
  dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
}

}  // namespace ly

namespace ly {
int stem_tc_u8(const uint8_t* img, int N, int H, int W, int Cout, const float* wk, const float* scale, const float* shift,
               float in_scale, __nv_bfloat16* out, long ld_out, cudaStream_t st);   // stem_tc.cu
}
using namespace ly;

extern "C" int ly_stem_fwd(const void* img, int img_is_u8, int N, int H, int W, int Cout, const float* wk,
                           const float* scale, const float* shift, float in_scale, void* out, long ld_out,
                           ly_stream_t st) {
  LY_REQUIRE(img && wk && out, "ly_stem_fwd: null pointer");
  LY_REQUIRE(N > 0 && H % 4 == 0 && W % 4 == 0 && H > 0 && W > 0, "ly_stem_fwd: H,W must be multiples of 4");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(img) % 16 == 0, "ly_stem_fwd: image not 16-byte aligned");
  int rc = chk(out, ld_out, Cout, "ly_stem_fwd(out)");
  if (rc) return rc;
  if (img_is_u8) {
    // uint8 pixels are exact in bf16: the tensor-core kernel (stem_tc.cu) replaces 48*Cout FP32 FMAs per pixel
    if (Cout % 8 == 0 && Cout <= 64 && W % 16 == 0)
      return stem_tc_u8((const uint8_t*)img, N, H, W, Cout, wk, scale, shift, in_scale, (__nv_bfloat16*)out, ld_out,
                        (cudaStream_t)st);
    return stem_dispatch<uint8_t>((const uint8_t*)img, N, H, W, Cout, wk, scale, shift, in_scale,
                                  (__nv_bfloat16*)out, ld_out, (cudaStream_t)st);
  }
  return stem_dispatch<float>((const float*)img, N, H, W, Cout, wk, scale, shift, in_scale, (__nv_bfloat16*)out,
                              ld_out, (cudaStream_t)st);
}

extern "C" int ly_stem_wgrad(const void* img, int img_is_u8, int N, int H, int W, int Cout, const void* dz,
                             long ld_dz, float in_scale, float* dwk, ly_stream_t st) {
  LY_REQUIRE(img && dz && dwk, "ly_stem_wgrad: null pointer");
  LY_REQUIRE(N > 0 && H % 4 == 0 && W % 4 == 0, "ly_stem_wgrad: H,W must be multiples of 4");
  if (img_is_u8)
    return stem_wgrad_dispatch<uint8_t>((const uint8_t*)img, N, H, W, Cout, (const __nv_bfloat16*)dz, ld_dz,
                                        in_scale, dwk, (cudaStream_t)st);
  return stem_wgrad_dispatch<float>((const float*)img, N, H, W, Cout, (const __nv_bfloat16*)dz, ld_dz, in_scale,
                                    dwk, (cudaStream_t)st);
}

extern "C" int ly_stem_im2col(const void* img, int img_is_u8, int N, int H, int W, void* P, ly_stream_t st) {
  LY_REQUIRE(img && P, "ly_stem_im2col: null pointer");
  LY_REQUIRE(N > 0 && H % 4 == 0 && W % 4 == 0 && H > 0 && W > 0, "ly_stem_im2col: H,W must be multiples of 4");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(img) % 16 == 0 && reinterpret_cast<uintptr_t>(P) % 16 == 0,
             "ly_stem_im2col: pointers must be 16-byte aligned");
  const long total = (long)N * (H / 4) * (W / 4) * 4;
  if (img_is_u8)
    LY_CUDA(launch(stem_im2col_kernel<uint8_t>, ceil_div(total, 256), 256, 0, (cudaStream_t)st, (const uint8_t*)img, N, H, W,
                   (__nv_bfloat16*)P));
  else
    LY_CUDA(launch(stem_im2col_kernel<float>, ceil_div(total, 256), 256, 0, (cudaStream_t)st, (const float*)img, N, H, W,
                   (__nv_bfloat16*)P));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_maxpool5_fwd(const void* x, long ldx, int N, int H, int W, int C, void* y, long ldy,
                               uint8_t* idx, ly_stream_t st) {
  int rc = chk(x, ldx, C, "ly_maxpool5_fwd(x)");
  if (rc) return rc;
  rc = chk(y, ldy, C, "ly_maxpool5_fwd(y)");
  if (rc) return rc;
  const int V = C / 8;
  const long total = (long)N * H * W * V;
  LY_CUDA(launch(maxpool5_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, (const __nv_bfloat16*)x, ldx, N, H, W, V,
                                                                      (__nv_bfloat16*)y, ldy, idx));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// y_k = k-fold 5x5 max pool of x for k = 1, 2, 3, written at y + (k-1)*ystep elements (three channel slices of one concat
// buffer with pixel stride ldy).  H*W <= 1600.
extern "C" int ly_sppf_pool3(const void* x, long ldx, int N, int H, int W, int C, void* y, long ldy, long ystep,
                             ly_stream_t st) {
  int rc = chk(x, ldx, C, "ly_sppf_pool3(x)");
  if (rc) return rc;
  rc = chk(y, ldy, C, "ly_sppf_pool3(y)");
  if (rc) return rc;
  LY_REQUIRE(ystep % 8 == 0 && ystep >= C && 3 * ystep <= ldy, "ly_sppf_pool3: ystep=%ld invalid", ystep);
  LY_REQUIRE((long)H * W <= 1600, "ly_sppf_pool3: H*W=%ld exceeds 1600", (long)H * W);
  const size_t smem = (size_t)4 * H * W * sizeof(uint4);
  static std::atomic<bool> attr_done[64];   // zero-initialised; set only after the attribute call succeeded
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(sppf_pool3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 1600 * 16));
    attr_done[dev & 63].store(true);
  }
  LY_CUDA(launch(sppf_pool3_kernel, (unsigned)(N * (C / 8)), 256, smem, (cudaStream_t)st, (const __nv_bfloat16*)x, ldx, H, W, C / 8,
                                                                             (__nv_bfloat16*)y, ldy, ystep));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_maxpool5_bwd(const void* dy, long lddy, const uint8_t* idx, int N, int H, int W, int C, void* dx,
                               long lddx, int accumulate, ly_stream_t st) {
  int rc = chk(dy, lddy, C, "ly_maxpool5_bwd(dy)");
  if (rc) return rc;
  rc = chk(dx, lddx, C, "ly_maxpool5_bwd(dx)");
  if (rc) return rc;
  LY_REQUIRE(idx != nullptr, "ly_maxpool5_bwd: null idx");
  const int V = C / 8;
  const long total = (long)N * H * W * V;
  LY_CUDA(launch(maxpool5_bwd_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, 
      (const __nv_bfloat16*)dy, lddy, idx, N, H, W, V, (__nv_bfloat16*)dx, lddx, accumulate));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_copy_upsample(const void* src, long lds, int N, int Hs, int Ws, int C, int up, void* dst,
                                long ldd, ly_stream_t st) {
  int rc = chk(src, lds, C, "ly_copy_upsample(src)");
  if (rc) return rc;
  rc = chk(dst, ldd, C, "ly_copy_upsample(dst)");
  if (rc) return rc;
  LY_REQUIRE(up == 1 || up == 2, "ly_copy_upsample: up must be 1 or 2");
  const int V = C / 8;
  const long total = (long)N * Hs * up * Ws * up * V;
  if (up == 2) {
    const long src_total = (long)N * Hs * Ws * V;
    LY_CUDA(launch(copy_up2_kernel, ceil_div(src_total, 512), 256, 0, (cudaStream_t)st, (const __nv_bfloat16*)src, lds, N, Hs, Ws, V,
                                                                           (__nv_bfloat16*)dst, ldd));
    LY_LAUNCH_CHECK();
    return LY_OK;
  }
  LY_CUDA(launch(copy_up_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, (const __nv_bfloat16*)src, lds, N, Hs, Ws, V,
                                                                     up, (__nv_bfloat16*)dst, ldd));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_downsum2(const void* src, long lds, int N, int H, int W, int C, void* dst, long ldd,
                           ly_stream_t st) {
  int rc = chk(src, lds, C, "ly_downsum2(src)");
  if (rc) return rc;
  rc = chk(dst, ldd, C, "ly_downsum2(dst)");
  if (rc) return rc;
  const int V = C / 8;
  const long total = (long)N * H * W * V;
  LY_CUDA(launch(downsum2_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, (const __nv_bfloat16*)src, lds, N, H, W, V,
                                                                      (__nv_bfloat16*)dst, ldd));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_gap(const void* x, long ldx, int N, int HW, int C, void* out, long ldo, ly_stream_t st) {
  int rc = chk(x, ldx, C, "ly_gap(x)");
  if (rc) return rc;
  rc = chk(out, ldo, C, "ly_gap(out)");
  if (rc) return rc;
  const int V = C / 8;
  dim3 grid(ceil_div(V, 32), N);
  LY_CUDA(launch(gap_kernel, grid, 512, 0, (cudaStream_t)st, (const __nv_bfloat16*)x, ldx, HW, V, (__nv_bfloat16*)out, ldo));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
