// CoordAtt (reference models/common.py:1583-1609) — the HBM-bound parts: directional average pooling and the
// gate application, forward and backward.  The tiny 1x1 convolutions / BatchNorm between them run through the
// generic conv / BN kernels on the pooled matrix.
//
// Pooled / gate matrix layout ("HW rows"): [N*H + N*W][C] bf16;  row n*H + h  <- mean over w of x[n,h,:,c]
//                                                                row N*H + n*W + w <- mean over h of x[n,:,w,c]
// so that conv_h works on the first N*H rows and conv_w on the last N*W rows, while conv1 + bn1 see all
// N*(H+W) positions jointly as in the reference (torch.cat along the spatial axis, common.py:1600-1603).
#include "ly_common.cuh"

namespace ly {

__device__ __forceinline__ uint4 ldq(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// Blocks [0, gw) : one thread per (n, w, v), loop over h  (perfectly coalesced column sums)
// Blocks [gw, ..): one thread per (n, h, v), loop over w  (row sums)
// mode 0: pooling      out = mean(x)
// mode 1: gate grads   out_w = Σ_h dy*x*a_h ; out_h = Σ_w dy*x*a_w
__global__ void __launch_bounds__(256) ca_reduce_kernel(const __nv_bfloat16* __restrict__ x, long ldx,
                                                        const __nv_bfloat16* __restrict__ dy, long lddy,
                                                        const __nv_bfloat16* __restrict__ gate, long ldg, int N,
                                                        int H, int W, int V, int gw, int mode,
                                                        __nv_bfloat16* __restrict__ out, long ldo) {
  pdl_enter();
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if ((int)blockIdx.x < gw) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long)N * W * V) return;
    const int v = (int)(t % V);
    const int w = (int)((t / V) % W);
    const int n = (int)(t / ((long)V * W));
    const long p0 = ((long)n * H) * W + w;
    for (int h = 0; h < H; ++h) {
      float f[8];
      unpack8(ldq(x + (p0 + (long)h * W) * ldx + v * 8), f);
      if (mode == 1) {
        float g[8], a[8];
        unpack8(ldq(dy + (p0 + (long)h * W) * lddy + v * 8), g);
        unpack8(ldq(gate + ((long)n * H + h) * ldg + v * 8), a);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(f[j] * g[j], a[j], acc[j]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    if (mode == 0) {
      const float inv = 1.f / (float)H;
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] *= inv;
    }
    *reinterpret_cast<uint4*>(out + ((long)N * H + (long)n * W + w) * ldo + v * 8) = pack8(acc);
  } else {
    const long t = (long)(blockIdx.x - gw) * blockDim.x + threadIdx.x;
    if (t >= (long)N * H * V) return;
    const int v = (int)(t % V);
    const int h = (int)((t / V) % H);
    const int n = (int)(t / ((long)V * H));
    const long p0 = ((long)n * H + h) * W;
    for (int w = 0; w < W; ++w) {
      float f[8];
      unpack8(ldq(x + (p0 + w) * ldx + v * 8), f);
      if (mode == 1) {
        float g[8], a[8];
        unpack8(ldq(dy + (p0 + w) * lddy + v * 8), g);
        unpack8(ldq(gate + ((long)N * H + (long)n * W + w) * ldg + v * 8), a);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(f[j] * g[j], a[j], acc[j]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    if (mode == 0) {
      const float inv = 1.f / (float)W;
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] *= inv;
    }
    *reinterpret_cast<uint4*>(out + ((long)n * H + h) * ldo + v * 8) = pack8(acc);
  }
}

// Pooling only (mode 0 of ca_reduce_kernel) with eight independent 16-byte loads in flight per thread: the reduction
// loops are short dependent chains otherwise and the kernel runs at a third of the HBM rate.
__global__ void __launch_bounds__(256) ca_pool_kernel(const __nv_bfloat16* __restrict__ x, long ldx, int N, int H,
                                                      int W, int V, int gw, __nv_bfloat16* __restrict__ out, long ldo) {
  pdl_enter();
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  const bool cols = (int)blockIdx.x < gw;
  const long t = (long)(cols ? blockIdx.x : blockIdx.x - gw) * blockDim.x + threadIdx.x;
  const int L = cols ? W : H;        // positions along the kept axis
  const int R = cols ? H : W;        // reduced axis
  if (t >= (long)N * L * V) return;
  const int v = (int)(t % V);
  const int l = (int)((t / V) % L);
  const int n = (int)(t / ((long)V * L));
  const long p0 = cols ? ((long)n * H) * W + l : ((long)n * H + l) * W;
  const long step = cols ? W : 1;
  const __nv_bfloat16* src = x + p0 * ldx + v * 8;
  const long sstep = step * ldx;
  int i = 0;
  for (; i + 8 <= R; i += 8) {
    uint4 u[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) u[k] = ldq(src + (long)(i + k) * sstep);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float f[8];
      unpack8(u[k], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  }
  for (; i < R; ++i) {
    float f[8];
    unpack8(ldq(src + (long)i * sstep), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += f[j];
  }
  const float inv = 1.f / (float)R;
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] *= inv;
  const long orow = cols ? (long)N * H + (long)n * W + l : (long)n * H + l;
  *reinterpret_cast<uint4*>(out + orow * ldo + v * 8) = pack8(acc);
}

// mode 0: out = x * a_h[n,h] * a_w[n,w]                         (forward, and dx1 = dy*a_h*a_w)
// mode 1: out = g_h[n,h]/W + g_w[n,w]/H                          (backward of the pooling)
// IT: index type.  The (pixel, vector) decomposition is four divisions per thread; in 64 bits they were most of this
// kernel's instructions (it ran instruction-bound at a third of the HBM rate), so the launcher picks 32-bit indices
// whenever the element count allows.
template <typename IT>
__global__ void __launch_bounds__(256) ca_pointwise_kernel(const __nv_bfloat16* __restrict__ x, long ldx,
                                                           const __nv_bfloat16* __restrict__ gate, long ldg, int N,
                                                           int H, int W, int V, int mode,
                                                           __nv_bfloat16* __restrict__ out, long ldo) {
  pdl_enter();
  const IT total = (IT)N * (IT)H * (IT)W * (IT)V;
  const IT t = (IT)blockIdx.x * (IT)blockDim.x + (IT)threadIdx.x;
  if (t >= total) return;
  const int v = (int)(t % (IT)V);
  const IT pixi = t / (IT)V;
  const int w = (int)(pixi % (IT)W);
  const IT rowi = pixi / (IT)W;
  const int h = (int)(rowi % (IT)H), n = (int)(rowi / (IT)H);
  const long pix = (long)pixi;
  float ah[8], aw[8], r[8];
  unpack8(ldq(gate + ((long)n * H + h) * ldg + v * 8), ah);
  unpack8(ldq(gate + ((long)N * H + (long)n * W + w) * ldg + v * 8), aw);
  if (mode == 0) {
    float f[8];
    unpack8(ldq(x + pix * ldx + v * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = f[j] * ah[j] * aw[j];
  } else {
    const float iw = 1.f / (float)W, ih = 1.f / (float)H;
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = ah[j] * iw + aw[j] * ih;
  }
  *reinterpret_cast<uint4*>(out + pix * ldo + v * 8) = pack8(r);
}

static int chk(const void* p, long ld, int C, const char* what) {
  LY_REQUIRE(p != nullptr, "%s: null pointer", what);
  LY_REQUIRE(C > 0 && C % 8 == 0 && ld >= C && ld % 8 == 0, "%s: C=%d ld=%ld invalid", what, C, ld);
  LY_REQUIRE(reinterpret_cast<uintptr_t>(p) % 16 == 0, "%s: pointer not 16-byte aligned", what);
  return LY_OK;
}

static int launch_reduce(const void* x, long ldx, const void* dy, long lddy, const void* gate, long ldg, int N,
                         int H, int W, int C, int mode, void* out, long ldo, cudaStream_t st) {
  const int V = C / 8;
  const int gw = ceil_div((long)N * W * V, 256);
  const int gh = ceil_div((long)N * H * V, 256);
  if (mode == 0) {
    LY_CUDA(launch(ca_pool_kernel, gw + gh, 256, 0, st, (const __nv_bfloat16*)x, ldx, N, H, W, V, gw, (__nv_bfloat16*)out, ldo));
    LY_LAUNCH_CHECK();
    return LY_OK;
  }
  LY_CUDA(launch(ca_reduce_kernel, gw + gh, 256, 0, st, (const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)dy, lddy,
                                            (const __nv_bfloat16*)gate, ldg, N, H, W, V, gw, mode,
                                            (__nv_bfloat16*)out, ldo));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

}  // namespace ly

using namespace ly;

extern "C" int ly_ca_pool(const void* x, long ldx, int N, int H, int W, int C, void* pooled, long ldp,
                          ly_stream_t st) {
  int rc = chk(x, ldx, C, "ly_ca_pool(x)");
  if (rc) return rc;
  rc = chk(pooled, ldp, C, "ly_ca_pool(pooled)");
  if (rc) return rc;
  return launch_reduce(x, ldx, nullptr, 0, nullptr, 0, N, H, W, C, 0, pooled, ldp, (cudaStream_t)st);
}

extern "C" int ly_ca_apply(const void* x, long ldx, const void* gates, long ldg, int N, int H, int W, int C,
                           void* out, long ldo, ly_stream_t st) {
  int rc = chk(x, ldx, C, "ly_ca_apply(x)");
  if (rc) return rc;
  rc = chk(gates, ldg, C, "ly_ca_apply(gates)");
  if (rc) return rc;
  rc = chk(out, ldo, C, "ly_ca_apply(out)");
  if (rc) return rc;
  const int V = C / 8;
  const long total = (long)N * H * W * V;
  if (total < (1L << 31))
    LY_CUDA(launch(ca_pointwise_kernel<unsigned>, ceil_div(total, 256), 256, 0, (cudaStream_t)st, 
        (const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)gates, ldg, N, H, W, V, 0, (__nv_bfloat16*)out, ldo));
  else
    LY_CUDA(launch(ca_pointwise_kernel<long>, ceil_div(total, 256), 256, 0, (cudaStream_t)st, 
        (const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)gates, ldg, N, H, W, V, 0, (__nv_bfloat16*)out, ldo));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_ca_gate_grad(const void* x, long ldx, const void* dy, long lddy, const void* gates, long ldg,
                               int N, int H, int W, int C, void* dgates, long ldd, ly_stream_t st) {
  int rc = chk(x, ldx, C, "ly_ca_gate_grad(x)");
  if (rc) return rc;
  rc = chk(dy, lddy, C, "ly_ca_gate_grad(dy)");
  if (rc) return rc;
  rc = chk(gates, ldg, C, "ly_ca_gate_grad(gates)");
  if (rc) return rc;
  rc = chk(dgates, ldd, C, "ly_ca_gate_grad(dgates)");
  if (rc) return rc;
  return launch_reduce(x, ldx, dy, lddy, gates, ldg, N, H, W, C, 1, dgates, ldd, (cudaStream_t)st);
}

extern "C" int ly_ca_pool_bwd(const void* dpooled, long ldp, int N, int H, int W, int C, void* dx, long ldx,
                              ly_stream_t st) {
  int rc = chk(dpooled, ldp, C, "ly_ca_pool_bwd(dpooled)");
  if (rc) return rc;
  rc = chk(dx, ldx, C, "ly_ca_pool_bwd(dx)");
  if (rc) return rc;
  const int V = C / 8;
  const long total = (long)N * H * W * V;
  if (total < (1L << 31))
    LY_CUDA(launch(ca_pointwise_kernel<unsigned>, ceil_div(total, 256), 256, 0, (cudaStream_t)st, 
        nullptr, 0, (const __nv_bfloat16*)dpooled, ldp, N, H, W, V, 1, (__nv_bfloat16*)dx, ldx));
  else
    LY_CUDA(launch(ca_pointwise_kernel<long>, ceil_div(total, 256), 256, 0, (cudaStream_t)st, 
        nullptr, 0, (const __nv_bfloat16*)dpooled, ldp, N, H, W, V, 1, (__nv_bfloat16*)dx, ldx));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
