// Weight layout kernels of the training path: fp32 OIHW parameters (the layout nn.Conv2d owns and the optimizer,
// EMA and DDP touch — SURVEY §8b / §7 H4) <-> the packed K-major bf16 / fp32 matrices the implicit-GEMM kernels read
// and write.  One launch per weight and direction instead of a chain of permute / pad / cat / cast operators.
//
// Packed K order (conv_gemm.cu, lead_yolo_b200/packing.py): for source s, for tap t (row-major; for a k = stride
// "patch" convolution a tap is a kernel row ky whose "channels" are the kw*Cs values (kx, c)), 64-wide chunks,
// zero padded to a multiple of 64.
#include "ly_common.cuh"

namespace ly {

struct SrcSplit {
  int n;
  int c[4];
};

__host__ __device__ inline int pad64(int v) { return (v + 63) / 64 * 64; }

// packed column of (input channel ci, ky, kx); returns -1 when out of range
__device__ __forceinline__ int packed_col(const SrcSplit& sp, int kh, int kw, int patch, int ci, int ky, int kx) {
  int col = 0, off = 0;
  for (int s = 0; s < sp.n; ++s) {
    const int cs = sp.c[s];
    if (ci < off + cs) {
      const int c = ci - off;
      if (patch) return col + ky * pad64(kw * cs) + kx * cs + c;
      return col + (ky * kw + kx) * pad64(cs) + c;
    }
    col += patch ? kh * pad64(kw * cs) : kh * kw * pad64(cs);
    off += cs;
  }
  return -1;
}

// forward pack: wp[o][Kp] bf16 <- w[o][ci][ky][kx]; one thread per packed element (zero padding included)
__global__ void pack_fwd_kernel(const float* __restrict__ w, int Cout, int Cin, int kh, int kw, int patch, SrcSplit sp,
                                int Kp, __nv_bfloat16* __restrict__ wp) {
  pdl_enter();
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)Cout * Kp) return;
  const int o = (int)(i / Kp);
  int col = (int)(i % Kp);
  float val = 0.f;
  int off = 0;
  for (int s = 0; s < sp.n; ++s) {
    const int cs = sp.c[s];
    const int segw = patch ? pad64(kw * cs) : pad64(cs);
    const int nseg = patch ? kh : kh * kw;
    if (col < nseg * segw) {
      const int t = col / segw, r = col % segw;
      if (patch) {
        if (r < kw * cs) val = w[(((long)o * Cin + off + r % cs) * kh + t) * kw + r / cs];
      } else if (r < cs) {
        val = w[(((long)o * Cin + off + r) * kh + t / kw) * kw + t % kw];
      }
      break;
    }
    col -= nseg * segw;
    off += cs;
  }
  wp[i] = __float2bfloat16(val);
}

// input-gradient weights of source channels [coff, coff+cs):
//   stride-1 conv: rows ci (cs of them), K = taps x pad64(cpad): wd[ci][t*P + co] = w[co][coff+ci][kh-1-ty][kw-1-tx]
//   patch conv   : rows (ky, kx, c) (kh*kw*cs), K = pad64(cpad):  wd[(ky*kw+kx)*cs + c][co] = w[co][coff+c][ky][kx]
__global__ void pack_dgrad_kernel(const float* __restrict__ w, int Cout, int Cin, int kh, int kw, int patch, int coff,
                                  int cs, int P, __nv_bfloat16* __restrict__ wd) {
  pdl_enter();
  const int rows = patch ? kh * kw * cs : cs;
  const int Kp = patch ? P : kh * kw * P;
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)rows * Kp) return;
  const int r = (int)(i / Kp), col = (int)(i % Kp);
  float val = 0.f;
  if (patch) {
    const int co = col;
    if (co < Cout) {
      const int t = r / cs, c = r % cs;
      val = w[(((long)co * Cin + coff + c) * kh + t / kw) * kw + t % kw];
    }
  } else {
    const int t = col / P, co = col % P;
    if (co < Cout) {
      const int ty = kh - 1 - t / kw, tx = kw - 1 - t % kw;
      val = w[(((long)co * Cin + coff + r) * kh + ty) * kw + tx];
    }
  }
  wd[i] = __float2bfloat16(val);
}

// weight gradient back to OIHW: dw[o][ci][ky][kx] (+)= dwp[o][packed_col]
__global__ void unpack_wgrad_kernel(const float* __restrict__ dwp, int Cout, int Cin, int kh, int kw, int patch,
                                    SrcSplit sp, int Kp, float* __restrict__ dw, int accumulate) {
  pdl_enter();
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)Cout * Cin * kh * kw) return;
  const int kx = (int)(i % kw);
  const int ky = (int)((i / kw) % kh);
  const int ci = (int)((i / ((long)kw * kh)) % Cin);
  const int o = (int)(i / ((long)kw * kh * Cin));
  const int col = packed_col(sp, kh, kw, patch, ci, ky, kx);
  const float g = dwp[(long)o * Kp + col];
  dw[i] = accumulate ? dw[i] + g : g;
}

static int split_of(int c0, int c1, int c2, int c3, SrcSplit* sp, int* cin) {
  const int c[4] = {c0, c1, c2, c3};
  sp->n = 0;
  *cin = 0;
  for (int s = 0; s < 4; ++s) {
    if (c[s] <= 0) break;
    sp->c[sp->n++] = c[s];
    *cin += c[s];
  }
  for (int s = sp->n; s < 4; ++s) {
    if (c[s] > 0) return -1;
    sp->c[s] = 0;
  }
  return sp->n > 0 ? 0 : -1;
}

static int packed_k_of(const SrcSplit& sp, int kh, int kw, int patch) {
  int k = 0;
  for (int s = 0; s < sp.n; ++s) k += patch ? kh * pad64(kw * sp.c[s]) : kh * kw * pad64(sp.c[s]);
  return k;
}

}  // namespace ly

using namespace ly;

extern "C" int ly_pack_weight(const float* w, int Cout, int kh, int kw, int patch, int c0, int c1, int c2, int c3,
                              void* wp, ly_stream_t st) {
  SrcSplit sp;
  int cin;
  LY_REQUIRE(w && wp && Cout > 0 && kh > 0 && kw > 0, "ly_pack_weight: bad arguments");
  LY_REQUIRE(split_of(c0, c1, c2, c3, &sp, &cin) == 0, "ly_pack_weight: bad source split (%d,%d,%d,%d)", c0, c1, c2, c3);
  const int Kp = packed_k_of(sp, kh, kw, patch);
  const long total = (long)Cout * Kp;
  LY_CUDA(launch(pack_fwd_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, w, Cout, cin, kh, kw, patch, sp, Kp,
                                                                       (__nv_bfloat16*)wp));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_pack_weight_dgrad(const float* w, int Cout, int Cin, int kh, int kw, int patch, int coff, int cs,
                                    void* wd, ly_stream_t st) {
  LY_REQUIRE(w && wd && Cout > 0 && Cin > 0 && kh > 0 && kw > 0 && cs > 0 && coff >= 0 && coff + cs <= Cin,
             "ly_pack_weight_dgrad: bad arguments");
  const int P = pad64((Cout + 7) / 8 * 8);
  const long total = patch ? (long)kh * kw * cs * P : (long)cs * kh * kw * P;
  LY_CUDA(launch(pack_dgrad_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, w, Cout, Cin, kh, kw, patch, coff, cs, P,
                                                                         (__nv_bfloat16*)wd));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_unpack_wgrad(const float* dwp, int Cout, int kh, int kw, int patch, int c0, int c1, int c2, int c3,
                               float* dw, int accumulate, ly_stream_t st) {
  SrcSplit sp;
  int cin;
  LY_REQUIRE(dwp && dw && Cout > 0 && kh > 0 && kw > 0, "ly_unpack_wgrad: bad arguments");
  LY_REQUIRE(split_of(c0, c1, c2, c3, &sp, &cin) == 0, "ly_unpack_wgrad: bad source split (%d,%d,%d,%d)", c0, c1, c2,
             c3);
  const int Kp = packed_k_of(sp, kh, kw, patch);
  const long total = (long)Cout * cin * kh * kw;
  LY_CUDA(launch(unpack_wgrad_kernel, ceil_div(total, 256), 256, 0, (cudaStream_t)st, dwp, Cout, cin, kh, kw, patch, sp, Kp, dw,
                                                                           accumulate));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
