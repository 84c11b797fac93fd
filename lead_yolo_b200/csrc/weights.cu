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
__device__ __forceinline__ void pack_fwd_elem(const float* __restrict__ w, int Cin, int kh, int kw, int patch,
                                              const SrcSplit& sp, int Kp, __nv_bfloat16* __restrict__ wp, long i) {
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

__global__ void pack_fwd_kernel(const float* __restrict__ w, int Cout, int Cin, int kh, int kw, int patch, SrcSplit sp,
                                int Kp, __nv_bfloat16* __restrict__ wp) {
  pdl_enter();
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)Cout * Kp) return;
  pack_fwd_elem(w, Cin, kh, kw, patch, sp, Kp, wp, i);
}

// input-gradient weights of source channels [coff, coff+cs):
//   stride-1 conv: rows ci (cs of them), K = taps x pad64(cpad): wd[ci][t*P + co] = w[co][coff+ci][kh-1-ty][kw-1-tx]
//   patch conv   : rows (ky, kx, c) (kh*kw*cs), K = pad64(cpad):  wd[(ky*kw+kx)*cs + c][co] = w[co][coff+c][ky][kx]
__device__ __forceinline__ void pack_dgrad_elem(const float* __restrict__ w, int Cout, int Cin, int kh, int kw, int patch,
                                                int coff, int cs, int P, __nv_bfloat16* __restrict__ wd, long i) {
  const int Kp = patch ? P : kh * kw * P;
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

__global__ void pack_dgrad_kernel(const float* __restrict__ w, int Cout, int Cin, int kh, int kw, int patch, int coff,
                                  int cs, int P, __nv_bfloat16* __restrict__ wd) {
  pdl_enter();
  const int rows = patch ? kh * kw * cs : cs;
  const int Kp = patch ? P : kh * kw * P;
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)rows * Kp) return;
  pack_dgrad_elem(w, Cout, Cin, kh, kw, patch, coff, cs, P, wd, i);
}

// Every parameter of a training step in every packed layout it is needed in, ONE launch: jobs[j] describes one
// (weight, layout) pair and owns the blocks [first_block, first_block of the next job).  A captured training step packs
// ~150 weights; as separate launches they cost ~3.5 us each on the replay's critical path.
__global__ void __launch_bounds__(256) pack_batch_kernel(const ly_pack_job* __restrict__ jobs, int njobs) {
  pdl_enter();
  int lo = 0, hi = njobs - 1;
  const long b = blockIdx.x;
  while (lo < hi) {   // last job whose first block is <= b
    const int mid = (lo + hi + 1) >> 1;
    if (jobs[mid].first_block <= b) lo = mid;
    else hi = mid - 1;
  }
  const ly_pack_job& j = jobs[lo];
  const long i = (b - j.first_block) * 256 + threadIdx.x;
  if (i >= j.total) return;
  if (j.kind == 0) {
    SrcSplit sp;
    sp.n = 0;
    for (int s = 0; s < 4; ++s) {
      sp.c[s] = j.c[s];
      if (j.c[s] > 0) sp.n = s + 1;
    }
    pack_fwd_elem((const float*)j.w, j.Cin, j.kh, j.kw, j.patch, sp, j.Kp, (__nv_bfloat16*)j.dst, i);
  } else {
    pack_dgrad_elem((const float*)j.w, j.Cout, j.Cin, j.kh, j.kw, j.patch, j.coff, j.cs, j.P, (__nv_bfloat16*)j.dst, i);
  }
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

extern "C" int ly_pack_batch(const ly_pack_job* jobs, int njobs, long total_blocks, ly_stream_t st) {
  LY_REQUIRE(jobs != nullptr && njobs > 0 && total_blocks > 0 && total_blocks < (1L << 31), "ly_pack_batch: bad arguments");
  LY_CUDA(launch(pack_batch_kernel, dim3((unsigned)total_blocks), 256, 0, (cudaStream_t)st, jobs, njobs));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
