// RFCBAMConv receptive-field feature generation on the tensor cores (inference, k = 3).
//
// generate = depthwise 3x3 (stride s) that turns every channel into 9 features (reference models/rfa.py:101-105).
// Per 16-channel group this is a small dense problem: a grouped 3x3 convolution 16 -> 144 channels whose weight
// matrix is block diagonal.  Each of the 9 taps is one tcgen05.mma (M = 128 pixels, N = 144, K = 16) on a 16-channel
// SW32 box fetched by TMA (for stride 2 from one of four parity planes of x); 15/16 of the multiplied weights are
// zeros, which is still several times faster than the FP32 pipe, and the zeros never leave the SM: the 9 weight tiles
// live in shared memory and only their 1296 diagonal values are rewritten when the channel group changes.
//
// Work item = (pixel tile, channel group), channel group fastest; every CTA owns a contiguous range of items.
//   warp 0      TMA producer: ring of tap boxes
//   warp 1      MMA issuer: 9 taps -> one of two TMEM accumulators
//   warps 2..9  epilogue: TMEM lane quadrant = warp % 4; warps 2..5 take t = 0..4, warps 6..9 take t = 5..8
//   warp 10     rewrites the diagonal of the (double-buffered) weight tiles for the next item
// Two passes over x, neither stores the ungated features (the reference materialises them three times):
//   STATS  g = relu(acc + shift); max / mean over channels per (pixel, t) kept in registers across the channel groups
//          of a pixel tile and merged into the attention input maps once per tile (features are >= 0, so an integer
//          atomicMax on the float bits is exact)
//   APPLY  A[m][t*C + c] = g * se[n][c] * att[n][3h + t/3][3w + t%3] -> the GEMM operand of RFCBAMConv.conv
#include <string.h>

#include "ly_common.cuh"
#include "ly_tc.cuh"

namespace ly {

constexpr int kRtStages = 6;          // one stage = the 3 tap boxes of a kernel row, filled under ONE mbarrier phase:
                                      // a lone producer lane pays ~250 cycles per barrier round trip but only ~128 per
                                      // extra box (tests/dev/tma_issue.cu), and the TMA unit itself moves 1 box row / cycle
constexpr int kRtA = 128 * 32;        // tap box: 128 pixels x 16 channels, SW32
constexpr int kRtB = 144 * 32;        // weight tile of one tap: 144 outputs x 16 inputs, SW32
constexpr int kRtB1 = 9 * kRtB;       // all 9 taps of one channel group
constexpr int kRtThreads = 352;

struct alignas(64) RfTcParams {
  // s = 1: a_map[0] = x as (C, W, 1, H, N).  s = 2: one map per input parity (ph, pw) = x[:, ph::2, pw::2, :], so a tap
  // is a dense box of its parity plane (a strided TMA traversal would walk 4x the rows).  Box (16, Wt, 1, Ht, 1), SW32.
  CUtensorMap a_map[4];
  int tiles_w, tiles_h, n_img, groups;
  int Wt, Ht, ho, wo, C, s;
  int items, per_cta;
  uint32_t a_bytes;
  const __nv_bfloat16* wc;  // [C/16][9 taps][144]: the diagonal values, BatchNorm scale folded in
  const float* shift;       // [C/16][144]
  const float* se;          // [N][C]                 (APPLY)
  const float* att;         // [N][ho*3][wo*3]        (APPLY)
  __nv_bfloat16* A;         // [M][9*C]               (APPLY)
  float* maps;              // [N][ho*3][wo*3][2]     (STATS; zero on entry)
};

__device__ __forceinline__ uint64_t rt_desc_sw32(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) |
         (6ull << 61);
}

__device__ __forceinline__ void st_global_256(void* ptr, const uint4& a, const uint4& b) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "r"(a.x), "r"(a.y), "r"(a.z),
               "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w)
               : "memory");
}

template <int MODE>  // 0 = STATS, 1 = APPLY
__global__ void __launch_bounds__(kRtThreads, 1) rf_tc_kernel(const __grid_constant__ RfTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[kRtStages], empty_bar[kRtStages], tfull[2], tempty[2], b1_full[2], b1_empty[2];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;
  uint8_t* sB = smem + kRtStages * 3 * kRtA;   // 2 x kRtB1

  // zero the weight tiles once: the off-diagonal entries stay zero for the whole kernel
  for (int i = threadIdx.x; i < 2 * kRtB1 / 16; i += kRtThreads) reinterpret_cast<uint4*>(sB)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (warp == 0 && lane < (p.s == 2 ? 4 : 1)) tma_prefetch_desc(&p.a_map[lane]);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kRtStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull[b], 1); mbar_init(&tempty[b], 8);
      mbar_init(&b1_full[b], 1); mbar_init(&b1_empty[b], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_s, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = tmem_base_s;

  const int first = (int)blockIdx.x * p.per_cta;
  const int mine = min(p.per_cta, p.items - first);

  if (warp == 0) {
    uint32_t it = 0;
    for (int i = 0; i < mine; ++i) {
      const int item = first + i;
      const int cg = item % p.groups, mt = item / p.groups;
      const int tw = mt % p.tiles_w, th = (mt / p.tiles_w) % p.tiles_h, n = mt / (p.tiles_w * p.tiles_h);
      for (int ti = 0; ti < 3; ++ti, ++it) {
        const int s = (int)(it % kRtStages);
        const uint32_t ph = (it / kRtStages) & 1u;
        mbar_wait(&empty_bar[s], ph ^ 1u);
        if (lane == 0) {
          mbar_arrive_expect_tx(&full_bar[s], 3u * p.a_bytes);
#pragma unroll
          for (int tj = 0; tj < 3; ++tj) {
            uint8_t* dst = sA + (s * 3 + tj) * kRtA;
            if (p.s == 2)   // input row 2h + ti - 1: parity (ti + 1) & 1, plane row h - (ti == 0)
              tma_load_5d(dst, &p.a_map[((ti + 1) & 1) * 2 + ((tj + 1) & 1)], &full_bar[s], cg * 16,
                          tw * p.Wt - (tj == 0), 0, th * p.Ht - (ti == 0), n);
            else
              tma_load_5d(dst, &p.a_map[0], &full_bar[s], cg * 16, tw * p.Wt + tj - 1, 0, th * p.Ht + ti - 1, n);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc_bf16_f32(128, 144, 0, 0);
    uint32_t it = 0;
    for (int i = 0; i < mine; ++i) {
      const int buf = i & 1;
      const uint32_t ph = ((uint32_t)i >> 1) & 1u;
      mbar_wait(&tempty[buf], ph ^ 1u);
      mbar_wait(&b1_full[buf], ph);
      tc_fence_after_sync();
      for (int ti = 0; ti < 3; ++ti, ++it) {
        const int s = (int)(it % kRtStages);
        mbar_wait(&full_bar[s], (it / kRtStages) & 1u);
        tc_fence_after_sync();
        if (lane == 0) {
#pragma unroll
          for (int tj = 0; tj < 3; ++tj)
            umma_bf16(tmem + (uint32_t)(buf * 256), rt_desc_sw32(smem_u32(sA + (s * 3 + tj) * kRtA)),
                      rt_desc_sw32(smem_u32(sB + buf * kRtB1 + (ti * 3 + tj) * kRtB)), idesc, (uint32_t)((ti | tj) != 0));
          umma_commit(&empty_bar[s]);
          if (ti == 2) { umma_commit(&tfull[buf]); umma_commit(&b1_empty[buf]); }
        }
        __syncwarp();
      }
    }
  } else if (warp == 10) {
    for (int i = 0; i < mine; ++i) {
      const int buf = i & 1;
      const int cg = (first + i) % p.groups;
      mbar_wait(&b1_empty[buf], (((uint32_t)i >> 1) & 1u) ^ 1u);
      // 1296 values = 162 vectors of 8; a vector shares its tap k and covers outputs n0..n0+7 (n0 % 8 == 0)
      const uint4* src = reinterpret_cast<const uint4*>(p.wc + (long)cg * 1296);
      const uint32_t dst = smem_u32(sB + buf * kRtB1);
      uint4 v[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const int vi = lane + 32 * j;
        v[j] = vi < 162 ? __ldg(src + vi) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const int vi = lane + 32 * j;
        if (vi < 162) {
          const int k = vi / 18, n0 = (vi - k * 18) * 8;
          const uint32_t w[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int n = n0 + e;
            const uint32_t off = (uint32_t)(k * kRtB + n * 32 + (((((n & 15) >> 3) ^ (n >> 2)) & 1) << 4) + e * 2);
            const uint16_t val = (uint16_t)((e & 1) ? (w[e >> 1] >> 16) : (w[e >> 1] & 0xFFFFu));
            asm volatile("st.shared.b16 [%0], %1;" ::"r"(dst + off), "h"(val) : "memory");
          }
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&b1_full[buf]);
    }
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int t0 = half * 5, nt = half ? 4 : 5;
    const int r = q * 32 + lane;
    const int hh = r / p.Wt, ww = r - hh * p.Wt;
    const float invC = 1.f / (float)p.C;
    float va[5], vb[5];   // STATS: running max / sum;  APPLY: va = attention values
    int cur_mt = -1, n = 0, h = 0, w = 0;
    bool valid = false;
    for (int i = 0; i < mine; ++i) {
      const int item = first + i;
      const int cg = item % p.groups, mt = item / p.groups;
      const int buf = i & 1;
      if (mt != cur_mt) {
        cur_mt = mt;
        const int tw = mt % p.tiles_w, th = (mt / p.tiles_w) % p.tiles_h;
        n = mt / (p.tiles_w * p.tiles_h);
        h = th * p.Ht + hh; w = tw * p.Wt + ww;
        valid = (r < p.Wt * p.Ht) && h < p.ho && w < p.wo;
#pragma unroll
        for (int tt = 0; tt < 5; ++tt) {
          va[tt] = 0.f; vb[tt] = 0.f;
          if (MODE == 1 && valid && tt < nt) {
            const int t = t0 + tt;
            va[tt] = __ldg(p.att + ((long)n * (p.ho * 3) + (h * 3 + t / 3)) * (long)(p.wo * 3) + (w * 3 + t % 3));
          }
        }
      }
      float sev[16];
      if (MODE == 1) {
        const float4* sp = reinterpret_cast<const float4*>(p.se + (long)n * p.C + cg * 16);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 v = __ldg(sp + j);
          sev[4 * j] = v.x; sev[4 * j + 1] = v.y; sev[4 * j + 2] = v.z; sev[4 * j + 3] = v.w;
        }
      }
      const long m = ((long)n * p.ho + h) * p.wo + w;
      mbar_wait(&tfull[buf], ((uint32_t)i >> 1) & 1u);
      tc_fence_after_sync();
      const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * 256);
      const float4* sh4 = reinterpret_cast<const float4*>(p.shift + cg * 144);
#pragma unroll
      for (int tt = 0; tt < 5; ++tt) {
        if (tt < nt) {
          const int t = t0 + tt;
          uint32_t acc[16];
          tmem_ld16(trow + (uint32_t)(t * 16), acc);
          tmem_ld_wait();
          float g[16];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 sv = __ldg(sh4 + t * 4 + j);
            g[4 * j] = fmaxf(__uint_as_float(acc[4 * j]) + sv.x, 0.f);
            g[4 * j + 1] = fmaxf(__uint_as_float(acc[4 * j + 1]) + sv.y, 0.f);
            g[4 * j + 2] = fmaxf(__uint_as_float(acc[4 * j + 2]) + sv.z, 0.f);
            g[4 * j + 3] = fmaxf(__uint_as_float(acc[4 * j + 3]) + sv.w, 0.f);
          }
          if (MODE == 0) {
            float mx = va[tt], sm = 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) { mx = fmaxf(mx, g[j]); sm += g[j]; }
            va[tt] = mx; vb[tt] += sm;
          } else {
            float lo[8], hi[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { lo[j] = g[j] * sev[j] * va[tt]; hi[j] = g[8 + j] * sev[8 + j] * va[tt]; }
            if (valid) st_global_256(p.A + m * 9L * p.C + (long)t * p.C + cg * 16, pack8(lo), pack8(hi));
          }
        }
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[buf]);
      if (MODE == 0) {
        const bool last = (i == mine - 1) || ((item + 1) / p.groups != mt);
        if (last && valid) {
#pragma unroll
          for (int tt = 0; tt < 5; ++tt) {
            if (tt < nt) {
              const int t = t0 + tt;
              const long pos = ((long)n * (p.ho * 3) + (h * 3 + t / 3)) * (long)(p.wo * 3) + (w * 3 + t % 3);
              atomicMax(reinterpret_cast<int*>(p.maps + pos * 2), __float_as_int(va[tt]));
              atomicAdd(p.maps + pos * 2 + 1, vb[tt] * invC);
            }
          }
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

typedef CUresult (*rt_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int rt_encode(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides,
                     const uint32_t* box, const uint32_t* estr) {
  static rt_encode_fn fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return fail(LY_EUNSUPPORTED, "cuTensorMapEncodeTiled not available");
    fn = (rt_encode_fn)f;
  }
  cuuint64_t gd[5], gs[4];
  cuuint32_t bd[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bd[i] = box[i]; es[i] = estr[i]; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides[i];
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bd, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(LY_ECUDA, "cuTensorMapEncodeTiled(SW32) failed: %d", (int)r);
  return LY_OK;
}

static int rf_tc_launch(int mode, const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo,
                        const void* wc, const float* shift, const float* se, const float* att, void* A, float* maps,
                        cudaStream_t stream) {
  LY_REQUIRE(x && wc && shift, "ly_rf_tc: null pointer");
  LY_REQUIRE(C % 16 == 0 && C > 0 && (s == 1 || s == 2) && ldx % 8 == 0 && ldx >= C, "ly_rf_tc: bad C/s/ld");
  LY_REQUIRE(ho == (H + 2 - 3) / s + 1 && wo == (W + 2 - 3) / s + 1, "ly_rf_tc: output size mismatch");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(wc) % 16 == 0 &&
                 reinterpret_cast<uintptr_t>(shift) % 16 == 0,
             "ly_rf_tc: alignment");
  RfTcParams p;
  memset(&p, 0, sizeof(p));
  // spatial tile of <= 128 output pixels; the strided TMA box spans s*Wt x s*Ht input positions (<= 256 each)
  int bw = 16, bh = 8;
  double best = -1;
  const int cand[6] = {8, 16, 32, 64, 128, wo <= 128 ? wo : 128};
  for (int i = 0; i < 6; ++i) {
    const int wt = cand[i];
    if (wt <= 0 || wt > 128) continue;
    int ht = 128 / wt;
    if (ht > ho) ht = ho;
    const double eff = (double)ho * wo / ((double)ceil_div(wo, wt) * ceil_div(ho, ht) * 128.0);
    if (eff > best + 1e-9) { best = eff; bw = wt; bh = ht; }
  }
  p.Wt = bw; p.Ht = bh;
  p.tiles_w = ceil_div(wo, bw); p.tiles_h = ceil_div(ho, bh); p.n_img = N;
  p.groups = C / 16;
  p.ho = ho; p.wo = wo; p.C = C; p.s = s;
  p.a_bytes = (uint32_t)(16 * bw * bh * 2);
  p.wc = (const __nv_bfloat16*)wc; p.shift = shift; p.se = se; p.att = att; p.A = (__nv_bfloat16*)A; p.maps = maps;
  for (int par = 0; par < (s == 2 ? 4 : 1); ++par) {
    const int ph = par >> 1, pw = par & 1;
    const uint64_t Hp = (uint64_t)(H - ph + s - 1) / s, Wp = (uint64_t)(W - pw + s - 1) / s;
    if (Hp == 0 || Wp == 0) { memcpy(&p.a_map[par], &p.a_map[0], sizeof(CUtensorMap)); continue; }  // 1-pixel inputs
    uint64_t dims[5] = {(uint64_t)C, Wp, 1, Hp, (uint64_t)N};
    uint64_t str[4] = {(uint64_t)s * ldx * 2, (uint64_t)W * ldx * 2, (uint64_t)s * W * ldx * 2,
                       (uint64_t)H * W * ldx * 2};
    uint32_t box[5] = {16, (uint32_t)bw, 1, (uint32_t)bh, 1};
    uint32_t es[5] = {1, 1, 1, 1, 1};
    int rc = rt_encode(&p.a_map[par], (const __nv_bfloat16*)x + ((long)ph * W + pw) * ldx, 5, dims, str, box, es);
    if (rc) return rc;
  }
  const long items = (long)p.tiles_w * p.tiles_h * N * p.groups;
  LY_REQUIRE(items < (1L << 30), "ly_rf_tc: problem too large");
  const int sms = device_sm_count();
  p.items = (int)items;
  p.per_cta = (int)ceil_div(items, (long)sms);
  const unsigned grid = (unsigned)ceil_div(items, (long)p.per_cta);
  const size_t smem = (size_t)kRtStages * 3 * kRtA + 2 * kRtB1 + 1024;
  static bool attr_done[64] = {false};
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    LY_CUDA(cudaFuncSetAttribute(rf_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LY_CUDA(cudaFuncSetAttribute(rf_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done[dev & 63] = true;
  }
  if (mode == 0) {
    LY_CUDA(cudaMemsetAsync(maps, 0, sizeof(float) * 2 * (size_t)N * ho * 3 * wo * 3, stream));
    rf_tc_kernel<0><<<grid, kRtThreads, smem, stream>>>(p);
  } else {
    rf_tc_kernel<1><<<grid, kRtThreads, smem, stream>>>(p);
  }
  LY_LAUNCH_CHECK();
  return LY_OK;
}

}  // namespace ly

using namespace ly;

// x: NHWC bf16 (pixel stride ldx); wc: bf16 [C/16][9 taps][9 t][16] depthwise generate weights with the BatchNorm scale
// folded in (value for output (t, c) and tap ij at wc[c/16][ij][t][c%16]); shift: fp32 [C/16][9 t][16].
// Pass 1: maps fp32 [N][3*ho][3*wo][2] = (max, mean) over channels of the generated features (zeroed by the call).
extern "C" int ly_rf_tc_stats(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                              const float* shift, float* maps, ly_stream_t st) {
  LY_REQUIRE(maps, "ly_rf_tc_stats: null pointer");
  return rf_tc_launch(0, x, ldx, N, H, W, C, s, ho, wo, wc, shift, nullptr, nullptr, nullptr, maps, (cudaStream_t)st);
}

// Pass 2: A bf16 [N*ho*wo][9*C], A[m][t*C + c] = relu(generate + shift) * se[n][c] * att[n][3h + t/3][3w + t%3].
extern "C" int ly_rf_tc_apply(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                              const float* shift, const float* se, const float* att, void* A, ly_stream_t st) {
  LY_REQUIRE(se && att && A && reinterpret_cast<uintptr_t>(A) % 32 == 0 && reinterpret_cast<uintptr_t>(se) % 16 == 0,
             "ly_rf_tc_apply: null or misaligned pointer");
  return rf_tc_launch(1, x, ldx, N, H, W, C, s, ho, wo, wc, shift, se, att, A, nullptr, (cudaStream_t)st);
}
