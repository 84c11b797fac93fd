// Stem (PatchEmbed_FasterNet.proj, models/common.py:1531,1537) on the tensor cores for uint8 images.
//
// The 4x4 stride-4 convolution 3 -> Cout is a GEMM with K = 48; on the FP32 pipe it costs 48*Cout FMAs per output pixel
// and was FMA-bound at ~5x the HBM time.  Here TMA streams raw uint8 image boxes (128 bytes x 32 rows x 3 channels = the
// input of 32 x 8 output pixels) into a shared-memory ring; converter threads turn them into the 128B-swizzled K-major
// fp16 A operand [pixel][(c*4+ky)*4+kx] (uint8 is exact in fp16, and 0x6400|b - 1024 converts on the integer/half
// pipes; K padded to 64); tcgen05.mma (kind::f16, fp16 x fp16 -> fp32, M=128, N=Cout, K=16 x 3) accumulates in TMEM; the
// epilogue applies in_scale and the shift and writes NHWC bf16.
// out = scale*conv(img*in_scale) + shift with scale folded into the fp16 weights.
//
//   persistent CTAs, 2 per SM; one "stage" = 256 output pixels (32 wide x 8 high) = two 128-row MMA tiles
//   warp 0       MMA issuer: 6 MMAs + ONE commit per stage (the commit frees the A slot and publishes both accumulators)
//   warps 1..8   converters (one output pixel per thread and stage): raw ring -> A slot
//   warps 9..16  epilogue (TMEM lane quadrant = warp % 4)
//   warp 17      TMA producer (one lane)
#include <string.h>

#include <cuda_fp16.h>

#include <atomic>

#include "ly_common.cuh"
#include "ly_tc.cuh"

namespace ly {

constexpr int kStThreads = 576;
constexpr int kStTile = 128 * 128;       // one A tile: 128 pixels x 64 K (bf16), SW128
constexpr int kStRaw = 3 * 32 * 128;     // one raw box: 3 channels x 32 image rows x 128 bytes
constexpr int kStRawSlots = 3;

struct alignas(64) StemTcParams {
  CUtensorMap img_map;  // uint8 (W, H, 3, N), box (128, 32, 3, 1)
  const float* wk;      // [48][Cout]
  const float* scale;   // [Cout] or null
  const float* shift;   // [Cout] or null
  __nv_bfloat16* out;
  long ld_out;
  float in_scale;
  int Ho, Wo, Cout, Npad;
  int tiles_w, tiles_h, stages;
};

__device__ __forceinline__ void tma_load_4d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

__global__ void __launch_bounds__(kStThreads, 2) stem_tc_kernel(const __grid_constant__ StemTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t raw_full[kStRawSlots], raw_empty[kStRawSlots], full_bar[2], done_bar[2], tempty_bar[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ float s_shift[64];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;                         // 2 slots x 2 tiles
  uint8_t* sB = smem + 4 * kStTile;           // [64][64] bf16, SW128
  uint8_t* sR = sB + 64 * 128;                // raw ring

  // zero A once (the K padding chunks 6,7 are never written again) and build the weight tile
  for (int i = threadIdx.x; i < 4 * kStTile / 16; i += kStThreads) reinterpret_cast<uint4*>(sA)[i] = make_uint4(0, 0, 0, 0);
  pdl_wait();
  for (int i = threadIdx.x; i < p.Npad * 64; i += kStThreads) {
    const int n = i >> 6, k = i & 63;
    float w = 0.f;
    if (n < p.Cout && k < 48) w = p.wk[k * p.Cout + n] * (p.scale != nullptr ? p.scale[n] : 1.f);
    *reinterpret_cast<__half*>(sB + n * 128 + ((((k >> 3) ^ (n & 7))) << 4) + (k & 7) * 2) = __float2half_rn(w);
  }
  if (threadIdx.x < 64) s_shift[threadIdx.x] = (p.shift != nullptr && (int)threadIdx.x < p.Cout) ? p.shift[threadIdx.x] : 0.f;
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(&full_bar[s], 256); mbar_init(&done_bar[s], 1); mbar_init(&tempty_bar[s], 8); }
    for (int s = 0; s < kStRawSlots; ++s) { mbar_init(&raw_full[s], 1); mbar_init(&raw_empty[s], 8); }
    fence_mbar_init();
  }
  if (warp == 17 && lane == 0) tma_prefetch_desc(&p.img_map);
  if (warp == 1) tmem_alloc(&tmem_base_s, 256);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_trigger();
  const uint32_t tmem = tmem_base_s;

  const int mine = ((int)blockIdx.x < p.stages) ? (p.stages - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int tiles_img = p.tiles_w * p.tiles_h;

  if (warp == 17) {
    if (lane == 0) {
      int rs = 0;
      uint32_t rph = 1;
      for (int i = 0; i < mine; ++i) {
        const int g = (int)blockIdx.x + i * (int)gridDim.x;
        const int n = g / tiles_img, tr = g - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
        mbar_wait(&raw_empty[rs], rph);
        mbar_arrive_expect_tx(&raw_full[rs], (uint32_t)kStRaw);
        tma_load_4d(sR + rs * kStRaw, &p.img_map, &raw_full[rs], tw * 128, th * 32, 0, n);
        if (++rs == kStRawSlots) { rs = 0; rph ^= 1u; }
      }
    }
  } else if (warp == 0) {
    if (lane == 0) {
      // kind::f16 with fp16 operands (format fields 0), fp32 accumulation, K-major A and B
      const uint32_t idesc = (1u << 4) | ((uint32_t)(p.Npad >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint64_t a0 = umma_smem_desc_sw128(smem_u32(sA), 16, 1024);
      const uint64_t b0 = umma_smem_desc_sw128(smem_u32(sB), 16, 1024);
      for (int i = 0; i < mine; ++i) {
        const int s = i & 1;
        const uint32_t ph = ((uint32_t)i >> 1) & 1u;
        mbar_wait(&tempty_bar[s], ph ^ 1u);
        mbar_wait(&full_bar[s], ph);
        tc_fence_after_sync();
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const uint64_t a = a0 + (uint64_t)((s * 2 + j) * (kStTile >> 4));
          const uint32_t d = tmem + (uint32_t)(s * 128 + j * 64);
          umma_bf16(d, a, b0, idesc, 0u);
          umma_bf16(d, a + 2, b0 + 2, idesc, 1u);
          umma_bf16(d, a + 4, b0 + 4, idesc, 1u);
        }
        umma_commit(&done_bar[s]);
      }
    }
  } else if (warp <= 8) {
    // ---------------- converters: raw uint8 box -> swizzled bf16 A rows ----------------
    const int t = threadIdx.x - 32;             // 0..255
    const int j = t >> 7, r = t & 127;
    const int ho_l = j * 4 + (r >> 5), wo_l = r & 31;
    int rs = 0;
    uint32_t rph = 0;
    for (int i = 0; i < mine; ++i) {
      const int s = i & 1;
      mbar_wait(&raw_full[rs], rph);
      const uint8_t* src = sR + rs * kStRaw + (ho_l * 4) * 128 + wo_l * 4;
      uchar4 v[12];
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int ky = 0; ky < 4; ++ky) v[c * 4 + ky] = *reinterpret_cast<const uchar4*>(src + (c * 32 + ky) * 128);
      __syncwarp();
      if (lane == 0) mbar_arrive(&raw_empty[rs]);          // the box is in registers: the slot can be refilled
      if (++rs == kStRawSlots) { rs = 0; rph ^= 1u; }
      mbar_wait(&done_bar[s], (((uint32_t)i >> 1) & 1u) ^ 1u);   // the MMAs that read this A slot two stages ago are done
      uint8_t* row = sA + (s * 2 + j) * kStTile + r * 128;
#pragma unroll
      for (int q = 0; q < 12; ++q) {
        // uint8 -> fp16 without the conversion pipe: 0x6400 | b is the half 1024 + b, minus 1024 (packed) gives b exactly
        const uint32_t u = *reinterpret_cast<const uint32_t*>(&v[q]);
        const __half2 k1024 = __half2half2(__ushort_as_half((unsigned short)0x6400));
        uint32_t lo = __byte_perm(u, 0x64646464u, 0x4140), hi = __byte_perm(u, 0x64646464u, 0x4342);
        const __half2 hlo = __hsub2(*reinterpret_cast<const __half2*>(&lo), k1024);
        const __half2 hhi = __hsub2(*reinterpret_cast<const __half2*>(&hi), k1024);
        uint2 w;
        w.x = *reinterpret_cast<const uint32_t*>(&hlo);
        w.y = *reinterpret_cast<const uint32_t*>(&hhi);
        *reinterpret_cast<uint2*>(row + ((((q >> 1) ^ (r & 7))) << 4) + (q & 1) * 8) = w;
      }
      fence_proxy_async_smem();
      mbar_arrive(&full_bar[s]);
    }
  } else {
    // ---------------- epilogue ----------------
    const int q = warp & 3;
    const int j = (warp - 9) >> 2;
    const int r = q * 32 + lane;
    const int ho_l = j * 4 + (r >> 5), wo_l = r & 31;
    for (int i = 0; i < mine; ++i) {
      const int g = (int)blockIdx.x + i * (int)gridDim.x;
      const int n = g / tiles_img, tr = g - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
      const int ho = th * 8 + ho_l, wo = tw * 32 + wo_l;
      const bool ok = ho < p.Ho && wo < p.Wo;
      const int s = i & 1;
      mbar_wait(&done_bar[s], ((uint32_t)i >> 1) & 1u);
      tc_fence_after_sync();
      const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(s * 128 + j * 64);
      __nv_bfloat16* o = p.out + (((long)n * p.Ho + ho) * p.Wo + wo) * p.ld_out;
#pragma unroll
      for (int cp = 0; cp < 2; ++cp) {            // 32 accumulator columns at a time (register budget: 2 CTAs / SM)
        uint32_t acc[2][16];
        if (cp * 32 < p.Npad) {
          tmem_ld16(trow + (uint32_t)(cp * 32), acc[0]);
          if (cp * 32 + 16 < p.Npad) tmem_ld16(trow + (uint32_t)(cp * 32 + 16), acc[1]);
          tmem_ld_wait();
        }
        if (cp == 1 || p.Npad <= 32) {            // last read of this accumulator: hand it back before the stores
          tc_fence_before_sync();
          __syncwarp();
          if (lane == 0 && (cp == 1) == (p.Npad > 32)) mbar_arrive(&tempty_bar[s]);
        }
        if (ok && cp * 32 < p.Npad) {
#pragma unroll
          for (int c = 0; c < 2; ++c) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int c0 = cp * 32 + c * 16 + h * 8;
              if (c0 < p.Cout) {
                float f[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = fmaf(__uint_as_float(acc[c][h * 8 + e]), p.in_scale, s_shift[c0 + e]);
                *reinterpret_cast<uint4*>(o + c0) = pack8(f);
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

typedef CUresult (*st_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// uint8 NCHW image -> NHWC bf16, out = scale * conv4x4s4(img * in_scale) + shift.  Cout % 8 == 0, Cout <= 64, W % 16 == 0.
int stem_tc_u8(const uint8_t* img, int N, int H, int W, int Cout, const float* wk, const float* scale, const float* shift,
               float in_scale, __nv_bfloat16* out, long ld_out, cudaStream_t st) {
  static st_encode_fn fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return fail(LY_EUNSUPPORTED, "cuTensorMapEncodeTiled not available");
    fn = (st_encode_fn)f;
  }
  StemTcParams p;
  memset(&p, 0, sizeof(p));
  p.wk = wk; p.scale = scale; p.shift = shift; p.out = out; p.ld_out = ld_out; p.in_scale = in_scale;
  p.Ho = H / 4; p.Wo = W / 4; p.Cout = Cout; p.Npad = (Cout + 15) / 16 * 16;
  p.tiles_w = ceil_div(p.Wo, 32); p.tiles_h = ceil_div(p.Ho, 8);
  const long stages = (long)p.tiles_w * p.tiles_h * N;
  LY_REQUIRE(stages < (1L << 30), "ly_stem_fwd: image batch too large");
  p.stages = (int)stages;
  {
    cuuint64_t gd[4] = {(cuuint64_t)W, (cuuint64_t)H, 3, (cuuint64_t)N};
    cuuint64_t gs[3] = {(cuuint64_t)W, (cuuint64_t)H * W, (cuuint64_t)3 * H * W};
    cuuint32_t bd[4] = {128, 32, 3, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = fn(&p.img_map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<uint8_t*>(img), gd, gs, bd, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(LY_ECUDA, "cuTensorMapEncodeTiled(uint8 image) failed: %d", (int)r);
  }
  const size_t smem = 4 * (size_t)kStTile + 64 * 128 + (size_t)kStRawSlots * kStRaw + 1024;
  static std::atomic<bool> attr_done[64];   // zero-initialised; set only after the attribute call succeeded
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(stem_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done[dev & 63].store(true);
  }
  const long cap = 2L * device_sm_count();
  LY_CUDA(launch(stem_tc_kernel, (unsigned)(stages < cap ? stages : cap), kStThreads, smem, st, p));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

}  // namespace ly
