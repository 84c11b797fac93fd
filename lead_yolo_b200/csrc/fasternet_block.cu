// Fused FasterNet block for inference (reference models/common.py:1432-1437, 1478-1482):
//
//     y = x + W2 relu(scale * (W1 [conv3x3(x[:, :Cp]) | x[:, Cp:]]) + shift)
//
// ONE kernel per block: a CTA owns a 128-pixel spatial tile; x is read once, y written once, the partial-conv output,
// the 2C-wide hidden tensor and its BN/ReLU never leave the SM.
//   warp 0   TMA: the x tile (64-channel SW128 box), nine 16-channel SW32 tap boxes (zero padding = TMA OOB fill) and
//            one bulk copy of the pre-swizzled weight image
//   warp 1   tcgen05 issuer: 9 MMAs (N=16,K=16) -> t1 in TMEM;  GEMM1 = x-part (K=64 chunk) + t1-part (K=16) -> u;
//            GEMM2 over the ReLU'd hidden tile -> y
//   warps 2-5  TMEM -> registers -> shared: t1 (bf16, SW32 A operand), hidden v = relu(scale*u+shift) (bf16, SW128 A
//            operand); final epilogue adds the residual from the x tile still in shared memory and stores y
// Several CTAs are resident per SM (72-85 KB smem each), so one CTA's TMA/epilogue phases overlap another's MMAs.
// Supported: C in {24, 40} (the HBM-bound FasterNet stages of LEAD-YOLO); other widths use the per-layer kernels.
#include <string.h>

#include "ly_common.cuh"
#include "ly_tc.cuh"
#include "ly_tmap.h"

namespace ly {

struct alignas(64) FnbParams {
  CUtensorMap x_map;    // (C, W, 1, H, N) box (64, Wt, 1, Ht, 1)  SW128
  CUtensorMap tap_map;  // (16, W, 1, H, N) box (16, Wt, 1, Ht, 1) SW32
  const __nv_bfloat16* x;
  __nv_bfloat16* y;
  const uint8_t* wimg;   // pre-swizzled weight image (see lead_yolo_b200/packing.py:fasternet_weight_image)
  const float* scale;    // [2C] folded BN
  const float* shift;
  int C, C2, N2;         // channels, hidden = 2C, GEMM2 N (C rounded up to 16)
  int H, W, Wt, Ht, tiles_w, tiles_h;
  int w_bytes;           // size of the weight image
  int off_w2, off_w1p, off_wp;   // byte offsets inside the image (W1x at 0)
  int v_chunks;          // 64-wide K chunks of the hidden tile
  int tmem_cols;
};

__device__ __forceinline__ uint64_t desc_sw32(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) |
         (6ull << 61);  // SWIZZLE_32B, 8-row groups 256 B apart
}
__device__ __forceinline__ void bulk_copy_g2s(void* smem, const void* gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem)),
               "l"(gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

constexpr int kXBytes = 128 * 128;      // x tile, SW128
constexpr int kTapBytes = 128 * 32;     // one tap box, SW32
constexpr int kOffTap = kXBytes;
constexpr int kOffT1 = kOffTap + 9 * kTapBytes;
constexpr int kOffW = kOffT1 + kTapBytes;   // weight image; the hidden tile V aliases the tap boxes

__global__ void __launch_bounds__(192) fasternet_block_kernel(const __grid_constant__ FnbParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t ld_bar, t1_full, t1_ready, u_full, v_ready, y_full;
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sX = smem;
  uint8_t* sTap = smem + kOffTap;
  uint8_t* sT1 = smem + kOffT1;
  uint8_t* sW = smem + kOffW;
  uint8_t* sV = sTap;  // alias: the tap boxes are dead once t1 has been computed

  if (warp == 1 && lane == 0) {
    mbar_init(&ld_bar, 1);
    mbar_init(&t1_full, 1);
    mbar_init(&t1_ready, 128);
    mbar_init(&u_full, 1);
    mbar_init(&v_ready, 128);
    mbar_init(&y_full, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_s, (uint32_t)p.tmem_cols);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = tmem_base_s;
  const uint32_t tm_t1 = tmem, tm_u = tmem + 16, tm_y = tmem + 16 + (uint32_t)p.C2;

  const int bx = blockIdx.x;
  const int tw = bx % p.tiles_w, th = (bx / p.tiles_w) % p.tiles_h, n = bx / (p.tiles_w * p.tiles_h);
  const int w0 = tw * p.Wt, h0 = th * p.Ht;

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&p.x_map);
      tma_prefetch_desc(&p.tap_map);
      const uint32_t box_x = (uint32_t)(64 * p.Wt * p.Ht * 2), box_t = (uint32_t)(16 * p.Wt * p.Ht * 2);
      mbar_arrive_expect_tx(&ld_bar, box_x + 9 * box_t + (uint32_t)p.w_bytes);
      bulk_copy_g2s(sW, p.wimg, (uint32_t)p.w_bytes, &ld_bar);
      tma_load_5d(sX, &p.x_map, &ld_bar, 0, w0, 0, h0, n);
#pragma unroll
      for (int t = 0; t < 9; ++t)
        tma_load_5d(sTap + t * kTapBytes, &p.tap_map, &ld_bar, 0, w0 + t % 3 - 1, 0, h0 + t / 3 - 1, n);
    }
  } else if (warp == 1) {
    // ---- partial 3x3 conv: t1[128 x 16] = sum_taps tap[128 x 16] * Wp_tap[16 x 16]^T
    mbar_wait(&ld_bar, 0);
    tc_fence_after_sync();
    if (lane == 0) {
      const uint32_t id16 = umma_idesc_bf16_f32(128, 16, 0, 0);
#pragma unroll
      for (int t = 0; t < 9; ++t)
        umma_bf16(tm_t1, desc_sw32(smem_u32(sTap + t * kTapBytes)), desc_sw32(smem_u32(sW + p.off_wp + t * 512)), id16,
                  (uint32_t)(t != 0));
      umma_commit(&t1_full);
    }
    __syncwarp();
    // ---- GEMM1: u[128 x 2C] = x[128 x 64] * W1x^T + t1[128 x 16] * W1p^T
    mbar_wait(&t1_ready, 0);
    tc_fence_after_sync();
    if (lane == 0) {
      const uint32_t id1 = umma_idesc_bf16_f32(128, p.C2, 0, 0);
      const uint32_t xb = smem_u32(sX), wb = smem_u32(sW);
      const int kx = (p.C + 15) / 16;
      for (int k = 0; k < kx; ++k)
        umma_bf16(tm_u, umma_smem_desc_sw128(xb + k * 32, 16, 1024), umma_smem_desc_sw128(wb + k * 32, 16, 1024), id1,
                  (uint32_t)(k != 0));
      umma_bf16(tm_u, desc_sw32(smem_u32(sT1)), desc_sw32(wb + p.off_w1p), id1, 1u);
      umma_commit(&u_full);
    }
    __syncwarp();
    // ---- GEMM2: y[128 x N2] = v[128 x 2C] * W2^T
    mbar_wait(&v_ready, 0);
    tc_fence_after_sync();
    if (lane == 0) {
      const uint32_t id2 = umma_idesc_bf16_f32(128, p.N2, 0, 0);
      const uint32_t vb = smem_u32(sV), wb = smem_u32(sW + p.off_w2);
      const int kv = (p.C2 + 15) / 16;
      for (int k = 0; k < kv; ++k) {
        const int kc = k >> 2, kk = k & 3;
        umma_bf16(tm_y, umma_smem_desc_sw128(vb + kc * kXBytes + kk * 32, 16, 1024),
                  umma_smem_desc_sw128(wb + kc * (p.N2 * 128) + kk * 32, 16, 1024), id2, (uint32_t)(k != 0));
      }
      umma_commit(&y_full);
    }
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int r = q * 32 + lane;                 // tile row = TMEM lane
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    // ---- t1: TMEM -> bf16 -> SW32 A operand
    mbar_wait(&t1_full, 0);
    tc_fence_after_sync();
    {
      uint32_t acc[16];
      tmem_ld16(tm_t1 + lane_addr, acc);
      tmem_ld_wait();
      float v[8];
#pragma unroll
      for (int g = 0; g < 2; ++g) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(acc[g * 8 + j]);
        *reinterpret_cast<uint4*>(sT1 + r * 32 + ((g ^ ((r >> 2) & 1)) << 4)) = pack8(v);
      }
    }
    tc_fence_before_sync();
    fence_proxy_async_smem();
    mbar_arrive(&t1_ready);
    // ---- hidden: v = relu(scale*u + shift) -> bf16 -> SW128 A operand (zero padded to the 64-wide K chunks)
    mbar_wait(&u_full, 0);
    tc_fence_after_sync();
    for (int c0 = 0; c0 < p.v_chunks * 64; c0 += 16) {
      uint32_t acc[16];
      if (c0 < p.C2) {
        tmem_ld16(tm_u + lane_addr + (uint32_t)c0, acc);
        tmem_ld_wait();
      }
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int c = c0 + g * 8;
        float v[8];
        if (c < p.C2) {
          const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.scale + c));
          const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.scale + c + 4));
          const float4 h0 = __ldg(reinterpret_cast<const float4*>(p.shift + c));
          const float4 h1 = __ldg(reinterpret_cast<const float4*>(p.shift + c + 4));
          const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
          const float sh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = fmaxf(fmaf(__uint_as_float(acc[g * 8 + j]), sc[j], sh[j]), 0.f);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = 0.f;
        }
        const int kc = c >> 6, ch = (c >> 3) & 7;
        *reinterpret_cast<uint4*>(sV + kc * kXBytes + r * 128 + ((ch ^ (r & 7)) << 4)) = pack8(v);
      }
    }
    tc_fence_before_sync();
    fence_proxy_async_smem();
    mbar_arrive(&v_ready);
    // ---- output: y = x + W2 v
    mbar_wait(&y_full, 0);
    tc_fence_after_sync();
    const int hh = r / p.Wt, ww = r - hh * p.Wt;
    const int h = h0 + hh, w = w0 + ww;
    const bool valid = (r < p.Wt * p.Ht) && h < p.H && w < p.W;
    const long pix = ((long)n * p.H + h) * p.W + w;
    for (int c0 = 0; c0 < p.N2; c0 += 16) {
      uint32_t acc[16];
      __syncwarp();
      tmem_ld16(tm_y + lane_addr + (uint32_t)c0, acc);
      tmem_ld_wait();
      if (!valid) continue;
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int c = c0 + g * 8;
        if (c >= p.C) break;
        float xr[8], v[8];
        unpack8(*reinterpret_cast<const uint4*>(sX + r * 128 + (((c >> 3) ^ (r & 7)) << 4)), xr);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(acc[g * 8 + j]) + xr[j];
        *reinterpret_cast<uint4*>(p.y + pix * p.C + c) = pack8(v);
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, (uint32_t)p.tmem_cols);
}

// TMA encode with a chosen swizzle (capi.cu offers SW128 only)
typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_sw32(CUtensorMap* out, const void* base, const uint64_t* dims, const uint64_t* strides,
                       const uint32_t* box) {
  static encode_fn fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return fail(LY_EUNSUPPORTED, "cuTensorMapEncodeTiled not available");
    fn = (encode_fn)f;
  }
  cuuint64_t gd[5], gs[4];
  cuuint32_t bd[5], es[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < 5; ++i) { gd[i] = dims[i]; bd[i] = box[i]; }
  for (int i = 0; i < 4; ++i) gs[i] = strides[i];
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), gd, gs, bd, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(LY_ECUDA, "cuTensorMapEncodeTiled(SW32) failed: %d", (int)r);
  return LY_OK;
}

}  // namespace ly

using namespace ly;

// x, y: dense NHWC bf16 [N][H][W][C], C in {24, 40}.  wimg: pre-swizzled weight image built by
// lead_yolo_b200.packing.fasternet_weight_image (sections W1x | W2 | W1p | Wp, offsets passed in).
extern "C" int ly_fasternet_block_infer(const void* x, void* y, int N, int H, int W, int C, const void* wimg,
                                        int w_bytes, int off_w2, int off_w1p, int off_wp, const float* scale,
                                        const float* shift, ly_stream_t st) {
  LY_REQUIRE(x && y && wimg && scale && shift && N > 0 && H > 0 && W > 0, "ly_fasternet_block_infer: bad args");
  LY_REQUIRE(C == 24 || C == 40, "ly_fasternet_block_infer: C=%d not supported (24 or 40)", C);
  LY_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0 &&
                 reinterpret_cast<uintptr_t>(wimg) % 16 == 0 && w_bytes % 16 == 0,
             "ly_fasternet_block_infer: alignment");
  FnbParams p;
  memset(&p, 0, sizeof(p));
  p.x = (const __nv_bfloat16*)x;
  p.y = (__nv_bfloat16*)y;
  p.wimg = (const uint8_t*)wimg;
  p.scale = scale;
  p.shift = shift;
  p.C = C;
  p.C2 = 2 * C;
  p.N2 = (C + 15) / 16 * 16;
  p.H = H;
  p.W = W;
  p.w_bytes = w_bytes;
  p.off_w2 = off_w2;
  p.off_w1p = off_w1p;
  p.off_wp = off_wp;
  p.v_chunks = (p.C2 + 63) / 64;
  LY_REQUIRE(p.v_chunks * kXBytes <= 9 * kTapBytes, "hidden tile does not fit the aliased tap region");
  const int need_cols = 16 + p.C2 + p.N2;
  p.tmem_cols = 32;
  while (p.tmem_cols < need_cols) p.tmem_cols *= 2;
  // spatial tile of 128 pixels
  int best_wt = 16, best_ht = 8;
  double best = -1;
  const int cand[6] = {8, 16, 32, 64, 128, W <= 128 ? W : 128};
  for (int i = 0; i < 6; ++i) {
    const int wt = cand[i];
    if (wt <= 0 || wt > 128) continue;
    int ht = 128 / wt;
    if (ht > H) ht = H;
    const double eff = (double)H * W / ((double)ceil_div(W, wt) * ceil_div(H, ht) * 128.0);
    if (eff > best + 1e-9) { best = eff; best_wt = wt; best_ht = ht; }
  }
  p.Wt = best_wt;
  p.Ht = best_ht;
  p.tiles_w = ceil_div(W, p.Wt);
  p.tiles_h = ceil_div(H, p.Ht);
  {
    uint64_t dims[5] = {(uint64_t)C, (uint64_t)W, 1, (uint64_t)H, (uint64_t)N};
    uint64_t str[4] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
    uint32_t box[5] = {64, (uint32_t)p.Wt, 1, (uint32_t)p.Ht, 1};
    int rc = tmap_encode_bf16(&p.x_map, x, 5, dims, str, box);
    if (rc) return rc;
    uint64_t dims2[5] = {16, (uint64_t)W, 1, (uint64_t)H, (uint64_t)N};
    uint32_t box2[5] = {16, (uint32_t)p.Wt, 1, (uint32_t)p.Ht, 1};
    rc = encode_sw32(&p.tap_map, x, dims2, str, box2);
    if (rc) return rc;
  }
  const size_t smem = (size_t)kOffW + (size_t)w_bytes + 1024;
  static bool attr_done[64] = {false};
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    LY_CUDA(cudaFuncSetAttribute(fasternet_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    attr_done[dev & 63] = true;
  }
  LY_REQUIRE(smem <= 112 * 1024, "ly_fasternet_block_infer: weight image too large (%d bytes)", w_bytes);
  const unsigned grid = (unsigned)(p.tiles_w * p.tiles_h * N);
  fasternet_block_kernel<<<grid, 192, smem, (cudaStream_t)st>>>(p);
  LY_LAUNCH_CHECK();
  return LY_OK;
}
