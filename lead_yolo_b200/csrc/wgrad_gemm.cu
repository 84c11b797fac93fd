// Weight gradient of the implicit-GEMM convolutions on tcgen05 tensor cores (sm_100a).
//
//   dWp[o][k] += Σ_pixels dZ[pixel][o] * Acol[pixel][k]          (fp32, packed-K layout of conv_gemm.cu)
//
// The contraction runs over pixels, so both operands are "MN-major" for the tensor core: a TMA box of
// (64 channels x P pixels) lands in 128B-swizzled shared memory exactly as the canonical MN-major SW128
// atom sequence (8 pixel rows x 128 B per atom).  One CTA owns a (128 output channels) x (up to 4 K chunks
// of 64) accumulator in TMEM, walks a strided subset of the pixel tiles (split over gridDim.x), and finally
// adds its partial tile into dWp with fp32 atomics (caller zero-initialises dWp).
#include <string.h>

#include <mutex>

#include <atomic>

#include "ly_common.cuh"
#include "ly_tc.cuh"
#include "ly_tmap.h"

namespace ly {

constexpr int kWMaxK = 96;
constexpr int kSlot = 128 * 128;  // one 64-channel x 128-pixel box
constexpr int kWMaxStages = 8;   // ring depth = 192 KB / stage size: 2 for (2 dZ + 4 activation) boxes, up to 6 for (1 + 1)

struct WEntry {
  int16_t src;
  int16_t c0;
  int8_t dw, dh, tz, pad_;
};

struct alignas(64) WgradParams {
  CUtensorMap dz_map;
  CUtensorMap a_map[4];
  int tiles_w, tiles_h, n_img;
  int Wt, Ht, P;  // P = Wt*Ht pixel rows per tile (multiple of 16)
  int num_k, Cout, Kp;
  int stages, ndz, slots;   // ring depth; dZ boxes per stage (1 when Cout <= 64); boxes per stage = ndz + min(4, num_k)
  float* dw;
  WEntry sched[kWMaxK];
};

__global__ void __launch_bounds__(192) wgrad_gemm_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[kWMaxStages];
  __shared__ uint64_t empty_bar[kWMaxStages];
  __shared__ uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  // A stage = the dZ boxes + the activation boxes of one pixel tile.  The ring is as deep as 192 KB allow: a tile is a
  // handful of MMAs, so what bounds a small-K layer is how many tiles' loads are in flight (two stages: ~2 us per tile).
  const int kStageBytes = p.slots * kSlot;

  const int ct = blockIdx.y;   // 128-wide output-channel tile
  const int kg = blockIdx.z;   // group of 4 K chunks
  const int k_begin = kg * 4;
  const int nch = min(4, p.num_k - k_begin);
  const int num_tiles = p.tiles_w * p.tiles_h * p.n_img;
  const int my_tiles = (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.dz_map);
    tma_prefetch_desc(&p.a_map[0]);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_s, 256);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_trigger();
  pdl_wait();
  const uint32_t tmem_base = tmem_base_s;
  const uint32_t box_bytes = (uint32_t)p.P * 128u;

  if (my_tiles > 0) {
    if (warp == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < my_tiles; ++i) {
        mbar_wait(&empty_bar[s], ph ^ 1u);
        if (lane == 0) {
          const int tile = (int)blockIdx.x + i * (int)gridDim.x;
          const int tw = tile % p.tiles_w;
          const int th = (tile / p.tiles_w) % p.tiles_h;
          const int n = tile / (p.tiles_w * p.tiles_h);
          const int w0 = tw * p.Wt, h0 = th * p.Ht;
          uint8_t* st = smem + s * kStageBytes;
          mbar_arrive_expect_tx(&full_bar[s], box_bytes * (uint32_t)(p.ndz + nch));
          tma_load_5d(st, &p.dz_map, &full_bar[s], ct * 128, w0, 0, h0, n);
          if (p.ndz == 2) tma_load_5d(st + kSlot, &p.dz_map, &full_bar[s], ct * 128 + 64, w0, 0, h0, n);
          for (int c = 0; c < nch; ++c) {
            const WEntry e = p.sched[k_begin + c];
            tma_load_5d(st + (p.ndz + c) * kSlot, &p.a_map[e.src], &full_bar[s], e.c0, w0 + e.dw, e.tz, h0 + e.dh, n);
          }
        }
        __syncwarp();
        if (++s == p.stages) { s = 0; ph ^= 1u; }
      }
    } else if (warp == 1) {
      const uint32_t idesc = umma_idesc_bf16_f32(128, nch * 64, 1, 1);
      const int ksteps = p.P / 16;
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < my_tiles; ++i) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after_sync();
        if (lane == 0) {
          // Cout <= 64: one dZ box; rows 64..127 of the M = 128 operand then run into the first activation box and fill
          // accumulator rows that are never stored
          const uint32_t a_base = smem_u32(smem + s * kStageBytes);
          const uint32_t b_base = a_base + p.ndz * kSlot;
          for (int k = 0; k < ksteps; ++k) {
            const uint64_t adesc = umma_smem_desc_sw128(a_base + k * 2048, kSlot, 1024);
            const uint64_t bdesc = umma_smem_desc_sw128(b_base + k * 2048, kSlot, 1024);
            umma_bf16(tmem_base, adesc, bdesc, idesc, (uint32_t)((i | k) != 0));
          }
          umma_commit(&empty_bar[s]);
          if (i == my_tiles - 1) umma_commit(&tmem_full_bar);
        }
        __syncwarp();
        if (++s == p.stages) { s = 0; ph ^= 1u; }
      }
    } else {
      const int q = warp & 3;
      const int o = ct * 128 + q * 32 + lane;
      mbar_wait(&tmem_full_bar, 0);
      tc_fence_after_sync();
      float* drow = p.dw + (long)o * p.Kp + (long)k_begin * 64;
      for (int c0 = 0; c0 < nch * 64; c0 += 16) {
        uint32_t acc[16];
        __syncwarp();
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, acc);
        tmem_ld_wait();
        if (o < p.Cout) {
          // 16-byte vector reductions (rows of dWp are 256-byte aligned: Kp is a multiple of 64)
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            atomicAdd(reinterpret_cast<float4*>(drow + c0 + j),
                      make_float4(__uint_as_float(acc[j]), __uint_as_float(acc[j + 1]), __uint_as_float(acc[j + 2]),
                                  __uint_as_float(acc[j + 3])));
        }
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 256);
}

static void pick_tile16(int Hout, int Wout, int* Wt, int* Ht) {
  double best = -1.0;
  *Wt = 16;
  *Ht = 8;
  const int cand[6] = {8, 16, 32, 64, 128, Wout <= 128 ? Wout : 128};
  for (int i = 0; i < 6; ++i) {
    const int wt = cand[i];
    if (wt <= 0 || wt > 128) continue;
    int ht = 128 / wt;
    while (ht > 1 && (wt * ht) % 16 != 0) --ht;
    if ((wt * ht) % 16 != 0) continue;
    const double eff = (double)Hout * Wout / ((double)ceil_div(Wout, wt) * ceil_div(Hout, ht) * 128.0);
    if (eff > best + 1e-9) {
      best = eff;
      *Wt = wt;
      *Ht = ht;
    }
  }
}

}  // namespace ly

using namespace ly;

// Same descriptor as ly_conv_fwd describes the forward convolution; dz is its output gradient
// (bf16 NHWC [N, Hout, Wout, Cout], pixel stride dz_ld); dw is fp32 [Cout][Kp], accumulated into.
extern "C" int ly_conv_wgrad(const ly_conv_desc* d, const void* dz, long dz_ld, float* dw, ly_stream_t stream) {
  LY_REQUIRE(d != nullptr && dz != nullptr && dw != nullptr, "ly_conv_wgrad: null pointer");
  LY_REQUIRE(d->n_src >= 1 && d->n_src <= 4, "ly_conv_wgrad: n_src=%d", d->n_src);
  LY_REQUIRE(d->Cout % 8 == 0 && dz_ld % 8 == 0 && dz_ld >= d->Cout, "ly_conv_wgrad: Cout=%d dz_ld=%ld invalid",
             d->Cout, dz_ld);
  bool flat = false, patch = false;
  int taps, Hout, Wout;
  if (d->kh == 1 && d->kw == 1 && d->stride == 1 && d->pad == 0) {
    flat = true; taps = 1; Hout = d->H; Wout = d->W;
  } else if (d->kh == 3 && d->kw == 3 && d->stride == 1 && d->pad == 1) {
    taps = 9; Hout = d->H; Wout = d->W;
  } else if (d->kh == d->kw && d->stride == d->kh && d->pad == 0 && (d->kh == 2 || d->kh == 4)) {
    LY_REQUIRE(d->H % d->kh == 0 && d->W % d->kw == 0, "ly_conv_wgrad: H,W not divisible by k");
    patch = true; taps = d->kh; Hout = d->H / d->kh; Wout = d->W / d->kw;
  } else {
    return fail(LY_EUNSUPPORTED, "ly_conv_wgrad: geometry k=%d s=%d p=%d unsupported", d->kh, d->stride, d->pad);
  }

  WgradParams p;
  memset(&p, 0, sizeof(p));
  if (flat) {
    const long M = (long)d->N * d->H * d->W;
    p.n_img = 1; p.tiles_h = 1; p.Wt = 128; p.Ht = 1;
    p.tiles_w = ceil_div(M, 128);
  } else {
    p.n_img = d->N;
    pick_tile16(Hout, Wout, &p.Wt, &p.Ht);
    p.tiles_w = ceil_div(Wout, p.Wt);
    p.tiles_h = ceil_div(Hout, p.Ht);
  }
  p.P = p.Wt * p.Ht;
  LY_REQUIRE(p.P % 16 == 0 && p.P <= 128, "ly_conv_wgrad: internal tile error P=%d", p.P);
  int rc;
  {
    uint64_t dims[5], str[4];
    uint32_t box[5] = {64, (uint32_t)p.Wt, 1, (uint32_t)p.Ht, 1};
    if (flat) {
      const uint64_t M = (uint64_t)d->N * d->H * d->W;
      dims[0] = (uint64_t)d->Cout; dims[1] = M; dims[2] = 1; dims[3] = 1; dims[4] = 1;
      str[0] = (uint64_t)dz_ld * 2; str[1] = M * dz_ld * 2; str[2] = str[1]; str[3] = str[1];
    } else {
      dims[0] = (uint64_t)d->Cout; dims[1] = (uint64_t)Wout; dims[2] = 1; dims[3] = (uint64_t)Hout;
      dims[4] = (uint64_t)d->N;
      str[0] = (uint64_t)dz_ld * 2; str[1] = (uint64_t)Wout * dz_ld * 2; str[2] = str[1];
      str[3] = (uint64_t)Hout * Wout * dz_ld * 2;
    }
    rc = tmap_encode_bf16(&p.dz_map, dz, 5, dims, str, box);
    if (rc != LY_OK) return rc;
  }
  int it = 0;
  for (int s = 0; s < d->n_src; ++s) {
    LY_REQUIRE(d->src[s] != nullptr && d->src_c[s] % 8 == 0 && d->src_ld[s] % 8 == 0 && d->src_ld[s] >= d->src_c[s],
               "ly_conv_wgrad: source %d invalid", s);
    const long ld = d->src_ld[s];
    uint64_t dims[5], str[4];
    uint32_t box[5] = {64, (uint32_t)p.Wt, 1, (uint32_t)p.Ht, 1};
    int ceff = d->src_c[s];
    if (flat) {
      const uint64_t M = (uint64_t)d->N * d->H * d->W;
      dims[0] = (uint64_t)ceff; dims[1] = M; dims[2] = 1; dims[3] = 1; dims[4] = 1;
      str[0] = (uint64_t)ld * 2; str[1] = M * ld * 2; str[2] = str[1]; str[3] = str[1];
    } else if (patch) {
      LY_REQUIRE(ld == d->src_c[s], "ly_conv_wgrad: patch conv needs a dense source");
      const int k = d->kh;
      ceff = d->src_c[s] * k;
      dims[0] = (uint64_t)ceff; dims[1] = (uint64_t)Wout; dims[2] = (uint64_t)k; dims[3] = (uint64_t)Hout;
      dims[4] = (uint64_t)d->N;
      str[0] = (uint64_t)ceff * 2; str[1] = (uint64_t)d->W * ld * 2; str[2] = (uint64_t)k * d->W * ld * 2;
      str[3] = (uint64_t)d->H * d->W * ld * 2;
    } else {
      dims[0] = (uint64_t)ceff; dims[1] = (uint64_t)d->W; dims[2] = 1; dims[3] = (uint64_t)d->H;
      dims[4] = (uint64_t)d->N;
      str[0] = (uint64_t)ld * 2; str[1] = (uint64_t)d->W * ld * 2; str[2] = str[1];
      str[3] = (uint64_t)d->H * d->W * ld * 2;
    }
    rc = tmap_encode_bf16(&p.a_map[s], d->src[s], 5, dims, str, box);
    if (rc != LY_OK) return rc;
    for (int t = 0; t < taps; ++t)
      for (int c0 = 0; c0 < ceff; c0 += 64) {
        LY_REQUIRE(it < kWMaxK, "ly_conv_wgrad: K too large");
        WEntry e;
        e.src = (int16_t)s; e.c0 = (int16_t)c0; e.pad_ = 0;
        if (patch) { e.dw = 0; e.dh = 0; e.tz = (int8_t)t; }
        else { e.dw = (int8_t)(t % d->kw - d->pad); e.dh = (int8_t)(t / d->kw - d->pad); e.tz = 0; }
        p.sched[it++] = e;
      }
  }
  p.num_k = it;
  p.Kp = it * 64;
  p.Cout = d->Cout;
  p.dw = dw;

  const int cts = ceil_div(d->Cout, 128);
  const int kgs = ceil_div(p.num_k, 4);
  const int num_tiles = p.tiles_w * p.tiles_h * p.n_img;
  // one CTA per SM (192 KB of shared memory each): exactly one wave, and half the atomic traffic of two
  int split = ceil_div((long)device_sm_count(), (long)cts * kgs);
  if (split > num_tiles) split = num_tiles;
  if (split < 1) split = 1;

  static std::atomic<bool> done[64];
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  p.ndz = d->Cout > 64 ? 2 : 1;
  p.slots = p.ndz + (p.num_k < 4 ? p.num_k : 4);
  p.stages = 12 / p.slots;                      // 12 boxes of 16 KB = 192 KB
  if (p.stages > kWMaxStages) p.stages = kWMaxStages;
  const size_t smem_bytes = (size_t)p.stages * p.slots * kSlot + 1024;
  if (!done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(wgrad_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 12 * kSlot + 1024));
    done[dev & 63].store(true);
  }
  dim3 grid((unsigned)split, (unsigned)cts, (unsigned)kgs);
  LY_CUDA(launch(wgrad_gemm_kernel, grid, 192, smem_bytes, (cudaStream_t)stream, p));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
