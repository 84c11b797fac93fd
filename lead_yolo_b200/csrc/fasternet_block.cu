// Fused FasterNet block for inference (reference models/common.py:1432-1437, 1478-1482):
//
//     y = x + W2 relu(scale * (W1 [conv3x3(x[:, :Cp]) | x[:, Cp:]]) + shift)
//
// ONE persistent kernel per block: a CTA walks 128-pixel spatial tiles; x is read once, y written once, the hidden
// tensor and its BN/ReLU never leave the SM.  The partial 3x3 conv is folded into the first GEMM on the host
// (Wc_tap = W1[:, :Cp] * Wp_tap, packing.py), so a tile is two MMA phases instead of three:
//     u = x * W1x^T + sum_taps tap * Wc_tap^T      (K = 64-chunk of x with its first Cp columns zeroed in W1x, 9 x K=16)
//     y = relu(scale*u + shift) * W2^T + x
// Tile rows use the pitch P = Wt + 2 of the tile's halo: accumulator row m = y*P + x (rows with x >= Wt are dead), so
// the nine taps are ONE 16-channel halo box (Ht+2 x P pixels, zero padding = TMA OOB fill) read through descriptors that
// start (dy+1)*P + (dx+1) rows further in (tests/dev/desc_shift.cu: the swizzle is a function of the absolute address).
// That is 160-256 TMA box rows per tile instead of 9 x 128 — the TMA unit moves one 32-byte row per cycle.
//   warp 0   TMA (one lane): weight image once; per tile the x box (64-channel SW128) and the halo box (SW32), both
//            double buffered and requested as soon as the MMAs two tiles back have consumed theirs
//   warp 1   tcgen05 issuer (one lane): phase A -> u in TMEM, commit; phase B -> y in TMEM, commit
//   warps 2-5  hidden group: u (TMEM) -> relu(scale*u+shift) -> bf16 SW128 A operand of phase B in shared memory
//   warps 6-9  output group: copies its row of the x box to registers (the box is then released), adds it to y (TMEM)
//            and stores; the two groups work on different phases of consecutive tiles at the same time
// Two CTAs per SM when the buffers fit (C = 24), so one CTA's epilogues overlap the other's MMAs.
// Supported by this kernel: C in {24, 40} (the HBM-bound FasterNet stages of LEAD-YOLO); C = 80 has its own kernel below;
// other widths use the per-layer kernels.
#include <string.h>

#include <atomic>

#include "ly_common.cuh"
#include "ly_tc.cuh"
#include "ly_tmap.h"

namespace ly {

// Per-role wait accounting for tests/dev/fnb_probe.cu; compiled out of the library.
#ifdef FNB_DEBUG
__device__ long long fnb_dbg[296][16];
#define FB_WAIT(slot, bar, par)                    \
  do {                                             \
    const long long t__ = clock64();               \
    mbar_wait(bar, par);                           \
    dbg_acc[slot] += clock64() - t__;              \
  } while (0)
#define FB_DBG_DECL long long dbg_acc[6] = {0, 0, 0, 0, 0, 0}; const long long dbg_t0 = clock64();
#define FB_DBG_FLUSH(base, nslots)                                                                     \
  if (blockIdx.x < 296) {                                                                              \
    for (int z = 0; z < nslots; ++z) fnb_dbg[blockIdx.x][base + z] = dbg_acc[z];                       \
    fnb_dbg[blockIdx.x][base + nslots] = clock64() - dbg_t0;                                           \
  }
__device__ __forceinline__ long long fb_gtime() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define FB_STAMP(slot) if (threadIdx.x == 0 && blockIdx.x < 296) fnb_dbg[blockIdx.x][slot] = fb_gtime();
#else
#define FB_WAIT(slot, bar, par) mbar_wait(bar, par)
#define FB_DBG_DECL
#define FB_DBG_FLUSH(base, nslots)
#define FB_STAMP(slot)
#endif

struct alignas(64) FnbParams {
  CUtensorMap x_map;    // (C, W, 1, H, N) box (64, P, 1, Ht, 1)  SW128
  CUtensorMap tap_map;  // (16, W, 1, H, N) box (16, P, 1, Ht+2, 1) SW32
  const __nv_bfloat16* x;
  __nv_bfloat16* y;
  const uint8_t* wimg;   // pre-swizzled weight image (see lead_yolo_b200/packing.py:fasternet_weight_image)
  const float* scale;    // [2C] folded BN
  const float* shift;
  int C, C2, N2;         // channels, hidden = 2C, GEMM2 N (C rounded up to 16)
  int H, W, Wt, Ht, P, tiles_w, tiles_h, tiles;
  int w_bytes;           // size of the weight image
  int off_w2, off_wc;    // byte offsets inside the image (W1x at 0)
  int v_chunks;          // 64-wide K chunks of the hidden tile
  int xbufs;             // 1 or 2 x buffers
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
constexpr int kHaloBytes = 288 * 32;    // halo box (<= 256 rows) + the rows a shifted 128-row read may run past it

__global__ void __launch_bounds__(320) fasternet_block_kernel(const __grid_constant__ FnbParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t w_bar, x_full[2], x_free[2], taps_full[2], u_full[2], v_ready, y_full, y_free;
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_sc[128], s_sh[128];   // folded BN of the hidden layer (2C <= 128 values)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sX = smem;                                   // xbufs x 16 KB
  uint8_t* sTap = sX + p.xbufs * kXBytes;               // 2 halo buffers
  uint8_t* sV = sTap + 2 * kHaloBytes;                  // v_chunks x 16 KB
  uint8_t* sW = sV + p.v_chunks * kXBytes;              // weight image

  if (warp == 1 && lane == 0) {
    mbar_init(&w_bar, 1);
    for (int b = 0; b < 2; ++b) { mbar_init(&x_full[b], 1); mbar_init(&x_free[b], 128); }
    mbar_init(&taps_full[0], 1);
    mbar_init(&taps_full[1], 1);
    mbar_init(&u_full[0], 1);
    mbar_init(&u_full[1], 1);
    mbar_init(&v_ready, 128);
    mbar_init(&y_full, 1);
    mbar_init(&y_free, 128);
    fence_mbar_init();
  }
  // the hidden tile's K padding (columns >= 2C of its last 64-wide chunk) is zeroed once and never written again
  for (int i = threadIdx.x; i < p.v_chunks * kXBytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(sV)[i] = make_uint4(0, 0, 0, 0);
  pdl_wait();
  for (int i = threadIdx.x; i < 128; i += blockDim.x) {
    s_sc[i] = i < p.C2 ? p.scale[i] : 0.f;
    s_sh[i] = i < p.C2 ? p.shift[i] : 0.f;
  }
  fence_proxy_async_smem();
  if (warp == 2) tmem_alloc(&tmem_base_s, (uint32_t)p.tmem_cols);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_trigger();
  const uint32_t tmem = tmem_base_s;
  // u is double buffered: phase A of tile i+1 runs while tile i is in its hidden epilogue / phase B / output epilogue
  const uint32_t tm_y = tmem + 2u * (uint32_t)p.C2;

  const int mine = ((int)blockIdx.x < p.tiles) ? (p.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int tiles_img = p.tiles_w * p.tiles_h;

  if (warp == 0) {
    if (lane == 0 && mine > 0) {
      FB_DBG_DECL
      tma_prefetch_desc(&p.x_map);
      tma_prefetch_desc(&p.tap_map);
      mbar_arrive_expect_tx(&w_bar, (uint32_t)p.w_bytes);
      bulk_copy_g2s(sW, p.wimg, (uint32_t)p.w_bytes, &w_bar);
      const uint32_t box_x = (uint32_t)(64 * p.P * p.Ht * 2), box_t = (uint32_t)(16 * p.P * (p.Ht + 2) * 2);
      for (int i = 0; i < mine; ++i) {
        const int t = (int)blockIdx.x + i * (int)gridDim.x;
        const int n = t / tiles_img, tr = t - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
        const int w0 = tw * p.Wt, h0 = th * p.Ht;
        const int b = p.xbufs == 2 ? (i & 1) : 0;
        const uint32_t use = p.xbufs == 2 ? ((uint32_t)i >> 1) : (uint32_t)i;
        // x buffer b: free once the output epilogue of its previous tile has read the residual
        FB_WAIT(0, &x_free[b], (use & 1u) ^ 1u);
        mbar_arrive_expect_tx(&x_full[b], box_x);
        tma_load_5d(sX + b * kXBytes, &p.x_map, &x_full[b], 0, w0, 0, h0, n);
        // halo buffer i & 1: free once the phase-A MMAs of tile i-2 have completed
        if (i > 1) FB_WAIT(1, &u_full[i & 1], (((uint32_t)i >> 1) - 1u) & 1u);
        mbar_arrive_expect_tx(&taps_full[i & 1], box_t);
        tma_load_5d(sTap + (i & 1) * kHaloBytes, &p.tap_map, &taps_full[i & 1], 0, w0 - 1, 0, h0 - 1, n);
      }
      FB_DBG_FLUSH(0, 2)
    }
  } else if (warp == 1) {
    if (lane == 0 && mine > 0) {
      FB_DBG_DECL
      const uint32_t id1 = umma_idesc_bf16_f32(128, p.C2, 0, 0);
      const uint32_t id2 = umma_idesc_bf16_f32(128, p.N2, 0, 0);
      const uint64_t w1x_d = umma_smem_desc_sw128(smem_u32(sW), 16, 1024);
      const uint64_t w2_d = umma_smem_desc_sw128(smem_u32(sW + p.off_w2), 16, 1024);
      const uint64_t wc_d = desc_sw32(smem_u32(sW + p.off_wc));
      const uint64_t tap_d = desc_sw32(smem_u32(sTap));
      const uint32_t P2 = (uint32_t)(p.P * 2);     // one halo row = 32 bytes = 2 descriptor address units
      const uint64_t v_d = umma_smem_desc_sw128(smem_u32(sV), 16, 1024);
      const uint32_t wc_step = (uint32_t)(p.C2 * 32) >> 4;
      const int kx = (p.C + 15) / 16, kv = (p.C2 + 15) / 16;
      mbar_wait(&w_bar, 0);
      // ---- phase A of tile i: u[i & 1][128 x 2C] = x * W1x^T + sum_taps tap * Wc_tap^T
      auto phase_a = [&](int i) {
        const int b = p.xbufs == 2 ? (i & 1) : 0;
        const uint32_t use = p.xbufs == 2 ? ((uint32_t)i >> 1) : (uint32_t)i;
        const uint32_t tm_u = tmem + (uint32_t)((i & 1) * p.C2);
        FB_WAIT(0, &x_full[b], use & 1u);
        tc_fence_after_sync();
        const uint64_t x_d = umma_smem_desc_sw128(smem_u32(sX + b * kXBytes), 16, 1024);
        for (int k = 0; k < kx; ++k) umma_bf16(tm_u, x_d + (uint64_t)(k * 2), w1x_d + (uint64_t)(k * 2), id1, (uint32_t)(k != 0));
        FB_WAIT(1, &taps_full[i & 1], ((uint32_t)i >> 1) & 1u);
        tc_fence_after_sync();
        const uint64_t halo_d = tap_d + (uint64_t)((i & 1) * (kHaloBytes >> 4));
#pragma unroll
        for (int k = 0; k < 9; ++k)   // tap (dy, dx) = the halo tile read (dy+1)*P + (dx+1) rows further in
          umma_bf16(tm_u, halo_d + (uint64_t)((uint32_t)(k / 3) * P2 + (uint32_t)(k % 3) * 2u),
                    wc_d + (uint64_t)((uint32_t)k * wc_step), id1, 1u);
        umma_commit(&u_full[i & 1]);
      };
      phase_a(0);
      for (int i = 0; i < mine; ++i) {
        const uint32_t ph = (uint32_t)i & 1u;
        // u[(i+1) & 1] was last read by the hidden epilogue of tile i-1, whose v_ready this lane has already seen; the
        // x buffer is handed back at the start of the hidden epilogue, so the look-ahead works with one buffer too.
        if (i + 1 < mine) phase_a(i + 1);
        // ---- phase B: y[128 x N2] = v[128 x 2C] * W2^T
        FB_WAIT(2, &v_ready, ph);
        FB_WAIT(3, &y_free, ph ^ 1u);     // the previous tile's output epilogue has drained y
        tc_fence_after_sync();
        for (int k = 0; k < kv; ++k) {
          const int kc = k >> 2, kk = k & 3;
          umma_bf16(tm_y, v_d + (uint64_t)(kc * (kXBytes >> 4) + kk * 2),
                    w2_d + (uint64_t)(kc * ((p.N2 * 128) >> 4) + kk * 2), id2, (uint32_t)(k != 0));
        }
        umma_commit(&y_full);
      }
      FB_DBG_FLUSH(3, 4)
    }
  } else if (warp < 6) {
    // ---------------- hidden group (warps 2..5): v = relu(scale*u + shift) -> bf16 -> SW128 A operand ----------------
    const int q = warp & 3;
    const int r = q * 32 + lane;                 // tile row = TMEM lane
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    FB_DBG_DECL
    for (int i = 0; i < mine; ++i) {
      FB_WAIT(0, &u_full[i & 1], ((uint32_t)i >> 1) & 1u);
      tc_fence_after_sync();
      // sV is also read by phase B of the previous tile: its completion (y_full) is implied by v_ready ordering — the
      // MMA lane issues phase B(i-1) before phase A(i+1)... but not before A(i); wait for y_full(i-1) explicitly.
      if (i > 0) FB_WAIT(1, &y_full, (uint32_t)(i - 1) & 1u);
      const uint32_t tm_u = tmem + (uint32_t)((i & 1) * p.C2);
      // columns >= C2 of the last K chunk were zeroed once at start.  Up to three 16-column loads per TMEM round trip;
      // scale / shift come from shared memory (staged once per CTA)
      for (int cb = 0; cb < p.C2; cb += 48) {
        uint32_t acc[3][16];
        tmem_ld16(tm_u + lane_addr + (uint32_t)cb, acc[0]);
        if (cb + 16 < p.C2) tmem_ld16(tm_u + lane_addr + (uint32_t)(cb + 16), acc[1]);
        if (cb + 32 < p.C2) tmem_ld16(tm_u + lane_addr + (uint32_t)(cb + 32), acc[2]);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 6; ++g) {
          const int c = cb + g * 8;
          if (c < p.C2) {
            const float4 s0 = *reinterpret_cast<const float4*>(s_sc + c), s1 = *reinterpret_cast<const float4*>(s_sc + c + 4);
            const float4 h0 = *reinterpret_cast<const float4*>(s_sh + c), h1 = *reinterpret_cast<const float4*>(s_sh + c + 4);
            const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
            const float sh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
            uint32_t w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {   // relu + convert + pack in one instruction
              const float lo = fmaf(__uint_as_float(acc[g >> 1][(g & 1) * 8 + 2 * j]), sc[2 * j], sh[2 * j]);
              const float hi = fmaf(__uint_as_float(acc[g >> 1][(g & 1) * 8 + 2 * j + 1]), sc[2 * j + 1], sh[2 * j + 1]);
              asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(w[j]) : "f"(hi), "f"(lo));
            }
            const int kc = c >> 6, ch = (c >> 3) & 7;
            *reinterpret_cast<uint4*>(sV + kc * kXBytes + r * 128 + ((ch ^ (r & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
          }
        }
      }
      tc_fence_before_sync();
      fence_proxy_async_smem();
      mbar_arrive(&v_ready);
    }
    if (threadIdx.x == 64) { FB_DBG_FLUSH(8, 2) }
  } else {
    // ---------------- output group (warps 6..9): y = x + W2 v ----------------
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const int hh = r / p.P, ww = r - hh * p.P;   // rows with ww >= Wt or hh >= Ht are dead
    FB_DBG_DECL
    for (int i = 0; i < mine; ++i) {
      const uint32_t ph = (uint32_t)i & 1u;
      const int b = p.xbufs == 2 ? (i & 1) : 0;
      const int t = (int)blockIdx.x + i * (int)gridDim.x;
      const int n = t / tiles_img, tr = t - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
      // once the phase-A MMAs are done with the x box, keep this row's residual in registers and hand the buffer back:
      // the x box two tiles ahead is then requested a whole tile earlier than "after the output epilogue"
      FB_WAIT(0, &u_full[i & 1], ((uint32_t)i >> 1) & 1u);
      uint4 xres[8];
      {
        const uint8_t* xrow = sX + b * kXBytes + r * 128;
#pragma unroll
        for (int cc = 0; cc < 8; ++cc)
          if (cc * 8 < p.C) xres[cc] = *reinterpret_cast<const uint4*>(xrow + ((cc ^ (r & 7)) << 4));
      }
      mbar_arrive(&x_free[b]);
      FB_WAIT(1, &y_full, ph);
      tc_fence_after_sync();
      const int h = th * p.Ht + hh, w = tw * p.Wt + ww;
      const bool valid = ww < p.Wt && hh < p.Ht && h < p.H && w < p.W;
      const long pix = ((long)n * p.H + h) * p.W + w;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        if (c0 >= p.N2) break;
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
          unpack8(xres[(c0 >> 3) + g], xr);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(acc[g * 8 + j]) + xr[j];
          *reinterpret_cast<uint4*>(p.y + pix * p.C + c) = pack8(v);
        }
      }
      tc_fence_before_sync();
      mbar_arrive(&y_free);
    }
    if (threadIdx.x == 192) { FB_DBG_FLUSH(11, 2) }
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

// ---------------------------------------------------------------------------------------------
// C = 80 (FasterNet stage 3: 40 x 40 maps at 640 input).  Same two-phase scheme with everything compile-time; the
// per-layer path spent three launches of ~6 tiles per SM on this block, most of it start-up and hand-off latency.
// Cp = 20 channels feed the 3x3, so the folded taps are two K=16 slices; channels 64..79 of x are a third K=16 (SW32)
// box.  Shared memory is the limit (weights 145 KB + hidden tile 48 KB): inputs are single buffered, pitch is fixed at
// P = 16 (tiles of 14 x 8 pixels), and the output group takes its residual from global memory (an L2 hit) instead of
// keeping the x box alive.
// ---------------------------------------------------------------------------------------------
namespace f80 {
constexpr int C = 80, C2 = 160, N2 = 80, P = 16, Wt = 14, Ht = 8;
constexpr int kOffW2 = 160 * 128;                  // W1x chunk 0 [160][64] SW128
constexpr int kOffW1b = kOffW2 + 3 * N2 * 128;     // W2: 3 chunks [80][64] SW128
constexpr int kOffWc = kOffW1b + 160 * 32;         // W1x channels 64..79 [160][16] SW32
constexpr int kWBytes = kOffWc + 18 * 160 * 32;    // Wc [slice][tap][160][16] SW32
constexpr int kHalo = 5632;                        // 160 box rows x 32 B + the two rows a shifted read runs past
constexpr int kSV = kXBytes;                       // after sX0
constexpr int kSW = kSV + 3 * kXBytes;
constexpr int kSX1 = kSW + kWBytes;
constexpr int kSHalo = kSX1 + 128 * 32;
constexpr int kSmem = kSHalo + 2 * kHalo + 1024;
constexpr uint32_t kTxBytes = 64 * P * Ht * 2 + 16 * P * Ht * 2 + 2 * 16 * P * (Ht + 2) * 2;
}  // namespace f80

struct alignas(64) Fnb80Params {
  CUtensorMap x_map;     // (C, W, 1, H, N) box (64, P, 1, Ht, 1)   SW128
  CUtensorMap x1_map;    // same tensor, box (16, P, 1, Ht, 1)      SW32
  CUtensorMap tap_map;   // same tensor, box (16, P, 1, Ht+2, 1)    SW32
  const __nv_bfloat16* x;
  __nv_bfloat16* y;
  const uint8_t* wimg;
  const float* scale;
  const float* shift;
  int H, W, tiles_w, tiles_h, tiles;
};

__global__ void __launch_bounds__(320, 1) fasternet_block80_kernel(const __grid_constant__ Fnb80Params p) {
  using namespace f80;
  FB_STAMP(14)
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t w_bar, in_full, u_full[2], v_ready, y_full, y_free;
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_sc[C2], s_sh[C2];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sX0 = smem;
  uint8_t* sV = smem + kSV;
  uint8_t* sW = smem + kSW;
  uint8_t* sX1 = smem + kSX1;
  uint8_t* sHalo = smem + kSHalo;

  if (warp == 1 && lane == 0) {
    mbar_init(&w_bar, 1);
    mbar_init(&in_full, 1);
    mbar_init(&u_full[0], 1);
    mbar_init(&u_full[1], 1);
    mbar_init(&v_ready, 128);
    mbar_init(&y_full, 1);
    mbar_init(&y_free, 128);
    fence_mbar_init();
  }
  // K padding of the hidden tile (columns 160..191 of chunk 2) and the halo overrun rows: zeroed once
  for (int i = threadIdx.x; i < 3 * kXBytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(sV)[i] = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < 2 * kHalo / 16; i += blockDim.x) reinterpret_cast<uint4*>(sHalo)[i] = make_uint4(0, 0, 0, 0);
  pdl_wait();
  for (int i = threadIdx.x; i < C2; i += blockDim.x) { s_sc[i] = p.scale[i]; s_sh[i] = p.shift[i]; }
  fence_proxy_async_smem();
  if (warp == 2) tmem_alloc(&tmem_base_s, 512u);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_trigger();
  const uint32_t tmem = tmem_base_s;
  const uint32_t tm_y = tmem + 2u * C2;

  const int mine = ((int)blockIdx.x < p.tiles) ? (p.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int tiles_img = p.tiles_w * p.tiles_h;

  if (warp == 0) {
    if (lane == 0 && mine > 0) {
      FB_DBG_DECL
      tma_prefetch_desc(&p.x_map);
      tma_prefetch_desc(&p.x1_map);
      tma_prefetch_desc(&p.tap_map);
      mbar_arrive_expect_tx(&w_bar, (uint32_t)kWBytes);
      bulk_copy_g2s(sW, p.wimg, (uint32_t)kWBytes, &w_bar);
      for (int i = 0; i < mine; ++i) {
        const int t = (int)blockIdx.x + i * (int)gridDim.x;
        const int n = t / tiles_img, tr = t - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
        const int w0 = tw * Wt, h0 = th * Ht;
        // single input buffer: free once the phase-A MMAs of the previous tile have completed
        if (i > 0) FB_WAIT(1, &u_full[(i - 1) & 1], ((uint32_t)(i - 1) >> 1) & 1u);
        mbar_arrive_expect_tx(&in_full, kTxBytes);
        tma_load_5d(sX0, &p.x_map, &in_full, 0, w0, 0, h0, n);
        tma_load_5d(sHalo, &p.tap_map, &in_full, 0, w0 - 1, 0, h0 - 1, n);
        tma_load_5d(sHalo + kHalo, &p.tap_map, &in_full, 16, w0 - 1, 0, h0 - 1, n);
        tma_load_5d(sX1, &p.x1_map, &in_full, 64, w0, 0, h0, n);
      }
      FB_DBG_FLUSH(0, 2)
    }
  } else if (warp == 1) {
    if (lane == 0 && mine > 0) {
      FB_DBG_DECL
      const uint32_t id1 = umma_idesc_bf16_f32(128, C2, 0, 0);
      const uint32_t id2 = umma_idesc_bf16_f32(128, N2, 0, 0);
      const uint64_t x0_d = umma_smem_desc_sw128(smem_u32(sX0), 16, 1024);
      const uint64_t w1a_d = umma_smem_desc_sw128(smem_u32(sW), 16, 1024);
      const uint64_t w2_d = umma_smem_desc_sw128(smem_u32(sW + kOffW2), 16, 1024);
      const uint64_t x1_d = desc_sw32(smem_u32(sX1));
      const uint64_t w1b_d = desc_sw32(smem_u32(sW + kOffW1b));
      const uint64_t wc_d = desc_sw32(smem_u32(sW + kOffWc));
      const uint64_t halo_d = desc_sw32(smem_u32(sHalo));
      const uint64_t v_d = umma_smem_desc_sw128(smem_u32(sV), 16, 1024);
      mbar_wait(&w_bar, 0);
      auto phase_a = [&](int i) {
        const uint32_t tm_u = tmem + (uint32_t)((i & 1) * C2);
        FB_WAIT(0, &in_full, (uint32_t)i & 1u);
        tc_fence_after_sync();
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16(tm_u, x0_d + (uint64_t)(k * 2), w1a_d + (uint64_t)(k * 2), id1, (uint32_t)(k != 0));
        umma_bf16(tm_u, x1_d, w1b_d, id1, 1u);
#pragma unroll
        for (int sl = 0; sl < 2; ++sl)
#pragma unroll
          for (int k = 0; k < 9; ++k)
            umma_bf16(tm_u, halo_d + (uint64_t)(sl * (kHalo >> 4) + ((k / 3) * P + (k % 3)) * 2),
                      wc_d + (uint64_t)((sl * 9 + k) * ((C2 * 32) >> 4)), id1, 1u);
        umma_commit(&u_full[i & 1]);
      };
      phase_a(0);
      for (int i = 0; i < mine; ++i) {
        const uint32_t ph = (uint32_t)i & 1u;
        if (i + 1 < mine) phase_a(i + 1);
        FB_WAIT(2, &v_ready, ph);
        FB_WAIT(3, &y_free, ph ^ 1u);
        tc_fence_after_sync();
#pragma unroll
        for (int k = 0; k < C2 / 16; ++k)
          umma_bf16(tm_y, v_d + (uint64_t)((k >> 2) * (kXBytes >> 4) + (k & 3) * 2),
                    w2_d + (uint64_t)((k >> 2) * ((N2 * 128) >> 4) + (k & 3) * 2), id2, (uint32_t)(k != 0));
        umma_commit(&y_full);
      }
      FB_DBG_FLUSH(3, 4)
    }
  } else if (warp < 6) {
    // hidden group: v = relu(scale*u + shift) -> bf16 -> SW128 A operand of phase B
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    FB_DBG_DECL
    for (int i = 0; i < mine; ++i) {
      FB_WAIT(0, &u_full[i & 1], ((uint32_t)i >> 1) & 1u);
      tc_fence_after_sync();
      const uint32_t tm_u = tmem + (uint32_t)((i & 1) * C2) + lane_addr;
      // two TMEM round trips of 80 columns; the first half is converted before sV is known to be free (phase B of the
      // previous tile still reads it), so that wait hides behind the math
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t acc[5][16];
#pragma unroll
        for (int b = 0; b < 5; ++b) tmem_ld16(tm_u + (uint32_t)(half * 80 + b * 16), acc[b]);
        tmem_ld_wait();
        uint32_t w[10][4];
#pragma unroll
        for (int g = 0; g < 10; ++g) {
          const int c = half * 80 + g * 8;
          const float4 s0 = *reinterpret_cast<const float4*>(s_sc + c), s1 = *reinterpret_cast<const float4*>(s_sc + c + 4);
          const float4 h0 = *reinterpret_cast<const float4*>(s_sh + c), h1 = *reinterpret_cast<const float4*>(s_sh + c + 4);
          const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
          const float sh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float lo = fmaf(__uint_as_float(acc[g >> 1][(g & 1) * 8 + 2 * j]), sc[2 * j], sh[2 * j]);
            const float hi = fmaf(__uint_as_float(acc[g >> 1][(g & 1) * 8 + 2 * j + 1]), sc[2 * j + 1], sh[2 * j + 1]);
            asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(w[g][j]) : "f"(hi), "f"(lo));
          }
        }
        if (half == 0 && i > 0) FB_WAIT(1, &y_full, (uint32_t)(i - 1) & 1u);   // phase B of the previous tile has read sV
#pragma unroll
        for (int g = 0; g < 10; ++g) {
          const int c = half * 80 + g * 8;
          const int kc = c >> 6, ch = (c >> 3) & 7;
          *reinterpret_cast<uint4*>(sV + kc * kXBytes + r * 128 + ((ch ^ (r & 7)) << 4)) = make_uint4(w[g][0], w[g][1], w[g][2], w[g][3]);
        }
      }
      tc_fence_before_sync();
      fence_proxy_async_smem();
      mbar_arrive(&v_ready);
    }
    if (threadIdx.x == 64) { FB_DBG_FLUSH(8, 2) }
  } else {
    // output group: y = x + W2 v
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const int hh = r / P, ww = r - hh * P;
    FB_DBG_DECL
    for (int i = 0; i < mine; ++i) {
      const uint32_t ph = (uint32_t)i & 1u;
      const int t = (int)blockIdx.x + i * (int)gridDim.x;
      const int n = t / tiles_img, tr = t - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
      const int h = th * Ht + hh, w = tw * Wt + ww;
      const bool valid = ww < Wt && h < p.H && w < p.W;
      const long pix = ((long)n * p.H + h) * p.W + w;
      uint4 xres[C / 8];
      if (valid) {
#pragma unroll
        for (int k = 0; k < C / 8; ++k) xres[k] = __ldg(reinterpret_cast<const uint4*>(p.x + pix * C) + k);
      }
      FB_WAIT(1, &y_full, ph);
      tc_fence_after_sync();
#pragma unroll
      for (int c0 = 0; c0 < N2; c0 += 16) {
        uint32_t acc[16];
        tmem_ld16(tm_y + lane_addr + (uint32_t)c0, acc);
        tmem_ld_wait();
        if (valid) {
          uint32_t o[8];
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            float xr[8], v[8];
            unpack8(xres[(c0 >> 3) + g], xr);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(acc[g * 8 + j]) + xr[j];
            const uint4 pk = pack8(v);
            o[g * 4 + 0] = pk.x; o[g * 4 + 1] = pk.y; o[g * 4 + 2] = pk.z; o[g * 4 + 3] = pk.w;
          }
          asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p.y + pix * C + c0), "r"(o[0]),
                       "r"(o[1]), "r"(o[2]), "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7])
                       : "memory");
        }
      }
      tc_fence_before_sync();
      mbar_arrive(&y_free);
    }
    if (threadIdx.x == 192) { FB_DBG_FLUSH(11, 2) }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512u);
  FB_STAMP(15)
}

}  // namespace ly

using namespace ly;


static int launch_fnb80(const void* x, void* y, int N, int H, int W, const void* wimg, int w_bytes, int off_w2,
                        int off_wc, const float* scale, const float* shift, ly_stream_t st) {
  using namespace f80;
  LY_REQUIRE(w_bytes == kWBytes && off_w2 == kOffW2 && off_wc == kOffWc,
             "ly_fasternet_block_infer: C=80 weight image layout mismatch (%d bytes, offsets %d / %d)", w_bytes, off_w2,
             off_wc);
  Fnb80Params p;
  memset(&p, 0, sizeof(p));
  p.x = (const __nv_bfloat16*)x;
  p.y = (__nv_bfloat16*)y;
  p.wimg = (const uint8_t*)wimg;
  p.scale = scale;
  p.shift = shift;
  p.H = H;
  p.W = W;
  p.tiles_w = ceil_div(W, Wt);
  p.tiles_h = ceil_div(H, Ht);
  p.tiles = p.tiles_w * p.tiles_h * N;
  uint64_t dims[5] = {(uint64_t)C, (uint64_t)W, 1, (uint64_t)H, (uint64_t)N};
  uint64_t str[4] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
  uint32_t box[5] = {64, P, 1, Ht, 1};
  int rc = tmap_encode_bf16(&p.x_map, x, 5, dims, str, box);
  if (rc) return rc;
  uint32_t box1[5] = {16, P, 1, Ht, 1};
  rc = encode_sw32(&p.x1_map, x, dims, str, box1);
  if (rc) return rc;
  uint64_t dims2[5] = {32, (uint64_t)W, 1, (uint64_t)H, (uint64_t)N};   // the taps see channels 0..31 only
  uint32_t box2[5] = {16, P, 1, Ht + 2, 1};
  rc = encode_sw32(&p.tap_map, x, dims2, str, box2);
  if (rc) return rc;
  static std::atomic<bool> attr_done[64];   // zero-initialised; set only after the attribute call succeeded
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(fasternet_block80_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    attr_done[dev & 63].store(true);
  }
  const int sms = device_sm_count();
  const unsigned grid = (unsigned)(p.tiles < sms ? p.tiles : sms);
  LY_CUDA(launch(fasternet_block80_kernel, grid, 320, kSmem, (cudaStream_t)st, p));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// x, y: dense NHWC bf16 [N][H][W][C], C in {24, 40, 80}.  wimg: pre-swizzled weight image built by
// lead_yolo_b200.packing.fasternet_weight_image (sections W1x | W2 | Wc, offsets passed in).
extern "C" int ly_fasternet_block_infer(const void* x, void* y, int N, int H, int W, int C, const void* wimg,
                                        int w_bytes, int off_w2, int off_wc, const float* scale, const float* shift,
                                        ly_stream_t st) {
  LY_REQUIRE(x && y && wimg && scale && shift && N > 0 && H > 0 && W > 0, "ly_fasternet_block_infer: bad args");
  LY_REQUIRE(C == 24 || C == 40 || C == 80, "ly_fasternet_block_infer: C=%d not supported (24, 40 or 80)", C);
  LY_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0 &&
                 reinterpret_cast<uintptr_t>(wimg) % 16 == 0 && w_bytes % 16 == 0,
             "ly_fasternet_block_infer: alignment");
  if (C == 80) {
    LY_REQUIRE(reinterpret_cast<uintptr_t>(y) % 32 == 0, "ly_fasternet_block_infer: y must be 32-byte aligned for C=80");
    return launch_fnb80(x, y, N, H, W, wimg, w_bytes, off_w2, off_wc, scale, shift, st);
  }
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
  p.off_wc = off_wc;
  p.v_chunks = (p.C2 + 63) / 64;
  p.tmem_cols = 32;
  while (p.tmem_cols < 2 * p.C2 + p.N2) p.tmem_cols *= 2;
  LY_REQUIRE(p.tmem_cols <= 256, "ly_fasternet_block_infer: accumulators do not fit 256 TMEM columns");
  LY_REQUIRE(off_w2 % 1024 == 0 && off_wc % 256 == 0, "ly_fasternet_block_infer: weight image sections misaligned");
  // spatial tile: Ht rows of pitch P = Wt + 2 inside the 128 accumulator rows
  int best_wt = 14, best_ht = 8;
  double best = -1;
  const int pitches[3] = {16, 32, 64};
  for (int i = 0; i < 3; ++i) {
    const int wt = pitches[i] - 2, ht = 128 / pitches[i];
    const double eff = (double)H * W / ((double)ceil_div(W, wt) * ceil_div(H, ht) * 128.0);
    if (eff > best + 1e-9) { best = eff; best_wt = wt; best_ht = ht; }
  }
  p.Wt = best_wt;
  p.Ht = best_ht;
  p.P = best_wt + 2;
  p.tiles_w = ceil_div(W, p.Wt);
  p.tiles_h = ceil_div(H, p.Ht);
  {
    uint64_t dims[5] = {(uint64_t)C, (uint64_t)W, 1, (uint64_t)H, (uint64_t)N};
    uint64_t str[4] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
    uint32_t box[5] = {64, (uint32_t)p.P, 1, (uint32_t)p.Ht, 1};
    int rc = tmap_encode_bf16(&p.x_map, x, 5, dims, str, box);
    if (rc) return rc;
    uint64_t dims2[5] = {16, (uint64_t)W, 1, (uint64_t)H, (uint64_t)N};
    uint32_t box2[5] = {16, (uint32_t)p.P, 1, (uint32_t)(p.Ht + 2), 1};
    rc = encode_sw32(&p.tap_map, x, dims2, str, box2);
    if (rc) return rc;
  }
  // two x buffers (prefetch across tiles) when two CTAs still fit an SM with them
  const size_t fixed = (size_t)2 * kHaloBytes + (size_t)p.v_chunks * kXBytes + (size_t)w_bytes + 1024;
  p.xbufs = (fixed + 2 * kXBytes <= 112 * 1024) ? 2 : 1;
  const size_t smem = fixed + (size_t)p.xbufs * kXBytes;
  static std::atomic<bool> attr_done[64];   // zero-initialised; set only after the attribute call succeeded
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(fasternet_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_done[dev & 63].store(true);
  }
  LY_REQUIRE(smem <= 200 * 1024, "ly_fasternet_block_infer: weight image too large (%d bytes)", w_bytes);
  p.tiles = p.tiles_w * p.tiles_h * N;
  const int per_sm = smem <= 112 * 1024 ? 2 : 1;
  const long cap = (long)per_sm * device_sm_count();
  const unsigned grid = (unsigned)(p.tiles < cap ? p.tiles : cap);
  LY_CUDA(launch(fasternet_block_kernel, grid, 320, smem, (cudaStream_t)st, p));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
