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
//   warps 2..17 epilogue: TMEM lane quadrant = warp % 4; four groups take t = {0,1,2}, {3,4}, {5,6}, {7,8}
//   warp 18     rewrites the diagonal of the (double-buffered) weight tiles for the next item
// Two passes over x, neither stores the ungated features (the reference materialises them three times):
//   STATS  g = relu(acc + shift); max / mean over channels per (pixel, t) kept in registers across the channel groups
//          of a pixel tile and merged into the attention input maps once per tile (features are >= 0, so an integer
//          atomicMax on the float bits is exact)
//   APPLY  A[m][(c/16)*144 + t*16 + c%16] = g * se[n][c] * att[n][3h + t/3][3w + t%3] -> the GEMM operand of
//          RFCBAMConv.conv, staged in shared memory and written back by TMA (288 contiguous bytes per pixel and item)
#include <string.h>

#include <atomic>

#include "ly_common.cuh"
#include "ly_tc.cuh"

namespace ly {

// Per-role wait accounting for tests/dev/rf_tc_probe.cu; compiled out of the library.
#ifdef RF_TC_DEBUG
__device__ long long rf_tc_dbg[148][16];
#define RT_WAIT(slot, bar, par)                    \
  do {                                             \
    const long long t__ = clock64();               \
    mbar_wait(bar, par);                           \
    dbg_acc[slot] += clock64() - t__;              \
  } while (0)
#define RT_DBG_DECL long long dbg_acc[4] = {0, 0, 0, 0}; const long long dbg_t0 = clock64();
#define RT_DBG_FLUSH(base, nslots)                                                                        \
  if (lane == 0 && blockIdx.x < 148) {                                                                    \
    for (int z = 0; z < nslots; ++z) rf_tc_dbg[blockIdx.x][base + z] = dbg_acc[z];                        \
    rf_tc_dbg[blockIdx.x][base + nslots] = clock64() - dbg_t0;                                            \
  }
#else
#define RT_WAIT(slot, bar, par) mbar_wait(bar, par)
#define RT_DBG_DECL
#define RT_DBG_FLUSH(base, nslots)
#endif

constexpr int kRtStagesMax = 6;       // one stage = the 3 tap boxes of a kernel row, filled under ONE mbarrier phase:
                                      // a lone producer lane pays ~250 cycles per barrier round trip but only ~128 per
                                      // extra box (tests/dev/tma_issue.cu), and the TMA unit itself moves 1 box row / cycle
constexpr int kRtA = 128 * 32;        // tap box: 128 pixels x 16 channels, SW32
constexpr int kRtB = 144 * 32;        // weight tile of one tap: 144 outputs x 16 inputs, SW32
constexpr int kRtB1 = 10 * kRtB;      // 9 taps + the shift tile of one channel group
constexpr int kRtStage = 128 * 288;   // APPLY: staged output of one item, [pixel][9 t][16] bf16, written back by TMA
constexpr int kRtThreads = 608;

struct alignas(64) RfTcParams {
  // s = 1: a_map[0] = x as (C, W, 1, H, N).  s = 2: one map per input parity (ph, pw) = x[:, ph::2, pw::2, :], so a tap
  // is a dense box of its parity plane (a strided TMA traversal would walk 4x the rows).  Box (16, Wt, 1, Ht, 1), SW32.
  CUtensorMap a_map[4];
  CUtensorMap o_map;   // APPLY: A as (9C, wo, ho, N), box (144, Wt, Ht, 1), no swizzle
  int tiles_w, tiles_h, n_img, groups;
  int Wt, Ht, ho, wo, C, s;
  int items, per_cta;
  int tap_slices;           // 1: the nine "taps" are the nine slices of dimension 2 of a_map[0] at the SAME pixels
                            // (input-gradient pass: A = dZ[m][t][16 channels], t = 0..8), not spatial neighbours
  uint32_t a_bytes;
  const __nv_bfloat16* wc;  // [C/16][9 taps][144]: the diagonal values, BatchNorm scale folded in
  const uint32_t* shift;    // [C/16][144]: BatchNorm shift as two bf16 (hi | lo << 16), hi + lo = shift to 16 bits
  const float* se;          // [N][C]                 (APPLY)
  const float* att;         // [N][ho*3][wo*3]        (APPLY)
  __nv_bfloat16* A;         // [M][9*C], column cg*144 + t*16 + c%16   (APPLY)
  float* maps;              // [N][ho*3][wo*3][2]     (STATS; zero on entry)
};

__host__ __device__ constexpr uint64_t rt_desc_sw32(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) |
         (6ull << 61);
}

__device__ __forceinline__ void st_global_256(void* ptr, const uint32_t (&w)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "r"(w[0]), "r"(w[1]), "r"(w[2]),
               "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}
// packed fp32 pairs (FMUL2 / FADD2) and relu + convert + pack in one instruction: the epilogue is issue-bound
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_mul(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint32_t relu_bf16x2(uint64_t v) {   // bf16(relu(lo)) | bf16(relu(hi)) << 16
  float lo, hi;
  f2_unpack(v, lo, hi);
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

template <int MODE>  // 0 = STATS, 1 = APPLY, 2 = RAW (training forward: generate + shift as bf16 [M][9][C], no gate)
__global__ void __launch_bounds__(kRtThreads, 1) rf_tc_kernel(const __grid_constant__ RfTcParams p) {
  constexpr int kRtStages = MODE == 1 ? 4 : 6;   // APPLY trades two ring stages for the output staging buffers
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[kRtStages], empty_bar[kRtStages], tfull[2], tempty[2], b1_full[2], b1_empty[2];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;                                  // ring: kRtStages x 3 tap boxes
  uint8_t* sOnes = smem + kRtStages * 3 * kRtA;        // [128][16] with columns 0 and 1 = 1.0: A operand of the shift MMA
  uint8_t* sB = sOnes + kRtA;                          // 2 x kRtB1
  uint8_t* sStage = sB + 2 * kRtB1;                    // APPLY: 2 x kRtStage

  // weight tiles: the off-diagonal entries stay zero for the whole kernel
  for (int i = threadIdx.x; i < (kRtA + 2 * kRtB1) / 16; i += kRtThreads)
    reinterpret_cast<uint4*>(sOnes)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  if (threadIdx.x < 128)   // columns 0,1 live in 16-byte chunk 0, stored at chunk (r >> 2) & 1 of row r
    *reinterpret_cast<uint32_t*>(sOnes + threadIdx.x * 32 + ((threadIdx.x >> 2) & 1) * 16) = 0x3F803F80u;
  fence_proxy_async_smem();
  if (warp == 0 && lane < (p.s == 2 ? 4 : 1)) tma_prefetch_desc(&p.a_map[lane]);
  if (MODE == 1 && warp == 2 && lane == 0) tma_prefetch_desc(&p.o_map);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kRtStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull[b], 1); mbar_init(&tempty[b], 16);
      mbar_init(&b1_full[b], 1); mbar_init(&b1_empty[b], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_s, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_trigger();
  pdl_wait();
  const uint32_t tmem = tmem_base_s;

  const int first = (int)blockIdx.x * p.per_cta;
  const int mine = min(p.per_cta, p.items - first);
  const int tiles_img = p.tiles_w * p.tiles_h;

  if (warp == 0) {
    if (lane == 0) {
      RT_DBG_DECL
      int cg = first % p.groups, mt = first / p.groups;
      uint32_t s = 0, ph = 1;   // ring slot and the parity to wait for on its empty barrier
      for (int i = 0; i < mine; ++i) {
        const int n = mt / tiles_img, tr = mt - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
        const int c0 = cg * 16;
        for (int ti = 0; ti < 3; ++ti) {
          RT_WAIT(0, &empty_bar[s], ph);
          mbar_arrive_expect_tx(&full_bar[s], 3u * p.a_bytes);
          uint8_t* dst = sA + s * 3 * kRtA;
          if (p.tap_slices) {
            const int row = th * p.Ht, col = tw * p.Wt;
            tma_load_5d(dst, &p.a_map[0], &full_bar[s], c0, col, ti * 3, row, n);
            tma_load_5d(dst + kRtA, &p.a_map[0], &full_bar[s], c0, col, ti * 3 + 1, row, n);
            tma_load_5d(dst + 2 * kRtA, &p.a_map[0], &full_bar[s], c0, col, ti * 3 + 2, row, n);
          } else if (p.s == 2) {   // input row 2h + ti - 1: parity (ti + 1) & 1, plane row h - (ti == 0)
            const int row = th * p.Ht - (ti == 0), col = tw * p.Wt;
            const CUtensorMap* m0 = &p.a_map[((ti + 1) & 1) * 2];
            tma_load_5d(dst, m0 + 1, &full_bar[s], c0, col - 1, 0, row, n);
            tma_load_5d(dst + kRtA, m0, &full_bar[s], c0, col, 0, row, n);
            tma_load_5d(dst + 2 * kRtA, m0 + 1, &full_bar[s], c0, col, 0, row, n);
          } else {
            const int row = th * p.Ht + ti - 1, col = tw * p.Wt;
            tma_load_5d(dst, &p.a_map[0], &full_bar[s], c0, col - 1, 0, row, n);
            tma_load_5d(dst + kRtA, &p.a_map[0], &full_bar[s], c0, col, 0, row, n);
            tma_load_5d(dst + 2 * kRtA, &p.a_map[0], &full_bar[s], c0, col + 1, 0, row, n);
          }
          if (++s == kRtStages) { s = 0; ph ^= 1u; }
        }
        if (++cg == p.groups) { cg = 0; ++mt; }
      }
      RT_DBG_FLUSH(0, 1)
    }
  } else if (warp == 1) {
    if (lane == 0) {
      RT_DBG_DECL
      const uint32_t idesc = umma_idesc_bf16_f32(128, 144, 0, 0);
      const uint64_t a_desc0 = rt_desc_sw32(smem_u32(sA)), b_desc0 = rt_desc_sw32(smem_u32(sB));
      const uint64_t ones_desc = rt_desc_sw32(smem_u32(sOnes));
      uint32_t s = 0, ph = 0;
      for (int i = 0; i < mine; ++i) {
        const int buf = i & 1;
        const uint32_t bph = ((uint32_t)i >> 1) & 1u;
        const uint32_t d_tmem = tmem + (uint32_t)(buf * 256);
        const uint64_t b_desc = b_desc0 + (uint64_t)(buf * (kRtB1 >> 4));
        RT_WAIT(0, &tempty[buf], bph ^ 1u);
        RT_WAIT(1, &b1_full[buf], bph);
        tc_fence_after_sync();
        umma_bf16(d_tmem, ones_desc, b_desc + (uint64_t)(9 * (kRtB >> 4)), idesc, 0u);   // acc = shift
        for (int ti = 0; ti < 3; ++ti) {
          RT_WAIT(2, &full_bar[s], ph);
          tc_fence_after_sync();
          const uint64_t a_desc = a_desc0 + (uint64_t)(s * (3 * kRtA >> 4));
#pragma unroll
          for (int tj = 0; tj < 3; ++tj)
            umma_bf16(d_tmem, a_desc + (uint64_t)(tj * (kRtA >> 4)), b_desc + (uint64_t)((ti * 3 + tj) * (kRtB >> 4)), idesc,
                      1u);
          umma_commit(&empty_bar[s]);
          if (++s == kRtStages) { s = 0; ph ^= 1u; }
        }
        umma_commit(&tfull[buf]);
        umma_commit(&b1_empty[buf]);
      }
      RT_DBG_FLUSH(2, 3)
    }
  } else if (warp == 18) {
    RT_DBG_DECL
    // 1296 diagonal values per channel group: value (tap k, output n) goes to row n, column n & 15 of tile k, i.e. byte
    //   k*kRtB + n*32 + 16*(((n >> 3) ^ (n >> 2)) & 1) + 2*(n & 7).  Its shared-memory bank depends only on n % 16 (16
    // distinct banks), so lane l takes n = l, l+32, ... of every tap: 2-way conflicts instead of the 32-way ones of a
    // "8 consecutive n per lane" split.  Shift tile: row n, columns 0,1 = (hi, lo): 4 bytes at 9*kRtB + n*32 + 16*((n>>2)&1).
    // The next group's values are fetched before waiting for the buffer.
    uint32_t off[5], soff[5];
#pragma unroll
    for (int m = 0; m < 5; ++m) {
      const int n = lane + 32 * m;
      off[m] = smem_u32(sB) + (uint32_t)(n * 32 + ((((n >> 3) ^ (n >> 2)) & 1) << 4) + (n & 7) * 2);
      soff[m] = smem_u32(sB) + (uint32_t)(9 * kRtB + n * 32 + ((n >> 2) & 1) * 16);
    }
    uint16_t v[9][5];
    uint32_t sh[5];
    auto fetch = [&](int cg) {
      const uint16_t* src = reinterpret_cast<const uint16_t*>(p.wc) + (long)cg * 1296 + lane;
#pragma unroll
      for (int k = 0; k < 9; ++k)
#pragma unroll
        for (int m = 0; m < 5; ++m) v[k][m] = (m < 4 || lane < 16) ? __ldg(src + k * 144 + 32 * m) : (uint16_t)0;
#pragma unroll
      for (int m = 0; m < 5; ++m) sh[m] = (m < 4 || lane < 16) ? __ldg(p.shift + cg * 144 + lane + 32 * m) : 0u;
    };
    int cg = first % p.groups;
    fetch(cg);
    for (int i = 0; i < mine; ++i) {
      const int buf = i & 1;
      RT_WAIT(0, &b1_empty[buf], (((uint32_t)i >> 1) & 1u) ^ 1u);
#pragma unroll
      for (int k = 0; k < 9; ++k)
#pragma unroll
        for (int m = 0; m < 5; ++m)
          if (m < 4 || lane < 16)
            asm volatile("st.shared.b16 [%0], %1;" ::"r"(off[m] + (uint32_t)(buf * kRtB1 + k * kRtB)), "h"(v[k][m]) : "memory");
#pragma unroll
      for (int m = 0; m < 5; ++m)
        if (m < 4 || lane < 16)
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(soff[m] + (uint32_t)(buf * kRtB1)), "r"(sh[m]) : "memory");
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&b1_full[buf]);
      if (++cg == p.groups) cg = 0;
      if (i + 1 < mine) fetch(cg);
    }
    RT_DBG_FLUSH(6, 1)
  } else {
    RT_DBG_DECL
    const int q = warp & 3;
    const int eg = (warp - 2) >> 2;                 // epilogue group: t = 0..2 | 3,4 | 5,6 | 7,8
    const int t0 = eg == 0 ? 0 : 1 + 2 * eg, nt = eg == 0 ? 3 : 2;
    const int r = q * 32 + lane;
    const int hh = r / p.Wt, ww = r - hh * p.Wt;
    const float invC = 1.f / (float)p.C;
    float va[3], vb[3];   // STATS: running max / sum;  APPLY: va = attention values
    int cg = first % p.groups, mt = first / p.groups;
    int n = 0, h = 0, w = 0;
    bool valid = false, fresh = true;
    const float* se_row = nullptr;
    int tile_w0 = 0, tile_h0 = 0;
    for (int i = 0; i < mine; ++i) {
      const int buf = i & 1;
      if (fresh) {
        fresh = false;
        n = mt / tiles_img;
        const int tr = mt - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
        tile_w0 = tw * p.Wt; tile_h0 = th * p.Ht;
        h = tile_h0 + hh; w = tile_w0 + ww;
        valid = (r < p.Wt * p.Ht) && h < p.ho && w < p.wo;
#pragma unroll
        for (int tt = 0; tt < 3; ++tt) {
          va[tt] = 0.f; vb[tt] = 0.f;
          if (MODE == 1 && valid && tt < nt) {
            const int t = t0 + tt;
            va[tt] = __ldg(p.att + ((long)n * (p.ho * 3) + (h * 3 + t / 3)) * (long)(p.wo * 3) + (w * 3 + t % 3));
          }
        }
        if (MODE == 1) se_row = p.se + (long)n * p.C;
      }
      uint64_t se2[8];
      if (MODE == 1) {
        const float4* sp = reinterpret_cast<const float4*>(se_row + cg * 16);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 e = __ldg(sp + j);
          se2[2 * j] = f2_pack(e.x, e.y); se2[2 * j + 1] = f2_pack(e.z, e.w);
        }
      }
      RT_WAIT(0, &tfull[buf], ((uint32_t)i >> 1) & 1u);
      tc_fence_after_sync();
      const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * 256 + t0 * 16);
      uint32_t acc[3][16];
      tmem_ld16(trow, acc[0]);
      tmem_ld16(trow + 16u, acc[1]);
      if (nt == 3) tmem_ld16(trow + 32u, acc[2]);
      tmem_ld_wait();
      // the accumulator is in registers: hand the TMEM buffer back before doing the arithmetic
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[buf]);
      uint8_t* stg = sStage + buf * kRtStage;
#pragma unroll
      for (int tt = 0; tt < 3; ++tt) {
        if (tt < nt) {
          if (MODE == 0) {   // acc = generate + shift; relu is implied for the max (running max starts at 0)
            float mx = va[tt];
            uint64_t s2 = f2_pack(0.f, 0.f);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float x0 = __uint_as_float(acc[tt][2 * j]), x1 = __uint_as_float(acc[tt][2 * j + 1]);
              mx = fmaxf(fmaxf(mx, x0), x1);
              s2 = f2_add(s2, f2_pack(fmaxf(x0, 0.f), fmaxf(x1, 0.f)));
            }
            float s_lo, s_hi;
            f2_unpack(s2, s_lo, s_hi);
            va[tt] = mx; vb[tt] += s_lo + s_hi;
          } else if (MODE == 2) {   // one full 32-byte sector per lane: Z[m][t][cg*16 .. +16]
            if (valid) {
              uint32_t wds[8];
#pragma unroll
              for (int j = 0; j < 8; ++j)
                wds[j] = pack_bf16x2(__uint_as_float(acc[tt][2 * j]), __uint_as_float(acc[tt][2 * j + 1]));
              const long m = ((long)n * p.ho + h) * p.wo + w;
              st_global_256(p.A + (m * 9 + (t0 + tt)) * (long)p.C + cg * 16, wds);
            }
          } else {           // relu(acc) * se * att = relu(acc * (se * att)): both gates are sigmoids (> 0)
            const uint64_t att2 = f2_pack(va[tt], va[tt]);
            uint32_t wds[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint64_t x2 = f2_pack(__uint_as_float(acc[tt][2 * j]), __uint_as_float(acc[tt][2 * j + 1]));
              wds[j] = relu_bf16x2(f2_mul(x2, f2_mul(se2[j], att2)));
            }
            uint4* dst = reinterpret_cast<uint4*>(stg + r * 288 + (t0 + tt) * 32);
            dst[0] = make_uint4(wds[0], wds[1], wds[2], wds[3]);
            dst[1] = make_uint4(wds[4], wds[5], wds[6], wds[7]);
          }
        }
      }
      if (MODE == 1) {
        // one barrier per item: the elected thread has seen the previous store finish reading its staging buffer (the
        // one the next item writes) before anybody passes, then sends this item's tile; TMA clips at the image border
        fence_proxy_async_smem();
        const bool elected = (warp == 2 && lane == 0);
        if (elected) tma_store_wait_read<0>();
        asm volatile("bar.sync 1, 512;" ::: "memory");
        if (elected) {
          tma_store_4d(&p.o_map, stg, cg * 144, tile_w0, tile_h0, n);
          tma_store_commit();
        }
      }
      const bool last = (i == mine - 1) || (cg + 1 == p.groups);
      if (MODE == 0 && last && valid) {
#pragma unroll
        for (int tt = 0; tt < 3; ++tt) {
          if (tt < nt) {
            const int t = t0 + tt;
            const long pos = ((long)n * (p.ho * 3) + (h * 3 + t / 3)) * (long)(p.wo * 3) + (w * 3 + t % 3);
            atomicMax(reinterpret_cast<int*>(p.maps + pos * 2), __float_as_int(va[tt]));
            atomicAdd(p.maps + pos * 2 + 1, vb[tt] * invC);
          }
        }
      }
      if (++cg == p.groups) { cg = 0; ++mt; fresh = true; }
    }
    if (MODE == 1 && warp == 2 && lane == 0) tma_store_wait_all<0>();
    if (warp == 2) { RT_DBG_FLUSH(8, 1) }
    if (warp == 14) { RT_DBG_FLUSH(10, 1) }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------------------------
// Weight gradient of the receptive-field generation (k = 3) on the tensor cores:
//     dW[t][ij][c] = Σ_pixels dZ[m][t][c] * x_tap_ij[m][c]
// Per 16-channel group this is the DIAGONAL (c == c') of D[(t,c)][(ij,c')] = dZ_gᵀ · Xtaps_g, a 144 x 144 product whose
// contraction runs over pixels — both operands are "MN-major" for the tensor core, and the nine SW32 boxes of a pixel
// tile ([128 pixels][16 channels], exactly what the forward loads) laid side by side ARE the canonical MN-major operand:
// 16-element atoms 4096 B apart (LBO), 8-pixel groups 256 B apart (SBO).  One CTA owns (group, slice of the pixel
// tiles), accumulates every tile into the same two TMEM accumulators (t = 0..7 → rows 0..127; t = 8 → rows 0..15 of the
// second) — 16 MMAs per tile — and writes its 81 x 16 diagonal values as one row segment of parts[slice][81*C];
// rf_wgrad_finalize adds the slices in order (bit-reproducible).  15/16 of the MMA work is off-diagonal waste and still
// a fraction of what the loads cost; the CUDA-core kernel spent 473 us on L17 reading the same 340 MB.
// ------------------------------------------------------------------------------------------------------------------
struct alignas(64) RfWgParams {
  CUtensorMap dz_map;   // dZ as (C, wo, 9, ho, N), box (16, Wt, 1, Ht, 1), SW32
  CUtensorMap x_map[4]; // as RfTcParams::a_map
  int tiles_w, tiles_h, n_img, groups, Wt, Ht, s, C;
  int tiles, slices;
  float* parts;         // [slices][81 * C]
};
constexpr int kWgStages = 2;
constexpr int kWgStage = 18 * kRtA;   // nine dZ slices + nine input taps of one pixel tile
constexpr int kWgThreads = 192;

__host__ __device__ constexpr uint64_t rt_desc_sw32_mn(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | (6ull << 61);
}

__global__ void __launch_bounds__(kWgThreads, 1) rf_wgrad_tc_kernel(const __grid_constant__ RfWgParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[kWgStages], empty_bar[kWgStages], tfull;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.dz_map);
    tma_prefetch_desc(&p.x_map[0]);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kWgStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&tfull, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_s, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_trigger();
  pdl_wait();
  const uint32_t tmem = tmem_base_s;

  const int g = (int)blockIdx.x % p.groups, sl = (int)blockIdx.x / p.groups;
  const int t_begin = (int)((long)p.tiles * sl / p.slices), t_end = (int)((long)p.tiles * (sl + 1) / p.slices);
  const int mine = t_end - t_begin;
  const int tiles_img = p.tiles_w * p.tiles_h;
  const uint32_t a_bytes = (uint32_t)(16 * p.Wt * p.Ht * 2);

  if (warp == 0) {
    if (lane == 0) {
      uint32_t s = 0, ph = 1;
      for (int i = 0; i < mine; ++i) {
        const int mt = t_begin + i;
        const int n = mt / tiles_img, tr = mt - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
        const int c0 = g * 16, row0 = th * p.Ht, col0 = tw * p.Wt;
        mbar_wait(&empty_bar[s], ph);
        mbar_arrive_expect_tx(&full_bar[s], 18u * a_bytes);
        uint8_t* st = smem + s * kWgStage;
        for (int t = 0; t < 9; ++t) tma_load_5d(st + t * kRtA, &p.dz_map, &full_bar[s], c0, col0, t, row0, n);
        uint8_t* sx = st + 9 * kRtA;
        for (int ti = 0; ti < 3; ++ti) {
          uint8_t* dst = sx + ti * 3 * kRtA;
          if (p.s == 2) {   // as the forward: input row 2h + ti - 1 lives in parity plane (ti + 1) & 1
            const int row = row0 - (ti == 0);
            const CUtensorMap* m0 = &p.x_map[((ti + 1) & 1) * 2];
            tma_load_5d(dst, m0 + 1, &full_bar[s], c0, col0 - 1, 0, row, n);
            tma_load_5d(dst + kRtA, m0, &full_bar[s], c0, col0, 0, row, n);
            tma_load_5d(dst + 2 * kRtA, m0 + 1, &full_bar[s], c0, col0, 0, row, n);
          } else {
            const int row = row0 + ti - 1;
            tma_load_5d(dst, &p.x_map[0], &full_bar[s], c0, col0 - 1, 0, row, n);
            tma_load_5d(dst + kRtA, &p.x_map[0], &full_bar[s], c0, col0, 0, row, n);
            tma_load_5d(dst + 2 * kRtA, &p.x_map[0], &full_bar[s], c0, col0 + 1, 0, row, n);
          }
        }
        if (++s == kWgStages) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && mine > 0) {
      const uint32_t idesc = umma_idesc_bf16_f32(128, 144, 1, 1);
      uint32_t s = 0, ph = 0;
      for (int i = 0; i < mine; ++i) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after_sync();
        const uint32_t base = smem_u32(smem + s * kWgStage);
        for (int k = 0; k < 8; ++k) {   // 16 pixels per MMA
          const uint64_t a1 = rt_desc_sw32_mn(base + k * 512, kRtA, 256);
          const uint64_t a2 = rt_desc_sw32_mn(base + 8 * kRtA + k * 512, kRtA, 256);   // t = 8 (+ 7 boxes of don't-care rows)
          const uint64_t b = rt_desc_sw32_mn(base + 9 * kRtA + k * 512, kRtA, 256);
          const uint32_t acc = (uint32_t)((i | k) != 0);
          umma_bf16(tmem, a1, b, idesc, acc);
          umma_bf16(tmem + 256u, a2, b, idesc, acc);
        }
        umma_commit(&empty_bar[s]);
        if (++s == kWgStages) { s = 0; ph ^= 1u; }
      }
      umma_commit(&tfull);
    }
  } else {
    const int q = warp & 3;
    const int r = q * 32 + lane, t = r >> 4, c = r & 15;
    float* prow = p.parts + (long)sl * 81 * p.C + g * 16 + c;
    if (mine > 0) {
      mbar_wait(&tfull, 0);
      tc_fence_after_sync();
    }
    for (int half = 0; half < 2; ++half) {
      if (half == 1 && q != 0) break;                 // second accumulator: rows 0..15 (t = 8) only
      for (int ij = 0; ij < 9; ++ij) {
        uint32_t acc[16];
        float v = 0.f;
        if (mine > 0) {
          tmem_ld16(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * 256 + ij * 16), acc);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j == c) v = __uint_as_float(acc[j]);
        }
        const int tt = half == 0 ? t : 8;
        if (half == 0 || lane < 16) prow[(long)(tt * 9 + ij) * p.C] = v;
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}


// ------------------------------------------------------------------------------------------------------------------
// FUSED pass 2 (inference): generate -> BN -> ReLU -> SE x spatial-attention gate -> the k x k / stride-k convolution of
// RFCBAMConv (models/rfa.py:121-129) -> its BatchNorm + ReLU, in ONE kernel.  The gated receptive-field matrix (9*C wide,
// the largest tensor of the module) never exists in HBM: an item's gated [128 pixels][144] tile is written by the epilogue
// warps into shared memory as nine SW32 K-major tiles and consumed from there as the A operand of a second accumulation
//     OUT[128 pixels][Cout] += G_item[128][144] * W2[group][Cout][144]^T            (tcgen05, fp32 in TMEM)
// over the C/16 channel groups of the pixel tile; W2 streams through a ring of bulk copies from a pre-swizzled image.
// Work unit = a whole pixel tile (all its channel groups stay on one CTA).  Single-buffered generate accumulator, weight
// tiles and gated tile: the second GEMM of item i is issued after the generate MMAs of item i+1, so it runs while the
// epilogue warps gate item i+1.
//   TMEM: generate accumulator at column 0 (144 columns), OUT at column 256 (Cout <= 256 columns).
// ------------------------------------------------------------------------------------------------------------------
struct alignas(64) RfFusedParams {
  CUtensorMap a_map[4];
  int tiles_w, tiles_h, n_img, groups;
  int Wt, Ht, ho, wo, C, s, Cout;
  int tiles, per_cta;
  int a_slots;               // tap buffers of one whole item (nine boxes) each: 1 or 2
  uint32_t a_bytes;
  const __nv_bfloat16* wc;
  const uint32_t* shift;
  const float* se;
  const float* att;
  const __nv_bfloat16* w2;   // [groups][9][Cout][16] bf16, every [Cout][16] tile SW32-swizzled
  const float* scale2;       // [Cout] folded conv BatchNorm
  const float* shift2;       // [Cout] (conv bias folded in)
  __nv_bfloat16* out;        // NHWC [N][ho][wo][Cout], pixel stride ld_out
  long ld_out;
};

// Barriers are per ITEM, not per ring stage: the MMA-issuing lane pays ~340 cycles per tcgen05.commit against ~80 per
// tcgen05.mma (tests/dev/mma_issue.cu), so an item costs it 19 MMAs and three commits (taps free + accumulator ready,
// W2 and gated tile free) — with one commit per three-box stage it was nine, and the lane was the bottleneck.
__global__ void __launch_bounds__(kRtThreads, 1) rf_fused_kernel(const __grid_constant__ RfFusedParams p) {
  extern __shared__ uint8_t smem_raw[];
  // gen_ready: 16 epilogue warps (accumulator of the previous item read) + the weight-tile writer (this item's diagonal in
  // place); g2_ready: 16 epilogue warps (gated tile written) + the W2 bulk copy (one arrive.expect_tx).  A lone issuing
  // lane pays ~250 cycles per mbarrier round trip, so what it waits on is merged wherever no arriver can run a phase ahead.
  __shared__ uint64_t a_full[2], a_empty[2], w_empty, tfull, gen_ready, g2_ready, out_full, out_empty;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  const int nA = p.a_slots;
  uint8_t* sA = smem;                                  // nA x 9 tap boxes
  uint8_t* sOnes = sA + nA * 9 * kRtA;
  uint8_t* sB = sOnes + kRtA;                          // generate weight tiles of the current group
  uint8_t* sStage = sB + kRtB1;                        // gated tile: 9 x [128][16] SW32
  uint8_t* sW2 = sStage + 9 * kRtA;                    // 9 x [Cout][16] SW32: the current group's slice of W2
  const uint32_t w2_tile = (uint32_t)p.Cout * 32u;

  for (int i = threadIdx.x; i < (kRtA + kRtB1) / 16; i += kRtThreads)
    reinterpret_cast<uint4*>(sOnes)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  if (threadIdx.x < 128)
    *reinterpret_cast<uint32_t*>(sOnes + threadIdx.x * 32 + ((threadIdx.x >> 2) & 1) * 16) = 0x3F803F80u;
  fence_proxy_async_smem();
  if (warp == 0 && lane < (p.s == 2 ? 4 : 1)) tma_prefetch_desc(&p.a_map[lane]);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    mbar_init(&w_empty, 1);
    mbar_init(&tfull, 1);
    mbar_init(&gen_ready, 17);
    mbar_init(&g2_ready, 17);
    mbar_init(&out_full, 1); mbar_init(&out_empty, 16);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_s, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_trigger();
  pdl_wait();
  const uint32_t tmem = tmem_base_s;

  const int tile0 = (int)blockIdx.x * p.per_cta;
  const int my_tiles = min(p.per_cta, p.tiles - tile0);
  const int mine = my_tiles > 0 ? my_tiles * p.groups : 0;   // items, channel group fastest
  const int tiles_img = p.tiles_w * p.tiles_h;

  if (warp == 0) {
    if (lane == 0) {
      // Two independent streams from one lane — an item's nine tap boxes (what the generate MMAs wait on) and its W2
      // slice (what the second GEMM waits on) — polled without blocking, so neither holds the other back.
      int cgA = 0, mtA = tile0, itemA = 0, cgB = 0, itemB = 0;
      long long t_idle = clock64();
      while (itemA < mine || itemB < mine) {
        bool moved = false;
        const int slot = nA == 2 ? (itemA & 1) : 0;
        const uint32_t useA = (uint32_t)(nA == 2 ? (itemA >> 1) : itemA);
        if (itemA < mine && mbar_try_wait(&a_empty[slot], (useA & 1u) ^ 1u)) {
          const int n = mtA / tiles_img, tr = mtA - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
          const int c0 = cgA * 16;
          uint64_t* bar = &a_full[slot];
          mbar_arrive_expect_tx(bar, 9u * p.a_bytes);
          uint8_t* dst = sA + slot * 9 * kRtA;
#pragma unroll
          for (int ti = 0; ti < 3; ++ti, dst += 3 * kRtA) {
            if (p.s == 2) {   // input row 2h + ti - 1: parity (ti + 1) & 1, plane row h - (ti == 0)
              const int row = th * p.Ht - (ti == 0), col = tw * p.Wt;
              const CUtensorMap* m0 = &p.a_map[((ti + 1) & 1) * 2];
              tma_load_5d(dst, m0 + 1, bar, c0, col - 1, 0, row, n);
              tma_load_5d(dst + kRtA, m0, bar, c0, col, 0, row, n);
              tma_load_5d(dst + 2 * kRtA, m0 + 1, bar, c0, col, 0, row, n);
            } else {
              const int row = th * p.Ht + ti - 1, col = tw * p.Wt;
              tma_load_5d(dst, &p.a_map[0], bar, c0, col - 1, 0, row, n);
              tma_load_5d(dst + kRtA, &p.a_map[0], bar, c0, col, 0, row, n);
              tma_load_5d(dst + 2 * kRtA, &p.a_map[0], bar, c0, col + 1, 0, row, n);
            }
          }
          ++itemA;
          if (++cgA == p.groups) { cgA = 0; ++mtA; }
          moved = true;
        }
        if (itemB < mine && mbar_try_wait(&w_empty, ((uint32_t)itemB & 1u) ^ 1u)) {
          const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.w2) + (size_t)cgB * 9u * w2_tile;
          mbar_arrive_expect_tx(&g2_ready, 9u * w2_tile);
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                           smem_u32(sW2)),
                       "l"(wsrc), "r"(9u * w2_tile), "r"(smem_u32(&g2_ready))
                       : "memory");
          ++itemB;
          if (++cgB == p.groups) cgB = 0;
          moved = true;
        }
        if (moved) t_idle = clock64();
        else if (clock64() - t_idle > 4000000000LL) __trap();   // a pipeline bug must surface as an error, not a hang
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // One lane issues everything: the generate MMAs of item i+1 go out before the second GEMM of item i, so the
      // tensor pipe works on them while the epilogue warps gate item i.  (Measured with tests/dev/rf_fused_probe.cu: a
      // second issuing lane for the second GEMM — in warp 18 or in a twentieth warp of its own — was SLOWER, 110 -> 125 us
      // on the 128-channel layer: the sixteen gating warps are the critical chain at ~2400 busy cycles per item, and the
      // single lane's 19 MMAs + 3 commits fit inside it.)
      const uint32_t idesc = umma_idesc_bf16_f32(128, 144, 0, 0);
      const uint32_t idesc2 = umma_idesc_bf16_f32(128, p.Cout, 0, 0);
      const uint64_t a_desc0 = rt_desc_sw32(smem_u32(sA)), b_desc = rt_desc_sw32(smem_u32(sB));
      const uint64_t ones_desc = rt_desc_sw32(smem_u32(sOnes));
      const uint64_t g_desc0 = rt_desc_sw32(smem_u32(sStage)), w_desc0 = rt_desc_sw32(smem_u32(sW2));
      const uint32_t d_gen = tmem, d_out = tmem + 256u;
      int cg2 = 0, tile = 0;
      RT_DBG_DECL
      for (int i = 0; i <= mine; ++i) {
        if (i < mine) {
          const int slot = nA == 2 ? (i & 1) : 0;
          const uint32_t use = (uint32_t)(nA == 2 ? (i >> 1) : i);
          RT_WAIT(0, &gen_ready, (uint32_t)i & 1u);
          RT_WAIT(1, &a_full[slot], use & 1u);
          tc_fence_after_sync();
          umma_bf16(d_gen, ones_desc, b_desc + (uint64_t)(9 * (kRtB >> 4)), idesc, 0u);   // acc = shift
          const uint64_t a_desc = a_desc0 + (uint64_t)(slot * (9 * kRtA >> 4));
#pragma unroll
          for (int t = 0; t < 9; ++t)
            umma_bf16(d_gen, a_desc + (uint64_t)(t * (kRtA >> 4)), b_desc + (uint64_t)(t * (kRtB >> 4)), idesc, 1u);
          umma_commit(&a_empty[slot]);
          umma_commit(&tfull);
        }
        if (i > 0) {   // second GEMM of item i-1: OUT (+)= gated tile x its W2 slice
          if (cg2 == 0) mbar_wait(&out_empty, ((uint32_t)tile & 1u) ^ 1u);   // previous tile's output has left TMEM
          RT_WAIT(2, &g2_ready, (uint32_t)(i - 1) & 1u);
          tc_fence_after_sync();
#pragma unroll
          for (int t = 0; t < 9; ++t)
            umma_bf16(d_out, g_desc0 + (uint64_t)(t * (kRtA >> 4)), w_desc0 + (uint64_t)((t * w2_tile) >> 4), idesc2,
                      (uint32_t)((cg2 | t) != 0));
          umma_commit(&w_empty);          // W2 slice AND gated tile are free again
          if (++cg2 == p.groups) {
            umma_commit(&out_full);
            cg2 = 0;
            ++tile;
          }
        }
      }
      RT_DBG_FLUSH(2, 3)
    }
  } else if (warp == 18) {
    RT_DBG_DECL
    uint32_t off[5], soff[5];
#pragma unroll
    for (int m = 0; m < 5; ++m) {
      const int n = lane + 32 * m;
      off[m] = smem_u32(sB) + (uint32_t)(n * 32 + ((((n >> 3) ^ (n >> 2)) & 1) << 4) + (n & 7) * 2);
      soff[m] = smem_u32(sB) + (uint32_t)(9 * kRtB + n * 32 + ((n >> 2) & 1) * 16);
    }
    uint16_t v[9][5];
    uint32_t sh[5];
    auto fetch = [&](int cg) {
      const uint16_t* src = reinterpret_cast<const uint16_t*>(p.wc) + (long)cg * 1296 + lane;
#pragma unroll
      for (int k = 0; k < 9; ++k)
#pragma unroll
        for (int m = 0; m < 5; ++m) v[k][m] = (m < 4 || lane < 16) ? __ldg(src + k * 144 + 32 * m) : (uint16_t)0;
#pragma unroll
      for (int m = 0; m < 5; ++m) sh[m] = (m < 4 || lane < 16) ? __ldg(p.shift + cg * 144 + lane + 32 * m) : 0u;
    };
    int cg = 0;
    if (mine > 0) fetch(0);
    for (int i = 0; i < mine; ++i) {
      if (i > 0) RT_WAIT(0, &tfull, ((uint32_t)i & 1u) ^ 1u);   // the generate MMAs of item i-1 have read the tiles
#pragma unroll
      for (int k = 0; k < 9; ++k)
#pragma unroll
        for (int m = 0; m < 5; ++m)
          if (m < 4 || lane < 16)
            asm volatile("st.shared.b16 [%0], %1;" ::"r"(off[m] + (uint32_t)(k * kRtB)), "h"(v[k][m]) : "memory");
#pragma unroll
      for (int m = 0; m < 5; ++m)
        if (m < 4 || lane < 16) asm volatile("st.shared.b32 [%0], %1;" ::"r"(soff[m]), "r"(sh[m]) : "memory");
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&gen_ready);
      if (++cg == p.groups) cg = 0;
      if (i + 1 < mine) fetch(cg);
    }
    RT_DBG_FLUSH(6, 1)
  } else {
    RT_DBG_DECL
    const int q = warp & 3;
    const int eg = (warp - 2) >> 2;                 // t = 0..2 | 3,4 | 5,6 | 7,8;  output columns eg * Cout/4 ..
    const int t0 = eg == 0 ? 0 : 1 + 2 * eg, nt = eg == 0 ? 3 : 2;
    const int r = q * 32 + lane;
    const int hh = r / p.Wt, ww = r - hh * p.Wt;
    const uint32_t swz = (uint32_t)((r >> 2) & 1);
    float va[3];
    int cg = 0, mt = tile0, tile = 0;
    int n = 0, h = 0, w = 0;
    bool valid = false, fresh = true;
    const float* se_row = nullptr;
    if (mine > 0 && lane == 0) mbar_arrive(&gen_ready);   // item 0: there is no previous accumulator to drain
    for (int i = 0; i < mine; ++i) {
      if (fresh) {
        fresh = false;
        n = mt / tiles_img;
        const int tr = mt - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
        h = th * p.Ht + hh; w = tw * p.Wt + ww;
        valid = (r < p.Wt * p.Ht) && h < p.ho && w < p.wo;
#pragma unroll
        for (int tt = 0; tt < 3; ++tt) {
          va[tt] = 0.f;
          if (valid && tt < nt) {
            const int t = t0 + tt;
            va[tt] = __ldg(p.att + ((long)n * (p.ho * 3) + (h * 3 + t / 3)) * (long)(p.wo * 3) + (w * 3 + t % 3));
          }
        }
        se_row = p.se + (long)n * p.C;
      }
      uint64_t se2[8];
      {
        const float4* sp = reinterpret_cast<const float4*>(se_row + cg * 16);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 e = __ldg(sp + j);
          se2[2 * j] = f2_pack(e.x, e.y); se2[2 * j + 1] = f2_pack(e.z, e.w);
        }
      }
      RT_WAIT(0, &tfull, (uint32_t)i & 1u);
      tc_fence_after_sync();
      const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(t0 * 16);
      uint32_t acc[3][16];
      tmem_ld16(trow, acc[0]);
      tmem_ld16(trow + 16u, acc[1]);
      if (nt == 3) tmem_ld16(trow + 32u, acc[2]);
      tmem_ld_wait();
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0 && i + 1 < mine) mbar_arrive(&gen_ready);   // the accumulator may be overwritten by item i+1
      uint32_t wds[3][8];
#pragma unroll
      for (int tt = 0; tt < 3; ++tt) {
        if (tt < nt) {
          const uint64_t att2 = f2_pack(va[tt], va[tt]);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint64_t x2 = f2_pack(__uint_as_float(acc[tt][2 * j]), __uint_as_float(acc[tt][2 * j + 1]));
            wds[tt][j] = relu_bf16x2(f2_mul(x2, f2_mul(se2[j], att2)));
          }
        }
      }
      // the second GEMM of the previous item has finished reading the gated tile (and its W2 slice)
      RT_WAIT(1, &w_empty, ((uint32_t)i & 1u) ^ 1u);
#pragma unroll
      for (int tt = 0; tt < 3; ++tt) {
        if (tt < nt) {
          uint8_t* row = sStage + (t0 + tt) * kRtA + r * 32;
          *reinterpret_cast<uint4*>(row + (swz << 4)) = make_uint4(wds[tt][0], wds[tt][1], wds[tt][2], wds[tt][3]);
          *reinterpret_cast<uint4*>(row + ((swz ^ 1u) << 4)) = make_uint4(wds[tt][4], wds[tt][5], wds[tt][6], wds[tt][7]);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&g2_ready);
      if (++cg == p.groups) {
        // the tile's OUT accumulator is complete: BatchNorm + ReLU, bf16, NHWC rows
        RT_WAIT(2, &out_full, (uint32_t)tile & 1u);
        tc_fence_after_sync();
        const int ncol = p.Cout >> 2;                // columns of this epilogue group: 32 or 64
        const uint32_t orow = tmem + ((uint32_t)(q * 32) << 16) + 256u + (uint32_t)(eg * ncol);
        __nv_bfloat16* dst = p.out + (((long)n * p.ho + h) * p.wo + w) * p.ld_out + eg * ncol;
        for (int c0 = 0; c0 < ncol; c0 += 32) {
          uint32_t o[2][16];
          tmem_ld16(orow + (uint32_t)c0, o[0]);
          tmem_ld16(orow + (uint32_t)c0 + 16u, o[1]);     // ncol is a multiple of 32 (Cout = 128 or 256)
          tmem_ld_wait();
          if (c0 + 32 >= ncol) {       // last read of this warp: hand the accumulator back
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&out_empty);
          }
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            if (valid) {
              const int cb = eg * ncol + c0 + half * 16;
              uint32_t wo8[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float2 sc = __ldg(reinterpret_cast<const float2*>(p.scale2 + cb) + j);
                const float2 sf = __ldg(reinterpret_cast<const float2*>(p.shift2 + cb) + j);
                const float y0 = fmaf(__uint_as_float(o[half][2 * j]), sc.x, sf.x);
                const float y1 = fmaf(__uint_as_float(o[half][2 * j + 1]), sc.y, sf.y);
                wo8[j] = relu_bf16x2(f2_pack(y0, y1));
              }
              st_global_256(dst + c0 + half * 16, wo8);
            }
          }
        }
        cg = 0; ++mt; ++tile; fresh = true;
      }
    }
    if (warp == 2) { RT_DBG_FLUSH(8, 3) }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

typedef CUresult (*rt_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int rt_encode(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides,
                     const uint32_t* box, const uint32_t* estr, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_32B) {
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
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(LY_ECUDA, "cuTensorMapEncodeTiled(rf_tc) failed: %d", (int)r);
  return LY_OK;
}

static int rf_tc_launch(int mode, const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo,
                        const void* wc, const void* shift, const float* se, const float* att, void* A, float* maps,
                        cudaStream_t stream) {
  LY_REQUIRE(x && wc && shift, "ly_rf_tc: null pointer");
  LY_REQUIRE(C % 16 == 0 && C > 0 && (s == 1 || s == 2) && ldx % 8 == 0 && ldx >= C, "ly_rf_tc: bad C/s/ld");
  LY_REQUIRE(ho == (H + 2 - 3) / s + 1 && wo == (W + 2 - 3) / s + 1, "ly_rf_tc: output size mismatch");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(wc) % 16 == 0 &&
                 reinterpret_cast<uintptr_t>(shift) % 4 == 0,
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
  p.wc = (const __nv_bfloat16*)wc; p.shift = (const uint32_t*)shift; p.se = se; p.att = att; p.A = (__nv_bfloat16*)A; p.maps = maps;
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
  if (mode == 1) {
    uint64_t dims[4] = {(uint64_t)9 * C, (uint64_t)wo, (uint64_t)ho, (uint64_t)N};
    uint64_t str[3] = {(uint64_t)9 * C * 2, (uint64_t)wo * 9 * C * 2, (uint64_t)ho * wo * 9 * C * 2};
    uint32_t box[4] = {144, (uint32_t)bw, (uint32_t)bh, 1};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = rt_encode(&p.o_map, A, 4, dims, str, box, es, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
  }
  const long items = (long)p.tiles_w * p.tiles_h * N * p.groups;
  LY_REQUIRE(items < (1L << 30), "ly_rf_tc: problem too large");
  const int sms = device_sm_count();
  p.items = (int)items;
  p.per_cta = (int)ceil_div(items, (long)sms);
  const unsigned grid = (unsigned)ceil_div(items, (long)p.per_cta);
  const size_t smem0 = (size_t)6 * 3 * kRtA + kRtA + 2 * kRtB1 + 1024;
  const size_t smem1 = (size_t)4 * 3 * kRtA + kRtA + 2 * kRtB1 + 2 * kRtStage + 1024;
  static std::atomic<bool> attr_done[64];   // zero-initialised; set only after the attribute call succeeded
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(rf_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem0));
    LY_CUDA(cudaFuncSetAttribute(rf_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    LY_CUDA(cudaFuncSetAttribute(rf_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem0));
    attr_done[dev & 63].store(true);
  }
  if (mode == 0) {
    LY_CUDA(cudaMemsetAsync(maps, 0, sizeof(float) * 2 * (size_t)N * ho * 3 * wo * 3, stream));
    LY_CUDA(launch(rf_tc_kernel<0>, grid, kRtThreads, smem0, stream, p));
  } else if (mode == 1) {
    LY_CUDA(launch(rf_tc_kernel<1>, grid, kRtThreads, smem1, stream, p));
  } else {
    LY_CUDA(launch(rf_tc_kernel<2>, grid, kRtThreads, smem0, stream, p));
  }
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// Input-gradient pass of the receptive-field generation: R[m][ij][c] = Σ_t wT[t][ij][c] * dZ[m][t][c] — the same nine
// K = 16 MMAs per item as the forward, with the nine feature slices of dZ (bf16 [N*ho*wo][9][C]) as the "taps".
static int rf_tc_launch_taps(const void* dZ, int N, int ho, int wo, int C, const void* wc, const void* shift, void* R,
                             cudaStream_t stream) {
  LY_REQUIRE(dZ && wc && shift && R, "ly_rf_tc_taps: null pointer");
  LY_REQUIRE(C % 16 == 0 && C > 0 && N > 0 && ho > 0 && wo > 0, "ly_rf_tc_taps: bad shape");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(dZ) % 16 == 0 && reinterpret_cast<uintptr_t>(R) % 32 == 0 &&
                 reinterpret_cast<uintptr_t>(wc) % 16 == 0,
             "ly_rf_tc_taps: alignment");
  RfTcParams p;
  memset(&p, 0, sizeof(p));
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
  p.ho = ho; p.wo = wo; p.C = C; p.s = 1;
  p.tap_slices = 1;
  p.a_bytes = (uint32_t)(16 * bw * bh * 2);
  p.wc = (const __nv_bfloat16*)wc; p.shift = (const uint32_t*)shift; p.A = (__nv_bfloat16*)R;
  {
    uint64_t dims[5] = {(uint64_t)C, (uint64_t)wo, 9, (uint64_t)ho, (uint64_t)N};
    uint64_t str[4] = {(uint64_t)9 * C * 2, (uint64_t)C * 2, (uint64_t)wo * 9 * C * 2, (uint64_t)ho * wo * 9 * C * 2};
    uint32_t box[5] = {16, (uint32_t)bw, 1, (uint32_t)bh, 1};
    uint32_t es[5] = {1, 1, 1, 1, 1};
    int rc = rt_encode(&p.a_map[0], dZ, 5, dims, str, box, es);
    if (rc) return rc;
  }
  const long items = (long)p.tiles_w * p.tiles_h * N * p.groups;
  LY_REQUIRE(items < (1L << 30), "ly_rf_tc_taps: problem too large");
  const int sms = device_sm_count();
  p.items = (int)items;
  p.per_cta = (int)ceil_div(items, (long)sms);
  const unsigned grid = (unsigned)ceil_div(items, (long)p.per_cta);
  const size_t smem0 = (size_t)6 * 3 * kRtA + kRtA + 2 * kRtB1 + 1024;
  static std::atomic<bool> attr_done[64];
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(rf_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem0));
    attr_done[dev & 63].store(true);
  }
  LY_CUDA(launch(rf_tc_kernel<2>, grid, kRtThreads, smem0, stream, p));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

// Launch of rf_wgrad_tc_kernel; `parts` holds at least `max_rows` rows of 81*C floats.  Returns the number of rows (slices)
// written through *rows_out; the caller adds them in order (rf_wgrad_finalize).
int rf_wgrad_tc_launch(const void* dZ, const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo,
                       float* parts, long max_rows, int* rows_out, cudaStream_t stream) {
  LY_REQUIRE(dZ && x && parts && rows_out, "rf_wgrad_tc: null pointer");
  LY_REQUIRE(C % 16 == 0 && C > 0 && (s == 1 || s == 2) && ldx % 8 == 0 && ldx >= C, "rf_wgrad_tc: bad C/s/ld");
  RfWgParams p;
  memset(&p, 0, sizeof(p));
  // 128-pixel tiles exactly (the pixel count is the K extent of the MMAs): power-of-two widths; overhang = zero fill
  int bw = 16;
  double best = -1;
  const int cand[5] = {8, 16, 32, 64, 128};
  for (int i = 0; i < 5; ++i) {
    const int wt = cand[i], ht = 128 / wt;
    const double eff = (double)ho * wo / ((double)ceil_div(wo, wt) * ceil_div(ho, ht) * 128.0);
    if (eff > best + 1e-9) { best = eff; bw = wt; }
  }
  const int bh = 128 / bw;
  p.Wt = bw; p.Ht = bh;
  p.tiles_w = ceil_div(wo, bw); p.tiles_h = ceil_div(ho, bh); p.n_img = N;
  p.groups = C / 16; p.s = s; p.C = C;
  const long tiles = (long)p.tiles_w * p.tiles_h * N;
  LY_REQUIRE(tiles < (1L << 30), "rf_wgrad_tc: problem too large");
  p.tiles = (int)tiles;
  long slices = device_sm_count() / p.groups;
  if (slices < 1) slices = 1;
  if (slices > tiles) slices = tiles;
  if (slices > max_rows) slices = max_rows;
  LY_REQUIRE(slices >= 1, "rf_wgrad_tc: no scratch rows");
  p.slices = (int)slices;
  p.parts = parts;
  uint32_t es[5] = {1, 1, 1, 1, 1};
  {
    uint64_t dims[5] = {(uint64_t)C, (uint64_t)wo, 9, (uint64_t)ho, (uint64_t)N};
    uint64_t str[4] = {(uint64_t)9 * C * 2, (uint64_t)C * 2, (uint64_t)wo * 9 * C * 2, (uint64_t)ho * wo * 9 * C * 2};
    uint32_t box[5] = {16, (uint32_t)bw, 1, (uint32_t)bh, 1};
    int rc = rt_encode(&p.dz_map, dZ, 5, dims, str, box, es);
    if (rc) return rc;
  }
  for (int par = 0; par < (s == 2 ? 4 : 1); ++par) {
    const int ph = par >> 1, pw = par & 1;
    const uint64_t Hp = (uint64_t)(H - ph + s - 1) / s, Wp = (uint64_t)(W - pw + s - 1) / s;
    if (Hp == 0 || Wp == 0) { memcpy(&p.x_map[par], &p.x_map[0], sizeof(CUtensorMap)); continue; }
    uint64_t dims[5] = {(uint64_t)C, Wp, 1, Hp, (uint64_t)N};
    uint64_t str[4] = {(uint64_t)s * ldx * 2, (uint64_t)W * ldx * 2, (uint64_t)s * W * ldx * 2, (uint64_t)H * W * ldx * 2};
    uint32_t box[5] = {16, (uint32_t)bw, 1, (uint32_t)bh, 1};
    int rc = rt_encode(&p.x_map[par], (const __nv_bfloat16*)x + ((long)ph * W + pw) * ldx, 5, dims, str, box, es);
    if (rc) return rc;
  }
  const size_t smem = (size_t)kWgStages * kWgStage + 1024;
  static std::atomic<bool> attr_done[64];
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!attr_done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(rf_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done[dev & 63].store(true);
  }
  LY_CUDA(launch(rf_wgrad_tc_kernel, dim3((unsigned)(slices * p.groups)), dim3(kWgThreads), smem, stream, p));
  LY_LAUNCH_CHECK();
  *rows_out = (int)slices;
  return LY_OK;
}

static int rf_fused_launch(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                           const void* shift, const float* se, const float* att, const void* w2, const float* scale2,
                           const float* shift2, int Cout, void* out, long ld_out, cudaStream_t stream) {
  LY_REQUIRE(x && wc && shift && se && att && w2 && scale2 && shift2 && out, "ly_rfcbam_fused_infer: null pointer");
  LY_REQUIRE(C % 16 == 0 && C > 0 && (s == 1 || s == 2) && ldx % 8 == 0 && ldx >= C, "ly_rfcbam_fused_infer: bad C/s/ld");
  LY_REQUIRE(Cout == 128 || Cout == 256, "ly_rfcbam_fused_infer: Cout=%d must be 128 or 256", Cout);
  LY_REQUIRE(ho == (H + 2 - 3) / s + 1 && wo == (W + 2 - 3) / s + 1, "ly_rfcbam_fused_infer: output size mismatch");
  LY_REQUIRE(ld_out % 16 == 0 && ld_out >= Cout && reinterpret_cast<uintptr_t>(out) % 32 == 0,
             "ly_rfcbam_fused_infer: out must be 32-byte aligned with a pixel stride that is a multiple of 16");
  LY_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(wc) % 16 == 0 &&
                 reinterpret_cast<uintptr_t>(w2) % 16 == 0 && reinterpret_cast<uintptr_t>(se) % 16 == 0 &&
                 reinterpret_cast<uintptr_t>(scale2) % 8 == 0 && reinterpret_cast<uintptr_t>(shift2) % 8 == 0,
             "ly_rfcbam_fused_infer: alignment");
  RfFusedParams p;
  memset(&p, 0, sizeof(p));
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
  p.ho = ho; p.wo = wo; p.C = C; p.s = s; p.Cout = Cout;
  p.a_bytes = (uint32_t)(16 * bw * bh * 2);
  p.wc = (const __nv_bfloat16*)wc; p.shift = (const uint32_t*)shift; p.se = se; p.att = att;
  p.w2 = (const __nv_bfloat16*)w2; p.scale2 = scale2; p.shift2 = shift2;
  p.out = (__nv_bfloat16*)out; p.ld_out = ld_out;
  for (int par = 0; par < (s == 2 ? 4 : 1); ++par) {
    const int ph = par >> 1, pw = par & 1;
    const uint64_t Hp = (uint64_t)(H - ph + s - 1) / s, Wp = (uint64_t)(W - pw + s - 1) / s;
    if (Hp == 0 || Wp == 0) { memcpy(&p.a_map[par], &p.a_map[0], sizeof(CUtensorMap)); continue; }
    uint64_t dims[5] = {(uint64_t)C, Wp, 1, Hp, (uint64_t)N};
    uint64_t str[4] = {(uint64_t)s * ldx * 2, (uint64_t)W * ldx * 2, (uint64_t)s * W * ldx * 2,
                       (uint64_t)H * W * ldx * 2};
    uint32_t box[5] = {16, (uint32_t)bw, 1, (uint32_t)bh, 1};
    uint32_t es[5] = {1, 1, 1, 1, 1};
    int rc = rt_encode(&p.a_map[par], (const __nv_bfloat16*)x + ((long)ph * W + pw) * ldx, 5, dims, str, box, es);
    if (rc) return rc;
  }
  const long tiles = (long)p.tiles_w * p.tiles_h * N;
  LY_REQUIRE(tiles * p.groups < (1L << 30), "ly_rfcbam_fused_infer: problem too large");
  p.tiles = (int)tiles;
  p.per_cta = (int)ceil_div(tiles, (long)device_sm_count());
  const unsigned grid = (unsigned)ceil_div(tiles, (long)p.per_cta);
  // two whole-item tap buffers when the W2 slice leaves room for them (Cout = 128), else one
  p.a_slots = Cout <= 128 ? 2 : 1;
  const size_t smem = (size_t)p.a_slots * 9 * kRtA + kRtA + kRtB1 + 9 * kRtA + (size_t)9 * Cout * 32 + 1024;
  static std::atomic<size_t> attr_set[64];
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (attr_set[dev & 63].load() < smem) {
    LY_CUDA(cudaFuncSetAttribute(rf_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set[dev & 63].store(smem);
  }
  LY_CUDA(launch(rf_fused_kernel, grid, kRtThreads, smem, stream, p));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

}  // namespace ly

using namespace ly;

// x: NHWC bf16 (pixel stride ldx); wc: bf16 [C/16][9 taps][9 t][16] depthwise generate weights with the BatchNorm scale
// folded in (value for output (t, c) and tap ij at wc[c/16][ij][t][c%16]); shift: uint32 [C/16][9 t][16], the BatchNorm
// shift as two bf16 (hi | lo << 16, hi + lo = shift): it enters the accumulator through a tenth MMA against a tile of ones.
// Pass 1: maps fp32 [N][3*ho][3*wo][2] = (max, mean) over channels of the generated features (zeroed by the call).
extern "C" int ly_rf_tc_stats(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                              const void* shift, float* maps, ly_stream_t st) {
  LY_REQUIRE(maps, "ly_rf_tc_stats: null pointer");
  return rf_tc_launch(0, x, ldx, N, H, W, C, s, ho, wo, wc, shift, nullptr, nullptr, nullptr, maps, (cudaStream_t)st);
}

// Pass 2: A bf16 [N*ho*wo][9*C], A[m][(c/16)*144 + t*16 + c%16] = relu(generate + shift) * se[n][c] * att[n][3h+t/3][3w+t%3]
// (channel-group-major columns: one item's 144 values per pixel are contiguous, so TMA writes them back in 288-byte rows).
extern "C" int ly_rf_tc_apply(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                              const void* shift, const float* se, const float* att, void* A, ly_stream_t st) {
  LY_REQUIRE(se && att && A && reinterpret_cast<uintptr_t>(A) % 16 == 0 && reinterpret_cast<uintptr_t>(se) % 16 == 0,
             "ly_rf_tc_apply: null or misaligned pointer");
  return rf_tc_launch(1, x, ldx, N, H, W, C, s, ho, wo, wc, shift, se, att, A, nullptr, (cudaStream_t)st);
}

// Training forward of the receptive-field generation (models/rfa.py:95-100, the grouped 3x3 convolution alone):
// Z bf16 [N*ho*wo][9][C], Z[m][t][c] = generate (+ shift; pass zeros for the raw convolution), the layout ly_rf_generate
// writes.  The diagonal weights are bf16 here (fp32 in ly_rf_generate): the difference is below the bf16 rounding of Z.
extern "C" int ly_rf_tc_generate(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                                 const void* shift, void* Z, ly_stream_t st) {
  LY_REQUIRE(Z && reinterpret_cast<uintptr_t>(Z) % 32 == 0, "ly_rf_tc_generate: null or misaligned pointer");
  return rf_tc_launch(2, x, ldx, N, H, W, C, s, ho, wo, wc, shift, nullptr, nullptr, Z, nullptr, (cudaStream_t)st);
}

// Backward of the generation with respect to its input, first half: R bf16 [N*ho*wo][9 (ij)][C],
// R[m][ij][c] = Σ_t wcT[t][ij][c] * dZ[m][t][c] (+ shift, pass zeros): the per-output-pixel gradient of the 3x3 input patch.
// wcT: bf16 [C/16][9 t][9 ij][16] (the forward's weight image with its two 9-axes swapped).  ly_rf_col2im folds R into dx.
extern "C" int ly_rf_tc_taps(const void* dZ, int N, int ho, int wo, int C, const void* wcT, const void* shift, void* R,
                             ly_stream_t st) {
  return rf_tc_launch_taps(dZ, N, ho, wo, C, wcT, shift, R, (cudaStream_t)st);
}

// Pass 2 fused with RFCBAMConv.conv (models/rfa.py:121-129): out bf16 NHWC [N][ho][wo][Cout] = relu(scale2 * (gated
// receptive-field matrix x W2) + shift2); the gated matrix never reaches HBM.  w2: bf16 [C/16][9 t][Cout][16], value
// conv.weight[o][g*16+j][t/3][t%3] at tile row o, column j, every [Cout][16] tile in the SW32 shared-memory image
// (16-byte chunk c of row o stored at chunk c ^ ((o >> 2) & 1)); scale2/shift2: conv BatchNorm folded with the conv bias.
extern "C" int ly_rfcbam_fused_infer(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo,
                                     const void* wc, const void* shift, const float* se, const float* att, const void* w2,
                                     const float* scale2, const float* shift2, int Cout, void* out, long ld_out,
                                     ly_stream_t st) {
  return rf_fused_launch(x, ldx, N, H, W, C, s, ho, wo, wc, shift, se, att, w2, scale2, shift2, Cout, out, ld_out,
                         (cudaStream_t)st);
}
