// Implicit-GEMM convolution on tcgen05 tensor cores (sm_100a).
//
//   persistent grid (one CTA per SM, static tile striding);  block = 608 threads:
//     warp 0       TMA producer, one lane (A boxes: 64 channels x Wt x Ht pixels of one tap, or one halo box per chunk
//                  for 3x3 convolutions with resident weights; B boxes: 64 x BN weights unless resident)
//     warp 1       MMA issuer, one lane (4 x tcgen05.mma K=16 per 64-wide K chunk, one commit per ring stage)
//     warps 2..17  epilogue: four groups of four warps (TMEM lane quadrant = warp % 4), each group drains a different
//                  tile: TMEM -> registers -> scale/shift/activation/residual -> TMA store (flat bf16) or global stores
//   A and B tiles live in 128B-swizzled shared memory, up to four accumulators (128 x BN fp32) in TMEM.
//   Zero padding of a 3x3 convolution and ragged channel counts are TMA out-of-bounds zero fill.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <type_traits>

#include <atomic>

#include "ly_common.cuh"
#include "ly_tc.cuh"
#include "ly_tmap.h"

namespace ly {

// Per-role wait accounting for tests/dev/conv_probe.cu; compiled out of the library.
#ifdef CONV_GEMM_DEBUG
__device__ long long conv_dbg[148][16];
#define CG_WAIT(slot, bar, par)                    \
  do {                                             \
    const long long t__ = clock64();               \
    mbar_wait(bar, par);                           \
    dbg_acc[slot] += clock64() - t__;              \
  } while (0)
#define CG_DBG_DECL long long dbg_acc[4] = {0, 0, 0, 0}; const long long dbg_t0 = clock64();
#define CG_DBG_FLUSH(base, nslots)                                                                      \
  if (lane == 0 && blockIdx.x < 148) {                                                                  \
    for (int z = 0; z < nslots; ++z) conv_dbg[blockIdx.x][base + z] = dbg_acc[z];                       \
    conv_dbg[blockIdx.x][base + nslots] = clock64() - dbg_t0;                                           \
  }
#else
#define CG_WAIT(slot, bar, par) mbar_wait(bar, par)
#define CG_DBG_DECL
#define CG_DBG_FLUSH(base, nslots)
#endif

constexpr int kMaxK = 96;      // K chunks of 64 per launch
constexpr int kMaxStages = 8;
constexpr int kMaxCout = 2560;  // scale/shift staged in shared memory
constexpr int kAStage = 128 * 128;  // 128 rows x 128 B

struct KEntry {
  int16_t src;  // which A tensor map
  int16_t c0;   // channel coordinate of the chunk
  int8_t dw, dh, tz;
  int8_t nk16;  // K = 16 slices of this chunk that hold channels (the rest of the 64-wide box is zero fill: not multiplied)
  int16_t kb, kstep;  // halo mode: weight chunk of tap 0 and the chunk distance between taps
};

struct alignas(64) ConvGemmParams {
  CUtensorMap a_map[4];
  CUtensorMap b_map;
  CUtensorMap o_map;   // output as [rows][Cout] for the TMA-store epilogue (flat 1x1 case)
  int tiles_w, tiles_h, n_img, tiles_n;
  int Wt, Ht, Hout, Wout;
  int num_k, Cout, BN, tmem_cols, stages, nbuf;
  int halo;        // 3x3 with resident weights: pitch P = Wt + 2 of the halo tile ONE box per 64-channel chunk serves all
                   // nine taps through row-shifted descriptors (tests/dev/desc_shift.cu); 0 = one box per tap
  int a_stage;     // bytes of one activation slot of the ring
  int num_kb;      // 64-wide weight chunks (= num_k unless halo)
  int kgroup;      // K chunks per ring stage: one mbarrier phase and ONE tcgen05.commit (~340 cycles of the issuing lane,
                   // tests/dev/mma_issue.cu) cover kgroup x 4 MMAs
  int tma_store;   // epilogue stages 32-row x 64-column slabs in shared memory and stores them with TMA
  int vec32_ok;    // bf16 output rows 32-byte aligned, no patch scatter: 16 columns go out as one 32-byte store
  int issuers;     // MMA-issuing lanes (1 or 2); the ring is split between them
  int det_bulk;    // Detect epilogue stages raw / z rows in the warp's slab and stores them with bulk copies
  int slab_bytes;  // shared-memory slab per epilogue warp (0: none)
  int b_resident;  // all weight chunks stay in shared memory for the whole kernel (single N tile, small K*N)
  int act, out_fp32, vec_ok, res_vec_ok;
  uint32_t a_bytes;
  int patch_k, patch_c;
  const float* scale;
  const float* shift;
  const __nv_bfloat16* residual;
  long res_ld;
  void* out;
  long out_ld;
  // Detect head (ly_detect_head_infer): the epilogue decodes boxes instead of storing the raw GEMM output
  float* det_z;          // [N][rows_total][no] decoded rows, or NULL (= ordinary convolution)
  float* det_raw;        // [N][na][ny][nx][no] raw logits, or NULL
  long det_rows_total, det_row_off;
  int det_na, det_no, det_ny, det_nx;
  float det_stride;
  const float* det_anchors;   // device, [na][2] (w, h) in stride units
  KEntry sched[kMaxK];
};

// act(scale*acc + shift).  For SiLU the kernel stages scale/2 and shift/2, so h = x/2 comes out of the one FFMA and
// silu(x) = x * sigmoid(x) = h + h * tanh(h): FFMA, MUFU, FFMA instead of five instructions per element.
template <int ACT>
__device__ __forceinline__ float ssa(uint32_t acc, float sc, float sh) {
  const float h = fmaf(__uint_as_float(acc), sc, sh);
  if (ACT == LY_ACT_SILU) {
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    return fmaf(h, t, h);
  }
  return act_apply_t<ACT>(h);
}

__device__ __forceinline__ void st_global_v8(void* ptr, const uint32_t (&w)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "r"(w[0]), "r"(w[1]), "r"(w[2]),
               "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}

// One epilogue thread = one output pixel (TMEM lane).  Columns [c_begin, BN) are drained 16 at a time, stride c_step.
template <int ACT>
__device__ __forceinline__ void epilogue_cols(const ConvGemmParams& p, uint32_t trow, bool valid, int n, int h, int w,
                                              long pix, int n0, const float* s_scale, const float* s_shift,
                                              int c_begin, int c_step) {
  const int kC = p.patch_k * p.patch_c;
  for (int c0 = c_begin; c0 < p.BN; c0 += c_step) {
    uint32_t acc[16];
    __syncwarp();
    tmem_ld16(trow + (uint32_t)c0, acc);
    tmem_ld_wait();
    if (!valid) continue;
    if (p.vec32_ok && c0 + 16 <= p.BN && n0 + c0 + 16 <= p.Cout) {
      // 16 columns = one full 32-byte sector per lane: half the L1TEX sector operations of two 16-byte stores
      const int c = n0 + c0;
      uint32_t w[8];
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const float4 sc0 = *reinterpret_cast<const float4*>(s_scale + c + g * 8);
        const float4 sc1 = *reinterpret_cast<const float4*>(s_scale + c + g * 8 + 4);
        const float4 sh0 = *reinterpret_cast<const float4*>(s_shift + c + g * 8);
        const float4 sh1 = *reinterpret_cast<const float4*>(s_shift + c + g * 8 + 4);
        const float sc[8] = {sc0.x, sc0.y, sc0.z, sc0.w, sc1.x, sc1.y, sc1.z, sc1.w};
        const float sh[8] = {sh0.x, sh0.y, sh0.z, sh0.w, sh1.x, sh1.y, sh1.z, sh1.w};
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = ssa<ACT>(acc[g * 8 + j], sc[j], sh[j]);
        if (p.residual != nullptr) {
          float rr[8];
          unpack8(*reinterpret_cast<const uint4*>(p.residual + pix * p.res_ld + c + g * 8), rr);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += rr[j];
        }
        const uint4 pk = pack8(v);
        w[g * 4 + 0] = pk.x; w[g * 4 + 1] = pk.y; w[g * 4 + 2] = pk.z; w[g * 4 + 3] = pk.w;
      }
      st_global_v8(reinterpret_cast<__nv_bfloat16*>(p.out) + pix * p.out_ld + c, w);
      continue;
    }
#pragma unroll
    for (int g = 0; g < 2; ++g) {
      const int cl = c0 + g * 8;       // column inside the tile
      const int c = n0 + cl;           // output channel
      if (cl >= p.BN || c >= p.Cout) break;
      const int nv = min(8, p.Cout - c);
      long o_off, r_off;
      if (p.patch_k > 0) {
        const int ky = c / kC, rem = c - ky * kC;
        const long fp = ((long)n * p.Hout * p.patch_k + (long)h * p.patch_k + ky) * ((long)p.Wout * p.patch_k) +
                        (long)w * p.patch_k;
        o_off = fp * p.out_ld + rem;
        r_off = fp * p.res_ld + rem;
      } else {
        o_off = pix * p.out_ld + c;
        r_off = pix * p.res_ld + c;
      }
      const float4 sc0 = *reinterpret_cast<const float4*>(s_scale + c);
      const float4 sc1 = *reinterpret_cast<const float4*>(s_scale + c + 4);
      const float4 sh0 = *reinterpret_cast<const float4*>(s_shift + c);
      const float4 sh1 = *reinterpret_cast<const float4*>(s_shift + c + 4);
      const float sc[8] = {sc0.x, sc0.y, sc0.z, sc0.w, sc1.x, sc1.y, sc1.z, sc1.w};
      const float sh[8] = {sh0.x, sh0.y, sh0.z, sh0.w, sh1.x, sh1.y, sh1.z, sh1.w};
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        v[j] = ssa<ACT>(acc[g * 8 + j], sc[j], sh[j]);
      if (p.residual != nullptr) {
        if (p.res_vec_ok && nv == 8) {
          float rr[8];
          unpack8(*reinterpret_cast<const uint4*>(p.residual + r_off), rr);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += rr[j];
        } else {
          for (int j = 0; j < nv; ++j) v[j] += __bfloat162float(p.residual[r_off + j]);
        }
      }
      if (p.out_fp32) {
        float* o = reinterpret_cast<float*>(p.out) + o_off;
        if (p.vec_ok && nv == 8) {
          reinterpret_cast<float4*>(o)[0] = make_float4(v[0], v[1], v[2], v[3]);
          reinterpret_cast<float4*>(o)[1] = make_float4(v[4], v[5], v[6], v[7]);
        } else {
          for (int j = 0; j < nv; ++j) o[j] = v[j];
        }
      } else {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + o_off;
        if (p.vec_ok && nv == 8) {
          *reinterpret_cast<uint4*>(o) = pack8(v);
        } else {
          for (int j = 0; j < nv; ++j) o[j] = __float2bfloat16(v[j]);
        }
      }
    }
  }
}

// Detect epilogue (reference models/yolo.py:95-125): the 1x1 head's na*no <= 32 outputs of one pixel sit in one thread;
// it adds the bias and writes the raw logits in the (bs, na, ny, nx, no) layout the loss consumes and the decoded row
// (xy = (2s - 0.5 + grid) * stride, wh = (2s)^2 * anchor, rest = s, s = sigmoid) of the concatenated output, both fp32.
// Consecutive lanes are consecutive x, so each store instruction covers a contiguous run of the output.
// NO > 0: `no` known at compile time, so (anchor, field) of every unrolled column and the decode branch are static —
// the runtime version is ~170 instructions per output value and made this epilogue the bottleneck of the head.
template <int NO>
__device__ __forceinline__ void epilogue_detect(const ConvGemmParams& p, uint32_t trow, bool valid, long pix,
                                                const float* s_shift) {
  uint32_t acc[32];
  {
    uint32_t lo[16], hi[16];
    tmem_ld16(trow, lo);
    tmem_ld16(trow + 16u, hi);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 16; ++j) { acc[j] = lo[j]; acc[16 + j] = hi[j]; }
  }
  if (!valid) return;
  const int ny = p.det_ny, nx = p.det_nx, no = NO > 0 ? NO : p.det_no;
  const long plane = (long)ny * nx;
  const long n = pix / plane;
  const int rem = (int)(pix - n * plane);
  const int y = rem / nx, x = rem - y * nx;
  const long a_step = plane * no;
  float* raw = p.det_raw != nullptr ? p.det_raw + ((n * p.det_na * ny + y) * nx + x) * (long)no : nullptr;
  float* z = p.det_z + (n * p.det_rows_total + p.det_row_off + (long)y * nx + x) * (long)no;
  const float gx = ((float)x - 0.5f) * p.det_stride, gy = ((float)y - 0.5f) * p.det_stride, s2 = 2.f * p.det_stride;
  int ar = 0, jr = 0;
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    if (c < p.Cout) {      // warp-uniform
      const int a = NO > 0 ? c / NO : ar, j = NO > 0 ? c % NO : jr;
      const float v = __uint_as_float(acc[c]) + s_shift[c];
      if (raw != nullptr) raw[a * a_step + j] = v;
      const float sg = __frcp_rn(1.f + __expf(-v));
      float o;
      if (j == 0) o = fmaf(sg, s2, gx);
      else if (j == 1) o = fmaf(sg, s2, gy);
      else if (j == 2) o = (sg * 2.f) * (sg * 2.f) * (__ldg(p.det_anchors + a * 2) * p.det_stride);
      else if (j == 3) o = (sg * 2.f) * (sg * 2.f) * (__ldg(p.det_anchors + a * 2 + 1) * p.det_stride);
      else o = sg;
      z[a * a_step + j] = o;
      if (NO == 0 && ++jr == no) { jr = 0; ++ar; }
    }
  }
}

// Detect epilogue for na = 3, no = 6 when a warp's 32 pixels are one contiguous run of an image (plane % 32 == 0): the
// lanes lay raw and z out in shared memory exactly as they sit in global memory ([anchor][pixel][6] floats, 768 B per
// anchor) and six lanes issue one bulk copy each.  The scalar version's 36 four-byte stores per lane touch 24 sectors
// per instruction and kept L1TEX busy for most of the kernel.
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void epilogue_detect_bulk(const ConvGemmParams& p, uint32_t trow, bool valid, long pix,
                                                     const float* s_shift, uint8_t* slab, int lane) {
  uint32_t acc[32];
  {
    uint32_t lo[16], hi[16];
    tmem_ld16(trow, lo);
    tmem_ld16(trow + 16u, hi);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 16; ++j) { acc[j] = lo[j]; acc[16 + j] = hi[j]; }
  }
  if (lane < 6) tma_store_wait_read<0>();   // the previous tile's rows have left the slab
  __syncwarp();
  if (!valid) return;                       // warp-uniform: the pixel count is a multiple of 32
  const int ny = p.det_ny, nx = p.det_nx;
  const long plane = (long)ny * nx;
  const long n = pix / plane;
  const int rem = (int)(pix - n * plane);
  const int y = rem / nx, x = rem - y * nx;
  const float gx = ((float)x - 0.5f) * p.det_stride, gy = ((float)y - 0.5f) * p.det_stride, s2 = 2.f * p.det_stride;
  float* zs = reinterpret_cast<float*>(slab) + lane * 6;
  float* rs = zs + 3 * 32 * 6;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float v[6], o[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      v[j] = __uint_as_float(acc[a * 6 + j]) + s_shift[a * 6 + j];
      o[j] = __frcp_rn(1.f + __expf(-v[j]));
    }
    o[0] = fmaf(o[0], s2, gx);
    o[1] = fmaf(o[1], s2, gy);
    o[2] = (o[2] * 2.f) * (o[2] * 2.f) * (__ldg(p.det_anchors + a * 2) * p.det_stride);
    o[3] = (o[3] * 2.f) * (o[3] * 2.f) * (__ldg(p.det_anchors + a * 2 + 1) * p.det_stride);
#pragma unroll
    for (int j = 0; j < 6; j += 2) {
      *reinterpret_cast<float2*>(rs + a * 192 + j) = make_float2(v[j], v[j + 1]);
      *reinterpret_cast<float2*>(zs + a * 192 + j) = make_float2(o[j], o[j + 1]);
    }
  }
  fence_proxy_async_smem();
  __syncwarp();
  if (lane < 6) {
    const int a = lane % 3;
    const long rem0 = rem - lane;   // first pixel of this warp inside its image
    const float* src = reinterpret_cast<const float*>(slab) + lane * 192;   // lanes 0-2: z, lanes 3-5: raw
    float* dst = lane < 3 ? p.det_z + (n * p.det_rows_total + p.det_row_off + a * plane + rem0) * 6
                          : p.det_raw + ((n * 3 + a) * plane + rem0) * 6;
    bulk_store(dst, src, 768u);
    tma_store_commit();
  }
}

// TMA-store epilogue (flat 1x1, bf16, no residual scatter): each warp converts its 32 rows x 64 columns into a
// 128B-swizzled shared-memory slab and one lane issues a bulk tensor store — full 128 B lines reach L2 instead of
// 16 B pieces of 32 B sectors.
template <int ACT>
__device__ __forceinline__ void epilogue_cols_tma(const ConvGemmParams& p, uint32_t trow, int row0, int n0,
                                                  const float* s_scale, const float* s_shift, uint8_t* slab, int lane,
                                                  long pix, bool valid) {
  for (int c64 = 0; c64 < p.BN; c64 += 64) {
    if (lane == 0) tma_store_wait_read<0>();   // previous slab has left shared memory
    __syncwarp();
#pragma unroll
    for (int sub = 0; sub < 4; ++sub) {
      const int c0 = c64 + sub * 16;
      if (c0 >= p.BN) break;
      uint32_t acc[16];
      tmem_ld16(trow + (uint32_t)c0, acc);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int c = n0 + c0 + g * 8;
        float v[8];
        if (c < p.Cout) {
          const float4 sc0 = *reinterpret_cast<const float4*>(s_scale + c);
          const float4 sc1 = *reinterpret_cast<const float4*>(s_scale + c + 4);
          const float4 sh0 = *reinterpret_cast<const float4*>(s_shift + c);
          const float4 sh1 = *reinterpret_cast<const float4*>(s_shift + c + 4);
          const float sc[8] = {sc0.x, sc0.y, sc0.z, sc0.w, sc1.x, sc1.y, sc1.z, sc1.w};
          const float sh[8] = {sh0.x, sh0.y, sh0.z, sh0.w, sh1.x, sh1.y, sh1.z, sh1.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = ssa<ACT>(acc[g * 8 + j], sc[j], sh[j]);
          if (p.residual != nullptr && valid) {
            float rr[8];
            unpack8(*reinterpret_cast<const uint4*>(p.residual + pix * p.res_ld + c), rr);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] += rr[j];
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = 0.f;
        }
        const int chunk = sub * 2 + g;
        *reinterpret_cast<uint4*>(slab + lane * 128 + ((chunk ^ (lane & 7)) << 4)) = pack8(v);
      }
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(&p.o_map, slab, n0 + c64, row0);
      tma_store_commit();
    }
  }
}

// Persistent, warp-specialised kernel: one CTA per SM walks tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...
//   warp 0      TMA producer   (smem ring of `stages` {A box, B box} slots, continuous across tiles)
//   warp 1      MMA issuer     (accumulates tile i into TMEM buffer i & 1)
//   warps 2..17 epilogue       four groups of 4 warps (one per TMEM lane quadrant); group g drains tiles g, g+4, ...
//                              from TMEM buffer i % nbuf while the MMA warp fills the other buffers, so up to four
//                              tiles' TMEM->global latencies overlap on one SM
constexpr int kEpiGroups = 4;                 // epilogue warp groups, each drains a different tile
constexpr int kIssuer2Warp = 2 + kEpiGroups * 4;   // the second MMA-issuing warp sits after the epilogue warps
constexpr int kThreads = 64 + kEpiGroups * 128 + 32;
constexpr int kMaxTmemBufs = 4;

__global__ void __launch_bounds__(kThreads, 1) conv_gemm_kernel(const __grid_constant__ ConvGemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[kMaxStages];
  __shared__ uint64_t empty_bar[kMaxStages];
  __shared__ uint64_t b_ready_bar;
  __shared__ uint64_t tmem_full_bar[kMaxTmemBufs];
  __shared__ uint64_t tmem_empty_bar[kMaxTmemBufs];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_scale[kMaxCout];
  __shared__ __align__(16) float s_shift[kMaxCout];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;
  uint8_t* sB = smem + p.stages * p.kgroup * p.a_stage;
  const int b_stage = p.BN * 128;
  // store slab per epilogue warp, after the A/B buffers (1024-byte aligned: every piece above is)
  uint8_t* slabs = sB + (p.b_resident ? p.num_kb : p.stages * p.kgroup) * b_stage;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.b_map);
    tma_prefetch_desc(&p.a_map[0]);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&b_ready_bar, 1);
    for (int b = 0; b < kMaxTmemBufs; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 4 * (kEpiGroups / p.nbuf));
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_s, (uint32_t)p.tmem_cols);
  pdl_wait();   // everything above touched only this CTA's own state
  // folded BatchNorm / bias of every output channel, padded so that vector reads past Cout stay in bounds
  const int cpad = (p.Cout + 7) & ~7;
  for (int i = threadIdx.x; i < cpad; i += blockDim.x) {
    const float half = p.act == LY_ACT_SILU ? 0.5f : 1.f;   // see ssa<>
    s_scale[i] = half * ((p.scale != nullptr && i < p.Cout) ? p.scale[i] : 1.f);
    s_shift[i] = half * ((p.shift != nullptr && i < p.Cout) ? p.shift[i] : 0.f);
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_trigger();   // TMEM is ours: the next kernel's CTAs may come in behind the tail of this one
  const uint32_t tmem_base = tmem_base_s;

  const int tiles_m = p.tiles_w * p.tiles_h * p.n_img;
  const int num_tiles = tiles_m * p.tiles_n;
  const int my_tiles = ((int)blockIdx.x < num_tiles) ? (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  // The two single-lane roles below share their SM sub-partition with four busy epilogue warps, so every instruction of
  // theirs costs several issue slots: indices advance incrementally (no division), descriptors are base + offset.
  if (warp == 0) {
    // ---------------- TMA producer ----------------
    if (lane == 0 && my_tiles > 0) {
      CG_DBG_DECL
      if (p.b_resident) {   // weights once per CTA, not once per tile
        mbar_arrive_expect_tx(&b_ready_bar, (uint32_t)(p.num_kb * b_stage));
        for (int k = 0; k < p.num_kb; ++k) tma_load_2d(sB + k * b_stage, &p.b_map, &b_ready_bar, k * 64, 0);
      }
      const int kg = p.kgroup;
      const uint32_t tx = (uint32_t)kg * (p.a_bytes + (p.b_resident ? 0u : (uint32_t)b_stage));
      const int tiles_img = p.tiles_w * p.tiles_h;
      int tile = (int)blockIdx.x;
      // one ring per MMA-issuing lane (tile i -> ring i % issuers): s/ph of this tile's ring, s2/ph2 of the other
      const int rs = p.issuers == 2 ? p.stages >> 1 : p.stages;
      int s = 0, s2 = rs;
      uint32_t ph = 1, ph2 = 1;   // parity to wait for on the empty barrier of slot s
      for (int i = 0; i < my_tiles; ++i, tile += (int)gridDim.x) {
        const int mt = tile / p.tiles_n, nt = tile - mt * p.tiles_n;
        const int n = mt / tiles_img, tr = mt - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
        const int w0 = tw * p.Wt, h0 = th * p.Ht, n0 = nt * p.BN;
        for (int k = 0; k < p.num_k; k += kg) {
          CG_WAIT(0, &empty_bar[s], ph);
          mbar_arrive_expect_tx(&full_bar[s], tx);
          for (int j = 0; j < kg; ++j) {
            const KEntry e = p.sched[k + j];
            tma_load_5d(sA + (s * kg + j) * p.a_stage, &p.a_map[e.src], &full_bar[s], e.c0, w0 + e.dw, e.tz, h0 + e.dh, n);
            if (!p.b_resident) tma_load_2d(sB + (s * kg + j) * b_stage, &p.b_map, &full_bar[s], (k + j) * 64, n0);
          }
          ++s;
          if (s == rs) { s = 0; ph ^= 1u; }                       // ring 0 wraps
          else if (s == p.stages) { s = rs; ph ^= 1u; }           // ring 1 (or the only ring) wraps
        }
        if (p.issuers == 2) {   // next tile: the other ring
          const int ts = s; s = s2; s2 = ts;
          const uint32_t tp = ph; ph = ph2; ph2 = tp;
        }
      }
      CG_DBG_FLUSH(0, 1)
    }
  } else if (warp == 1 || warp == kIssuer2Warp) {
    // ---------------- MMA issuers ----------------
    // Two lanes on different SM sub-partitions take alternate tiles when p.issuers == 2 (tests/dev/mma_issue.cu: one
    // lane sustains an MMA every ~77 clk whatever N is, two lanes together ~50 clk at N = 64).  Each lane owns half of
    // the ring: an mbarrier parity wait is only safe for a consumer that sees every phase of its barrier.
    // `which` must be a compile-time constant: anything derived from threadIdx leaves the uniform datapath, and the
    // descriptors of every tcgen05.mma would then be moved from vector to uniform registers one by one (R2UR).
    auto issue = [&](auto which_c) {
    constexpr int which = decltype(which_c)::value;
    if (lane == 0 && my_tiles > which && which < p.issuers) {
      CG_DBG_DECL
      const uint32_t idesc = umma_idesc_bf16_f32(128, p.BN, 0, 0);
      const uint64_t a_desc0 = umma_smem_desc_sw128(smem_u32(sA), 16, 1024);
      const uint64_t b_desc0 = umma_smem_desc_sw128(smem_u32(sB), 16, 1024);
      const uint32_t a_step = (uint32_t)(p.a_stage >> 4), b_step = (uint32_t)(b_stage >> 4);
      if (p.b_resident) mbar_wait(&b_ready_bar, 0);
      const int kg = p.kgroup;
      const int istep = p.issuers;
      const int s_lo = (which && istep == 2) ? p.stages >> 1 : 0;           // this lane's ring: slots [s_lo, s_hi)
      const int s_hi = (!which && istep == 2) ? p.stages >> 1 : p.stages;
      int s = s_lo, buf = which;
      uint32_t ph = 0, use_par = 1;   // use_par: parity to wait for on tmem_empty_bar[buf]
      for (int i = which; i < my_tiles; i += istep) {
        CG_WAIT(0, &tmem_empty_bar[buf], use_par);   // epilogue has drained this accumulator
        tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + (uint32_t)(buf * p.BN);
        for (int k = 0; k < p.num_k; k += kg) {
          CG_WAIT(1, &full_bar[s], ph);
          tc_fence_after_sync();
          if (p.halo) {
            // nine taps from the one halo tile of this chunk: tap (dy, dx) starts (dy+1)*P + (dx+1) rows further in
            const KEntry e = p.sched[k];
            const uint64_t ahalo = a_desc0 + (uint64_t)((uint32_t)s * a_step);
            for (int t = 0; t < 9; ++t) {
              const int ty = t / 3, tx_ = t - ty * 3;
              const uint64_t adesc = ahalo + (uint64_t)((uint32_t)(ty * p.halo + tx_) * 8u);   // 128-byte rows
              const uint64_t bdesc = b_desc0 + (uint64_t)((uint32_t)(e.kb + t * e.kstep) * b_step);
              umma_bf16(tmem_d, adesc, bdesc, idesc, (uint32_t)((k | t) != 0));
              if (e.nk16 > 1) umma_bf16(tmem_d, adesc + 2, bdesc + 2, idesc, 1u);
              if (e.nk16 > 2) umma_bf16(tmem_d, adesc + 4, bdesc + 4, idesc, 1u);
              if (e.nk16 > 3) umma_bf16(tmem_d, adesc + 6, bdesc + 6, idesc, 1u);
            }
          } else
          for (int j = 0; j < kg; ++j) {
            const uint64_t adesc = a_desc0 + (uint64_t)((uint32_t)(s * kg + j) * a_step);
            const uint64_t bdesc = b_desc0 + (uint64_t)((uint32_t)(p.b_resident ? k + j : s * kg + j) * b_step);
            umma_bf16(tmem_d, adesc, bdesc, idesc, (uint32_t)((k | j) != 0));
            umma_bf16(tmem_d, adesc + 2, bdesc + 2, idesc, 1u);   // + 32 bytes = next 16 K elements inside the swizzle atom
            umma_bf16(tmem_d, adesc + 4, bdesc + 4, idesc, 1u);
            umma_bf16(tmem_d, adesc + 6, bdesc + 6, idesc, 1u);
          }
          umma_commit(&empty_bar[s]);
          if (++s == s_hi) { s = s_lo; ph ^= 1u; }
        }
        umma_commit(&tmem_full_bar[buf]);
        buf += istep;
        if (buf >= p.nbuf) { buf -= p.nbuf; use_par ^= 1u; }
      }
      if (which == 0) { CG_DBG_FLUSH(2, 2) }
    }
    };
    if (warp == 1) issue(std::integral_constant<int, 0>{});
    else issue(std::integral_constant<int, 1>{});
  } else {
    // ---------------- epilogue ----------------
    CG_DBG_DECL
    const int q = warp & 3;             // TMEM lane quadrant this warp may read
    // Group g serves TMEM buffer g % nbuf; when there are fewer buffers than groups (wide tiles), the groups that
    // share a buffer split its columns.  Each mbarrier is then waited on for consecutive phases only.
    const int grp = (warp - 2) >> 2;
    const int slot = grp % p.nbuf;
    const int part = grp / p.nbuf;
    const int nparts = kEpiGroups / p.nbuf;
    const int r = q * 32 + lane;
    const int pitch = p.halo ? p.halo : p.Wt;   // halo tiles keep the halo's pitch; columns >= Wt of a row are dead
    const int hh = r / pitch;
    const int ww = r - hh * pitch;
    for (int i = slot; i < my_tiles; i += p.nbuf) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int nt = tile % p.tiles_n, mt = tile / p.tiles_n;
      const int tw = mt % p.tiles_w, th = (mt / p.tiles_w) % p.tiles_h, n = mt / (p.tiles_w * p.tiles_h);
      const int h = th * p.Ht + hh, w = tw * p.Wt + ww, n0 = nt * p.BN;
      const bool valid = (ww < p.Wt) && (hh < p.Ht) && (h < p.Hout) && (w < p.Wout);
      const long pix = ((long)n * p.Hout + h) * p.Wout + w;
      const int buf = i % p.nbuf;
      const uint32_t use = (uint32_t)(i / p.nbuf);
      CG_WAIT(0, &tmem_full_bar[buf], use & 1u);
      tc_fence_after_sync();
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * p.BN);
      const int cb = part * 16, cs = nparts * 16;
      if (p.tma_store) {
        // flat rows: tile row r <-> pixel w; the whole tile's columns go through this warp's slab
        uint8_t* slab = slabs + (warp - 2) * p.slab_bytes;
        const int row0 = tw * p.Wt + q * 32;
        if (part == 0) {
          switch (p.act) {
            case LY_ACT_RELU: epilogue_cols_tma<LY_ACT_RELU>(p, trow, row0, n0, s_scale, s_shift, slab, lane, pix, valid); break;
            case LY_ACT_SILU: epilogue_cols_tma<LY_ACT_SILU>(p, trow, row0, n0, s_scale, s_shift, slab, lane, pix, valid); break;
            case LY_ACT_HSWISH: epilogue_cols_tma<LY_ACT_HSWISH>(p, trow, row0, n0, s_scale, s_shift, slab, lane, pix, valid); break;
            case LY_ACT_SIGMOID: epilogue_cols_tma<LY_ACT_SIGMOID>(p, trow, row0, n0, s_scale, s_shift, slab, lane, pix, valid); break;
            default: epilogue_cols_tma<LY_ACT_NONE>(p, trow, row0, n0, s_scale, s_shift, slab, lane, pix, valid); break;
          }
        }
      } else if (p.det_z != nullptr) {
        if (part == 0) {
          if (p.det_bulk) epilogue_detect_bulk(p, trow, valid, pix, s_shift, slabs + (warp - 2) * p.slab_bytes, lane);
          else if (p.det_no == 6) epilogue_detect<6>(p, trow, valid, pix, s_shift);
          else epilogue_detect<0>(p, trow, valid, pix, s_shift);
        }
      } else
      switch (p.act) {  // activation resolved once per tile: the column loop is branch-free
        case LY_ACT_RELU: epilogue_cols<LY_ACT_RELU>(p, trow, valid, n, h, w, pix, n0, s_scale, s_shift, cb, cs); break;
        case LY_ACT_SILU: epilogue_cols<LY_ACT_SILU>(p, trow, valid, n, h, w, pix, n0, s_scale, s_shift, cb, cs); break;
        case LY_ACT_HSWISH: epilogue_cols<LY_ACT_HSWISH>(p, trow, valid, n, h, w, pix, n0, s_scale, s_shift, cb, cs); break;
        case LY_ACT_SIGMOID: epilogue_cols<LY_ACT_SIGMOID>(p, trow, valid, n, h, w, pix, n0, s_scale, s_shift, cb, cs); break;
        default: epilogue_cols<LY_ACT_NONE>(p, trow, valid, n, h, w, pix, n0, s_scale, s_shift, cb, cs); break;
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
    }
    if ((p.tma_store && lane == 0) || (p.det_bulk && lane < 6)) tma_store_wait_all<0>();   // stores land before the CTA retires
    if (warp == 2) { CG_DBG_FLUSH(5, 1) }
    if (warp == 6) { CG_DBG_FLUSH(7, 1) }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
}


// ---------------------------------------------------------------------------------------------
// 3x3 convolutions whose weights do not fit in one SM's shared memory but whose HALF does (128 -> 128, 64 -> 256 ...):
// CTA PAIRS, cta_group::2.
// With one box per tap these layers move 590 KB from L2 per 128-pixel tile through a two-stage ring and the issuing
// lane waits on the ring most of the time (tensor pipe 40 % active, ncu).  Here two CTAs of a cluster (one TPC) share
// one accumulation M = 256 x N = Cout: each CTA keeps HALF of the weight matrix resident (Cout/2 rows x 9·Cin, 147 KB
// for 128 -> 128) and feeds its own 128 pixels from one halo box per 64-channel chunk (row-shifted descriptors, as in the
// resident-weight mode above), so a tile pair costs 2 barrier waits, 72 MMAs and 3 commits on ONE lane for both SMs and
// no weight traffic at all: 50 -> 38 us on the 40x40 layers.  (Measured and not kept: the same scheme for 1x1 layers
// with N = 256 — 256 -> 256 at 40x40 runs 34 us either way; those are bound by the fill and drain of six tiles per SM.)
//   both CTAs   warp 0: TMA lane (own halo boxes; completes on the LEADER's barrier), warps 2..17: epilogue of own rows
//   leader      warp 1: MMA lane, tcgen05.mma.cta_group::2; commits are multicast to both CTAs' barriers
// Every load here reads this GPU's memory; none is a multicast load.
constexpr int kC2Threads = 64 + kEpiGroups * 128;
constexpr int kC2MaxCout = 256;
constexpr int kC2Slots = 8;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t* dst_smem, uint32_t ncols) {   // same warp id in both CTAs
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16_2cta(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs once the MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3)
               : "memory");
}
// TMA loads into this CTA's shared memory whose bytes complete on the barrier at the same offset in the LEADER
// (bit 24 of a shared::cluster address is the CTA's rank inside the pair)
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;
__device__ __forceinline__ void tma_load_2d_2cta(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d_2cta(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                                 int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, "
      "%6, %7}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1),
      "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {   // barrier at this offset in CTA `cta`
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}"
      ::"r"(smem_u32(bar)), "r"(cta)
      : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {   // arrivals may come from the peer CTA
  const long long t0 = clock64();
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return;
    if (clock64() - t0 > 4000000000LL) __trap();
  }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kC2Threads, 1)
    conv_cta2_kernel(const __grid_constant__ ConvGemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t a_full[kC2Slots], a_empty[kC2Slots], b_ready_bar;
  __shared__ uint64_t tmem_full_bar[4], tmem_empty_bar[4];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_scale[kC2MaxCout];
  __shared__ __align__(16) float s_shift[kC2MaxCout];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;                                  // `stages` halo boxes
  uint8_t* sB = smem + p.stages * p.a_stage;           // num_kb x [BN/2][64] weight tiles: this CTA's half of the rows
  const int b_half = (p.BN >> 1) * 128;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.b_map);
    tma_prefetch_desc(&p.a_map[0]);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    mbar_init(&b_ready_bar, 1);
    for (int b = 0; b < 4; ++b) {   // drained by (4 warps per group) x (groups per accumulator) x (2 CTAs)
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 8 * (kEpiGroups / p.nbuf));
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_2cta(&tmem_base_s, (uint32_t)p.tmem_cols);
  pdl_wait();
  const int cpad = (p.Cout + 7) & ~7;
  for (int i = threadIdx.x; i < cpad; i += blockDim.x) {
    const float half = p.act == LY_ACT_SILU ? 0.5f : 1.f;   // see ssa<>
    s_scale[i] = half * ((p.scale != nullptr && i < p.Cout) ? p.scale[i] : 1.f);
    s_shift[i] = half * ((p.shift != nullptr && i < p.Cout) ? p.shift[i] : 0.f);
  }
  tc_fence_before_sync();
  __syncthreads();
  cluster_sync_all();   // the peer's barriers exist before anything is sent to them
  tc_fence_after_sync();
  pdl_trigger();
  const uint32_t tmem_base = tmem_base_s;

  const int tiles_img = p.tiles_w * p.tiles_h;
  const int tiles_m = tiles_img * p.n_img;
  const int units = (tiles_m + 1) >> 1;                 // unit u = tiles 2u (leader) and 2u + 1 (peer)
  const int cid = (int)blockIdx.x >> 1, ncl = (int)gridDim.x >> 1;
  const int mine = cid < units ? (units - cid + ncl - 1) / ncl : 0;

  if (warp == 0) {
    // ---------------- TMA producer (both CTAs) ----------------
    if (lane == 0 && mine > 0) {
      if (rank == 0) mbar_arrive_expect_tx(&b_ready_bar, (uint32_t)(2 * p.num_kb * b_half));
      for (int k = 0; k < p.num_kb; ++k)
        tma_load_2d_2cta(sB + k * b_half, &p.b_map, &b_ready_bar, k * 64, (int)rank * (p.BN >> 1));
      int s = 0;
      uint32_t ph = 1;
      int u = cid;
      for (int i = 0; i < mine; ++i, u += ncl) {
        const int m = 2 * u + (int)rank;
        int n = p.n_img, h0 = 0, w0 = 0;                // a unit's missing second tile: out of bounds = zero fill
        if (m < tiles_m) {
          n = m / tiles_img;
          const int tr = m - n * tiles_img, th = tr / p.tiles_w;
          h0 = th * p.Ht;
          w0 = (tr - th * p.tiles_w) * p.Wt;
        }
        for (int k = 0; k < p.num_k; ++k) {
          const KEntry e = p.sched[k];
          mbar_wait(&a_empty[s], ph);
          if (rank == 0) mbar_arrive_expect_tx(&a_full[s], 2u * p.a_bytes);
          tma_load_5d_2cta(sA + s * p.a_stage, &p.a_map[e.src], &a_full[s], e.c0, w0 + e.dw, e.tz, h0 + e.dh, n);
          if (++s == p.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer (leader only) ----------------
    if (lane == 0 && rank == 0 && mine > 0) {
      const uint32_t idesc = umma_idesc_bf16_f32(256, p.BN, 0, 0);
      const uint64_t a_desc0 = umma_smem_desc_sw128(smem_u32(sA), 16, 1024);
      const uint64_t b_desc0 = umma_smem_desc_sw128(smem_u32(sB), 16, 1024);
      const uint32_t a_step = (uint32_t)(p.a_stage >> 4), b_step = (uint32_t)(b_half >> 4);
      mbar_wait_cluster(&b_ready_bar, 0);
      int s = 0;
      uint32_t ph = 0;
      int buf = 0;
      uint32_t use_par = 1;   // parity to wait for on tmem_empty_bar[buf]
      for (int i = 0; i < mine; ++i) {
        mbar_wait_cluster(&tmem_empty_bar[buf], use_par);
        tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + (uint32_t)(buf * p.BN);
        for (int k = 0; k < p.num_k; ++k) {
          mbar_wait_cluster(&a_full[s], ph);
          tc_fence_after_sync();
          // nine taps from the one halo box of this chunk: tap (dy, dx) starts (dy+1)*P + (dx+1) rows further in
          const KEntry e = p.sched[k];
          const uint64_t ahalo = a_desc0 + (uint64_t)((uint32_t)s * a_step);
          for (int t = 0; t < 9; ++t) {
            const int ty = t / 3, tx = t - ty * 3;
            const uint64_t adesc = ahalo + (uint64_t)((uint32_t)(ty * p.halo + tx) * 8u);   // 128-byte rows
            const uint64_t bdesc = b_desc0 + (uint64_t)((uint32_t)(e.kb + t * e.kstep) * b_step);
            umma_bf16_2cta(tmem_d, adesc, bdesc, idesc, (uint32_t)((k | t) != 0));
            if (e.nk16 > 1) umma_bf16_2cta(tmem_d, adesc + 2, bdesc + 2, idesc, 1u);
            if (e.nk16 > 2) umma_bf16_2cta(tmem_d, adesc + 4, bdesc + 4, idesc, 1u);
            if (e.nk16 > 3) umma_bf16_2cta(tmem_d, adesc + 6, bdesc + 6, idesc, 1u);
          }
          umma_commit_2cta(&a_empty[s]);
          if (++s == p.stages) { s = 0; ph ^= 1u; }
        }
        umma_commit_2cta(&tmem_full_bar[buf]);
        if (++buf == p.nbuf) { buf = 0; use_par ^= 1u; }
      }
    }
  } else {
    // ---------------- epilogue (both CTAs, own 128 rows) ----------------
    const int q = warp & 3;
    const int grp = (warp - 2) >> 2;
    const int slot = grp % p.nbuf;         // TMEM buffer this group drains; groups that share one split its columns
    const int part = grp / p.nbuf, nparts = kEpiGroups / p.nbuf;
    const int r = q * 32 + lane;
    const int hh = r / p.halo, ww = r - hh * p.halo;
    for (int i = slot; i < mine; i += p.nbuf) {
      const int m = 2 * (cid + i * ncl) + (int)rank;
      const int n = m / tiles_img, tr = m - n * tiles_img, th = tr / p.tiles_w, tw = tr - th * p.tiles_w;
      const int h = th * p.Ht + hh, w = tw * p.Wt + ww;
      const bool valid = (m < tiles_m) && (ww < p.Wt) && (hh < p.Ht) && (h < p.Hout) && (w < p.Wout);
      const long pix = ((long)n * p.Hout + h) * p.Wout + w;
      mbar_wait(&tmem_full_bar[slot], (uint32_t)(i / p.nbuf) & 1u);
      tc_fence_after_sync();
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(slot * p.BN);
      const int cb = part * 16, cs = nparts * 16;
      switch (p.act) {
        case LY_ACT_RELU: epilogue_cols<LY_ACT_RELU>(p, trow, valid, n, h, w, pix, 0, s_scale, s_shift, cb, cs); break;
        case LY_ACT_SILU: epilogue_cols<LY_ACT_SILU>(p, trow, valid, n, h, w, pix, 0, s_scale, s_shift, cb, cs); break;
        case LY_ACT_HSWISH: epilogue_cols<LY_ACT_HSWISH>(p, trow, valid, n, h, w, pix, 0, s_scale, s_shift, cb, cs); break;
        case LY_ACT_SIGMOID: epilogue_cols<LY_ACT_SIGMOID>(p, trow, valid, n, h, w, pix, 0, s_scale, s_shift, cb, cs); break;
        default: epilogue_cols<LY_ACT_NONE>(p, trow, valid, n, h, w, pix, 0, s_scale, s_shift, cb, cs); break;
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tmem_empty_bar[slot], 0u);   // the leader's lane waits for both CTAs
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  cluster_sync_all();   // no CTA of the pair leaves while the other may still signal it or read its shared memory
  if (warp == 2) tmem_dealloc_2cta(tmem_base, (uint32_t)p.tmem_cols);
}

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
struct ConvGeom {
  bool flat;       // 1x1: rows are the flattened pixels
  bool patch;      // kernel == stride, no padding
  int taps;        // K taps per source
  int Hout, Wout;
};

static int conv_geometry(const ly_conv_desc* d, ConvGeom* g) {
  LY_REQUIRE(d != nullptr, "null descriptor");
  LY_REQUIRE(d->n_src >= 1 && d->n_src <= 4, "n_src=%d out of range", d->n_src);
  LY_REQUIRE(d->N > 0 && d->H > 0 && d->W > 0 && d->Cout > 0, "bad sizes N=%d H=%d W=%d Cout=%d", d->N, d->H,
             d->W, d->Cout);
  g->flat = g->patch = false;
  if (d->kh == 1 && d->kw == 1 && d->stride == 1 && d->pad == 0) {
    g->flat = true;
    g->taps = 1;
    g->Hout = d->H;
    g->Wout = d->W;
  } else if (d->kh == 3 && d->kw == 3 && d->stride == 1 && d->pad == 1) {
    g->taps = 9;
    g->Hout = d->H;
    g->Wout = d->W;
  } else if (d->kh == d->kw && d->stride == d->kh && d->pad == 0 && (d->kh == 2 || d->kh == 4)) {
    LY_REQUIRE(d->H % d->kh == 0 && d->W % d->kw == 0, "patch conv needs H,W divisible by k=%d", d->kh);
    g->patch = true;
    g->taps = d->kh;
    g->Hout = d->H / d->kh;
    g->Wout = d->W / d->kw;
  } else {
    return fail(LY_EUNSUPPORTED, "conv geometry k=%dx%d s=%d p=%d not supported", d->kh, d->kw, d->stride,
                d->pad);
  }
  for (int s = 0; s < d->n_src; ++s) {
    LY_REQUIRE(d->src[s] != nullptr, "source %d is null", s);
    LY_REQUIRE(d->src_c[s] > 0 && d->src_c[s] % 8 == 0, "source %d: channels %d must be a multiple of 8", s,
               d->src_c[s]);
    LY_REQUIRE(d->src_ld[s] >= d->src_c[s] && d->src_ld[s] % 8 == 0, "source %d: ld %ld invalid", s,
               d->src_ld[s]);
    if (g->patch) LY_REQUIRE(d->src_ld[s] == d->src_c[s], "patch conv needs a dense source (ld == C)");
  }
  return LY_OK;
}

static int conv_chunks(const ly_conv_desc* d, const ConvGeom& g) {
  int n = 0;
  for (int s = 0; s < d->n_src; ++s) {
    const int ceff = g.patch ? d->src_c[s] * d->kw : d->src_c[s];
    n += g.taps * ceil_div(ceff, 64);
  }
  return n;
}

static void pick_tile(int Hout, int Wout, int* Wt, int* Ht) {
  double best = -1.0;
  int cand[7] = {8, 16, 32, 64, 128, Wout <= 128 ? Wout : 128, 0};
  for (int i = 0; i < 6; ++i) {
    int wt = cand[i];
    if (wt <= 0 || wt > 128) continue;
    int ht = 128 / wt;
    if (ht > Hout) ht = Hout;
    if (ht < 1) ht = 1;
    double eff = (double)Hout * Wout / ((double)ceil_div(Wout, wt) * ceil_div(Hout, ht) * 128.0);
    if (eff > best + 1e-9 || (eff > best - 1e-9 && wt > *Wt)) {
      best = eff;
      *Wt = wt;
      *Ht = ht;
    }
  }
}

static int set_smem_attr_once() {
  // per device; the flag is set only after success, so a failed attempt is reported to every caller and retried
  static std::atomic<bool> done[64];
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (!done[dev & 63].load()) {
    LY_CUDA(cudaFuncSetAttribute(conv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    done[dev & 63].store(true);
  }
  return LY_OK;
}

}  // namespace ly

using namespace ly;

extern "C" int ly_conv_num_chunks(const ly_conv_desc* d) {
  ConvGeom g;
  int rc = conv_geometry(d, &g);
  if (rc != LY_OK) return rc;
  return conv_chunks(d, g);
}

namespace ly {
struct DetectArgs {
  int na, no;
  float stride;
  const float* anchors;        // device, na x (w, h) in stride units
  float* raw;
  float* z;
  long rows_total, row_off;
};
}  // namespace ly

static int conv_launch(const ly_conv_desc* d, const ly::DetectArgs* det, ly_stream_t stream) {
  ConvGeom g;
  int rc = conv_geometry(d, &g);
  if (rc != LY_OK) return rc;
  LY_REQUIRE(d->w_packed != nullptr && d->out != nullptr, "null weight/out pointer");
  const int num_k = conv_chunks(d, g);
  LY_REQUIRE(num_k <= kMaxK, "K too large: %d chunks of 64 (max %d)", num_k, kMaxK);
  if (d->patch_k > 0) {
    LY_REQUIRE(g.flat, "patch scatter output needs a 1x1 GEMM");
    LY_REQUIRE(d->patch_c % 8 == 0 && d->Cout == d->patch_k * d->patch_k * d->patch_c,
               "patch scatter: Cout=%d != k*k*c", d->Cout);
    LY_REQUIRE(d->out_ld == d->patch_c && (d->residual == nullptr || d->res_ld == d->patch_c),
               "patch scatter needs dense output");
  }

  static_assert(sizeof(ConvGemmParams) <= 4000, "kernel parameter block too large");
  ConvGemmParams p;
  memset(&p, 0, sizeof(p));

  // ---- tile geometry
  const bool flat_rows = g.flat;
  if (flat_rows) {
    const long M = (long)d->N * d->H * d->W;
    p.n_img = 1;
    p.Hout = 1;
    p.Wout = (int)M;
    p.Wt = 128;
    p.Ht = 1;
    if (d->patch_k > 0) {
      // scatter addressing needs (n, h, w): tile inside one image row-block instead
      p.n_img = d->N;
      p.Hout = d->H;
      p.Wout = d->W;
      pick_tile(p.Hout, p.Wout, &p.Wt, &p.Ht);
    }
  } else {
    p.n_img = d->N;
    p.Hout = g.Hout;
    p.Wout = g.Wout;
    pick_tile(p.Hout, p.Wout, &p.Wt, &p.Ht);
  }
  // 3x3 with weights that stay resident in shared memory: ONE halo box per 64-channel chunk feeds all nine taps
  // (row-shifted descriptors), which cuts the L2 -> SM activation traffic ~6x.  The tile keeps the halo's pitch.
  p.a_stage = kAStage;
  p.num_kb = num_k;
  bool cta2 = false;
  constexpr long kC2Budget = 224 * 1024 - 1024;   // dynamic shared memory of conv_cta2_kernel after alignment
  {
    static const int no_halo = getenv("LY_CONV_NO_HALO") ? 1 : 0;   // dev knob
    const int cr = (d->Cout + 15) / 16 * 16;
    if (!no_halo && g.taps == 9 && !g.patch && cr <= 256 && (long)num_k * cr * 128 <= 100 * 1024) {
      double best = -1;
      const int pitches[3] = {16, 32, 64};
      for (int i = 0; i < 3; ++i) {
        const int wt = pitches[i] - 2, ht = 128 / pitches[i];
        const double eff = (double)p.Hout * p.Wout / ((double)ceil_div(p.Wout, wt) * ceil_div(p.Hout, ht) * 128.0);
        if (eff > best + 1e-9) { best = eff; p.halo = pitches[i]; }
      }
      p.Wt = p.halo - 2;
      p.Ht = 128 / p.halo;
      p.a_stage = ((130 + 2 * p.halo) * 128 + 1023) / 1024 * 1024;   // box rows + what a shifted 128-row read runs past
    }
    // weights too large for one SM but half of them fit: CTA pairs (conv_cta2_kernel), each CTA keeps Cout/2 rows
    static const int no_cta2 = getenv("LY_CONV_NO_CTA2") ? 1 : 0;   // dev knob
    const long b_half_all = (long)num_k * (cr / 2) * 128;
    if (!p.halo && !no_cta2 && !g.patch && det == nullptr && d->patch_k == 0 && cr >= 32 && cr <= kC2MaxCout &&
        d->Cout <= kC2MaxCout && (long)num_k * cr * 128 > 100 * 1024) {
      if (g.taps == 9) {
        double best = -1;
        const int pitches[3] = {16, 32, 64};
        for (int i = 0; i < 3; ++i) {
          const int wt = pitches[i] - 2, ht = 128 / pitches[i];
          const long a_st = ((130 + 2 * pitches[i]) * 128 + 1023) / 1024 * 1024;
          if (b_half_all + 2 * a_st > kC2Budget) continue;
          const double eff = (double)p.Hout * p.Wout / ((double)ceil_div(p.Wout, wt) * ceil_div(p.Hout, ht) * 128.0);
          if (eff > best + 1e-9) { best = eff; p.halo = pitches[i]; }
        }
        if (p.halo) {
          p.Wt = p.halo - 2;
          p.Ht = 128 / p.halo;
          p.a_stage = ((130 + 2 * p.halo) * 128 + 1023) / 1024 * 1024;
          cta2 = true;
        }
      }
    }
  }
  p.tiles_w = ceil_div(p.Wout, p.Wt);
  p.tiles_h = ceil_div(p.Hout, p.Ht);
  p.a_bytes = p.halo ? (uint32_t)(64 * p.halo * (p.Ht + 2) * 2) : (uint32_t)(64 * p.Wt * p.Ht * 2);

  // ---- A tensor maps + K schedule
  int it = 0, kb_base = 0;
  for (int s = 0; s < d->n_src; ++s) {
    const long ld = d->src_ld[s];
    uint64_t dims[5], str[4];
    uint32_t box[5] = {64, (uint32_t)(p.halo ? p.halo : p.Wt), 1, (uint32_t)(p.halo ? p.Ht + 2 : p.Ht), 1};
    int ceff = d->src_c[s];
    if (flat_rows && d->patch_k == 0) {
      const uint64_t M = (uint64_t)d->N * d->H * d->W;
      dims[0] = (uint64_t)d->src_c[s]; dims[1] = M; dims[2] = 1; dims[3] = 1; dims[4] = 1;
      str[0] = (uint64_t)ld * 2; str[1] = M * ld * 2; str[2] = str[1]; str[3] = str[1];
    } else if (g.patch) {
      const int k = d->kh;
      ceff = d->src_c[s] * k;
      dims[0] = (uint64_t)ceff; dims[1] = (uint64_t)g.Wout; dims[2] = (uint64_t)k; dims[3] = (uint64_t)g.Hout;
      dims[4] = (uint64_t)d->N;
      str[0] = (uint64_t)ceff * 2;
      str[1] = (uint64_t)d->W * ld * 2;
      str[2] = (uint64_t)k * d->W * ld * 2;
      str[3] = (uint64_t)d->H * d->W * ld * 2;
    } else {
      dims[0] = (uint64_t)d->src_c[s]; dims[1] = (uint64_t)d->W; dims[2] = 1; dims[3] = (uint64_t)d->H;
      dims[4] = (uint64_t)d->N;
      str[0] = (uint64_t)ld * 2;
      str[1] = (uint64_t)d->W * ld * 2;
      str[2] = str[1];
      str[3] = (uint64_t)d->H * d->W * ld * 2;
    }
    rc = tmap_encode_bf16(&p.a_map[s], d->src[s], 5, dims, str, box);
    if (rc != LY_OK) return rc;
    if (p.halo) {   // one entry per chunk; its nine weight chunks sit kstep apart in the packed K order (tap-major)
      const int nch = ceil_div(ceff, 64);
      for (int c = 0; c < nch; ++c) {
        KEntry e;
        e.src = (int16_t)s; e.c0 = (int16_t)(c * 64); e.dw = -1; e.dh = -1; e.tz = 0;
        e.nk16 = (int8_t)ceil_div(ceff - c * 64 < 64 ? ceff - c * 64 : 64, 16);
        e.kb = (int16_t)(kb_base + c); e.kstep = (int16_t)nch;
        p.sched[it++] = e;
      }
      kb_base += 9 * nch;
      continue;
    }
    for (int t = 0; t < g.taps; ++t) {
      for (int c0 = 0; c0 < ceff; c0 += 64) {
        KEntry e;
        e.src = (int16_t)s;
        e.c0 = (int16_t)c0;
        e.nk16 = 4;
        e.kb = 0; e.kstep = 0;
        if (g.patch) {
          e.dw = 0; e.dh = 0; e.tz = (int8_t)t;
        } else {
          e.dw = (int8_t)(t % d->kw - d->pad);
          e.dh = (int8_t)(t / d->kw - d->pad);
          e.tz = 0;
        }
        p.sched[it++] = e;
      }
    }
  }
  p.num_k = it;

  // ---- N tiling, B map
  const int cout_r = (d->Cout + 15) / 16 * 16;
  const int ny = ceil_div(cout_r, 256);
  p.BN = (ceil_div(cout_r, ny) + 15) / 16 * 16;
  p.Cout = d->Cout;
  p.tiles_n = ny;
  LY_REQUIRE(d->Cout <= kMaxCout, "Cout=%d exceeds %d", d->Cout, kMaxCout);
  p.nbuf = p.BN <= 128 ? 4 : 2;   // accumulators in flight; must divide kEpiGroups
  p.tmem_cols = 32;
  while (p.tmem_cols < p.nbuf * p.BN) p.tmem_cols *= 2;
  {
    uint64_t dims[2] = {(uint64_t)num_k * 64, (uint64_t)d->Cout};
    uint64_t str[1] = {(uint64_t)num_k * 64 * 2};
    uint32_t box[2] = {64, (uint32_t)(cta2 ? p.BN / 2 : p.BN)};   // CTA pairs: each CTA loads its half of the rows
    rc = tmap_encode_bf16(&p.b_map, d->w_packed, 2, dims, str, box);
    if (rc != LY_OK) return rc;
  }
  // one persistent CTA per SM.  When the whole packed weight matrix of this launch is small (single N tile), keep it
  // resident in shared memory: the TMA ring then carries activations only and the L2 -> SM traffic per tile halves.
  const uintptr_t oa = reinterpret_cast<uintptr_t>(d->out);
  p.vec_ok = d->out_fp32 ? ((oa % 16 == 0) && (d->out_ld % 4 == 0)) : ((oa % 16 == 0) && (d->out_ld % 8 == 0));
  p.res_vec_ok = d->residual != nullptr && (reinterpret_cast<uintptr_t>(d->residual) % 16 == 0) &&
                 (d->res_ld % 8 == 0);
  p.vec32_ok = (!d->out_fp32 && d->patch_k == 0 && oa % 32 == 0 && d->out_ld % 16 == 0 &&
                (d->residual == nullptr || p.res_vec_ok)) ? 1 : 0;
  // TMA-store epilogue: flat rows, bf16, dense-enough output, no scatter; needs nparts == 1 (one group per tile).
  // (Wide tiles, two groups per accumulator: the 64 KB of slabs cost ring stages and the layers got slower — measured.)
  p.tma_store = (flat_rows && d->patch_k == 0 && !d->out_fp32 && p.vec_ok && p.nbuf == kEpiGroups &&
                 (d->residual == nullptr || p.res_vec_ok)) ? 1 : 0;
  if (det != nullptr) {
    // bulk-copy epilogue: na = 3, no = 6, every warp's 32 pixels inside one image, 16-byte aligned runs
    const long plane = (long)d->H * d->W;
    p.tma_store = 0;
    p.det_bulk = (flat_rows && det->na == 3 && det->no == 6 && det->raw != nullptr && plane % 32 == 0 &&
                  det->rows_total % 2 == 0 && det->row_off % 2 == 0 && reinterpret_cast<uintptr_t>(det->z) % 16 == 0 &&
                  reinterpret_cast<uintptr_t>(det->raw) % 16 == 0 && p.nbuf == kEpiGroups) ? 1 : 0;
  }
  p.slab_bytes = p.tma_store ? 4096 : (p.det_bulk ? 2 * 3 * 32 * 6 * 4 : 0);
  if (p.tma_store) {
    const uint64_t M = (uint64_t)d->N * d->H * d->W;
    uint64_t dims[2] = {(uint64_t)d->Cout, M};
    uint64_t str[1] = {(uint64_t)d->out_ld * 2};
    uint32_t box[2] = {64, 32};
    rc = tmap_encode_bf16(&p.o_map, d->out, 2, dims, str, box);
    if (rc != LY_OK) return rc;
  }
  const int budget = 196 * 1024 - kEpiGroups * 4 * p.slab_bytes;
  const long b_all = (long)num_k * p.BN * 128;
  p.b_resident = (ny == 1 && b_all <= 100 * 1024) ? 1 : 0;
  // K chunks per ring stage: the largest divisor of num_k (<= 4) that still leaves a ring of >= 3 stages (>= 2 when
  // one chunk alone is big).  Fewer, larger stages amortise the commit / mbarrier round trips of the single-lane roles.
  if (cta2) {
    const long b_half_all = (long)num_k * (p.BN / 2) * 128;
    p.b_resident = 1;
    p.tma_store = 0;
    p.slab_bytes = 0;
    p.stages = (int)((kC2Budget - b_half_all) / p.a_stage);
    if (p.stages > kC2Slots) p.stages = kC2Slots;
    p.kgroup = 1;
    LY_REQUIRE(p.stages >= 2 && ny == 1 && p.nbuf * p.BN <= 512, "CTA-pair mode: bad configuration (BN=%d)", p.BN);
    p.issuers = 1;
    p.act = d->act; p.scale = d->scale; p.shift = d->shift;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(d->residual);
    p.res_ld = d->res_ld; p.out = d->out; p.out_ld = d->out_ld; p.out_fp32 = d->out_fp32;
    const long tiles_c = (long)p.tiles_w * p.tiles_h * p.n_img;
    static const bool verbose_c = getenv("LY_CONV_VERBOSE") != nullptr;
    if (verbose_c)
      fprintf(stderr, "[conv cta2] BN=%d chunks=%d weight chunks=%d stages=%d halo=%d a_stage=%d tiles=%ld\n", p.BN,
              p.num_k, p.num_kb, p.stages, p.halo, p.a_stage, tiles_c);
    static std::atomic<bool> c2_attr[64];
    int dev = 0;
    LY_CUDA(cudaGetDevice(&dev));
    if (!c2_attr[dev & 63].load()) {
      LY_CUDA(cudaFuncSetAttribute(conv_cta2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
      c2_attr[dev & 63].store(true);
    }
    const size_t smem_c = (size_t)p.stages * p.a_stage + (size_t)b_half_all + 1024;
    const long units = (tiles_c + 1) / 2;
    const long ncl = units < device_sm_count() / 2 ? units : device_sm_count() / 2;
    conv_cta2_kernel<<<dim3((unsigned)(2 * ncl)), dim3(kC2Threads), smem_c, (cudaStream_t)stream>>>(p);   // cluster (2,1,1)
    LY_LAUNCH_CHECK();
    return LY_OK;
  }
  LY_REQUIRE(!p.halo || p.b_resident, "halo tiles need resident weights");
  const long chunk_bytes = p.a_stage + (p.b_resident ? 0 : p.BN * 128);
  const long ring_budget = budget - (p.b_resident ? b_all : 0);
  p.kgroup = 1;
  static const int g_max = getenv("LY_CONV_KGROUP_MAX") ? atoi(getenv("LY_CONV_KGROUP_MAX")) : 4;     // dev knobs
  static const int min_stages = getenv("LY_CONV_MIN_STAGES") ? atoi(getenv("LY_CONV_MIN_STAGES")) : 2;
  for (int g = g_max; g >= 2 && !p.halo; --g) {
    if (num_k % g == 0 && ring_budget / (g * chunk_bytes) >= min_stages) { p.kgroup = g; break; }
  }
  p.stages = (int)(ring_budget / (p.kgroup * chunk_bytes));
  if (p.stages > kMaxStages) p.stages = kMaxStages;
  LY_REQUIRE(p.stages >= 2, "shared-memory ring too small (BN=%d)", p.BN);
  // two issuing lanes when each can have a ring that holds a whole tile (>= 2 stages) and one lane alone is
  // issue-bound (BN <= 128, which also means four accumulators).  With rings shorter than a tile the [128,40]->128
  // layer went from 53 to 71 us — measured, so the rule stays strict.
  static const bool one_issuer = getenv("LY_CONV_ONE_ISSUER") != nullptr;   // dev knob
  p.issuers = (!one_issuer && p.stages >= 4 && p.BN <= 128 && p.stages / 2 >= p.num_k / p.kgroup) ? 2 : 1;
  if (p.issuers == 2) p.stages &= ~1;
  static const bool verbose = getenv("LY_CONV_VERBOSE") != nullptr;   // dev knob
  if (verbose)
    fprintf(stderr, "[conv] BN=%d num_k=%d kgroup=%d stages=%d issuers=%d halo=%d b_resident=%d a_stage=%d\n", p.BN, num_k,
            p.kgroup, p.stages, p.issuers, p.halo, p.b_resident, p.a_stage);
  const size_t smem_bytes = (size_t)kEpiGroups * 4 * p.slab_bytes + (p.b_resident ? (size_t)b_all : 0) +
                            (size_t)p.stages * p.kgroup * chunk_bytes + 1024;

  // ---- epilogue
  p.act = d->act;
  p.scale = d->scale;
  p.shift = d->shift;
  p.residual = reinterpret_cast<const __nv_bfloat16*>(d->residual);
  p.res_ld = d->res_ld;
  p.out = d->out;
  p.out_ld = d->out_ld;
  p.out_fp32 = d->out_fp32;
  p.patch_k = d->patch_k;
  p.patch_c = d->patch_c;
  if (det != nullptr) {
    LY_REQUIRE(flat_rows && d->patch_k == 0 && p.tiles_n == 1 && d->Cout == det->na * det->no && d->Cout <= 32 &&
                   det->na <= 4 && d->residual == nullptr,
               "detect head: needs a 1x1 convolution with na*no <= 32 outputs (got %d)", d->Cout);
    p.det_z = det->z; p.det_raw = det->raw;
    p.det_rows_total = det->rows_total; p.det_row_off = det->row_off;
    p.det_na = det->na; p.det_no = det->no; p.det_ny = d->H; p.det_nx = d->W;
    p.det_stride = det->stride;
    p.det_anchors = det->anchors;
  }

  rc = set_smem_attr_once();
  if (rc != LY_OK) return rc;
  const long num_tiles = (long)p.tiles_w * p.tiles_h * p.n_img * ny;
  const int sms = device_sm_count();
  dim3 grid((unsigned)(num_tiles < sms ? num_tiles : sms), 1, 1);
  LY_CUDA(launch(conv_gemm_kernel, grid, kThreads, smem_bytes, (cudaStream_t)stream, p));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_conv_fwd(const ly_conv_desc* d, ly_stream_t stream) { return conv_launch(d, nullptr, stream); }

// Detect head of one level at inference (reference models/yolo.py:79,90,95-125): 1x1 convolution Cin -> na*no with bias
// on the tensor cores whose epilogue writes raw [N][na][H][W][no] (may be NULL) and the decoded rows
// z[N][rows_total][no] at row_off + a*H*W + y*W + x, both fp32.  x: NHWC bf16 (pixel stride ldx), w_packed as for
// ly_conv_fwd, bias fp32 [na*no], anchors fp32 [na][2] in stride units (device).  na*no <= 32, na <= 4.
extern "C" int ly_detect_head_infer(const void* x, long ldx, int N, int H, int W, int Cin, const void* w_packed,
                                    const float* bias, int na, int no, float stride, const float* anchors, float* raw,
                                    float* z, long rows_total, long row_off, ly_stream_t stream) {
  LY_REQUIRE(x && w_packed && bias && anchors && z && na > 0 && no >= 5, "ly_detect_head_infer: bad args");
  LY_REQUIRE(row_off >= 0 && row_off + (long)na * H * W <= rows_total, "ly_detect_head_infer: rows out of range");
  ly_conv_desc d;
  memset(&d, 0, sizeof(d));
  d.n_src = 1; d.src[0] = x; d.src_c[0] = Cin; d.src_ld[0] = ldx;
  d.N = N; d.H = H; d.W = W; d.kh = d.kw = 1; d.stride = 1; d.pad = 0;
  d.Cout = na * no; d.w_packed = w_packed; d.shift = bias; d.act = LY_ACT_NONE;
  d.out = z; d.out_ld = na * no; d.out_fp32 = 1;   // `out` is not written in detect mode
  ly::DetectArgs det = {na, no, stride, anchors, raw, z, rows_total, row_off};
  return conv_launch(&d, &det, stream);
}
