// Experiment: can a UMMA shared-memory descriptor start at an arbitrary ROW of a TMA-written swizzled tile?
// (needed to read the 9 taps of a 3x3 convolution from ONE halo tile).  A = [256 rows][KC channels] loaded by TMA with
// SWIZZLE_32B (KC=16) or SWIZZLE_128B (KC=64); B = identity; D[m][n] must equal A[m + shift][n] for every shift.
#include <cstdio>
#include <vector>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "ly_tc.cuh"
using namespace ly;

__device__ __forceinline__ uint64_t desc_sw32(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) | (6ull << 61);
}

template <int KC>
__global__ void __launch_bounds__(128) k(const __grid_constant__ CUtensorMap amap, int shift, int use_base_off, float* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar, done;
  __shared__ uint32_t tm;
  uint8_t* sA = smem;                       // 256 rows x KC*2 bytes
  uint8_t* sB = smem + 256 * KC * 2;        // identity [KC][KC], same swizzle
  const int warp = threadIdx.x >> 5;
  constexpr int RB = KC * 2;                // row bytes
  for (int i = threadIdx.x; i < KC * KC; i += 128) {
    const int n = i / KC, kk = i % KC;
    const int chunk = kk >> 3, sw = (RB == 32) ? ((n >> 2) & 1) : (n & 7);
    *reinterpret_cast<__nv_bfloat16*>(sB + n * RB + ((chunk ^ sw) << 4) + (kk & 7) * 2) = __float2bfloat16(n == kk ? 1.f : 0.f);
  }
  fence_proxy_async_smem();
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&done, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(&tm, 64);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bar, 256 * RB);
    tma_load_2d(sA, &amap, &bar, 0, 0);
    mbar_wait(&bar, 0);
    tc_fence_after_sync();
    const uint32_t idesc = umma_idesc_bf16_f32(128, KC, 0, 0);
    const uint32_t a_addr = smem_u32(sA) + (uint32_t)(shift * RB);
    for (int kk = 0; kk < KC / 16; ++kk) {
      uint64_t ad, bd;
      if (RB == 32) { ad = desc_sw32(a_addr); bd = desc_sw32(smem_u32(sB)); }
      else {
        ad = umma_smem_desc_sw128(a_addr + kk * 32, 16, 1024); bd = umma_smem_desc_sw128(smem_u32(sB) + kk * 32, 16, 1024);
        if (use_base_off) ad |= (uint64_t)((a_addr >> 7) & 7u) << 49;
      }
      umma_bf16(tm, ad, bd, idesc, (uint32_t)(kk != 0));
    }
    umma_commit(&done);
  }
  mbar_wait(&done, 0);
  tc_fence_after_sync();
  const int r = threadIdx.x;
  for (int c0 = 0; c0 < KC; c0 += 16) {
    uint32_t acc[16];
    tmem_ld16(tm + ((uint32_t)(warp * 32) << 16) + c0, acc);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) out[r * KC + c0 + j] = __uint_as_float(acc[j]);
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 64);
}

template <int KC>
int run() {
  const int R = 256;
  std::vector<__nv_bfloat16> h(R * KC);
  for (int r = 0; r < R; ++r) for (int c = 0; c < KC; ++c) h[r * KC + c] = __float2bfloat16((float)(r + (c % 8) * 0.125f - (c / 8) * 256));
  void* a; cudaMalloc(&a, R * KC * 2); cudaMemcpy(a, h.data(), R * KC * 2, cudaMemcpyHostToDevice);
  float* out; cudaMalloc(&out, 128 * KC * 4);
  CUtensorMap m;
  cuuint64_t gd[2] = {(cuuint64_t)KC, (cuuint64_t)R}; cuuint64_t gs[1] = {(cuuint64_t)KC * 2};
  cuuint32_t bd[2] = {(cuuint32_t)KC, 256}; cuuint32_t es[2] = {1, 1};
  if (cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a, gd, gs, bd, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             KC == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) { printf("encode failed\n"); return 1; }
  const size_t smem = 256 * KC * 2 + KC * KC * 2 + 2048;
  cudaFuncSetAttribute(k<KC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int bo = 0; bo < (KC == 64 ? 2 : 1); ++bo) {
    printf("KC=%d base_offset_field=%d: shifts OK:", KC, bo);
    for (int shift = 0; shift <= 40; ++shift) {
      k<KC><<<1, 128, smem>>>(m, shift, bo, out);
      if (cudaDeviceSynchronize() != cudaSuccess) { printf(" [kernel failed at %d: %s]\n", shift, cudaGetErrorString(cudaGetLastError())); return 1; }
      std::vector<float> o(128 * KC);
      cudaMemcpy(o.data(), out, 128 * KC * 4, cudaMemcpyDeviceToHost);
      bool ok = true;
      for (int r = 0; r < 128 && ok; ++r) for (int c = 0; c < KC; ++c)
        if (o[r * KC + c] != __bfloat162float(h[(r + shift) * KC + c])) { ok = false; break; }
      printf(" %d%s", shift, ok ? "" : "x");
    }
    printf("\n");
  }
  return 0;
}

int main() { if (run<16>()) return 1; return run<64>(); }
