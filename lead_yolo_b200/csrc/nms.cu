// Confidence filter + class-aware non-maximum suppression on the decoded predictions, on the device and with static
// shapes — reference utils/general.py:884-994 (non_max_suppression, best-class branch: multi_label=False, no
// apriori labels, no masks, no merge) around torchvision.ops.nms.  The reference loops over images on the host, builds
// variable-length candidate lists with boolean masks and stops after a wall-clock time limit (general.py:919,990-992);
// here one launch compacts the candidates of every image and one block per image sorts and suppresses them, so the
// output is a fixed (bs, max_det, 6) tensor + counts and nothing depends on the host clock.
//
// pred: fp32 [bs][rows][no] = (cx, cy, w, h, obj, cls...) as Detect writes it.  Candidate = obj > conf_thres and
// obj * max_c cls_c > conf_thres.  Order = confidence descending, ties by ascending row (argsort in the reference is
// unstable; this is one of its valid orders).  Suppression = torchvision's rule: IoU > iou_thres against a kept box
// of higher rank, boxes shifted by class * max_wh unless agnostic.
#include <atomic>

#include "ly_common.cuh"

namespace ly {

__global__ void __launch_bounds__(256) nms_candidates_kernel(const float* __restrict__ pred, int bs, int rows, int no,
                                                             float conf_thres, unsigned long long* __restrict__ keys,
                                                             int* __restrict__ ncand, int cap) {
  pdl_enter();
  const int img = blockIdx.y;
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const float* p = pred + ((long)img * rows + r) * no;
  const float obj = p[4];
  if (!(obj > conf_thres)) return;
  float best = p[5] * obj;
  for (int c = 1; c < no - 5; ++c) best = fmaxf(best, p[5 + c] * obj);
  if (!(best > conf_thres)) return;
  const int slot = atomicAdd(&ncand[img], 1);
  if (slot < cap)
    keys[(long)img * cap + slot] = ((unsigned long long)__float_as_uint(best) << 32) | (0xFFFFFFFFu - (unsigned)r);
}

struct Box {
  float x1, y1, x2, y2;
};

__device__ __forceinline__ bool iou_over(const Box& a, const Box& b, float thr) {
  const float left = fmaxf(a.x1, b.x1), right = fminf(a.x2, b.x2);
  const float top = fmaxf(a.y1, b.y1), bottom = fminf(a.y2, b.y2);
  const float w = fmaxf(right - left, 0.f), h = fmaxf(bottom - top, 0.f);
  const float inter = w * h;
  const float sa = (a.x2 - a.x1) * (a.y2 - a.y1), sb = (b.x2 - b.x1) * (b.y2 - b.y1);
  return inter / (sa + sb - inter) > thr;
}

// one block per image; dynamic shared memory: keys [cap2] (8 B) | boxes [cap2] (16 B) | class [cap2] (2 B) | removed
// bitmap [cap2/32].  The boxes are stored WITH their class offset; the serial keep loop (one thread) reads only shared
// memory — when it re-read the prediction rows from global memory, its ~40 dependent round trips per image were most
// of the kernel's 58 us.
__global__ void __launch_bounds__(1024) nms_select_kernel(const float* __restrict__ pred, int rows, int no,
                                                          const unsigned long long* __restrict__ keys,
                                                          const int* __restrict__ ncand, int cap, int cap2,
                                                          float iou_thres, float max_wh, int max_nms, int max_det,
                                                          float* __restrict__ det, int* __restrict__ det_count) {
  pdl_enter();
  extern __shared__ __align__(16) unsigned char smraw[];
  unsigned long long* sk = reinterpret_cast<unsigned long long*>(smraw);
  Box* sb = reinterpret_cast<Box*>(smraw + (size_t)cap2 * 8);
  unsigned short* scls = reinterpret_cast<unsigned short*>(smraw + (size_t)cap2 * 24);
  unsigned* removed = reinterpret_cast<unsigned*>(smraw + (size_t)cap2 * 26);
  short* keep_idx = reinterpret_cast<short*>(smraw + (size_t)cap2 * 26 + (size_t)cap2 / 8 + 16);   // [max_det]
  __shared__ unsigned long long cm[64];
  __shared__ unsigned long long keepmask;
  __shared__ int nkept;

  const int img = blockIdx.x;
  const int tid = threadIdx.x, nth = blockDim.x;
  int n = min(ncand[img], cap);
  int n2 = 64;
  while (n2 < n) n2 <<= 1;
  for (int i = tid; i < n2; i += nth) sk[i] = i < n ? keys[(long)img * cap + i] : 0ull;
  for (int i = tid; i < n2 / 32; i += nth) removed[i] = 0u;
  if (tid == 0) nkept = 0;
  __syncthreads();
  // bitonic sort, descending
  for (int k = 2; k <= n2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < n2; i += nth) {
        const int l = i ^ j;
        if (l > i) {
          const unsigned long long a = sk[i], b = sk[l];
          const bool desc = (i & k) == 0;
          if (desc ? a < b : a > b) {
            sk[i] = b;
            sk[l] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  n = min(n, max_nms);
  const float* pimg = pred + (long)img * rows * no;
  for (int i = tid; i < n; i += nth) {
    const unsigned r = 0xFFFFFFFFu - (unsigned)(sk[i] & 0xFFFFFFFFull);
    const float* p = pimg + (long)r * no;
    int cls = 0;
    float best = p[5];
    for (int c = 1; c < no - 5; ++c)
      if (p[5 + c] > best) { best = p[5 + c]; cls = c; }
    const float off = (float)cls * max_wh;
    const float hw = p[2] / 2, hh = p[3] / 2;
    sb[i] = Box{p[0] - hw + off, p[1] - hh + off, p[0] + hw + off, p[1] + hh + off};
    scls[i] = (unsigned short)cls;
  }
  __syncthreads();
  for (int c0 = 0; c0 < n; c0 += 64) {
    const int cn = min(64, n - c0);
    if (tid < 64) {
      // row `tid` of the chunk's own suppression matrix (bits j > tid)
      unsigned long long m = 0ull;
      if (tid < cn) {
        const Box a = sb[c0 + tid];
        for (int j = tid + 1; j < cn; ++j)
          if (iou_over(a, sb[c0 + j], iou_thres)) m |= 1ull << j;
      }
      cm[tid] = m;
    }
    __syncthreads();
    if (tid == 0) {
      unsigned long long rem = (unsigned long long)removed[c0 / 32] | ((unsigned long long)removed[c0 / 32 + 1] << 32);
      unsigned long long km = 0ull;
      int kept = nkept;
      for (int i = 0; i < cn && kept < max_det; ++i) {
        if (!((rem >> i) & 1ull)) {
          km |= 1ull << i;
          rem |= cm[i];
          keep_idx[kept] = (short)(c0 + i);   // written out by the whole block after the loop
          ++kept;
        }
      }
      nkept = kept;
      keepmask = km;
    }
    __syncthreads();
    if (nkept >= max_det) break;
    const unsigned long long km = keepmask;
    if (km != 0ull) {
      for (int j = c0 + 64 + tid; j < n; j += nth) {
        if ((removed[j >> 5] >> (j & 31)) & 1u) continue;
        const Box b = sb[j];
        unsigned long long m = km;
        while (m) {
          const int i = __ffsll((long long)m) - 1;
          m &= m - 1;
          if (iou_over(sb[c0 + i], b, iou_thres)) {
            atomicOr(&removed[j >> 5], 1u << (j & 31));
            break;
          }
        }
      }
    }
    __syncthreads();
  }
  __syncthreads();
  // the kept boxes, in rank order: one thread each (the un-shifted coordinates come from the prediction row again,
  // bit for bit what the reference returns)
  for (int k = tid; k < nkept; k += nth) {
    const int idx = keep_idx[k];
    const unsigned long long key = sk[idx];
    const unsigned r = 0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull);
    const float* p = pimg + (long)r * no;
    const float hw = p[2] / 2, hh = p[3] / 2;
    float* o = det + ((long)img * max_det + k) * 6;
    o[0] = p[0] - hw; o[1] = p[1] - hh; o[2] = p[0] + hw; o[3] = p[1] + hh;
    o[4] = __uint_as_float((unsigned)(key >> 32));
    o[5] = (float)scls[idx];
  }
  if (tid == 0) det_count[img] = nkept;
}

}  // namespace ly

using namespace ly;

extern "C" long ly_nms_workspace_bytes(int bs, int max_candidates) {
  if (bs <= 0 || max_candidates <= 0) return -1;
  return (long)bs * max_candidates * 8 + (long)bs * 4;
}

extern "C" int ly_detect_nms(const float* pred, int bs, int rows, int no, float conf_thres, float iou_thres,
                             int agnostic, int max_det, int max_candidates, void* workspace, float* det,
                             int* det_count, int* n_candidates, ly_stream_t st) {
  LY_REQUIRE(pred && workspace && det && det_count, "ly_detect_nms: null pointer");
  LY_REQUIRE(bs > 0 && rows > 0 && no >= 6, "ly_detect_nms: bad shape bs=%d rows=%d no=%d", bs, rows, no);
  LY_REQUIRE(max_det > 0 && max_candidates >= 64 && max_candidates <= 8192,
             "ly_detect_nms: max_candidates=%d must be in [64, 8192] (shared-memory sort)", max_candidates);
  int cap2 = 64;
  while (cap2 < max_candidates) cap2 <<= 1;
  LY_REQUIRE(max_det <= 4096, "ly_detect_nms: max_det=%d exceeds 4096", max_det);
  LY_REQUIRE(no - 5 <= 65535, "ly_detect_nms: too many classes");
  const size_t smem = (size_t)cap2 * 26 + (size_t)cap2 / 8 + 16 + (size_t)max_det * 2 + 16;
  static std::atomic<size_t> attr_set[64];
  int dev = 0;
  LY_CUDA(cudaGetDevice(&dev));
  if (attr_set[dev & 63].load() < smem) {
    LY_CUDA(cudaFuncSetAttribute(nms_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set[dev & 63].store(smem);
  }
  cudaStream_t s = (cudaStream_t)st;
  unsigned long long* keys = (unsigned long long*)workspace;
  int* ncand = (int*)((char*)workspace + (size_t)bs * max_candidates * 8);
  LY_CUDA(cudaMemsetAsync(ncand, 0, (size_t)bs * 4, s));
  LY_CUDA(cudaMemsetAsync(det, 0, (size_t)bs * max_det * 6 * 4, s));
  dim3 g1((unsigned)ceil_div(rows, 256), (unsigned)bs);
  LY_CUDA(launch(nms_candidates_kernel, g1, 256, 0, s, pred, bs, rows, no, conf_thres, keys, ncand, max_candidates));
  LY_LAUNCH_CHECK();
  LY_CUDA(launch(nms_select_kernel, bs, 1024, smem, s, pred, rows, no, keys, ncand, max_candidates, cap2, iou_thres,
                                           agnostic ? 0.f : 7680.f, 30000, max_det, det, det_count));
  LY_LAUNCH_CHECK();
  if (n_candidates != nullptr)
    LY_CUDA(cudaMemcpyAsync(n_candidates, ncand, (size_t)bs * 4, cudaMemcpyDeviceToDevice, s));
  return LY_OK;
}
