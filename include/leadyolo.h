/* leadyolo.h — C-ABI of libleadyolo.so: hand-written sm_100a kernels for the LEAD-YOLO module hot path.
 *
 * The reference (qingqing-zijin/LEAD-YOLO) is pure Python/PyTorch and has no FFI of its own; each entry point
 * below cites the reference nn.Module code (file:line under the reference tree) whose arithmetic it replaces.
 * The reference-side binding is the ctypes stub in lead_yolo_b200/_lib.py.
 *
 * Conventions
 *  - Every pointer is a DEVICE pointer on the current CUDA device. Activations are NHWC ("channels_last")
 *    bf16 unless stated; parameters/statistics are fp32.  `ld` = distance between consecutive pixels, in
 *    elements (lets a tensor be a channel slice of a wider concat buffer).
 *  - The caller owns every buffer (inputs, outputs, saved statistics, workspaces).  Unless a function says
 *    otherwise the library allocates nothing, never synchronises, and launches on `stream`.
 *  - Return 0 on success or a negative ly_status; ly_last_error() gives a thread-local message.
 *  - There is no CPU fallback: a device that is not sm_100 yields LY_EUNSUPPORTED.
 */
#ifndef LEADYOLO_H_
#define LEADYOLO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* ly_stream_t; /* cudaStream_t */

enum ly_status { LY_OK = 0, LY_EINVAL = -1, LY_EUNSUPPORTED = -2, LY_ECUDA = -3 };
enum ly_act { LY_ACT_NONE = 0, LY_ACT_RELU = 1, LY_ACT_SILU = 2, LY_ACT_HSWISH = 3, LY_ACT_SIGMOID = 4 };

const char* ly_last_error(void);
int ly_version(void);
/* 0 when the current device can run the kernels (compute capability 10.x), else LY_EUNSUPPORTED. */
int ly_device_check(void);

/* ------------------------------------------------------------------------------------------------------
 * Dense convolution as an implicit GEMM on tcgen05 tensor cores (TMA-fed, fp32 accumulation in TMEM).
 * Replaces nn.Conv2d (+ folded BatchNorm + activation) at: models/common.py:1894-1910 (Conv),
 * 1461-1464 (MLPBlock.mlp 1x1s), 1531/1549 (patch embed / merging), models/rfa.py:110 (RFCBAMConv.conv),
 * models/yolo.py:79,90 (Detect heads), and — with transposed/flipped weights — their input gradients.
 *
 *   out[p, o] = act( scale[o] * sum_{s,t,c} A_s[p (+) tap t, c] * Wp[o, k(s,t,c)] + shift[o] ) (+ residual[p, o])
 *
 * A sources: up to 4 NHWC bf16 tensors sharing (N,H,W), concatenated along K ("virtual concat").
 * Geometry: (kh,kw,stride,pad) in {(1,1,1,0), (3,3,1,1), (k,k,k,0) with k in {2,4}}.
 * Packed weights Wp: bf16 [Cout][Kp], Kp = 64 * ly_conv_num_chunks(); K order: for source, for tap
 * (row-major), for 64-channel chunk (zero padded).  For (k,k,k,0) a "tap" is a kernel row ky and the
 * "channels" are the kw*C contiguous values (kx, c).  lead_yolo_b200.packing mirrors this order.
 * ------------------------------------------------------------------------------------------------------ */
typedef struct {
  int n_src;
  const void* src[4]; /* bf16, first channel of the slice; 16-byte aligned */
  int src_c[4];       /* channels taken from each source (multiple of 8) */
  long src_ld[4];     /* pixel stride in elements (multiple of 8) */
  int N, H, W;        /* input spatial size */
  int kh, kw, stride, pad;
  int Cout;
  const void* w_packed; /* bf16 [Cout][Kp] */
  const float* scale;   /* [Cout] or NULL (=1) */
  const float* shift;   /* [Cout] or NULL (=0) */
  int act;              /* ly_act */
  const void* residual; /* bf16 [pixels][res_ld] added after the activation, or NULL */
  long res_ld;
  void* out;   /* bf16 (or fp32 when out_fp32) */
  long out_ld; /* pixel stride of out in elements */
  int out_fp32;
  /* patch_k > 0: scatter the GEMM columns (ky, kx, c) back onto a patch grid — the input gradient of a
   * (k,k,k,0) convolution.  out is then the full-resolution NHWC tensor with patch_c channels and
   * out_ld its pixel stride; Cout must equal patch_k*patch_k*patch_c. */
  int patch_k, patch_c;
} ly_conv_desc;

int ly_conv_num_chunks(const ly_conv_desc* d); /* number of 64-wide K chunks (Kp / 64); <0 on error */
int ly_conv_fwd(const ly_conv_desc* d, ly_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* LEADYOLO_H_ */
