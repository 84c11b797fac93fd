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

/* Weight gradient of the convolution described by `d` (autograd of nn.Conv2d; the reference relies on PyTorch
 * autograd).  dz: bf16 NHWC output gradient [N,Hout,Wout,Cout] with pixel stride dz_ld.
 * dw: fp32 [Cout][Kp] in the packed-K layout; partial sums are ADDED (zero it first).  tcgen05, MN-major. */
int ly_conv_wgrad(const ly_conv_desc* d, const void* dz, long dz_ld, float* dw, ly_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * BatchNorm2d / activation streaming kernels on an NHWC bf16 matrix [M pixels][C], C % 8 == 0.
 * Reference: nn.BatchNorm2d at models/common.py:91,1462,1533,1551,1591, models/rfa.py:104,111 with
 * eps/momentum from utils/torch_utils.py:217-219; activations common.py:84 (SiLU), 1463 (ReLU), 1565-1580.
 * ------------------------------------------------------------------------------------------------------ */
/* sum[c] += Σ_m x[m][c];  sq[c] += Σ_m x[m][c]^2 (sq may be NULL).  Accumulates: zero the outputs first. */
int ly_col_stats(const void* x, long ld, long M, int C, float* sum, float* sq, ly_stream_t stream);
/* Training-mode statistics -> mean, rstd, scale = gamma*rstd, shift = beta - mean*scale; updates the running
 * statistics in place (unbiased variance, momentum) when running_mean/var are non-NULL. */
int ly_bn_finalize(const float* sum, const float* sq, int C, long n, float eps, float momentum, const float* gamma,
                   const float* beta, float* running_mean, float* running_var, float* mean_out, float* rstd_out,
                   float* scale_out, float* shift_out, ly_stream_t stream);
/* Inference fold: scale = gamma/sqrt(running_var+eps), shift = beta - running_mean*scale (+ conv_bias*scale). */
int ly_bn_fold(int C, float eps, const float* gamma, const float* beta, const float* running_mean,
               const float* running_var, const float* conv_bias, float* mean_out, float* rstd_out, float* scale_out,
               float* shift_out, ly_stream_t stream);
/* y = act(x*scale[c] + shift[c]) */
int ly_scale_shift_act(const void* x, long ldx, long M, int C, const float* scale, const float* shift, int act,
                       void* y, long ldy, ly_stream_t stream);
/* Backward of y = act(scale*z + shift): s1 = Σ dact, s2 = Σ dact*xhat (= dbeta, dgamma); dz per BatchNorm
 * backward (training: batch statistics; otherwise dz = scale*dact).
 * parts == NULL: the sums are ADDED into s1/s2 with float atomics (zero them first; order-dependent rounding).
 * parts != NULL: fp32 scratch [ly_reduce_blocks(M, C)][2*C]; every block stores its partial sums and a small second
 * launch adds them in block order into s1/s2 (written, no zeroing needed) — bit-reproducible, which matters because
 * s1/n and s2/n are subtracted from EVERY element of dz.
 * sync != NULL (with parts): two ZEROED 32-bit words; for large tensors (>= 48 Mi elements) the three phases then run as
 * ONE cooperatively launched kernel separated by grid barriers (same arithmetic, same order); NULL keeps the launches. */
int ly_reduce_blocks(long M, int C);
int ly_bn_act_bwd(const void* dy, long lddy, const void* z, long ldz, long M, int C, const float* scale,
                  const float* shift, const float* mean, const float* rstd, int act, int training, float* s1,
                  float* s2, float* parts, unsigned* sync, void* dz, long lddz, ly_stream_t stream);
/* dz = dy * act'(.) with the derivative taken from the activation OUTPUT y (none / relu / sigmoid). */
int ly_act_bwd_from_out(const void* dy, long lddy, const void* y, long ldy, long M, int C, int act, void* dz,
                        long lddz, ly_stream_t stream);

/* Training-mode BatchNorm2d forward in two launches (nn.BatchNorm2d at the sites listed above, batch statistics):
 * ly_bn_stats accumulates, per channel, the sums about a pivot K = x[0][c] into acc = [C] fp32 K | [C] int64 Σ(x-K)·2^24
 * | [C] int64 Σ(x-K)²·2^20, i.e. 5*C floats of space (zero acc first).  Shifted sums keep the variance from cancelling
 * when |mean| >> std; the cross-block step is 64-bit fixed point, so the statistics — and the training forward — are
 * bit-reproducible from run to run (valid for Σ(x-K)² < 8e12).  ly_bn_apply_train derives
 * mean / biased variance from acc, writes y = act(gamma*(z-mean)*rstd + beta), saves stats = fp32 [4][C]
 * (mean, rstd, scale, shift) for ly_bn_act_bwd and updates running_mean / running_var (unbiased variance, `momentum`)
 * in place when they are non-NULL. */
int ly_bn_stats(const void* x, long ld, long M, int C, float* acc, ly_stream_t stream);
int ly_bn_apply_train(const void* z, long ldz, long M, int C, const float* acc, const float* gamma, const float* beta,
                      float eps, float momentum, float* running_mean, float* running_var, float* stats, int act,
                      void* y, long ldy, ly_stream_t stream);


/* ------------------------------------------------------------------------------------------------------
 * Weight layouts of the training path.  nn.Conv2d keeps fp32 OIHW parameters (optimizer, EMA, DDP and checkpoints
 * touch them: utils/torch_utils.py:318-346, 404-432, train.py:387-403); the GEMM kernels read packed bf16.  One launch
 * per weight and direction.  c0..c3: channels of the K-concatenated sources (0 = unused), patch = 1 for kernel == stride.
 *   ly_pack_weight        w [Cout][Cin][kh][kw] -> wp bf16 [Cout][Kp]                         (forward operand)
 *   ly_pack_weight_dgrad  -> weights of the input-gradient GEMM for source channels [coff, coff+cs):
 *                            stride 1: [cs][kh*kw*P], flipped taps, K = output channel; patch: [kh*kw*cs][P];
 *                            P = Cout rounded up to a multiple of 64 (zero filled)
 *   ly_unpack_wgrad       dwp fp32 [>=Cout][Kp] (ly_conv_wgrad's output) -> dw [Cout][Cin][kh][kw], added when accumulate
 * ------------------------------------------------------------------------------------------------------ */
int ly_pack_weight(const float* w, int Cout, int kh, int kw, int patch, int c0, int c1, int c2, int c3, void* wp,
                   ly_stream_t stream);
int ly_pack_weight_dgrad(const float* w, int Cout, int Cin, int kh, int kw, int patch, int coff, int cs, void* wd,
                         ly_stream_t stream);
int ly_unpack_wgrad(const float* dwp, int Cout, int kh, int kw, int patch, int c0, int c1, int c2, int c3, float* dw,
                    int accumulate, ly_stream_t stream);
/* ly_pack_weight / ly_pack_weight_dgrad for MANY weights in one launch (a captured training step packs every parameter
 * once per replay).  jobs: DEVICE array; kind 0 = forward layout (c = source split, Kp = packed K), kind 1 = input-gradient
 * layout (coff, cs, P = pad64 of the 8-aligned Cout); total = elements of dst; first_block = prefix sum of
 * ceil(total / 256) over the jobs before this one; total_blocks = that sum over all jobs. */
typedef struct ly_pack_job {
  const void* w;     /* fp32 OIHW parameter */
  void* dst;         /* bf16 packed matrix */
  int kind, Cout, Cin, kh, kw, patch;
  int c[4];
  int coff, cs, P, Kp;
  long total;
  long first_block;
} ly_pack_job;
int ly_pack_batch(const ly_pack_job* jobs, int njobs, long total_blocks, ly_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Stem: PatchEmbed_FasterNet.proj on the raw image (models/common.py:1531,1537): 4x4/s4 conv, 3 -> Cout,
 * NCHW fp32 (or uint8 with in_scale = 1/255, train.py:295) -> NHWC bf16; out = scale*conv + shift.
 * wk: fp32 [48][Cout], k = (c*4+ky)*4+kx.  ly_stem_wgrad ADDS the weight gradient into dwk (same layout).
 * ------------------------------------------------------------------------------------------------------ */
int ly_stem_fwd(const void* img, int img_is_u8, int N, int H, int W, int Cout, const float* wk, const float* scale,
                const float* shift, float in_scale, void* out, long ld_out, ly_stream_t stream);
int ly_stem_wgrad(const void* img, int img_is_u8, int N, int H, int W, int Cout, const void* dz, long ld_dz,
                  float in_scale, float* dwk, ly_stream_t stream);
/* Patch matrix of the stem: P bf16 [N*(H/4)*(W/4)][64], column c*16+ky*4+kx = img[n][c][4y+ky][4x+kx] (raw values, no
 * in_scale; uint8 is exact in bf16), columns 48..63 zero.  With it the stem's weight gradient is ly_conv_wgrad of a 1x1
 * convolution over P (tensor cores) instead of ly_stem_wgrad (CUDA cores): dW[o][c][ky][kx] = in_scale * dWp[o][c*16+ky*4+kx]. */
int ly_stem_im2col(const void* img, int img_is_u8, int N, int H, int W, void* P, ly_stream_t stream);

/* Fused FasterNet block, inference (models/common.py:1432-1437,1478-1482): y = x + W2 relu(scale*(W1 [pconv3x3(x[:,:C/4])
 * | x[:,C/4:]]) + shift) in ONE persistent kernel; x and y are dense NHWC bf16 [N][H][W][C], C in {24, 40, 80} (y 32-byte aligned for C = 80).  The 2C-wide
 * hidden tensor stays in TMEM / shared memory; the partial conv is folded into the first GEMM (Wc_tap = W1[:, :C/4] Wp_tap).
 * wimg: pre-swizzled bf16 weight image with sections W1x | W2 | Wc at byte offsets 0 / off_w2 / off_wc
 * (lead_yolo_b200.packing.fasternet_weight_image; C = 80 has a fixed layout W1x[:, :64] | W2 | W1x[:, 64:] | Wc, checked
 * against w_bytes / off_w2 / off_wc); scale/shift: folded mlp.1 BatchNorm [2C]. */
int ly_fasternet_block_infer(const void* x, void* y, int N, int H, int W, int C, const void* wimg, int w_bytes,
                             int off_w2, int off_wc, const float* scale, const float* shift, ly_stream_t stream);

/* SPPF pooling (models/common.py:356,364-367): 5x5 / stride 1 / pad 2 max pool on an NHWC channel slice.
 * idx (optional, uint8 [pixels][C]) records the arg-max tap for ly_maxpool5_bwd. */
int ly_maxpool5_fwd(const void* x, long ldx, int N, int H, int W, int C, void* y, long ldy, uint8_t* idx,
                    ly_stream_t stream);
/* SPPF at inference (models/common.py:364-367): the three cascaded 5x5 pools in one launch; y_k (k = 1, 2, 3 pools deep) is
 * written at y + (k-1)*ystep elements, i.e. into three channel slices of the concat buffer.  H*W <= 1600. */
int ly_sppf_pool3(const void* x, long ldx, int N, int H, int W, int C, void* y, long ldy, long ystep,
                  ly_stream_t stream);
int ly_maxpool5_bwd(const void* dy, long lddy, const uint8_t* idx, int N, int H, int W, int C, void* dx, long lddx,
                    int accumulate, ly_stream_t stream);

/* nn.Upsample(None, 2, 'nearest') / Concat (LEAD-YOLO.yaml:29-34, models/common.py:531-538): copy an NHWC
 * tensor into a channel slice of a wider buffer, optionally replicating 2x2 (up = 2); and its backward. */
int ly_copy_upsample(const void* src, long lds, int N, int Hs, int Ws, int C, int up, void* dst, long ldd,
                     ly_stream_t stream);
int ly_downsum2(const void* src, long lds, int N, int H, int W, int C, void* dst, long ldd, ly_stream_t stream);

/* Detect post-processing of one level (models/yolo.py:95-125), fp32: p = conv output [N][ny][nx][na*no] (NHWC);
 * raw [N][na][ny][nx][no] (the permuted view training consumes) and/or the decoded rows of this level inside
 * z [N][rows_total][no] at row_off + a*ny*nx + y*nx + x: xy = (2s - 0.5 + grid)*stride, wh = (2s)^2 * anchor*stride,
 * rest = s, s = sigmoid.  anchors: [na][2] in stride units (Detect.anchors[i]). */
int ly_detect_decode(const float* p, int N, int ny, int nx, int na, int no, float stride, const float* anchors,
                     float* raw, float* z, long rows_total, long row_off, ly_stream_t stream);
/* The same with the 1x1 head convolution in front (models/yolo.py:79,90): ONE tensor-core kernel per level whose epilogue
 * adds the bias and writes raw (may be NULL) and the decoded rows — the head's output never exists in its conv layout.
 * x: NHWC bf16; w_packed as for ly_conv_fwd ([na*no][Kp]); bias fp32 [na*no]; anchors fp32 [na][2] in stride units.
 * na*no <= 32, na <= 4. */
int ly_detect_head_infer(const void* x, long ldx, int N, int H, int W, int Cin, const void* w_packed, const float* bias,
                         int na, int no, float stride, const float* anchors, float* raw, float* z, long rows_total,
                         long row_off, ly_stream_t stream);

/* rfa.SE global average pool (models/rfa.py:80,90): bf16 [N][C] <- mean over HW. */
int ly_gap(const void* x, long ldx, int N, int HW, int C, void* out, long ldo, ly_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * CoordAtt (models/common.py:1595-1609).  "HW rows" matrix: [N*H + N*W][C]; row n*H+h = mean_w x[n,h,:,c],
 * row N*H + n*W + w = mean_h x[n,:,w,c].  Gates use the same row layout.
 * ------------------------------------------------------------------------------------------------------ */
int ly_ca_pool(const void* x, long ldx, int N, int H, int W, int C, void* pooled, long ldp, ly_stream_t stream);
/* Inference: everything between pooling and gating in one kernel (common.py:1601-1608).  pooled/gates: "HW rows"
 * matrices with `rows` rows of which the first nh use conv_h (wh [C][mip], bh [C]) and the rest conv_w (ww, bw);
 * w1 [mip][C] = conv1.weight; scale/shift [mip] = bn1 folded with conv1.bias; mip <= 32. */
int ly_ca_mid_infer(const void* pooled, long ldp, long rows, long nh, int C, int mip, const float* w1,
                    const float* scale, const float* shift, const float* wh, const float* bh, const float* ww,
                    const float* bw, void* gates, long ldg, ly_stream_t stream);
int ly_ca_apply(const void* x, long ldx, const void* gates, long ldg, int N, int H, int W, int C, void* out,
                long ldo, ly_stream_t stream); /* out = x * a_h[n,h,c] * a_w[n,w,c] */
int ly_ca_gate_grad(const void* x, long ldx, const void* dy, long lddy, const void* gates, long ldg, int N, int H,
                    int W, int C, void* dgates, long ldd, ly_stream_t stream);
int ly_ca_pool_bwd(const void* dpooled, long ldp, int N, int H, int W, int C, void* dx, long ldx,
                   ly_stream_t stream); /* dx = dP_h[n,h,c]/W + dP_w[n,w,c]/H */

/* ------------------------------------------------------------------------------------------------------
 * RFCBAMConv (models/rfa.py:95-129).  G/A: bf16 [N*ho*wo][k*k*C], column t*C+c (t = n1*k+n2);
 * maps fp32 [N][ho*k][wo*k][2] (max, mean over c); amax int16 [N][ho*k][wo*k]; att fp32 [N][ho*k][wo*k];
 * wgk fp32 [k*k][k*k][C] = generate.0.weight[c*k*k+t, 0, i, j] at [t][i*k+j][c]; wa fp32 [2][3][3].
 * ------------------------------------------------------------------------------------------------------ */
int ly_rf_generate(const void* x, long ldx, int N, int H, int W, int C, int k, int s, int ho, int wo,
                   const float* wgk, const float* scale, const float* shift, int act, void* G, ly_stream_t stream);
int ly_rf_attention(const void* G, int N, int ho, int wo, int C, int k, const float* wa, float* maps, int16_t* amax,
                    float* att, ly_stream_t stream);
/* No counterpart in the reference: a variant of ly_rf_attention for callers that cannot allocate device
 * memory when they call (a server whose memory arena is handed out at start-up, a real-time loop where
 * cudaMalloc's device-wide synchronisation is off-limits).  The max/mean maps between the two kernels live in a
 * library-owned static device array of LY_RF_NOALLOC_MAX_FLOATS = 2,097,152 floats (8 MiB, held by every GPU that
 * loads the library until the process exits) instead of a caller buffer; N*ho*k*wo*k*2 floats must fit or the call
 * returns LY_EINVAL.  The first call on each device looks the array's address up (make it before any stream
 * capture); no call allocates or frees.  All calls on a device share the array: calls that may overlap on the GPU
 * race.  Results are identical to ly_rf_attention. */
int ly_rf_attention_noalloc(const void* G, int N, int ho, int wo, int C, int k, const float* wa, int16_t* amax,
                            float* att, ly_stream_t stream);
int ly_rf_gate(const void* G, int N, int ho, int wo, int C, int k, const void* se, long ldse, const float* att,
               void* A, ly_stream_t stream); /* A = G * se[n,c] * att */
/* Inference front half of RFCBAMConv (rfa.py:113-128) without materialising G: pass 1 recomputes the receptive-field
 * features and emits only the channel max/mean maps and the SE sums, a tiny kernel evaluates the SE gate
 * (w1 [R][C], w2 [C][R] = se.fc weights) and the 3x3 spatial attention, pass 2 recomputes the features and writes the
 * gated GEMM operand A once.  Workspaces (fp32, caller-allocated): maps [N][ho*k][wo*k][2], gap_sum [N][C],
 * se [N][C], att [N][ho*k][wo*k].  scale/shift: folded generate BatchNorm in G-column order (t*C + c).  C <= 256. */
int ly_rfcbam_front_infer(const void* x, long ldx, int N, int H, int W, int C, int k, int s, int ho, int wo,
                          const float* wgk, const float* scale, const float* shift, const float* w1, const float* w2,
                          int R, const float* wa, float* maps, float* gap_sum, float* se, float* att, void* A,
                          ly_stream_t stream);
/* k = 3 inference pipeline with the feature generation on the tensor cores (csrc/rf_generate_tc.cu): the depthwise
 * 3x3 of rfa.py:101-105 runs as a grouped 3x3 convolution (16 -> 144 channels per group) with one tcgen05.mma per tap.
 * Two passes over x, the ungated features are never stored:
 *   ly_rf_tc_stats  maps fp32 [N][3ho][3wo][2] = (max, mean) over channels of relu(generate + shift); zeroed by the call
 *   ly_rf_tc_apply  A bf16 [N*ho*wo][9*C], A[m][t*C+c] = relu(generate + shift) * se[n][c] * att[n][3h+t/3][3w+t%3]
 * wc: bf16 [C/16][9 taps][9 t][16] generate weights with the BatchNorm scale folded in; shift: uint32 [C/16][9 t][16],
 * the BatchNorm shift split into two bf16 (hi | lo << 16, hi + lo = shift).  C % 16 == 0, s in {1, 2}; A 32-byte aligned. */
int ly_rf_tc_stats(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                   const void* shift, float* maps, ly_stream_t stream);
int ly_rf_tc_apply(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                   const void* shift, const float* se, const float* att, void* A, ly_stream_t stream);
/* Training forward of the receptive-field generation on the same kernel (models/rfa.py:95-100, the grouped 3x3 alone):
 * Z bf16 [N*ho*wo][9][C], Z[m][t*C+c] = generate + shift — the layout and the meaning of ly_rf_generate's output; pass a
 * zero shift for the raw convolution that feeds the batch-statistics BatchNorm.  wc as above with scale 1 (bf16 weights:
 * the difference to ly_rf_generate's fp32 weights is below the bf16 rounding of Z).  Z 32-byte aligned. */
int ly_rf_tc_generate(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                      const void* shift, void* Z, ly_stream_t stream);
/* Backward of that generation with respect to x, on the same kernel (k = 3): ly_rf_tc_taps computes the gradient of every
 * output pixel's 3x3 input patch, R bf16 [N*ho*wo][9 (ij)][C], R[m][ij][c] = Σ_t wcT[t][ij][c] * dZ[m][t][c] (+ shift: pass
 * zeros); wcT bf16 [C/16][9 t][9 ij][16] = the forward image with its two 9-axes swapped; dZ bf16 [N*ho*wo][9][C].
 * ly_rf_col2im then folds the patches into dx[n,y,x,c] = Σ R[m(h,w)][ij][c] over h*s+i-1 == y, w*s+j-1 == x (+ extra[n][c],
 * may be NULL).  Together they replace the dx half of ly_rf_generate_bwd (fp32 weights on the CUDA cores). */
int ly_rf_tc_taps(const void* dZ, int N, int ho, int wo, int C, const void* wcT, const void* shift, void* R,
                  ly_stream_t stream);
int ly_rf_col2im(const void* R, int N, int H, int W, int C, int s, int ho, int wo, const float* extra, void* dx, long lddx,
                 ly_stream_t stream);
/* ly_rf_tc_apply fused with RFCBAMConv.conv + its BatchNorm + ReLU (models/rfa.py:121-129): the gated receptive-field
 * matrix (9*C wide) never reaches HBM — each [128 pixels][144] tile goes from the gating epilogue through shared memory
 * straight into a second tcgen05 accumulation over the channel groups.  out: bf16 NHWC [N][ho][wo][Cout] (pixel stride
 * ld_out, multiple of 16; 32-byte aligned).  w2: bf16 [C/16][9 t][Cout][16] = conv.weight[o][g*16+j][t/3][t%3], every
 * [Cout][16] tile in the SW32 shared-memory image (lead_yolo_b200.packing.rf_fused_weight_image); scale2/shift2 fp32
 * [Cout]: conv.1 BatchNorm folded with conv.0.bias.  C % 16 == 0, Cout in {128, 256}, s in {1, 2}. */
int ly_rfcbam_fused_infer(const void* x, long ldx, int N, int H, int W, int C, int s, int ho, int wo, const void* wc,
                          const void* shift, const float* se, const float* att, const void* w2, const float* scale2,
                          const float* shift2, int Cout, void* out, long ld_out, ly_stream_t stream);
/* gap_sum[n][c] = sum_hw x (zeroed by the call);  se = sigmoid(W2 relu(W1 gap_sum/HW)), att = sigmoid(conv3x3(maps)). */
int ly_rf_gap_sum(const void* x, long ldx, int N, int HW, int C, float* gap_sum, ly_stream_t stream);
/* Training-path average pool of rfa.SE (models/rfa.py:80,90), bit-reproducible: out bf16 [N][C] = mean over HW;
 * parts: fp32 scratch [LY_GAP_PARTS][N][C] (per-block partial sums, added in index order).  C <= 512. */
#define LY_GAP_PARTS 16
int ly_gap_mean(const void* x, long ldx, int N, int HW, int C, float* parts, void* out, ly_stream_t stream);
int ly_rf_se_att(const float* gap_sum, int N, int C, int HW, const float* w1, const float* w2, int R, const float* maps,
                 int HK, int WK, const float* wa, float* se, float* att, ly_stream_t stream);
/* dse: fp32 [LY_RF_GATE_PARTS][N][C] per-block partial sums, WRITTEN (not accumulated); the gradient of the SE gate is
 * their sum over the first axis in index order (deterministic — the value is broadcast over every pixel afterwards). */
#define LY_RF_GATE_PARTS 16
int ly_rf_gate_bwd(const void* dA, const void* G, int N, int ho, int wo, int C, int k, const void* se, long ldse,
                   const float* att, void* dG, float* dse, float* datt, ly_stream_t stream);
int ly_rf_attention_bwd(const float* datt, const float* att, const float* maps, const int16_t* amax, int N, int ho,
                        int wo, int C, int k, const float* wa, float* dmaps, float* dwa, void* dG,
                        ly_stream_t stream); /* dG += route(dmax) + dmean/C ; dwa += */
/* dx (may be NULL) = input gradient (+ extra[n][c], the SE pooling gradient already divided by H*W, may be NULL);
 * dwgk (may be NULL) += weight gradient in the wgk layout.  wscratch (may be NULL): ly_rf_wgrad_scratch_floats(C, k) floats
 * — with it every pixel lane stores its partial sums and a second launch adds them in order (no atomics, reproducible);
 * without it the sums are added with float atomics. */
long ly_rf_wgrad_scratch_floats(int C, int k);
int ly_rf_generate_bwd(const void* dZ, const void* x, long ldx, int N, int H, int W, int C, int k, int s, int ho,
                       int wo, const float* wgk, const float* extra, void* dx, long lddx, float* dwgk,
                       float* wscratch, long wscratch_floats, ly_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Detection loss (utils/loss.py:121-268 ComputeLoss.__call__ / build_targets with the EIoU box term of
 * utils/metrics.py:293-354), fp32, static shapes (CUDA-graph capturable).  p[l]: raw head map of level l,
 * logical (bs, na, ny, nx, nc+5) with element strides sb/sa/sy/sx/so; dp[l]: its gradient, same strides (backward
 * only; every logical element is written).  targets: [nt][6] = (image, class, x, y, w, h) normalised, rows with
 * image < 0 are padding.  anchors: [nl][na][2] in grid units (Detect.anchors).  Focal loss (fl_gamma > 0),
 * autobalance and sort_obj_iou are not supported.  workspace: ly_yolo_loss_workspace_bytes() bytes, written by
 * _fwd and read by _bwd.  out: [4] = (lbox + lobj + lcls) * bs | lbox | lobj | lcls (weighted by hyp box/obj/cls).
 * grad_scale: DEVICE scalar, the upstream gradient of out[0].
 * ------------------------------------------------------------------------------------------------------ */
typedef struct {
  int nl, na, nc, bs, nt;
  int ny[3], nx[3];
  const float* p[3];
  float* dp[3];
  long sb[3], sa[3], sy[3], sx[3], so[3];
  const float* anchors;
  const float* targets;
  float balance[3]; /* loss.py:112: [4.0, 1.0, 0.4] */
  float anchor_t, box, obj, cls, cls_pw, obj_pw, cp, cn; /* hyp + smooth_BCE targets */
} ly_loss_desc;
long ly_yolo_loss_workspace_bytes(const ly_loss_desc* d);
int ly_yolo_loss_fwd(const ly_loss_desc* d, void* workspace, float* out, ly_stream_t stream);
int ly_yolo_loss_bwd(const ly_loss_desc* d, const void* workspace, const float* grad_scale, ly_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Confidence filter + class-aware NMS on the decoded predictions (utils/general.py:884-994 non_max_suppression,
 * best-class branch, around torchvision.ops.nms), static shapes, no wall-clock time limit.
 * pred: fp32 [bs][rows][no]; det: fp32 [bs][max_det][6] = (x1, y1, x2, y2, conf, cls), rows past det_count[i] are 0;
 * n_candidates (may be NULL): candidates above conf_thres per image — when it exceeds max_candidates (<= 8192) the
 * surplus was dropped in arrival order and the result for that image is not the reference's.
 * workspace: ly_nms_workspace_bytes(bs, max_candidates) bytes.
 * ------------------------------------------------------------------------------------------------------ */
long ly_nms_workspace_bytes(int bs, int max_candidates);
int ly_detect_nms(const float* pred, int bs, int rows, int no, float conf_thres, float iou_thres, int agnostic,
                  int max_det, int max_candidates, void* workspace, float* det, int* det_count, int* n_candidates,
                  ly_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Optimiser step on flat fp32 buffers (train.py:330-341; utils/torch_utils.py:318-346 smart_optimizer = SGD with
 * Nesterov momentum, groups [0,n0) decayed weights | [n0,n1) BatchNorm weights | [n1,n) biases; 404-432 ModelEMA):
 * gradient scaling + norm clipping + SGD + EMA in one pass.  hp: DEVICE fp32 [8] = lr0, lr1, lr2, momentum,
 * weight_decay, max_norm (<= 0: off), ema_decay, grad_scale.  grad_sq: DEVICE scalar Σ g² (ly_sq_sum; ADDS into out),
 * NULL = no clipping.  ema may be NULL.  ly_ema_update: ema = d*ema + (1-d)*src (BatchNorm buffers).
 * ------------------------------------------------------------------------------------------------------ */
int ly_sq_sum(const float* g, long n, float* out, ly_stream_t stream);
int ly_sgd_ema_step(float* params, const float* grads, float* momentum_buf, float* ema, long n, long n0, long n1,
                    const float* hp, const float* grad_sq, ly_stream_t stream);
int ly_ema_update(float* ema, const float* src, long n, const float* hp, ly_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* LEADYOLO_H_ */
