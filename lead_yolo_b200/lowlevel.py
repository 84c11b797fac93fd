"""Raw (no autograd) Python wrappers over the C-ABI.  Tensors are logical NCHW torch tensors stored
channels_last (NHWC in memory); a channel slice of a wider NHWC buffer is accepted (pixel stride `ld`)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import ConvDesc, LeadYoloError, check, lib, ptr, stream_ptr

CL = torch.channels_last
COUNT = [0]        # kernels launched through the C-ABI (bench.py: gpu_launches)
PROFILE = None     # when a list: (entry point, start event, end event, algorithmic bytes, flops) per call
_LAUNCHES = {"ly_gap_mean": 2, "ly_yolo_loss_fwd": 4, "ly_yolo_loss_bwd": 2, "ly_detect_nms": 2, "ly_rfcbam_front_infer": 3, "ly_bn_act_bwd": 3, "ly_rf_attention": 2, "ly_rf_attention_bwd": 2, "ly_rf_generate_bwd": 3}


class _Prof:
    __slots__ = ("name", "nbytes", "flops", "e0")

    def __init__(self, name, nbytes=0.0, flops=0.0):
        self.name, self.nbytes, self.flops = name, nbytes, flops

    def __enter__(self):
        COUNT[0] += _LAUNCHES.get(self.name, 1)
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *a):
        if PROFILE is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            PROFILE.append((self.name, self.e0, e1, self.nbytes, self.flops))
        return False


ACT = {"none": 0, "relu": 1, "silu": 2, "hswish": 3, "sigmoid": 4}


def empty_cl(n, c, h, w, dtype=torch.bfloat16, device="cuda"):
    return torch.empty((n, c, h, w), dtype=dtype, device=device, memory_format=CL)


def zeros_cl(n, c, h, w, dtype=torch.bfloat16, device="cuda"):
    return torch.zeros((n, c, h, w), dtype=dtype, device=device, memory_format=CL)


def pixel_ld(x: torch.Tensor):
    """Pixel stride (elements) of an NHWC-in-memory view, or None if x is not such a view."""
    n, c, h, w = x.shape
    sn, sc, sh, sw = x.stride()
    if w > 1:
        ld = sw
    elif h > 1:
        ld = sh
    elif n > 1:
        ld = sn
    else:
        ld = c
    ok = (c == 1 or sc == 1) and (w == 1 or sw == ld) and (h == 1 or sh == w * ld) and \
         (n == 1 or sn == h * w * ld) and ld >= c
    return ld if ok else None


def as_nhwc(x: torch.Tensor, dtype=torch.bfloat16):
    """Return (tensor, ld): x in `dtype`, NHWC in memory (copying only when needed)."""
    if x.dtype != dtype:
        x = x.to(dtype)
    ld = pixel_ld(x)
    if ld is None or (x.data_ptr() % 16) != 0 or (ld % 8) != 0:
        x = x.contiguous(memory_format=CL)
        ld = pixel_ld(x)
        if ld is None:  # degenerate strides of size-1 dims
            x = x.clone(memory_format=CL)
            ld = x.shape[1]
    return x, ld


def conv_gemm(srcs, w_packed, cout, k=1, stride=1, pad=0, scale=None, shift=None, act="none",
              residual=None, out=None, out_fp32=False, patch_scatter=None):
    """out = act(scale * conv(cat(srcs), W) + shift) (+ residual).  See ly_conv_fwd in include/leadyolo.h.

    srcs: list of logical-NCHW bf16 tensors (NHWC memory) with equal (N,H,W).
    patch_scatter=(k, c): treat the GEMM columns as (ky,kx,c) and scatter to an (N, c, H*k, W*k) tensor.
    """
    d = ConvDesc()
    keep = []
    n, _, h, w = srcs[0].shape
    d.n_src = len(srcs)
    for i, s in enumerate(srcs):
        s, ld = as_nhwc(s)
        keep.append(s)
        assert s.shape[0] == n and s.shape[2] == h and s.shape[3] == w
        d.src[i] = s.data_ptr()
        d.src_c[i] = s.shape[1]
        d.src_ld[i] = ld
    d.N, d.H, d.W = n, h, w
    d.kh = d.kw = k
    d.stride, d.pad = stride, pad
    d.Cout = cout
    assert w_packed.dtype == torch.bfloat16 and w_packed.is_contiguous()
    d.w_packed = w_packed.data_ptr()
    if scale is not None:
        assert scale.dtype == torch.float32 and scale.numel() == cout
        d.scale = scale.data_ptr()
    if shift is not None:
        assert shift.dtype == torch.float32 and shift.numel() == cout
        d.shift = shift.data_ptr()
    d.act = ACT[act]
    ho, wo = (h, w) if stride == 1 else (h // stride, w // stride)
    if patch_scatter is not None:
        pk, pc = patch_scatter
        oshape = (n, pc, h * pk, w * pk)
        d.patch_k, d.patch_c = pk, pc
    else:
        oshape = (n, cout, ho, wo)
    if out is None:
        out = torch.empty(oshape, dtype=torch.float32 if out_fp32 else torch.bfloat16, device=srcs[0].device,
                          memory_format=CL)
    else:
        assert tuple(out.shape) == oshape, (tuple(out.shape), oshape)
    old = pixel_ld(out)
    assert old is not None, "out must be NHWC in memory"
    d.out = out.data_ptr()
    d.out_ld = old
    d.out_fp32 = 1 if out.dtype == torch.float32 else 0
    if residual is not None:
        residual, rld = as_nhwc(residual)
        keep.append(residual)
        assert tuple(residual.shape) == oshape
        d.residual = residual.data_ptr()
        d.res_ld = rld
    m_out = out.numel() // out.shape[1]
    cin = sum(int(x.shape[1]) for x in keep[:len(srcs)])
    nbytes = sum(x.numel() * 2 for x in keep) + out.numel() * out.element_size() + w_packed.numel() * 2
    with _Prof("ly_conv_fwd", nbytes, 2.0 * (n * ho * wo) * cin * k * k * cout):
        check(lib().ly_conv_fwd(C.byref(d), stream_ptr()), "ly_conv_fwd")
    return out


# ---------------------------------------------------------------------------------------------------------
# Generic typed call into the C-ABI.  Signature letters: p = device pointer (torch tensor or None),
# i = int, l = long, f = float, s = stream (appended automatically).
# ---------------------------------------------------------------------------------------------------------
_SIGS = {
    "ly_col_stats": "pllipp",
    "ly_bn_finalize": "ppilffpppppppp",
    "ly_bn_fold": "ifppppppppp",
    "ly_scale_shift_act": "pllippipl",
    "ly_bn_act_bwd": "plpllippppiipppppl",
    "ly_bn_stats": "pllip",
    "ly_bn_apply_train": "pllipppffpppipl",
    "ly_pack_weight": "piiiiiiiip",
    "ly_pack_weight_dgrad": "piiiiiiip",
    "ly_unpack_wgrad": "piiiiiiiipi",
    "ly_pack_batch": "pil",
    "ly_sq_sum": "plp",
    "ly_sgd_ema_step": "pppplllpp",
    "ly_ema_update": "pplp",
    "ly_act_bwd_from_out": "plplliipl",
    "ly_stem_fwd": "piiiiipppfpl",
    "ly_stem_wgrad": "piiiiiplfp",
    "ly_stem_im2col": "piiiip",
    "ly_fasternet_block_infer": "ppiiiipiiipp",
    "ly_maxpool5_fwd": "pliiiiplp",
    "ly_maxpool5_bwd": "plpiiiipli",
    "ly_sppf_pool3": "pliiiipll",
    "ly_copy_upsample": "pliiiiipl",
    "ly_downsum2": "pliiiipl",
    "ly_gap": "pliiipl",
    "ly_detect_decode": "piiiiifpppll",
    "ly_detect_head_infer": "pliiiippiifpppll",
    "ly_ca_pool": "pliiiipl",
    "ly_ca_mid_infer": "plllii" + "ppppppp" + "pl",
    "ly_ca_apply": "plpliiiipl",
    "ly_ca_gate_grad": "plplpliiiipl",
    "ly_ca_pool_bwd": "pliiiipl",
    "ly_rf_generate": "pliiiiiiiipppip",
    "ly_rf_attention": "piiiiipppp",
    "ly_rf_attention_noalloc": "piiiiippp",
    "ly_rf_gate": "piiiiiplpp",
    "ly_rfcbam_front_infer": "pliiiiiiiipppppipppppp",
    "ly_rf_tc_stats": "pliiiiiiippp",
    "ly_rf_tc_apply": "pliiiiiiippppp",
    "ly_rf_tc_generate": "pliiiiiiippp",
    "ly_rf_tc_taps": "piiiippp",
    "ly_rf_col2im": "piiiiiiippl",
    "ly_rfcbam_fused_infer": "pliiiiiiipppppppipl",
    "ly_rf_gap_sum": "pliiip",
    "ly_gap_mean": "pliiipp",
    "ly_rf_se_att": "piiippipiippp",
    "ly_rf_gate_bwd": "ppiiiiiplpppp",
    "ly_rf_attention_bwd": "ppppiiiiipppp",
    "ly_rf_generate_bwd": "ppliiiiiiiippplppl",
}
_CT = {"p": C.c_void_p, "i": C.c_int, "l": C.c_long, "f": C.c_float}
_prepared = {}


def call(name, *args):
    fn = _prepared.get(name)
    sig = _SIGS[name]
    if fn is None:
        fn = getattr(lib(), name)
        fn.argtypes = [_CT[ch] for ch in sig] + [C.c_void_p]
        fn.restype = C.c_int
        _prepared[name] = fn
    assert len(args) == len(sig), f"{name}: expected {len(sig)} args, got {len(args)}"
    conv = []
    for ch, a in zip(sig, args):
        if ch == "p":
            conv.append(None if a is None else a.data_ptr())
        elif ch == "f":
            conv.append(float(a))
        else:
            conv.append(int(a))
    nbytes = 0.0
    if PROFILE is not None:
        nbytes = float(sum(a.numel() * a.element_size() for a in args if isinstance(a, torch.Tensor)))
    with _Prof(name, nbytes, 0.0):
        check(fn(*conv, torch.cuda.current_stream().cuda_stream), name)


def rf_wgrad_scratch(c, k, device):
    """fp32 scratch of ly_rf_generate_bwd's weight-gradient pass (per-lane partial sums)."""
    fn = lib().ly_rf_wgrad_scratch_floats
    fn.argtypes = [C.c_int, C.c_int]
    fn.restype = C.c_long
    return torch.empty(int(fn(c, k)), dtype=torch.float32, device=device)


def reduce_blocks(m, c):
    """Blocks the library's reducing kernels use for an [m][c] matrix (size of their partial-sum scratch)."""
    fn = lib().ly_reduce_blocks
    fn.argtypes = [C.c_long, C.c_int]
    fn.restype = C.c_int
    return int(fn(m, c))


def conv_wgrad(srcs, dz, cout, k=1, stride=1, pad=0, zeros=None, out=None):
    """fp32 [Cout, Kp] packed weight gradient of conv(cat(srcs)) given the output gradient dz (bf16 NHWC).
    `zeros(n)` supplies the zero-initialised accumulator (the training arena); default torch.zeros.  `out`: an already
    ZEROED fp32 buffer of Cout*Kp elements to accumulate into instead."""
    from .packing import packed_k
    d = ConvDesc()
    keep = []
    n, _, h, w = srcs[0].shape
    d.n_src = len(srcs)
    for i, s in enumerate(srcs):
        s, ld = as_nhwc(s)
        keep.append(s)
        d.src[i] = s.data_ptr()
        d.src_c[i] = s.shape[1]
        d.src_ld[i] = ld
    d.N, d.H, d.W = n, h, w
    d.kh = d.kw = k
    d.stride, d.pad = stride, pad
    d.Cout = cout
    dz, dld = as_nhwc(dz)
    kp = packed_k([s.shape[1] for s in srcs], k, k, patch=(stride == k and k > 1))
    if out is not None:
        assert out.numel() == cout * kp and out.dtype == torch.float32 and out.is_contiguous()
        dw = out.view(cout, kp)
    elif zeros is None:
        dw = torch.zeros((cout, kp), dtype=torch.float32, device=dz.device)
    else:
        dw = zeros(cout * kp, dz.device).view(cout, kp)
    fn = lib().ly_conv_wgrad
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    cin = sum(int(x.shape[1]) for x in keep)
    nbytes = sum(x.numel() * 2 for x in keep) + dz.numel() * 2 + dw.numel() * 4
    with _Prof("ly_conv_wgrad", nbytes, 2.0 * (dz.numel() // dz.shape[1]) * cin * k * k * cout):
        check(fn(C.byref(d), dz.data_ptr(), dld, dw.data_ptr(), torch.cuda.current_stream().cuda_stream),
              "ly_conv_wgrad")
    return dw
