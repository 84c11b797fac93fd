"""Raw (no autograd) Python wrappers over the C-ABI.  Tensors are logical NCHW torch tensors stored
channels_last (NHWC in memory); a channel slice of a wider NHWC buffer is accepted (pixel stride `ld`)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import ConvDesc, check, lib, ptr, stream_ptr

CL = torch.channels_last
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
    check(lib().ly_conv_fwd(C.byref(d), stream_ptr()), "ly_conv_fwd")
    return out
