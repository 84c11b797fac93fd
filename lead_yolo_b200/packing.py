"""Weight packing for the implicit-GEMM kernels (mirrors the K schedule of csrc/conv_gemm.cu).

Packed layout: bf16 [Cout][Kp]; K order = for source, for tap (row-major), for 64-channel chunk (zero
padded to 64).  For a patch convolution (kernel == stride) a tap is a kernel row ky and its "channels"
are the kw*C contiguous values (kx, c) of an NHWC pixel run.
"""
from __future__ import annotations

import torch


def _pad64(blk: torch.Tensor) -> torch.Tensor:
    k = blk.shape[1]
    kp = (k + 63) // 64 * 64
    if kp == k:
        return blk
    return torch.nn.functional.pad(blk, (0, kp - k))


def k_segments(src_channels, kh, kw, patch):
    """[(source, tap, width)] in packed order; each occupies ceil(width/64)*64 packed columns."""
    segs = []
    for s, cs in enumerate(src_channels):
        if patch:
            segs += [(s, ky, kw * cs) for ky in range(kh)]
        else:
            segs += [(s, t, cs) for t in range(kh * kw)]
    return segs


def packed_k(src_channels, kh, kw, patch) -> int:
    return sum((w + 63) // 64 * 64 for _, _, w in k_segments(src_channels, kh, kw, patch))


def pack_conv_weight(w: torch.Tensor, src_channels=None, patch: bool = False) -> torch.Tensor:
    """w: [Cout, Cin, kh, kw] (any float dtype) -> bf16 [Cout, Kp]."""
    cout, cin, kh, kw = w.shape
    if src_channels is None:
        src_channels = [cin]
    assert sum(src_channels) == cin, (src_channels, cin)
    w = w.detach().float()
    cols, off = [], 0
    for cs in src_channels:
        ws = w[:, off:off + cs]
        off += cs
        if patch:
            for ky in range(kh):
                cols.append(_pad64(ws[:, :, ky, :].permute(0, 2, 1).reshape(cout, kw * cs)))
        else:
            for t in range(kh * kw):
                cols.append(_pad64(ws[:, :, t // kw, t % kw]))
    return torch.cat(cols, 1).to(torch.bfloat16).contiguous()


def unpack_conv_weight(wp: torch.Tensor, cout: int, src_channels, kh: int, kw: int, patch: bool = False):
    """Inverse of pack_conv_weight for an fp32 [Cout, Kp] matrix (weight gradients) -> [Cout, Cin, kh, kw]."""
    cin = sum(src_channels)
    out = wp.new_zeros((cout, cin, kh, kw), dtype=torch.float32)
    col, off = 0, 0
    for cs in src_channels:
        if patch:
            for ky in range(kh):
                width = kw * cs
                blk = wp[:, col:col + width].float().reshape(cout, kw, cs).permute(0, 2, 1)
                out[:, off:off + cs, ky, :] = blk
                col += (width + 63) // 64 * 64
        else:
            for t in range(kh * kw):
                out[:, off:off + cs, t // kw, t % kw] = wp[:, col:col + cs].float()
                col += (cs + 63) // 64 * 64
        off += cs
    return out
