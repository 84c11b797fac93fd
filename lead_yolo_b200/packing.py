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


def dgrad_weight_matrix(w: torch.Tensor, coff: int, cs: int, patch: bool) -> torch.Tensor:
    """Packed bf16 weights of the GEMM that maps dZ to the gradient of source channels [coff, coff+cs) — the host
    statement of what csrc/weights.cu:pack_dgrad_kernel writes (the tests pin the kernel to it).  The output
    channel count is zero-padded to a multiple of 8 first (Detect's 18 -> 24)."""
    cout, _, kh, kw = w.shape
    cpad = (cout + 7) // 8 * 8
    ws = torch.nn.functional.pad(w.detach().float()[:, coff:coff + cs], (0, 0, 0, 0, 0, 0, 0, cpad - cout))
    if patch:
        wd = ws.permute(2, 3, 1, 0).reshape(kh * kw * cs, cpad, 1, 1)      # rows (ky, kx, c), K = output channel
    elif kh == 1:
        wd = ws.transpose(0, 1)
    else:
        wd = ws.flip(2, 3).transpose(0, 1)
    return pack_conv_weight(wd, [cpad])


# ---------------------------------------------------------------------------------------------------------
# Pre-swizzled shared-memory images (csrc/fasternet_block.cu): K-major tiles exactly as tcgen05 reads them, so
# the kernel fetches all weights of a block with ONE bulk copy.
#   SW128 tile: rows of 64 bf16 (128 B); 16-byte chunk j of row r is stored at chunk j ^ (r & 7)
#   SW32  tile: rows of 16 bf16 (32 B);  16-byte chunk j of row r is stored at chunk j ^ ((r >> 2) & 1)
# ---------------------------------------------------------------------------------------------------------
def _swizzle_rows(mat: torch.Tensor, chunks: int, fn) -> torch.Tensor:
    rows = mat.shape[0]
    m = mat.reshape(rows, chunks, 8)
    r = torch.arange(rows, device=mat.device).view(-1, 1)
    j = torch.arange(chunks, device=mat.device).view(1, -1)
    dst = (j ^ fn(r)).expand(rows, chunks)
    out = torch.empty_like(m)
    out.scatter_(1, dst.unsqueeze(-1).expand(rows, chunks, 8), m)
    return out.reshape(-1)


def swizzle128(mat: torch.Tensor) -> torch.Tensor:
    assert mat.shape[1] == 64 and mat.shape[0] % 8 == 0
    return _swizzle_rows(mat, 8, lambda r: r & 7)


def swizzle32(mat: torch.Tensor) -> torch.Tensor:
    assert mat.shape[1] == 16 and mat.shape[0] % 8 == 0
    return _swizzle_rows(mat, 2, lambda r: (r >> 2) & 1)


def fasternet_weight_image(w1: torch.Tensor, w2: torch.Tensor, wp: torch.Tensor):
    """(image bf16 1-D, off_w2, off_wc in bytes) for MLPBlock weights mlp.0 (2C,C,1,1), mlp.3 (C,2C,1,1) and
    partial_conv3 (Cp,Cp,3,3).  Sections: W1x [2C][64] | W2 [chunks][N2][64] | Wc [9][2C][16] with the partial conv folded
    into the first GEMM: Wc_tap = W1[:, :Cp] @ Wp[:, :, tap] (fp32 product, rounded to bf16 once)."""
    c2, c = w1.shape[0], w1.shape[1]
    cp = wp.shape[0]
    if c == 80:
        return _fasternet_weight_image80(w1, w2, wp)
    assert cp <= 16 and c <= 64 and c2 % 8 == 0
    n2 = (c + 15) // 16 * 16
    chunks = (c2 + 63) // 64
    f = lambda t: t.detach().float()  # noqa: E731
    pad = torch.nn.functional.pad
    w1m = f(w1)[:, :, 0, 0]
    w1x = w1m.clone()
    w1x[:, :cp] = 0
    parts = [swizzle128(pad(w1x, (0, 64 - c)).to(torch.bfloat16))]
    w2m = pad(f(w2)[:, :, 0, 0], (0, chunks * 64 - c2, 0, n2 - c))
    for kc in range(chunks):
        parts.append(swizzle128(w2m[:, kc * 64:(kc + 1) * 64].to(torch.bfloat16).contiguous()))
    off_w2 = parts[0].numel() * 2
    off_wc = off_w2 + sum(p.numel() for p in parts[1:]) * 2
    wpf = f(wp)
    for t in range(9):
        wc = w1m[:, :cp] @ wpf[:, :, t // 3, t % 3]                      # [2C][Cp]
        parts.append(swizzle32(pad(wc, (0, 16 - cp)).to(torch.bfloat16)))
    return torch.cat(parts).contiguous(), off_w2, off_wc


def _fasternet_weight_image80(w1: torch.Tensor, w2: torch.Tensor, wp: torch.Tensor):
    """C = 80 (csrc/fasternet_block.cu, namespace f80): W1x[:, :64] SW128 | W2 3 chunks [80][64] SW128 |
    W1x[:, 64:80] SW32 | Wc [slice 0..1][tap 0..8][160][16] SW32 (Cp = 20 channels = two K=16 slices)."""
    c2, c = w1.shape[0], w1.shape[1]
    cp = wp.shape[0]
    assert c == 80 and c2 == 160 and cp == 20
    f = lambda t: t.detach().float()  # noqa: E731
    pad = torch.nn.functional.pad
    bf = torch.bfloat16
    w1m = f(w1)[:, :, 0, 0]
    w1x = w1m.clone()
    w1x[:, :cp] = 0
    parts = [swizzle128(w1x[:, :64].to(bf).contiguous())]
    off_w2 = parts[0].numel() * 2
    w2m = pad(f(w2)[:, :, 0, 0], (0, 192 - c2))
    for kc in range(3):
        parts.append(swizzle128(w2m[:, kc * 64:(kc + 1) * 64].to(bf).contiguous()))
    parts.append(swizzle32(w1x[:, 64:80].to(bf).contiguous()))
    off_wc = sum(p.numel() for p in parts) * 2
    wpf = f(wp)
    wcs = [pad(w1m[:, :cp] @ wpf[:, :, t // 3, t % 3], (0, 32 - cp)) for t in range(9)]     # [160][32] each
    for sl in range(2):
        for t in range(9):
            parts.append(swizzle32(wcs[t][:, sl * 16:(sl + 1) * 16].to(bf).contiguous()))
    return torch.cat(parts).contiguous(), off_w2, off_wc


def rf_fused_weight_image(w: torch.Tensor, c: int) -> torch.Tensor:
    """RFCBAMConv.conv weight [Cout, C, 3, 3] -> bf16 [C/16][9 t][Cout][16] with every [Cout][16] tile in the SW32
    shared-memory image (csrc/rf_generate_tc.cu: rf_fused_kernel streams it with plain bulk copies)."""
    cout = w.shape[0]
    assert w.shape[1] == c and c % 16 == 0 and tuple(w.shape[2:]) == (3, 3) and cout % 8 == 0
    m = w.detach().float().permute(1, 2, 3, 0).reshape(c // 16, 16, 9, cout)          # [g][j][t][o]
    m = m.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)                          # [g][t][o][j]
    tiles = [swizzle32(m[g, t]) for g in range(c // 16) for t in range(9)]
    return torch.cat(tiles).contiguous()
