"""The C = 80 fused FasterNet block kernel (csrc/fasternet_block.cu, namespace f80) hard-codes the layout of its weight
image; this pins lead_yolo_b200.packing._fasternet_weight_image80 to that layout element by element (CPU only)."""
import torch

from lead_yolo_b200.packing import fasternet_weight_image


def test_fnb80_weight_image_layout():
    g = torch.Generator().manual_seed(0)
    c, c2, cp = 80, 160, 20
    w1 = torch.randn(c2, c, 1, 1, generator=g)
    w2 = torch.randn(c, c2, 1, 1, generator=g)
    wp = torch.randn(cp, cp, 3, 3, generator=g)
    img, off_w2, off_wc = fasternet_weight_image(w1, w2, wp)
    # the constants of namespace f80
    assert off_w2 == 160 * 128
    off_w1b = off_w2 + 3 * 80 * 128
    assert off_wc == off_w1b + 160 * 32
    assert img.numel() * 2 == off_wc + 18 * 160 * 32 and img.dtype == torch.bfloat16
    bf = lambda t: t.to(torch.bfloat16)  # noqa: E731

    def sw128(base, r, col):      # element (r, col) of a [rows][64] SW128 tile starting at element offset `base`
        return img[base + r * 64 + (((col >> 3) ^ (r & 7)) << 3) + (col & 7)]

    def sw32(base, r, col):       # element (r, col) of a [rows][16] SW32 tile
        return img[base + r * 16 + (((col >> 3) ^ ((r >> 2) & 1)) << 3) + (col & 7)]

    w1m, w2m = w1[:, :, 0, 0], w2[:, :, 0, 0]
    for r, col in [(0, 0), (5, 19), (5, 20), (77, 43), (159, 63)]:
        want = bf(w1m[r, col]) if col >= cp else torch.tensor(0, dtype=torch.bfloat16)   # first Cp columns are folded into Wc
        assert sw128(0, r, col) == want, (r, col)
    for r, k in [(0, 0), (79, 63), (13, 64), (40, 127), (7, 128), (79, 159)]:
        assert sw128(off_w2 // 2 + (k // 64) * 80 * 64, r, k % 64) == bf(w2m[r, k]), (r, k)
    for r in (0, 33, 79):         # K padding of the last W2 chunk
        assert sw128(off_w2 // 2 + 2 * 80 * 64, r, 40) == 0
    for r, col in [(0, 64), (3, 71), (4, 72), (159, 79)]:
        assert sw32(off_w1b // 2, r, col - 64) == bf(w1m[r, col]), (r, col)
    for t in (0, 4, 8):
        wc = w1m[:, :cp] @ wp[:, :, t // 3, t % 3]      # [160][20]
        for r, col in [(0, 0), (9, 15), (100, 16), (159, 19)]:
            sl = col // 16
            got = sw32(off_wc // 2 + (sl * 9 + t) * 160 * 16, r, col % 16)
            assert got == bf(wc[r, col]), (t, r, col)
        assert sw32(off_wc // 2 + (9 + t) * 160 * 16, 50, 7) == 0      # channels 20..31 of slice 1 are padding
