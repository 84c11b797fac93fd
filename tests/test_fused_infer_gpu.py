"""Inference-only fused kernels at RAGGED sizes (tiles that hang over the image border, odd H/W): each is compared with
the oracle (fp32 restatement of the reference module, oracle/restate.py) or with a plain torch expression.
The model-level tests only see 160/80/40/20-pixel maps; the halo tiles (pitch Wt + 2) deserve their own edge cases."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from helpers import TOL_BF16, half_normal, oracle_params, randomise_bn, rel_l2   # noqa: E402
from oracle import restate as R                                                 # noqa: E402


def _mods():
    from lead_yolo_b200 import modules as M
    return M


@pytest.mark.parametrize("shape", [(2, 64, 37, 45), (1, 64, 8, 130), (3, 32, 19, 7), (2, 128, 40, 40), (1, 128, 37, 45),
                                   (1, 256, 33, 30)])
def test_conv3x3_halo_ragged(shape):
    """3x3 Conv with resident weights takes the halo path (one box per K chunk, row-shifted taps)."""
    M = _mods()
    n, c, h, w = shape
    torch.manual_seed(3)
    mod = M.Conv(c, c, 3, 1).cuda().eval()
    randomise_bn(mod, 5)
    x = half_normal(shape, 11, signed=True)
    with torch.no_grad():
        y = mod(x)
        ref = R.conv_block(oracle_params(mod), "m", x.float(), R.Ctx(False), k=3)
    assert y.shape == ref.shape
    assert rel_l2(y.float(), ref) < TOL_BF16


@pytest.mark.parametrize("case", [(2, 128, 128, 40, 40, 3), (1, 128, 128, 37, 45, 3), (1, 128, 128, 5, 3, 3),
                                  (3, 128, 128, 15, 14, 3), (5, 160, 96, 20, 20, 3), (2, 72, 112, 33, 30, 3),
                                  (64, 128, 128, 40, 40, 3), (2, 64, 256, 20, 20, 3), (3, 48, 208, 9, 31, 3),
                                  (1, 256, 256, 33, 30, 3), (5, 136, 200, 20, 20, 3), (1, 64, 320, 5, 3, 3)])
def test_conv_cta2_ragged(case):
    """3x3 convolutions whose weights fit in shared memory only when split over a CTA pair: conv_cta2_kernel (cta_group::2,
    M = 256, each CTA keeps Cout/2 weight rows resident, halo boxes).  Single tiles and odd tile counts (the peer CTA of
    the last unit has no tile), ragged channel chunks, four and two accumulators (Cout <= 128 / <= 256), the bench
    geometry; the last three cases are too large even for a pair and stay on the streaming path.  Forward in eval mode
    and input gradient (the dgrad convolution takes the same kernels) against the oracle."""
    M = _mods()
    n, c1, c2, h, w, k = case
    torch.manual_seed(3)
    mod = M.Conv(c1, c2, k, 1).cuda().eval()
    randomise_bn(mod, 5)
    x = half_normal((n, c1, h, w), 11, signed=True)
    with torch.no_grad():
        y = mod(x)
        ref = R.conv_block(oracle_params(mod), "m", x.float(), R.Ctx(False), k=k)
    assert y.shape == ref.shape
    assert rel_l2(y.float(), ref) < TOL_BF16
    # gradient with respect to the input (eval-mode BN: a fixed affine map, so the oracle's autograd is the reference)
    xg = x.clone().requires_grad_(True)
    g = half_normal(tuple(ref.shape), 12, signed=True)
    (mod(xg).float() * g.float()).sum().backward()
    xr = x.float().clone().requires_grad_(True)
    (R.conv_block(oracle_params(mod), "m", xr, R.Ctx(False), k=k) * g.float()).sum().backward()
    assert rel_l2(xg.grad.float(), xr.grad) < TOL_BF16


@pytest.mark.parametrize("shape", [(2, 24, 37, 45), (1, 40, 9, 70), (2, 24, 16, 14), (2, 80, 37, 45), (3, 80, 40, 40),
                                   (1, 80, 5, 3)])
def test_fasternet_block_ragged(shape):
    M = _mods()
    torch.manual_seed(4)
    mod = M.BasicStage(shape[1], 1).cuda().eval()
    randomise_bn(mod, 6)
    x = half_normal(shape, 12, signed=True)
    with torch.no_grad():
        y = mod(x)
        ref = R.basic_stage(oracle_params(mod), "m", x.float(), R.Ctx(False))
    assert rel_l2(y.float(), ref) < TOL_BF16


@pytest.mark.parametrize("hw", [(20, 20), (13, 31), (5, 3)])
def test_sppf_pool3_exact(hw):
    """ly_sppf_pool3 = three cascaded 5x5 max pools: bit-exact against torch on bf16 data."""
    from lead_yolo_b200 import lowlevel as ll
    h, w = hw
    n, c = 3, 32
    x = half_normal((n, c, h, w), 21, signed=True)
    buf = ll.empty_cl(n, 4 * c, h, w)
    buf[:, :c] = x
    ll.call("ly_sppf_pool3", buf[:, :c], 4 * c, n, h, w, c, buf[:, c:2 * c], 4 * c, c)
    ref = x.float()
    for k in range(3):
        ref = torch.nn.functional.max_pool2d(ref, 5, 1, 2)
        assert torch.equal(buf[:, (k + 1) * c:(k + 2) * c].float(), ref), f"pool depth {k + 1}"


def test_stem_u8_ragged():
    """Tensor-core stem on a uint8 image whose patch grid is not a multiple of the 32 x 8 stage."""
    M = _mods()
    torch.manual_seed(5)
    mod = M.PatchEmbed_FasterNet(3, 24, 4, 4).cuda().eval()
    randomise_bn(mod, 7)
    g = torch.Generator().manual_seed(9)
    img = torch.randint(0, 256, (2, 3, 52, 208), generator=g, dtype=torch.uint8).cuda()
    with torch.no_grad():
        y = mod(img)
        ref = R.patch_embed(oracle_params(mod), "m", img.float() / 255, R.Ctx(False))
    assert rel_l2(y.float(), ref) < TOL_BF16
