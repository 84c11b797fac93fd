"""Tensor-core training kernels against their CUDA-core counterparts at RAGGED sizes (tiles overhanging the image, odd
H / W): the receptive-field generation (ly_rf_tc_generate vs ly_rf_generate), its input gradient (ly_rf_tc_taps +
ly_rf_col2im vs the dx half of ly_rf_generate_bwd), its weight gradient (rf_wgrad_tc_kernel vs rf_generate_wgrad_kernel)
and the stem's weight gradient (ly_stem_im2col + ly_conv_wgrad vs
ly_stem_wgrad).  The CUDA-core kernels are themselves held to the oracle by the module tests (tests/test_modules_gpu.py);
weights are pre-rounded to bf16 here so that both sides multiply the same numbers."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from helpers import rel_l2   # noqa: E402

CL = torch.channels_last


def _nhwc(t):
    return t.to(torch.bfloat16).contiguous(memory_format=CL)


@pytest.mark.parametrize("case", [(3, 32, 37, 37, 2), (2, 128, 40, 40, 2), (1, 16, 5, 9, 2), (2, 48, 19, 23, 1),
                                  (1, 256, 20, 20, 2)])
def test_rf_generate_tc_matches_cuda_cores(case):
    from lead_yolo_b200 import lowlevel as ll, ops
    n, c, h, w, s = case
    g = torch.Generator(device="cuda").manual_seed(5)
    x = _nhwc(torch.randn(n, c, h, w, device="cuda", generator=g))
    wg = torch.randn(c * 9, 1, 3, 3, device="cuda", generator=g).bfloat16().float()
    wgk = ops.rf_pack_generate_weight(wg, c, 3)
    ho, wo = ops.rf_out_size(h, w, 3, s)
    z_tc = ops._rf_generate_raw(x, wgk, 3, s, ho, wo)
    z_cc = ll.empty_cl(n, 9 * c, ho, wo)
    ll.call("ly_rf_generate", x, c, n, h, w, c, 3, s, ho, wo, wgk, None, None, ll.ACT["none"], z_cc)
    assert rel_l2(z_tc.float(), z_cc.float()) < 3e-3          # both bf16 roundings of the same fp32 sums

    # input gradient: patch gradients on the tensor cores + fold, with the SE pooling term riding along
    dz = _nhwc(torch.randn(n, 9 * c, ho, wo, device="cuda", generator=g))
    extra = torch.randn(n, c, device="cuda", generator=g)
    dx_tc, _ = ops._rf_generate_bwd(dz, x, wgk, 3, s, ho, wo, extra, True, False)
    dx_cc = ll.empty_cl(n, c, h, w)
    ll.call("ly_rf_generate_bwd", dz, x, c, n, h, w, c, 3, s, ho, wo, wgk, extra, dx_cc, c, None, None, 0)
    # the tensor-core path rounds the nine patch gradients to bf16 before folding up to four of them
    assert rel_l2(dx_tc.float(), dx_cc.float()) < 6e-3

    # weight gradient: tensor cores (scratch rows given) against the CUDA-core kernel (no scratch: atomics)
    dw_tc = torch.zeros_like(wgk)
    scr = ll.rf_wgrad_scratch(c, 3, "cuda")
    ll.call("ly_rf_generate_bwd", dz, x, c, n, h, w, c, 3, s, ho, wo, wgk, None, None, c, dw_tc, scr, scr.numel())
    dw_cc = torch.zeros_like(wgk)
    ll.call("ly_rf_generate_bwd", dz, x, c, n, h, w, c, 3, s, ho, wo, wgk, None, None, c, dw_cc, None, 0)
    assert rel_l2(dw_tc, dw_cc) < 1e-4


@pytest.mark.parametrize("u8", [True, False])
@pytest.mark.parametrize("shape", [(2, 52, 64), (1, 128, 128), (3, 4, 16)])
def test_stem_wgrad_tensor_cores(shape, u8):
    from lead_yolo_b200 import lowlevel as ll, ops
    n, h, w = shape
    cout = 24
    g = torch.Generator(device="cuda").manual_seed(9)
    if u8:
        img = torch.randint(0, 256, (n, 3, h, w), device="cuda", generator=g, dtype=torch.uint8)
        in_scale = 1.0 / 255.0
    else:
        img = torch.rand(n, 3, h, w, device="cuda", generator=g).bfloat16().float()     # exact in the bf16 patch matrix
        in_scale = 1.0
    dz = _nhwc(torch.randn(n, cout, h // 4, w // 4, device="cuda", generator=g))
    dwk = torch.zeros(48, cout, device="cuda")
    ll.call("ly_stem_wgrad", img, 1 if u8 else 0, n, h, w, cout, dz, cout, in_scale, dwk)
    ref = dwk.t().reshape(cout, 3, 4, 4)
    pm = ll.empty_cl(n, 64, h // 4, w // 4)
    ll.call("ly_stem_im2col", img, 1 if u8 else 0, n, h, w, pm)
    assert float(pm[:, 48:].float().abs().max()) == 0.0
    dwp = ll.conv_wgrad([pm], dz, cout, k=1)
    got = dwp[:, :48].reshape(cout, 3, 4, 4) * in_scale
    assert rel_l2(got, ref) < 1e-4
