"""ly_rf_attention_noalloc — the entry point for callers that cannot allocate device memory at call time (a
server that gave away its whole memory arena at start-up, a real-time loop where cudaMalloc's device-wide
synchronisation is not allowed).  It keeps the max/mean maps of RFCBAMConv's spatial attention in an 8 MiB static
device array inside the library instead of a caller-provided buffer.

Checked here, through the C-ABI: its outputs equal those of the ordinary ly_rf_attention on two calls in a row
(with different inputs, so contents left behind by the first call would show), and sizes beyond the built-in
scratch are refused with a clear message.  This file sorts last on purpose: a stop-at-first-failure run has then
already executed every other test.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS = 2 * 1024 * 1024


def _inputs(seed, n, ho, wo, c, k):
    g = torch.Generator(device="cpu").manual_seed(seed)
    G = torch.randn(n * ho * wo, k * k * c, generator=g).abs().cuda().to(torch.bfloat16).contiguous()
    wa = (torch.randn(18, generator=g) * 0.3).cuda()
    return G, wa


def _run(name, G, wa, n, ho, wo, c, k, with_maps):
    from lead_yolo_b200 import lowlevel as ll
    amax = torch.full((n, ho * k, wo * k), -1, dtype=torch.int16, device="cuda")
    att = torch.full((n, ho * k, wo * k), float("nan"), dtype=torch.float32, device="cuda")
    if with_maps:
        maps = torch.empty((n, ho * k, wo * k, 2), dtype=torch.float32, device="cuda")
        ll.call(name, G, n, ho, wo, c, k, wa, maps, amax, att)
    else:
        ll.call(name, G, n, ho, wo, c, k, wa, amax, att)
    torch.cuda.synchronize()
    return att, amax


@pytest.mark.parametrize("k", [1, 3])
def test_noalloc_matches_regular_entry_point_twice(k):
    n, ho, wo, c = 3, 10, 12, 64
    for seed in (1, 2):                       # two calls in a row, different data
        G, wa = _inputs(seed, n, ho, wo, c, k)
        att_ref, amax_ref = _run("ly_rf_attention", G, wa, n, ho, wo, c, k, True)
        att, amax = _run("ly_rf_attention_noalloc", G, wa, n, ho, wo, c, k, False)
        assert torch.isfinite(att).all()
        assert torch.equal(att, att_ref)      # same kernels, same arithmetic: bit-identical
        assert torch.equal(amax, amax_ref)


def test_noalloc_rejects_oversize():
    from lead_yolo_b200 import lowlevel as ll
    from lead_yolo_b200._lib import LeadYoloError
    n, ho, wo, c, k = 64, 48, 48, 8, 3        # 64*144*144*2 floats > 2 Mi floats
    assert n * ho * k * wo * k * 2 > MAX_FLOATS
    G = torch.zeros(8, 8, dtype=torch.bfloat16, device="cuda")   # never read: the size check comes first
    wa = torch.zeros(18, device="cuda")
    att = torch.zeros(8, device="cuda")
    with pytest.raises(LeadYoloError, match="exceeds the built-in scratch"):
        ll.call("ly_rf_attention_noalloc", G, n, ho, wo, c, k, wa, None, att)
