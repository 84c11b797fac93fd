"""GPU parity of the training-step kernels that sit either side of the module path: weight layout kernels, the
pivot-shifted BatchNorm statistics, the detection loss (vs the real-reference fixtures and the oracle restatement),
device NMS (vs the real-reference fixtures and the oracle), and the fused SGD + EMA step (vs torch.optim.SGD)."""
import os

import pytest
import torch

from helpers import rel_l2
from oracle import restate as R

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = "cuda"


def _gold():
    return torch.load(os.path.join(ROOT, "tests", "golden", "loss_nms.pt"))


# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("cout,srcs,k,patch", [(24, [48, 24], 3, False), (40, [24], 2, True), (18, [128], 1, False),
                                               (160, [8, 80], 1, False), (24, [3 * 8], 4, True), (256, [128], 3, False)])
def test_weight_layout_kernels_bit_exact(cout, srcs, k, patch):
    from lead_yolo_b200 import lowlevel as ll, ops
    from lead_yolo_b200.packing import dgrad_weight_matrix, pack_conv_weight, packed_k, unpack_conv_weight
    torch.manual_seed(0)
    w = torch.randn(cout, sum(srcs), k, k, device=DEV)
    assert torch.equal(ops.packed_weight(w, srcs, k, patch), pack_conv_weight(w, srcs, patch=patch))
    off = 0
    for cs in srcs:
        got = ops.packed_weight_dgrad(w, cout, sum(srcs), k, patch, off, cs)
        assert torch.equal(got, dgrad_weight_matrix(w, off, cs, patch)), (off, cs)
        off += cs
    cpad = (cout + 7) // 8 * 8
    dwp = torch.randn(cpad, packed_k(srcs, k, k, patch), device=DEV)
    dw = torch.empty(cout, sum(srcs), k, k, device=DEV)
    c4 = srcs + [0] * (4 - len(srcs))
    ll.call("ly_unpack_wgrad", dwp, cout, k, k, int(patch), *c4, dw, 0)
    assert torch.equal(dw, unpack_conv_weight(dwp, cpad, srcs, k, k, patch)[:cout])


def test_bn_statistics_do_not_cancel_for_large_means():
    """E[x²] - mean² in fp32 loses the variance when |mean| >> std; the pivot-shifted sums do not."""
    from lead_yolo_b200 import ops
    import torch.nn as nn
    torch.manual_seed(1)
    n, c, h, w = 8, 32, 40, 40
    x = (torch.randn(n, c, h, w, device=DEV) * 0.5 + 60.0).bfloat16().contiguous(memory_format=torch.channels_last)
    bn = nn.BatchNorm2d(c, eps=1e-3, momentum=0.03).to(DEV).train()
    y = ops.bn_act(x, bn, "none")
    xd = x.double()
    mean, var = xd.mean((0, 2, 3)), xd.var((0, 2, 3), unbiased=False)
    ref = (xd - mean.view(1, -1, 1, 1)) / torch.sqrt(var.view(1, -1, 1, 1) + 1e-3)
    assert rel_l2(y.double(), ref) < 6e-3          # bf16 output rounding only
    m = n * h * w
    assert torch.allclose(bn.running_mean.double(), 0.03 * mean, rtol=1e-5)
    assert torch.allclose(bn.running_var.double(), 0.97 + 0.03 * var * m / (m - 1), rtol=1e-4)


def test_bn_fold_cache_sees_training_updates():
    """eval -> train forward -> eval must not reuse the stale folded scale/shift."""
    from lead_yolo_b200 import modules as M
    torch.manual_seed(0)
    m = M.Conv(32, 32, 1, 1).to(DEV)
    x = torch.randn(4, 32, 16, 16, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        m.eval()
        y0 = m(x).float()
        m.train()
        for _ in range(3):
            m(x * 3 + 1)
        m.eval()
        y1 = m(x).float()
    P = {"m." + k: v.detach().float().cpu() for k, v in m.state_dict().items()}
    ref = R.conv_block(P, "m", x.float().cpu(), R.Ctx(False, eps=m.bn.eps), k=1)
    assert rel_l2(y1.cpu(), ref) < 1e-2
    assert rel_l2(y0.cpu(), ref) > 5e-2            # the statistics really moved


# ---------------------------------------------------------------------------------------------------------------
def _loss_ours(p, targets, anchors, hyp, nc):
    from lead_yolo_b200.loss import YoloLossFn, smooth_BCE
    cp, cn = smooth_BCE(0.0)
    cfg = {"na": 3, "nc": nc, "balance": [4.0, 1.0, 0.4], "anchor_t": hyp["anchor_t"], "box": hyp["box"],
           "obj": hyp["obj"], "cls": hyp["cls"], "cls_pw": hyp["cls_pw"], "obj_pw": hyp["obj_pw"], "cp": cp, "cn": cn}
    out = YoloLossFn.apply(cfg, anchors, targets, *p)
    return out[0:1], out[1:4].detach()


@pytest.mark.parametrize("name", ["nc1_128", "nc3_160", "nc1_dups"])
def test_loss_matches_reference_fixture(name):
    c = _gold()["loss"][name]
    p = [t.to(DEV).requires_grad_(True) for t in c["p"]]
    pad = torch.full((7, 6), -1.0)
    targets = torch.cat([c["targets"][:2], pad, c["targets"][2:]]).to(DEV)        # padding rows anywhere
    loss, items = _loss_ours(p, targets, c["anchors"].to(DEV), c["hyp"], c["nc"])
    (loss * 3.0).sum().backward()
    assert torch.allclose(loss.cpu(), c["loss"], rtol=2e-5, atol=1e-6), (loss, c["loss"])
    assert torch.allclose(items.cpu(), c["items"], rtol=2e-5, atol=1e-6)
    for a, b in zip(p, c["dp"]):
        assert rel_l2(a.grad.cpu(), 3.0 * b) < 1e-5


def test_loss_full_size_strided_maps_vs_oracle():
    """bs=64, 640x640 geometry, head maps as strided views of an NHWC buffer (what Detect returns in training)."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from make_golden_post import ANCHORS, synthetic_targets
    bs, nc = 64, 1
    strides = (8, 16, 32)
    anchors = ANCHORS.view(3, 3, 2) / torch.tensor(strides).view(3, 1, 1)
    g = torch.Generator().manual_seed(5)
    base = [torch.randn((bs, 640 // s, 640 // s, 3 * (nc + 5)), generator=g) for s in strides]
    targets = synthetic_targets(bs, 11, nc)
    hyp = R.scaled_hyp(3, nc, 640)

    def views(ts, dev):
        out = []
        for t in ts:
            t = t.to(dev).requires_grad_(True)
            b, ny, nx, _ = t.shape
            out.append((t, t.view(b, ny, nx, 3, nc + 5).permute(0, 3, 1, 2, 4)))
        return out
    ours = views(base, DEV)
    ref = views(base, "cpu")
    loss, items = _loss_ours([v for _, v in ours], targets.to(DEV), anchors.to(DEV), hyp, nc)
    loss.sum().backward()
    rl, ri = R.compute_loss([v for _, v in ref], targets, anchors, hyp, nc)
    rl.backward()
    assert torch.allclose(loss.cpu(), rl, rtol=1e-4), (loss, rl)
    assert torch.allclose(items.cpu(), ri, rtol=1e-4, atol=1e-7)
    for (a, _), (b, _) in zip(ours, ref):
        assert rel_l2(a.grad.cpu(), b.grad) < 1e-4


# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["nc1", "nc3"])
def test_nms_matches_reference_fixture(name):
    from lead_yolo_b200.postprocess import non_max_suppression
    c = _gold()["nms"][name]
    out = non_max_suppression(c["pred"].to(DEV), 0.25, 0.45)
    for a, b in zip(out, c["out"]):
        assert a.shape == b.shape, (a.shape, b.shape)
        assert torch.equal(a.cpu(), b)


def test_nms_full_size_vs_oracle_and_static_outputs():
    """bs=8 x 25200 rows with planted clusters; kept boxes identical to the oracle, in order; padded rows zero."""
    from lead_yolo_b200.postprocess import nms_static, non_max_suppression
    bs, rows = 8, 25200
    g = torch.Generator().manual_seed(3)
    pred = torch.zeros(bs, rows, 6)
    centres = torch.rand((bs, 40, 2), generator=g) * 560 + 40
    idx = torch.randint(0, 40, (bs, rows), generator=g)
    pred[..., 0:2] = torch.gather(centres, 1, idx[..., None].expand(-1, -1, 2)) + torch.randn((bs, rows, 2), generator=g) * 5
    pred[..., 2:4] = torch.rand((bs, rows, 2), generator=g) * 50 + 15
    pred[..., 4] = torch.rand((bs, rows), generator=g) ** 12          # ~1500 candidates above 0.25 per image
    pred[..., 5] = torch.rand((bs, rows), generator=g) * 0.3 + 0.7
    ref = R.non_max_suppression(pred, 0.25, 0.45)
    out = non_max_suppression(pred.to(DEV), 0.25, 0.45)
    det, count, ncand = nms_static(pred.to(DEV), 0.25, 0.45)
    for i, (a, b) in enumerate(zip(out, ref)):
        assert a.shape == b.shape and torch.equal(a.cpu(), b), i
        assert int(count[i]) == b.shape[0] and float(det[i, b.shape[0]:].abs().sum()) == 0.0
        assert 300 < int(ncand[i]) < 4096
    # empty image and an image with a single candidate
    pred2 = pred[:2].clone()
    pred2[0, :, 4] = 0.0
    pred2[1, :, 4] = 0.0
    pred2[1, 77, 4] = 0.9
    out2 = non_max_suppression(pred2.to(DEV), 0.25, 0.45)
    assert out2[0].shape == (0, 6) and out2[1].shape == (1, 6)


# ---------------------------------------------------------------------------------------------------------------
def test_fused_sgd_ema_matches_torch():
    from lead_yolo_b200 import lowlevel as ll
    torch.manual_seed(0)
    n0, n1, n = 4000, 4400, 5004
    p = torch.randn(n, device=DEV)
    p_ref = [p[:n0].clone().requires_grad_(True), p[n0:n1].clone().requires_grad_(True), p[n1:].clone().requires_grad_(True)]
    opt = torch.optim.SGD([{"params": [p_ref[0]], "weight_decay": 5e-4, "lr": 0.01},
                           {"params": [p_ref[1]], "weight_decay": 0.0, "lr": 0.02},
                           {"params": [p_ref[2]], "weight_decay": 0.0, "lr": 0.1}], lr=0.01, momentum=0.937,
                          nesterov=True)
    mom = torch.zeros(n, device=DEV)
    ema = p.clone()
    ema_ref = p.clone()
    d = 0.9
    hp = torch.tensor([0.01, 0.02, 0.1, 0.937, 5e-4, 10.0, d, 0.5], device=DEV)
    for it in range(3):
        g = torch.randn(n, device=DEV) * (40.0 if it == 1 else 1.0)       # step 1 gets clipped
        gsq = torch.zeros(1, device=DEV)
        ll.call("ly_sq_sum", g, n, gsq)
        assert torch.allclose(gsq, (g.double() ** 2).sum().float(), rtol=1e-5)
        ll.call("ly_sgd_ema_step", p, g, mom, ema, n, n0, n1, hp, gsq)
        for q, sl in zip(p_ref, (slice(0, n0), slice(n0, n1), slice(n1, n))):
            q.grad = (g[sl] * 0.5).clone()
        torch.nn.utils.clip_grad_norm_(p_ref, max_norm=10.0)
        opt.step()
        ema_ref = d * ema_ref + (1 - d) * torch.cat([q.detach() for q in p_ref])
        assert torch.allclose(p, torch.cat([q.detach() for q in p_ref]), rtol=1e-5, atol=1e-6)
        assert torch.allclose(ema, ema_ref, rtol=1e-5, atol=1e-6)


def test_training_forward_is_bit_reproducible_and_gradients_agree():
    """Two identical training steps: raw heads and loss bit-identical (fixed-point BatchNorm statistics, ordered SE
    pooling), gradients equal up to the fp32 atomics of the weight-gradient kernels."""
    from lead_yolo_b200.loss import ComputeLoss
    from lead_yolo_b200.model import build_lead_yolo
    from lead_yolo_b200.train import HYP_DEFAULT, scale_hyp
    g = torch.Generator().manual_seed(9)
    img = torch.randint(0, 256, (4, 3, 128, 128), generator=g, dtype=torch.uint8).cuda()
    tg = torch.tensor([[i, 0, 0.3 + 0.1 * i, 0.5, 0.2, 0.25] for i in range(4)], device=DEV)
    outs = []
    for _ in range(4):
        m = build_lead_yolo("lead", 1, DEV, seed=1).train()
        m.hyp = scale_hyp(HYP_DEFAULT, 3, 1, 128)
        raw = m(img)
        loss, _ = ComputeLoss(m)(raw, tg)
        loss.backward()
        outs.append(([r.detach().clone() for r in raw], loss.detach().clone(),
                     torch.cat([p.grad.flatten() for p in m.parameters()])))
    for raw, loss, grad in outs[1:]:
        assert all(torch.equal(a, b) for a, b in zip(raw, outs[0][0]))
        assert torch.equal(loss, outs[0][1])
        assert ((grad - outs[0][2]).norm() / outs[0][2].norm()).item() < 1e-4


def test_pack_batch_matches_single_launches():
    """ly_pack_batch (every weight of a captured training step in one launch) writes exactly what ly_pack_weight /
    ly_pack_weight_dgrad write one weight at a time: forward layouts with one and several K-concatenated sources, 3x3 and
    k = stride patch convolutions, input-gradient layouts of channel slices."""
    from lead_yolo_b200 import lowlevel as ll, ops
    g = torch.Generator(device="cuda").manual_seed(3)
    ws = {
        "a": torch.randn(128, 336, 1, 1, device="cuda", generator=g, requires_grad=True),
        "b": torch.randn(64, 64, 3, 3, device="cuda", generator=g, requires_grad=True),
        "c": torch.randn(80, 40, 2, 2, device="cuda", generator=g, requires_grad=True),
        "d": torch.randn(18, 128, 1, 1, device="cuda", generator=g, requires_grad=True),
    }
    calls = [
        ("f", "a", ([256, 80], 1, False)), ("f", "b", ([64], 3, False)), ("f", "c", ([40], 2, True)), ("f", "d", ([128], 1, False)),
        ("d", "a", (128, 336, 1, False, 0, 256)), ("d", "a", (128, 336, 1, False, 256, 80)), ("d", "b", (64, 64, 3, False, 0, 64)),
        ("d", "c", (80, 40, 2, True, 0, 40)), ("d", "d", (18, 128, 1, False, 0, 128)),
    ]

    def run_all():
        out = []
        for kind, name, a in calls:
            out.append(ops.packed_weight(ws[name], *a) if kind == "f" else ops.packed_weight_dgrad(ws[name], *a))
        return out

    ops.GRAD_SINK = {w.data_ptr(): w for w in ws.values()}      # marks the weights as parameters of a running step
    ops.BYPASS_PARAM_CACHE = True
    ops.PACK_RECORD = []
    try:
        single = [t.clone() for t in run_all()]
        table, njobs, blocks, plan = ops.pack_batch_table(ops.PACK_RECORD, "cuda")
    finally:
        ops.PACK_RECORD = None
    assert njobs == len(calls)
    for buf in plan.values():
        buf.fill_(float("nan"))
    ll.call("ly_pack_batch", table, njobs, blocks)
    ops.PACK_PLAN = plan
    try:
        batched = run_all()
    finally:
        ops.PACK_PLAN = None
        ops.GRAD_SINK = {}
        ops.BYPASS_PARAM_CACHE = False
    for s, b in zip(single, batched):
        assert torch.equal(s.view(torch.int16), b.view(torch.int16))
