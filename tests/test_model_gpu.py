"""Whole-network parity on the GPU: LEAD-YOLO built from the drop-in modules vs (a) the real reference's outputs
committed in tests/golden/model_small.pt, (b) the oracle restatement at 640x640 incl. post-NMS box parity
(north_star: identical box count, IoU > 0.99), (c) a training-mode forward/backward."""
import os

import pytest
import torch

from cases import BN_SEED, IN_SEED, INIT_SEED
from helpers import TOL_BF16, oracle_params, randomise_bn, rel_l2
from oracle import restate as R

pytestmark = pytest.mark.gpu
GOLD_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _model(variant="lead"):
    from lead_yolo_b200.model import build_lead_yolo
    m = build_lead_yolo(variant, 1, "cuda", seed=INIT_SEED)
    randomise_bn(m, BN_SEED)
    return m


def _P(model):
    return {k[2:]: v for k, v in oracle_params(model, prefix="m").items()}


def test_param_count_and_state_dict_layout():
    m = _model()
    assert sum(p.numel() for p in m.parameters()) == 3_135_478     # SURVEY §0 / authors' log
    assert len(m.state_dict()) == 319


def test_small_input_vs_real_reference():
    gold = torch.load(os.path.join(GOLD_DIR, "model_small.pt"), map_location="cuda", weights_only=True)
    m = _model()
    assert gold["n_params"] == sum(p.numel() for p in m.parameters())
    img = gold["img"]
    m.eval()
    with torch.no_grad():
        dec, raw = m(img.float() / 255)
        dec8, _ = m(img)                       # uint8 image, /255 folded into the stem kernel
    for a, b in zip(raw, gold["raw"]):
        assert rel_l2(a, b) <= 2e-2            # 24 layers deep, bf16 storage between layers
    assert rel_l2(dec, gold["decoded"]) <= 2e-2
    assert rel_l2(dec8, dec) <= 1e-3
    m.train()
    with torch.no_grad():
        raw_tr = m(img.float() / 255)
    for a, b in zip(raw_tr, gold["raw_train"]):
        assert rel_l2(a, b) <= 5e-2            # batch statistics over as few as 4 positions at 64x64


def _sar_like(n, s, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.empty(n, 1, s, s).exponential_(1 / 40.0, generator=g).clamp_(0, 255)
    for i in range(n):
        for _ in range(3):
            cx, cy = torch.randint(40, s - 40, (2,), generator=g).tolist()
            w, h = torch.randint(10, 60, (2,), generator=g).tolist()
            x[i, 0, cy - h // 2:cy + h // 2, cx - w // 2:cx + w // 2] = torch.empty(1).uniform_(180, 255, generator=g)
    return x.expand(n, 3, s, s).round().to(torch.uint8).contiguous()


@pytest.mark.parametrize("variant", ["lead", "c3"])
def test_640_eval_and_nms_box_parity(variant):
    import torchvision
    m = _model(variant)
    # random-init objectness is ~0.02: raise the Detect objectness bias so that O(100) boxes pass conf 0.25
    with torch.no_grad():
        for mi in m.model[-1].m:
            mi.bias.view(3, -1)[:, 4] += 3.5
    img = _sar_like(2, 640, IN_SEED).cuda()
    m.eval()
    P = _P(m)
    with torch.no_grad():
        dec, raw = m(img)
        dec_ref, raw_ref, _ = R.model_forward(P, img.float() / 255, R.Ctx(False), variant=variant)
    assert dec.shape == dec_ref.shape == (2, 25200, 6)
    for a, b in zip(raw, raw_ref):
        assert rel_l2(a, b) <= 2e-2
    # Outputs are index-aligned (row j = same anchor/cell).  Box geometry and confidence parity for every candidate
    # the reference passes to NMS:
    n_total = 0
    for i in range(2):
        conf, conf_ref = dec[i][:, 4] * dec[i][:, 5], dec_ref[i][:, 4] * dec_ref[i][:, 5]
        cand = conf_ref > 0.25
        assert (conf - conf_ref).abs().max() < 2e-2
        xyxy = [torch.cat((d[:, :2] - d[:, 2:4] / 2, d[:, :2] + d[:, 2:4] / 2), 1) for d in (dec[i], dec_ref[i])]
        iou = torchvision.ops.box_iou(xyxy[0][cand], xyxy[1][cand]).diagonal()
        assert iou.min() > 0.99, f"same-candidate IoU min {iou.min():.4f}"
        # The post-NMS COUNT criterion (identical count, IoU > 0.99 per box) is asserted exactly on a head with
        # well-separated detections in tests/test_parity_full_gpu.py::test_planted_detections_identical_count_and_boxes:
        # this random-init head scores its candidates within bf16 noise of each other and of the thresholds.
        n_total += int(cand.sum())
    assert n_total > 0, "parity run produced no boxes: threshold/bias choice is not exercising NMS"


def test_nms_exact_count_on_top_candidates():
    """north_star box criterion (identical post-NMS count, IoU > 0.99 per box).  A random-init head scores all 25200
    candidates almost equally (adjacent scores differ by less than the bf16 noise), so which near-duplicate wins a
    tie is arbitrary in ANY bf16 run; to exercise the criterion deterministically, NMS runs on the reference's 60
    strongest candidates per image with the reference's ranking for both box sets — only box geometry decides."""
    import torchvision
    m = _model()
    img = _sar_like(2, 640, IN_SEED + 5).cuda()
    m.eval()
    P = _P(m)
    with torch.no_grad():
        dec, _ = m(img)
        dec_ref, _, _ = R.model_forward(P, img.float() / 255, R.Ctx(False))
    for i in range(2):
        conf_ref = dec_ref[i][:, 4] * dec_ref[i][:, 5]
        score, order = conf_ref.sort(descending=True)
        score, order = score[:60], order[:60]
        rank = torch.arange(60, 0, -1, device=score.device, dtype=torch.float32)   # strict ranking, no ties
        keeps, kept_boxes = [], []
        for d in (dec[i], dec_ref[i]):
            c = d[order]
            xyxy = torch.cat((c[:, :2] - c[:, 2:4] / 2, c[:, :2] + c[:, 2:4] / 2), 1)
            k = torchvision.ops.nms(xyxy, rank, 0.45)
            keeps.append(sorted(k.tolist()))
            kept_boxes.append(xyxy[k.sort()[0]])
        assert len(keeps[0]) == len(keeps[1]) > 0, f"post-NMS box count {len(keeps[0])} vs reference {len(keeps[1])}"
        assert keeps[0] == keeps[1]
        assert torchvision.ops.box_iou(kept_boxes[0], kept_boxes[1]).diagonal().min() > 0.99


def test_train_step_vs_oracle():
    """Training-mode forward/backward of the whole network at 640x640 against the fp32 oracle.  The floor for a
    bf16 implementation is measured live: the same oracle run under bf16 autocast (what the reference's own AMP
    path does, train.py:316) — our gradient must be as close to fp32 as that, within a small margin."""
    m = _model()
    m.train()
    img = _sar_like(2, 640, IN_SEED + 1).cuda().float() / 255

    def oracle_run(autocast):
        P = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running" not in k and
                 "anchors" not in k else v.clone()) for k, v in _P(m).items()}
        ctx = R.Ctx(True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            _, raw_ref, _ = R.model_forward(P, img, ctx, decode=False)
        g = torch.Generator(device="cpu").manual_seed(11)
        rs = [torch.randn(r.shape, generator=g).cuda() for r in raw_ref]
        sum((a.float() * r).sum() for a, r in zip(raw_ref, rs)).backward()
        grads = torch.cat([P[k].grad.flatten().float() for k, _ in m.named_parameters()])
        return [r.detach().float() for r in raw_ref], grads, ctx, rs

    raw32, g32, ctx, rs = oracle_run(False)
    raw16, g16, _, _ = oracle_run(True)
    raw = m(img)
    sum((a * r).sum() for a, r in zip(raw, rs)).backward()
    cos = lambda a, b: torch.nn.functional.cosine_similarity(a, b, dim=0).item()  # noqa: E731
    for a, b, c in zip(raw, raw32, raw16):
        assert rel_l2(a, b) <= max(3e-2, 2.0 * rel_l2(c, b))
    missing = [k for k, p in m.named_parameters() if p.grad is None]
    assert not missing, f"parameters without gradient (DDP static_graph needs all): {missing[:5]}"
    ga = torch.cat([p.grad.flatten().float() for _, p in m.named_parameters()])
    c_mine, c_ref16 = cos(ga, g32), cos(g16, g32)
    assert c_mine >= min(0.98, c_ref16 - 0.03), f"cosine(grad, fp32 oracle) = {c_mine:.4f}; bf16 oracle: {c_ref16:.4f}"
    assert 0.8 < (ga.norm() / g32.norm()).item() < 1.25
    sd = m.state_dict()
    worst = max(rel_l2(sd[k], v) for k, v in ctx.updates.items())
    assert worst <= 3e-2, f"running statistics rel-L2 {worst:.3e}"
