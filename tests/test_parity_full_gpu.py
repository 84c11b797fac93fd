"""GPU parity where round 1 was soft ('Harden parity'):

  * the north_star box criterion — IDENTICAL post-NMS count and IoU > 0.99 per box at conf 0.25 / IoU 0.45 — on a Detect
    head with planted, well-separated detections.  A random-init head scores all 25200 cells almost equally, so which
    near-duplicate survives is noise in ANY bf16 run; here the head is constructed from the oracle's own features so that
    its detections are separated from the thresholds and from each other by more than the measured bf16 error, which the
    test verifies before it asserts anything (the construction is checked, not assumed);
  * every module at the geometry the bench runs (bs = 64, LEAD-YOLO's real channel counts / resolutions) against the
    oracle evaluated in fp32 on the same GPU (TF32 off) — multi-wave persistent scheduling, 32-bit index switches;
  * configs[4]: 1024x1024, 4 images, whole model, eval;  configs[3]: the C3 ablation graph, training forward/backward.
"""
import pytest
import torch

from cases import BN_SEED, IN_SEED, INIT_SEED
from helpers import oracle_params, randomise_bn, rel_l2
from oracle import restate as R

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def _no_tf32():
    a, b = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = a, b


def _model(variant="lead"):
    from lead_yolo_b200.model import build_lead_yolo
    m = build_lead_yolo(variant, 1, DEV, seed=INIT_SEED)
    randomise_bn(m, BN_SEED)
    return m


def _P(model, device="cpu"):
    return {k[2:]: v.to(device) for k, v in oracle_params(model, prefix="m").items()}


def _sar_like(n, s, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.empty(n, 1, s, s).exponential_(1 / 40.0, generator=g).clamp_(0, 255)
    for i in range(n):
        for _ in range(4):
            cx, cy = torch.randint(60, s - 60, (2,), generator=g).tolist()
            w, h = torch.randint(10, 60, (2,), generator=g).tolist()
            x[i, 0, cy - h // 2:cy + h // 2, cx - w // 2:cx + w // 2] = torch.empty(1).uniform_(180, 255, generator=g)
    return x.expand(n, 3, s, s).round().to(torch.uint8).contiguous()


# ---------------------------------------------------------------------------------------------------------------
# planted detections
# ---------------------------------------------------------------------------------------------------------------
STRIDES = (8.0, 16.0, 32.0)


def _nms_rows(dec, conf_thres=0.25, iou_thres=0.45):
    """Reference NMS (oracle restatement) on one image's decoded rows -> kept ROW indices in rank order."""
    conf = dec[:, 4] * dec[:, 5]
    cand = ((dec[:, 4] > conf_thres) & (conf > conf_thres)).nonzero().flatten()
    if cand.numel() == 0:
        return cand
    order = torch.sort(conf[cand], descending=True, stable=True)[1]
    cand = cand[order]
    d = dec[cand]
    box = torch.stack((d[:, 0] - d[:, 2] / 2, d[:, 1] - d[:, 3] / 2, d[:, 0] + d[:, 2] / 2, d[:, 1] + d[:, 3] / 2), 1)
    return cand[R.nms_greedy(box, conf[cand], iou_thres)[:300]]


def _plant_head(P, feats, level, anchor, j, theta, k):
    """Detect weights for one level: objectness of `anchor` = k * (feature_j - theta), every other objectness -15;
    class logit +6; wh logits 0 (box = anchor size), xy logits a faint image-dependent signal."""
    w = P[f"model.23.m.{level}.weight"]
    b = P[f"model.23.m.{level}.bias"]
    w = w.clone().view(3, 6, -1)
    b = b.clone().view(3, 6)
    w[:, 0:2] *= 0.05
    b[:, 0:2] = 0.0
    w[:, 2:] = 0.0
    b[:, 2:4] = 0.0
    b[:, 4] = -15.0
    b[:, 5] = 6.0
    if j is not None:
        w[anchor, 4, j] = k
        b[anchor, 4] = -k * theta
    P[f"model.23.m.{level}.weight"] = w.view(18, -1, 1, 1)
    P[f"model.23.m.{level}.bias"] = b.view(18)


def _choose(feat, feat_gpu, lo=3, hi=60):
    """Candidate (channel j, threshold theta, separation) triples, best first: theta sits mid-gap between two consecutive
    order statistics (lo..hi detections) and separation = min over ALL cells of |f_i - theta| / err_i, err_i = that
    cell's own measured bf16 error |GPU feature - oracle feature|.  Also returns err (c, cells)."""
    n, c, h, w = feat.shape
    flat = feat.permute(1, 0, 2, 3).reshape(c, -1)
    err = (feat_gpu.permute(1, 0, 2, 3).reshape(c, -1) - flat).abs()
    err = err + 0.05 * err.mean(1, keepdim=True) + 1e-7          # floor: a cell that happened to round exactly
    vals, _ = flat.sort(1, descending=True)
    cands = []
    for r in range(lo, hi + 1):
        theta = (vals[:, r - 1] + vals[:, r]) / 2                                   # (c,)
        sep = ((flat - theta[:, None]).abs() / err).min(1)[0]                        # (c,)
        for j in sep.topk(3)[1].tolist():
            cands.append((float(sep[j]), j, float(theta[j])))
    cands.sort(reverse=True)
    return [(j, th, sp) for sp, j, th in cands[:24]], err


def _nms_is_robust(P, feats, level, anchor, k, err, trials=8):
    """Oracle side only: the kept set of this level's detections does not change when every objectness logit moves by up
    to +-2 k err_i (twice the measured feature error) and every box centre by +-0.3 px."""
    raw = R.detect_raw(P, "model.23", feats)[level:level + 1]
    Pl = {"model.23.anchors": P["model.23.anchors"][level:level + 1]}
    g = torch.Generator().manual_seed(0)
    base = None
    for t in range(trials + 1):
        r = [raw[0].clone()]
        if t:
            sign = torch.randint(0, 2, err.shape, generator=g).float() * 2 - 1
            r[0][:, anchor, :, :, 4] += 2.0 * k * err * sign
        dec = R.detect_decode(Pl, "model.23", r, strides=(STRIDES[level],))
        if t:
            dec[..., 0:2] += (torch.rand(dec[..., 0:2].shape, generator=g) - 0.5) * 0.6
        kept = [sorted(_nms_rows(dec[i]).tolist()) for i in range(dec.shape[0])]
        if base is None:
            base = kept
            if sum(len(x) for x in kept) < 3:
                return False
        elif kept != base:
            return False
    return True


def test_planted_detections_identical_count_and_boxes():
    from lead_yolo_b200.postprocess import non_max_suppression
    m = _model().eval()
    img = _sar_like(2, 640, IN_SEED + 11)
    x = img.float() / 255
    P = _P(m)
    _, _, ys = R.model_forward(P, x, R.Ctx(False))
    feats = [ys[16], ys[19], ys[22]]
    got = {}
    hooks = [m.model[i].register_forward_hook(lambda mod, inp, out, i=i: got.__setitem__(i, out.detach().float().cpu()))
             for i in (16, 19, 22)]
    with torch.no_grad():
        m(img.to(DEV))
    for hk in hooks:
        hk.remove()
    feats_gpu = [got[16], got[19], got[22]]
    # P3 / anchor 0 (10x13 boxes: neighbours do not overlap) and P5 / anchor 2 (373x326 boxes: everything near
    # overlaps, so the suppression decisions are exercised)
    plan = []
    for level, anchor in ((0, 0), (1, None), (2, 2)):
        if anchor is None:
            _plant_head(P, feats, level, 0, None, 0.0, 0.0)
            continue
        cands, err = _choose(feats[level], feats_gpu[level])
        chosen = None
        for j, theta, sep in cands:
            if sep < 3.0:
                break
            fj = feats[level][:, j]
            gap = float((fj - theta).abs().min()) * 2
            k = 8.0 / gap                                          # >= 4 logits either side of the threshold
            Pt = dict(P)
            _plant_head(Pt, feats, level, anchor, j, theta, k)
            if _nms_is_robust(Pt, feats, level, anchor, k, err[j].view(fj.shape)):
                chosen = (j, theta, k, gap, sep)
                break
        assert chosen is not None, f"level {level}: no head whose NMS outcome survives the measured bf16 error"
        j, theta, k, gap, sep = chosen
        _plant_head(P, feats, level, anchor, j, theta, k)
        plan.append((level, anchor, j, theta, k, gap, sep))
    raw_ref = R.detect_raw(P, "model.23", feats)
    dec_ref = R.detect_decode(P, "model.23", raw_ref)
    # same head on the product model
    with torch.no_grad():
        for lvl, mi in enumerate(m.model[-1].m):
            mi.weight.copy_(P[f"model.23.m.{lvl}.weight"])
            mi.bias.copy_(P[f"model.23.m.{lvl}.bias"])
        dec, raw = m(img.to(DEV))
    # 1. the construction holds on the GPU path: measured objectness-logit error << the planted margin (4 logits)
    for (level, anchor, j, theta, k, gap, sep) in plan:
        lo_ref = raw_ref[level][:, anchor, :, :, 4]
        lo_gpu = raw[level][:, anchor, :, :, 4].float().cpu()
        assert bool(((lo_ref > 0) == (lo_gpu > 0)).all()), f"level {level}: a planted objectness logit changed sign"
        near = lo_ref.abs() < 12
        assert (lo_gpu - lo_ref)[near].abs().max() < 3.0, f"level {level}: logit error eats the 4-logit margin (sep {sep:.1f})"
    # 2. identical post-NMS count, every box matched with IoU > 0.99 (north_star), through the device NMS
    import torchvision
    out = non_max_suppression(dec, 0.25, 0.45)
    total = 0
    for i in range(2):
        keep_ref = _nms_rows(dec_ref[i])
        ref = dec_ref[i][keep_ref]
        ref_box = torch.stack((ref[:, 0] - ref[:, 2] / 2, ref[:, 1] - ref[:, 3] / 2, ref[:, 0] + ref[:, 2] / 2,
                               ref[:, 1] + ref[:, 3] / 2), 1)
        ours = out[i].cpu()
        assert ours.shape[0] == ref.shape[0], f"image {i}: post-NMS count {ours.shape[0]} vs reference {ref.shape[0]}"
        iou = torchvision.ops.box_iou(ref_box, ours[:, :4])
        best, arg = iou.max(1)
        assert best.min() > 0.99, f"image {i}: worst matched IoU {best.min():.4f}"
        assert arg.unique().numel() == arg.numel()               # one-to-one
        assert (ours[arg, 4] - ref[:, 4] * ref[:, 5]).abs().max() < 2e-2
        total += ref.shape[0]
    assert total >= 8, f"only {total} planted detections survived: the construction is not exercising NMS"


# ---------------------------------------------------------------------------------------------------------------
# modules at the bench geometry (bs = 64), eval, oracle in fp32 on the GPU
# ---------------------------------------------------------------------------------------------------------------
FULL = {
    # name: (class, ctor args, oracle key, input shape, signed input)
    "L1_stage24": ("BasicStage", (24, 1), "BasicStage", (64, 24, 160, 160), True),
    "L2_merge24_40": ("PatchMerging_FasterNet", (24, 40, 2, 2), "PatchMerging", (64, 24, 160, 160), True),
    "L3_stage40": ("BasicStage", (40, 1), "BasicStage", (64, 40, 80, 80), True),
    "L5_stage80": ("BasicStage", (80, 1), "BasicStage", (64, 80, 40, 40), True),
    "L7_stage160": ("BasicStage", (160, 1), "BasicStage", (64, 160, 20, 20), True),
    "L8_sppf": ("SPPF", (160, 160, 5), "SPPF", (64, 160, 20, 20), True),
    "L9_rfcbam_k1": ("RFCBAMConv", (160, 256, 1, 1), "RFCBAM", (64, 160, 20, 20), False),
    "L12_c3ca": ("C3_CA", (336, 256, 1, False), "C3_CA", (64, 336, 40, 40), False),
    "L13_rfcbam_k1": ("RFCBAMConv", (256, 128, 1, 1), "RFCBAM", (64, 256, 40, 40), False),
    "L16_c3ca": ("C3_CA", (168, 128, 1, False), "C3_CA", (64, 168, 80, 80), False),
    "L17_rfcbam_k3s2": ("RFCBAMConv", (128, 128, 3, 2), "RFCBAM", (64, 128, 80, 80), False),
    "L19_c3ca": ("C3_CA", (256, 256, 1, False), "C3_CA", (64, 256, 40, 40), False),
    "L20_rfcbam_k3s2": ("RFCBAMConv", (256, 256, 3, 2), "RFCBAM", (64, 256, 40, 40), False),
    "L22_c3ca": ("C3_CA", (512, 512, 1, False), "C3_CA", (64, 512, 20, 20), False),
    "L16_c3": ("C3", (168, 128, 1, False), "C3", (64, 168, 80, 80), False),
    "coordatt64_80": ("CoordAtt", (64, 64, 32), "CoordAtt", (64, 64, 80, 80), False),
}


@pytest.mark.parametrize("name", sorted(FULL))
def test_module_eval_parity_at_bench_geometry(name):
    from helpers import ORACLE_FN, half_normal
    from lead_yolo_b200 import modules as M
    cls, args, okey, shape, signed = FULL[name]
    torch.manual_seed(INIT_SEED)
    mod = getattr(M, cls)(*args).to(DEV)
    randomise_bn(mod, BN_SEED)
    mod.eval()
    x = half_normal(shape, IN_SEED, signed=signed)
    P = {k: v.to(DEV) for k, v in oracle_params(mod).items()}
    if okey == "RFCBAM":
        P["__stride__m"] = args[3] if len(args) > 3 else 1
    with torch.no_grad():
        y = mod(x)
        ref = ORACLE_FN[okey](P, x.float(), R.Ctx(False))
    assert y.shape == ref.shape
    assert rel_l2(y, ref) <= 1e-2, f"{name}: rel-L2 {rel_l2(y, ref):.3e}"


def test_1024_whole_model_eval_vs_oracle():
    """configs[4]: 1024x1024, the 4-image per-GPU share of bs = 32 (maps 256 / 128 / 64 / 32)."""
    m = _model().eval()
    img = _sar_like(4, 1024, IN_SEED + 3).to(DEV)
    P = _P(m, DEV)
    with torch.no_grad():
        dec, raw = m(img)
        dec_ref, raw_ref, _ = R.model_forward(P, img.float() / 255, R.Ctx(False))
    assert dec.shape == dec_ref.shape == (4, 3 * (128 * 128 + 64 * 64 + 32 * 32), 6)
    for a, b in zip(raw, raw_ref):
        assert rel_l2(a, b) <= 2e-2
    assert rel_l2(dec, dec_ref) <= 2e-2


def test_c3_ablation_train_step_vs_oracle():
    """configs[3]: yolov5s-FasterNet+RFCBAMConv.yaml (plain C3 instead of C3_CA), training forward/backward at 640x640."""
    m = _model("c3").train()
    img = _sar_like(4, 640, IN_SEED + 4).to(DEV)
    x = img.float() / 255
    P = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running" not in k and "anchors" not in k else v)
         for k, v in _P(m, DEV).items()}
    raw = m(x)
    loss = sum((r.float() ** 2).mean() for r in raw)
    loss.backward()
    _, raw_ref, _ = R.model_forward(P, x, R.Ctx(True), variant="c3", decode=False)
    loss_ref = sum((r ** 2).mean() for r in raw_ref)
    loss_ref.backward()
    for a, b in zip(raw, raw_ref):
        assert rel_l2(a, b) <= 3e-2
    assert abs(float(loss) - float(loss_ref)) <= 3e-2 * abs(float(loss_ref))
    # gradients: direction and scale over all parameters (the criterion of test_model_gpu.test_train_step_vs_oracle:
    # a whole-network bf16 gradient sits at cosine ~0.98-0.99 against fp32)
    ga, gr = [], []
    for k, p in m.named_parameters():
        assert p.grad is not None and P[k].grad is not None, k
        ga.append(p.grad.flatten().float())
        gr.append(P[k].grad.flatten().float())
    ga, gr = torch.cat(ga), torch.cat(gr)
    cos = torch.nn.functional.cosine_similarity(ga, gr, dim=0).item()
    assert cos >= 0.97, f"cosine(grad, fp32 oracle) = {cos:.4f}"
    assert 0.8 < (ga.norm() / gr.norm()).item() < 1.25
