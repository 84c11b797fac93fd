"""TEST INFRASTRUCTURE ONLY (checker, never shipped or measured as the product).

Functional restatement, in plain PyTorch on whatever dtype the parameters have (fp32 / fp64), of the
LEAD-YOLO module arithmetic.  Each function takes a flat parameter dict `P` (a state_dict, tensors may
require grad) plus a key prefix, and cites the reference lines it follows.  Backward comes from torch
autograd through these functions.  Pinned against the imported reference classes by
tests/test_oracle_vs_reference.py (run in the build container) and by the committed fixtures under
tests/golden/ (made by oracle/make_golden.py from the real reference).

The reference's own tests hold no golden vectors for this path: parity is pinned to
outputs of the reference itself, run here.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-3       # utils/torch_utils.py:217-219 (initialize_weights overrides nn.BatchNorm2d defaults)
BN_MOMENTUM = 0.03


class Ctx:
    """training=True -> batch statistics; new running stats are recorded in `updates` (not applied)."""

    def __init__(self, training=False, eps=BN_EPS, momentum=BN_MOMENTUM):
        self.training = training
        self.eps = eps
        self.momentum = momentum
        self.updates = {}


def batchnorm(P, pre, x, ctx):
    """nn.BatchNorm2d semantics: biased var to normalise, unbiased into running_var."""
    g, b = P[pre + ".weight"], P[pre + ".bias"]
    if ctx.training:
        dims = (0, 2, 3)
        mu = x.mean(dims)
        var = x.var(dims, unbiased=False)
        n = x.numel() // x.shape[1]
        ctx.updates[pre + ".running_mean"] = (1 - ctx.momentum) * P[pre + ".running_mean"] + ctx.momentum * mu.detach()
        ctx.updates[pre + ".running_var"] = (1 - ctx.momentum) * P[pre + ".running_var"] + \
            ctx.momentum * var.detach() * n / max(n - 1, 1)
    else:
        mu, var = P[pre + ".running_mean"], P[pre + ".running_var"]
    v = lambda t: t.view(1, -1, 1, 1)
    return (x - v(mu)) * torch.rsqrt(v(var) + ctx.eps) * v(g) + v(b)


def conv_block(P, pre, x, ctx, k=1, s=1, act=True):
    """Conv = SiLU(BN(conv)) — models/common.py:1894-1907; bias-free, pad k//2."""
    y = F.conv2d(x, P[pre + ".conv.weight"], None, s, k // 2)
    y = batchnorm(P, pre + ".bn", y, ctx)
    return F.silu(y) if act else y


def faster_block(P, pre, x, ctx):
    """MLPBlock: x + W2 relu(BN(W1 [conv3x3(x[:, :C/4]) | x[:, C/4:]])) — common.py:1432-1437, 1478-1482."""
    wp = P[pre + ".spatial_mixing.partial_conv3.weight"]
    cp = wp.shape[0]
    t = torch.cat((F.conv2d(x[:, :cp], wp, None, 1, 1), x[:, cp:]), 1)
    u = F.conv2d(t, P[pre + ".mlp.0.weight"])
    v = F.relu(batchnorm(P, pre + ".mlp.1", u, ctx))
    return x + F.conv2d(v, P[pre + ".mlp.3.weight"])


def basic_stage(P, pre, x, ctx):
    """BasicStage = sequence of MLPBlocks — common.py:1493-1525."""
    i = 0
    while f"{pre}.blocks.{i}.mlp.0.weight" in P:
        x = faster_block(P, f"{pre}.blocks.{i}", x, ctx)
        i += 1
    return x


def patch_embed(P, pre, x, ctx):
    """PatchEmbed_FasterNet: BN(conv k=s) — common.py:1537-1539."""
    w = P[pre + ".proj.weight"]
    return batchnorm(P, pre + ".norm", F.conv2d(x, w, None, w.shape[-1]), ctx)


def patch_merge(P, pre, x, ctx):
    """PatchMerging_FasterNet: BN(conv k=s) — common.py:1555-1557."""
    w = P[pre + ".reduction.weight"]
    return batchnorm(P, pre + ".norm", F.conv2d(x, w, None, w.shape[-1]), ctx)


def h_swish(t):
    """t * relu6(t+3)/6 — common.py:1565-1580."""
    return t * F.relu6(t + 3.0) / 6.0


def coord_att(P, pre, x, ctx):
    """CoordAtt — common.py:1595-1609."""
    n, c, h, w = x.shape
    ph = x.mean(3, keepdim=True)                       # (n,c,h,1)
    pw = x.mean(2, keepdim=True).permute(0, 1, 3, 2)   # (n,c,w,1)
    y = torch.cat((ph, pw), 2)
    y = F.conv2d(y, P[pre + ".conv1.weight"], P[pre + ".conv1.bias"])
    y = h_swish(batchnorm(P, pre + ".bn1", y, ctx))
    yh, yw = y[:, :, :h], y[:, :, h:].permute(0, 1, 3, 2)
    ah = torch.sigmoid(F.conv2d(yh, P[pre + ".conv_h.weight"], P[pre + ".conv_h.bias"]))
    aw = torch.sigmoid(F.conv2d(yw, P[pre + ".conv_w.weight"], P[pre + ".conv_w.bias"]))
    return x * aw * ah


def ca_bottleneck(P, pre, x, ctx, add=False):
    """CA_Bottleneck: ca(cv2_3x3(cv1_1x1(x))) (+x) — common.py:1612-1623."""
    y = coord_att(P, pre + ".ca", conv_block(P, pre + ".cv2", conv_block(P, pre + ".cv1", x, ctx), ctx, k=3), ctx)
    return x + y if add else y


def c3_ca(P, pre, x, ctx, shortcut=False):
    """C3_CA: cv3(cat(m(cv1 x), cv2 x)) — common.py:1626-1637."""
    a = conv_block(P, pre + ".cv1", x, ctx)
    i = 0
    while f"{pre}.m.{i}.cv1.conv.weight" in P:
        a = ca_bottleneck(P, f"{pre}.m.{i}", a, ctx, add=shortcut)
        i += 1
    return conv_block(P, pre + ".cv3", torch.cat((a, conv_block(P, pre + ".cv2", x, ctx)), 1), ctx)


def bottleneck(P, pre, x, ctx, add=False):
    """Bottleneck: cv2_3x3(cv1_1x1(x)) (+x) — common.py:195-207."""
    y = conv_block(P, pre + ".cv2", conv_block(P, pre + ".cv1", x, ctx), ctx, k=3)
    return x + y if add else y


def c3(P, pre, x, ctx, shortcut=False):
    """C3 — common.py:254-268."""
    a = conv_block(P, pre + ".cv1", x, ctx)
    i = 0
    while f"{pre}.m.{i}.cv1.conv.weight" in P:
        a = bottleneck(P, f"{pre}.m.{i}", a, ctx, add=shortcut)
        i += 1
    return conv_block(P, pre + ".cv3", torch.cat((a, conv_block(P, pre + ".cv2", x, ctx)), 1), ctx)


def sppf(P, pre, x, ctx):
    """SPPF: three chained 5x5/s1/p2 max pools — common.py:359-367."""
    a = conv_block(P, pre + ".cv1", x, ctx)
    y1 = F.max_pool2d(a, 5, 1, 2)
    y2 = F.max_pool2d(y1, 5, 1, 2)
    y3 = F.max_pool2d(y2, 5, 1, 2)
    return conv_block(P, pre + ".cv2", torch.cat((a, y1, y2, y3), 1), ctx)


def se_gate(P, pre, x):
    """rfa.SE: sigmoid(W2 relu(W1 mean_hw(x))) — models/rfa.py:88-92 (hidden width is the literal 16)."""
    y = x.mean((2, 3))
    y = torch.sigmoid(F.linear(F.relu(F.linear(y, P[pre + ".fc.0.weight"])), P[pre + ".fc.2.weight"]))
    return y[:, :, None, None]


def rfcbam(P, pre, x, ctx):
    """RFCBAMConv — models/rfa.py:113-129."""
    b, c = x.shape[:2]
    wg = P[pre + ".generate.0.weight"]            # (c*k*k, 1, k, k), groups=c
    k = wg.shape[-1]
    wc = P[pre + ".conv.0.weight"]
    # stride of the generate conv is not stored on the module; derive it from the conv geometry:
    s = P["__stride__" + pre] if ("__stride__" + pre) in P else 1
    g = F.conv2d(x, wg, None, s, k // 2, 1, c)
    g = F.relu(batchnorm(P, pre + ".generate.1", g, ctx))
    h, w = g.shape[2:]
    # channel c*k*k + n1*k + n2 -> position (h*k+n1, w*k+n2)
    u = g.view(b, c, k, k, h, w).permute(0, 1, 4, 2, 5, 3).reshape(b, c, h * k, w * k)
    gated = u * se_gate(P, pre + ".se", x)
    mx = u.max(1, keepdim=True)[0]
    mn = u.mean(1, keepdim=True)
    att = torch.sigmoid(F.conv2d(torch.cat((mx, mn), 1), P[pre + ".get_weight.0.weight"], None, 1, 1))
    y = F.conv2d(gated * att, wc, P[pre + ".conv.0.bias"], k)
    return F.relu(batchnorm(P, pre + ".conv.1", y, ctx))


def detect_raw(P, pre, feats, na=3):
    """Detect heads: 1x1 conv + reshape to (bs, na, ny, nx, no) — models/yolo.py:90-95."""
    outs = []
    for i, f in enumerate(feats):
        y = F.conv2d(f, P[f"{pre}.m.{i}.weight"], P[f"{pre}.m.{i}.bias"])
        bs, ch, ny, nx = y.shape
        outs.append(y.view(bs, na, ch // na, ny, nx).permute(0, 1, 3, 4, 2).contiguous())
    return outs


def detect_decode(P, pre, raw, strides=(8.0, 16.0, 32.0)):
    """Inference decode — models/yolo.py:98-125, 128-150: xy=(2σ-0.5+grid)*stride, wh=(2σ)^2*anchor."""
    z = []
    for i, x in enumerate(raw):
        bs, na, ny, nx, no = x.shape
        d, t = x.device, x.dtype
        yv, xv = torch.meshgrid(torch.arange(ny, device=d, dtype=t), torch.arange(nx, device=d, dtype=t),
                                indexing="ij")
        grid = torch.stack((xv, yv), 2).expand(1, na, ny, nx, 2) - 0.5
        anchor_grid = (P[pre + ".anchors"][i].to(t) * strides[i]).view(1, na, 1, 1, 2)
        s = x.sigmoid()
        xy = (s[..., 0:2] * 2 + grid) * strides[i]
        wh = (s[..., 2:4] * 2) ** 2 * anchor_grid
        z.append(torch.cat((xy, wh, s[..., 4:]), 4).view(bs, na * nx * ny, no))
    return torch.cat(z, 1)


# Layer table of models/LEAD-YOLO.yaml after width 0.5 / depth 0.33 scaling; "c3" variant
# = models/yolov5s-FasterNet+RFCBAMConv.yaml.  (kind, from, extra)
def layer_table(variant="lead"):
    csp = "c3_ca" if variant == "lead" else "c3"
    return [
        ("patch_embed", -1, None), ("stage", -1, 1), ("patch_merge", -1, None), ("stage", -1, 1),
        ("patch_merge", -1, None), ("stage", -1, 3), ("patch_merge", -1, None), ("stage", -1, 1),
        ("sppf", -1, None), ("rfcbam", -1, 1), ("up", -1, None), ("cat", (-1, 5), None), (csp, -1, None),
        ("rfcbam", -1, 1), ("up", -1, None), ("cat", (-1, 3), None), (csp, -1, None), ("rfcbam", -1, 2),
        ("cat", (-1, 13), None), (csp, -1, None), ("rfcbam", -1, 2), ("cat", (-1, 9), None), (csp, -1, None),
        ("detect", (16, 19, 22), None),
    ]


def model_forward(P, x, ctx, variant="lead", decode=True, upto=None):
    """Whole-network forward mirroring BaseModel._forward_once (models/yolo.py:179-195).
    Returns (decoded or None, raw heads, per-layer outputs)."""
    ys = []
    P = dict(P)
    for i, (kind, frm, extra) in enumerate(layer_table(variant)):
        pre = f"model.{i}"
        if frm != -1:
            x = [x if j == -1 else ys[j] for j in frm]
        if kind == "patch_embed":
            x = patch_embed(P, pre, x, ctx)
        elif kind == "patch_merge":
            x = patch_merge(P, pre, x, ctx)
        elif kind == "stage":
            if extra == 1:
                x = basic_stage(P, pre, x, ctx)
            else:
                for r in range(extra):
                    x = basic_stage(P, f"{pre}.{r}", x, ctx)
        elif kind == "sppf":
            x = sppf(P, pre, x, ctx)
        elif kind == "rfcbam":
            P["__stride__" + pre] = extra
            x = rfcbam(P, pre, x, ctx)
        elif kind == "up":
            x = F.interpolate(x, scale_factor=2.0, mode="nearest")
        elif kind == "cat":
            x = torch.cat(x, 1)
        elif kind == "c3_ca":
            x = c3_ca(P, pre, x, ctx)
        elif kind == "c3":
            x = c3(P, pre, x, ctx)
        elif kind == "detect":
            raw = detect_raw(P, pre, x)
            ys.append(raw)
            dec = detect_decode(P, pre, raw) if decode else None
            return dec, raw, ys
        ys.append(x)
        if upto is not None and i == upto:
            return None, None, ys
    raise AssertionError("unreachable")


# ------------------------------------------------------------------------------------------------------------
# Detection loss and post-processing (the callers either side of the module path)
# ------------------------------------------------------------------------------------------------------------
HYP_SCRATCH_LOW = {  # data/hyps/hyp.scratch-low.yaml:14-21 (loss part), before the train.py:238-242 scaling
    "box": 0.05, "cls": 0.5, "cls_pw": 1.0, "obj": 1.0, "obj_pw": 1.0, "anchor_t": 4.0, "fl_gamma": 0.0,
    "label_smoothing": 0.0}


def scaled_hyp(nl=3, nc=1, imgsz=640, hyp=None):
    """train.py:238-242: box *= 3/nl; cls *= nc/80 * 3/nl; obj *= (imgsz/640)**2 * 3/nl; label_smoothing = 0."""
    h = dict(HYP_SCRATCH_LOW if hyp is None else hyp)
    h["box"] *= 3 / nl
    h["cls"] *= nc / 80 * 3 / nl
    h["obj"] *= (imgsz / 640) ** 2 * 3 / nl
    return h


def bbox_eiou(box1, box2, eps=1e-7):
    """bbox_iou(box1, box2, xywh=True, EIoU=True), utils/metrics.py:309-347 with alpha = 1."""
    (x1, y1, w1, h1), (x2, y2, w2, h2) = box1.chunk(4, -1), box2.chunk(4, -1)
    b1_x1, b1_x2, b1_y1, b1_y2 = x1 - w1 / 2, x1 + w1 / 2, y1 - h1 / 2, y1 + h1 / 2
    b2_x1, b2_x2, b2_y1, b2_y2 = x2 - w2 / 2, x2 + w2 / 2, y2 - h2 / 2, y2 + h2 / 2
    inter = (b1_x2.minimum(b2_x2) - b1_x1.maximum(b2_x1)).clamp(0) * \
            (b1_y2.minimum(b2_y2) - b1_y1.maximum(b2_y1)).clamp(0)
    union = w1 * h1 + w2 * h2 - inter + eps
    iou = inter / (union + eps)
    cw = b1_x2.maximum(b2_x2) - b1_x1.minimum(b2_x1)
    ch = b1_y2.maximum(b2_y2) - b1_y1.minimum(b2_y1)
    c2 = cw ** 2 + ch ** 2 + eps
    rho2 = ((b2_x1 + b2_x2 - b1_x1 - b1_x2) ** 2 + (b2_y1 + b2_y2 - b1_y1 - b1_y2) ** 2) / 4
    rho_w2 = ((b2_x2 - b2_x1) - (b1_x2 - b1_x1)) ** 2
    rho_h2 = ((b2_y2 - b2_y1) - (b1_y2 - b1_y1)) ** 2
    return iou - (rho2 / c2 + rho_w2 / (cw ** 2 + eps) + rho_h2 / (ch ** 2 + eps))


def build_targets(p, targets, anchors, anchor_t):
    """ComputeLoss.build_targets, utils/loss.py:195-268.  Padding rows (image < 0) are dropped first."""
    targets = targets[targets[:, 0] >= 0]
    na, nt = anchors.shape[1], targets.shape[0]
    tcls, tbox, indices, anch = [], [], [], []
    gain = torch.ones(7)
    ai = torch.arange(na).float().view(na, 1).repeat(1, nt)
    targets = torch.cat((targets.repeat(na, 1, 1), ai[..., None]), 2)
    g = 0.5
    off = torch.tensor([[0, 0], [1, 0], [0, 1], [-1, 0], [0, -1]]).float() * g
    for i in range(len(p)):
        anc, shape = anchors[i], p[i].shape
        gain[2:6] = torch.tensor(shape)[[3, 2, 3, 2]]
        t = targets * gain
        if nt:
            r = t[..., 4:6] / anc[:, None]
            j = torch.max(r, 1 / r).max(2)[0] < anchor_t
            t = t[j]
            gxy = t[:, 2:4]
            gxi = gain[[2, 3]] - gxy
            j, k = ((gxy % 1 < g) & (gxy > 1)).T
            l, m = ((gxi % 1 < g) & (gxi > 1)).T
            j = torch.stack((torch.ones_like(j), j, k, l, m))
            t = t.repeat((5, 1, 1))[j]
            offsets = (torch.zeros_like(gxy)[None] + off[:, None])[j]
        else:
            t = targets[0]
            offsets = 0
        bc, gxy, gwh, a = t.chunk(4, 1)
        a, (b, c) = a.long().view(-1), bc.long().T
        gij = (gxy - offsets).long()
        gi, gj = gij.T
        gj, gi = gj.clamp(0, shape[2] - 1), gi.clamp(0, shape[3] - 1)
        indices.append((b, a, gj, gi))
        tbox.append(torch.cat((gxy - torch.stack((gi, gj), 1), gwh), 1))   # the reference clamps gij in place first
        anch.append(anc[a])
        tcls.append(c)
    return tcls, tbox, indices, anch


def compute_loss(p, targets, anchors, hyp, nc=1, balance=(4.0, 1.0, 0.4), cp=1.0, cn=0.0):
    """ComputeLoss.__call__, utils/loss.py:121-191 (EIoU box term, gr = 1, no focal loss, no autobalance).
    p: list of (bs, na, ny, nx, nc+5); returns ((lbox+lobj+lcls)*bs, (lbox, lobj, lcls))."""
    lcls, lbox, lobj = torch.zeros(1), torch.zeros(1), torch.zeros(1)
    tcls, tbox, indices, anch = build_targets(p, targets, anchors, hyp["anchor_t"])
    for i, pi in enumerate(p):
        b, a, gj, gi = indices[i]
        tobj = torch.zeros(pi.shape[:4], dtype=pi.dtype)
        n = b.shape[0]
        if n:
            pxy, pwh, _, pcls = pi[b, a, gj, gi].split((2, 2, 1, nc), 1)
            pxy = pxy.sigmoid() * 2 - 0.5
            pwh = (pwh.sigmoid() * 2) ** 2 * anch[i]
            iou = bbox_eiou(torch.cat((pxy, pwh), 1), tbox[i]).squeeze(-1)
            lbox = lbox + (1.0 - iou).mean()
            tobj[b, a, gj, gi] = iou.detach().clamp(0).type(tobj.dtype)   # duplicates: the last entry wins (CPU)
            if nc > 1:
                t = torch.full_like(pcls, cn)
                t[range(n), tcls[i]] = cp
                lcls = lcls + F.binary_cross_entropy_with_logits(pcls, t, pos_weight=torch.tensor([hyp["cls_pw"]]))
        obji = F.binary_cross_entropy_with_logits(pi[..., 4], tobj, pos_weight=torch.tensor([hyp["obj_pw"]]))
        lobj = lobj + obji * balance[i]
    lbox, lobj, lcls = lbox * hyp["box"], lobj * hyp["obj"], lcls * hyp["cls"]
    bs = p[0].shape[0]
    return (lbox + lobj + lcls) * bs, torch.cat((lbox, lobj, lcls)).detach()


def nms_greedy(boxes, scores, iou_thres):
    """torchvision.ops.nms (CPU kernel nms_kernel_impl): visit boxes by descending score (ties: lower index), suppress
    those with IoU > threshold against a kept box; fp32 arithmetic in the kernel's order of operations."""
    order = torch.sort(scores, descending=True, stable=True)[1]
    b = boxes[order].float()
    area = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
    dead = torch.zeros(b.shape[0], dtype=torch.bool)
    keep = []
    for i in range(b.shape[0]):
        if dead[i]:
            continue
        keep.append(int(order[i]))
        w = (torch.minimum(b[i, 2], b[i + 1:, 2]) - torch.maximum(b[i, 0], b[i + 1:, 0])).clamp(min=0)
        h = (torch.minimum(b[i, 3], b[i + 1:, 3]) - torch.maximum(b[i, 1], b[i + 1:, 1])).clamp(min=0)
        inter = w * h
        dead[i + 1:] |= inter / (area[i] + area[i + 1:] - inter) > iou_thres
    return torch.tensor(keep, dtype=torch.long)


def non_max_suppression(prediction, conf_thres=0.25, iou_thres=0.45, agnostic=False, max_det=300):
    """utils/general.py:884-994, best-class branch (multi_label=False, no labels / masks / merge), without the
    wall-clock time limit.  Returns a list of (n, 6) tensors (xyxy, conf, cls) per image."""
    max_wh, max_nms = 7680, 30000
    xc = prediction[..., 4] > conf_thres
    out = [torch.zeros((0, 6))] * prediction.shape[0]
    for xi, x in enumerate(prediction):
        x = x[xc[xi]].clone()
        if not x.shape[0]:
            continue
        x[:, 5:] *= x[:, 4:5]
        box = torch.stack((x[:, 0] - x[:, 2] / 2, x[:, 1] - x[:, 3] / 2, x[:, 0] + x[:, 2] / 2, x[:, 1] + x[:, 3] / 2), 1)
        conf, j = x[:, 5:].max(1, keepdim=True)
        x = torch.cat((box, conf, j.float()), 1)[conf.view(-1) > conf_thres]
        if not x.shape[0]:
            continue
        x = x[x[:, 4].argsort(descending=True, stable=True)[:max_nms]]
        c = x[:, 5:6] * (0 if agnostic else max_wh)
        i = nms_greedy(x[:, :4] + c, x[:, 4], iou_thres)[:max_det]
        out[xi] = x[i]
    return out
