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
