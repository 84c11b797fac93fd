"""Detection loss of the training step — drop-in for the reference's `utils.loss.ComputeLoss` (utils/loss.py:91-268)
with the EIoU box term (utils/metrics.py:293-354), computed by the fp32 kernels of csrc/yolo_loss.cu.

Same constructor and call signature: `ComputeLoss(model)(p, targets) -> (loss, items)` with `p` the list of raw head
maps `(bs, na, ny, nx, nc+5)` the Detect training branch returns and `targets` an `(nt, 6)` tensor
`(image, class, x, y, w, h)`.  Unlike the reference, shapes are static: rows of `targets` whose image index is
negative are padding, nothing is filtered on the host and no value is read back, so the call can sit inside a captured
CUDA graph.  Focal loss (`fl_gamma > 0`), `autobalance` and `sort_obj_iou` have no kernel and raise.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lowlevel as ll
from ._lib import LeadYoloError, LossDesc, check, lib


def _desc(cfg, p, dp, anchors, targets):
    d = LossDesc()
    d.nl, d.na, d.nc, d.bs, d.nt = len(p), cfg["na"], cfg["nc"], int(p[0].shape[0]), int(targets.shape[0])
    for i, pi in enumerate(p):
        bs, na, ny, nx, no = pi.shape
        if na != cfg["na"] or no != cfg["nc"] + 5 or bs != d.bs or pi.dtype != torch.float32:
            raise LeadYoloError(f"ComputeLoss: head map {i} has shape {tuple(pi.shape)} / dtype {pi.dtype}; expected "
                                f"(bs, {cfg['na']}, ny, nx, {cfg['nc'] + 5}) fp32")
        d.ny[i], d.nx[i] = ny, nx
        d.p[i] = pi.data_ptr()
        d.sb[i], d.sa[i], d.sy[i], d.sx[i], d.so[i] = pi.stride()
        if dp is not None:
            assert dp[i].stride() == pi.stride()
            d.dp[i] = dp[i].data_ptr()
        d.balance[i] = cfg["balance"][i]
    d.anchors = anchors.data_ptr()
    d.targets = targets.data_ptr()
    for k in ("anchor_t", "box", "obj", "cls", "cls_pw", "obj_pw", "cp", "cn"):
        setattr(d, k, float(cfg[k]))
    return d


class YoloLossFn(torch.autograd.Function):
    """out[4] = (lbox + lobj + lcls) * bs | lbox | lobj | lcls; gradient w.r.t. the head maps only (through out[0])."""

    @staticmethod
    def forward(ctx, cfg, anchors, targets, *p):
        p = [t if t.dtype == torch.float32 else t.float() for t in p]
        targets = targets.detach().float().contiguous()
        if targets.shape[0] == 0:
            targets = torch.full((1, 6), -1.0, device=p[0].device)
        anchors = anchors.detach().float().contiguous()
        d = _desc(cfg, p, None, anchors, targets)
        fn = lib().ly_yolo_loss_workspace_bytes
        fn.restype = C.c_long
        nbytes = fn(C.byref(d))
        if nbytes < 0:
            check(-1, "ly_yolo_loss_workspace_bytes")
        ws = torch.empty(nbytes, dtype=torch.uint8, device=p[0].device)
        out = torch.empty(4, dtype=torch.float32, device=p[0].device)
        with ll._Prof("ly_yolo_loss_fwd"):
            check(lib().ly_yolo_loss_fwd(C.byref(d), C.c_void_p(ws.data_ptr()), C.c_void_p(out.data_ptr()),
                                         ll.stream_ptr()), "ly_yolo_loss_fwd")
        ctx.cfg = cfg
        ctx.save_for_backward(anchors, targets, ws, *p)
        return out

    @staticmethod
    def backward(ctx, gout):
        anchors, targets, ws, *p = ctx.saved_tensors
        gs = gout[0:1].float().contiguous()
        dp = [torch.empty_strided(t.shape, t.stride(), dtype=torch.float32, device=t.device) for t in p]
        d = _desc(ctx.cfg, p, dp, anchors, targets)
        with ll._Prof("ly_yolo_loss_bwd"):
            check(lib().ly_yolo_loss_bwd(C.byref(d), C.c_void_p(ws.data_ptr()), C.c_void_p(gs.data_ptr()),
                                         ll.stream_ptr()), "ly_yolo_loss_bwd")
        return (None, None, None, *dp)


def smooth_BCE(eps=0.1):
    """positive / negative BCE targets (utils/loss.py:12-14)."""
    return 1.0 - 0.5 * eps, 0.5 * eps


class ComputeLoss:
    """utils/loss.py:91-119 constructor contract: reads `model.hyp` and the Detect module `model.model[-1]`."""
    sort_obj_iou = False

    def __init__(self, model, autobalance=False):
        inner = model.module if hasattr(model, "module") and hasattr(model.module, "model") else model
        h = inner.hyp
        m = inner.model[-1]
        if h.get("fl_gamma", 0.0) > 0:
            raise LeadYoloError("ComputeLoss: focal loss (fl_gamma > 0) has no kernel in lead_yolo_b200")
        if autobalance or self.sort_obj_iou:
            raise LeadYoloError("ComputeLoss: autobalance / sort_obj_iou have no kernel in lead_yolo_b200")
        if m.nl > 3:
            raise LeadYoloError("ComputeLoss: at most 3 detection levels")
        cp, cn = smooth_BCE(eps=h.get("label_smoothing", 0.0))
        self.balance = {3: [4.0, 1.0, 0.4]}.get(m.nl, [4.0, 1.0, 0.25, 0.06, 0.02])[:m.nl]
        self.na, self.nc, self.nl, self.anchors = m.na, m.nc, m.nl, m.anchors
        self.hyp, self.gr, self.autobalance = h, 1.0, False
        self.device = next(model.parameters()).device
        self.cfg = {"na": m.na, "nc": m.nc, "balance": self.balance, "anchor_t": h["anchor_t"], "box": h["box"],
                    "obj": h["obj"], "cls": h["cls"], "cls_pw": h["cls_pw"], "obj_pw": h["obj_pw"], "cp": cp, "cn": cn}

    def __call__(self, p, targets):
        out = YoloLossFn.apply(self.cfg, self.anchors, targets, *p)
        return out[0:1], out[1:4].detach()
