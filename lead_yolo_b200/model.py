"""Model assembly for the hot path: the LEAD-YOLO graph built from the drop-in modules.

Two ways in:
  * `install(ref_common, ref_rfa, ref_yolo)` rebinds the class names inside the reference's own module
    namespaces, so the reference's `models/yolo.py:parse_model()` (397-492) builds `LEAD-YOLO.yaml` out of these
    classes and reference checkpoints un-pickle into them (SURVEY §8b "Registry").
  * `build_lead_yolo()` is a self-contained builder (the reference tree is not present on the GPU box): it
    restates the layer table of models/LEAD-YOLO.yaml:14-43 after width 0.5 / depth 0.33 scaling
    (models/yolo.py:432,451) and wires layers exactly like BaseModel._forward_once (yolo.py:179-195).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from . import modules as M

ANCHORS = ((10, 13, 16, 30, 33, 23), (30, 61, 62, 45, 59, 119), (116, 90, 156, 198, 373, 326))  # LEAD-YOLO.yaml:8-11

DROP_IN = ("Conv", "Bottleneck", "C3", "SPPF", "Concat", "Partial_conv3", "MLPBlock", "BasicStage",
           "PatchEmbed_FasterNet", "PatchMerging_FasterNet", "h_sigmoid", "h_swish", "CoordAtt", "CA_Bottleneck",
           "C3_CA")
DROP_IN_RFA = ("SE", "RFCBAMConv")


def install(ref_common, ref_rfa, ref_yolo):
    """Rebind the hot-path class names in the reference's namespaces (star-imports copied the bindings into
    models.yolo, yolo.py:16,25, so all three modules are patched).  Returns the dict of replaced originals."""
    saved = {}
    for name in DROP_IN:
        cls = getattr(M, name)
        saved[("common", name)] = getattr(ref_common, name, None)
        setattr(ref_common, name, cls)
        if hasattr(ref_yolo, name):
            saved[("yolo", name)] = getattr(ref_yolo, name)
            setattr(ref_yolo, name, cls)
    for name in DROP_IN_RFA:
        cls = getattr(M, name)
        saved[("rfa", name)] = getattr(ref_rfa, name, None)
        setattr(ref_rfa, name, cls)
    # models.yolo resolves RFCBAMConv from `from models.rfa import *`; SE in yolo's namespace is common's SE
    # (star-import order), which is not on this path and stays untouched.
    saved[("yolo", "RFCBAMConv")] = getattr(ref_yolo, "RFCBAMConv", None)
    ref_yolo.RFCBAMConv = M.RFCBAMConv
    saved[("yolo", "Detect")] = getattr(ref_yolo, "Detect", None)
    ref_yolo.Detect = M.Detect
    return saved


def uninstall(ref_common, ref_rfa, ref_yolo, saved):
    mods = {"common": ref_common, "rfa": ref_rfa, "yolo": ref_yolo}
    for (where, name), cls in saved.items():
        if cls is not None:
            setattr(mods[where], name, cls)


def _layer_specs(variant: str, nc: int):
    csp = M.C3_CA if variant == "lead" else M.C3
    # (from, constructor) — channel numbers are models/LEAD-YOLO.yaml scaled by width_multiple 0.5 and rounded
    # up to a multiple of 8 (utils/general.py:669-673); BasicStage repeats are n*0.33 rounded (1/1/3/1).
    return [
        (-1, lambda: M.PatchEmbed_FasterNet(3, 24, 4, 4)),                                    # 0  P1/4
        (-1, lambda: M.BasicStage(24, 1)),                                                    # 1
        (-1, lambda: M.PatchMerging_FasterNet(24, 40, 2, 2)),                                 # 2  P2/8
        (-1, lambda: M.BasicStage(40, 1)),                                                    # 3
        (-1, lambda: M.PatchMerging_FasterNet(40, 80, 2, 2)),                                 # 4
        (-1, lambda: nn.Sequential(*(M.BasicStage(80, 1) for _ in range(3)))),               # 5
        (-1, lambda: M.PatchMerging_FasterNet(80, 160, 2, 2)),                                # 6
        (-1, lambda: M.BasicStage(160, 1)),                                                   # 7
        (-1, lambda: M.SPPF(160, 160, 5)),                                                    # 8
        (-1, lambda: M.RFCBAMConv(160, 256, 1, 1)),                                           # 9
        (-1, lambda: M.Upsample(None, 2, "nearest")),                                         # 10
        ([-1, 5], lambda: M.Concat(1)),                                                       # 11
        (-1, lambda: csp(336, 256, 1, False)),                                                # 12
        (-1, lambda: M.RFCBAMConv(256, 128, 1, 1)),                                           # 13
        (-1, lambda: M.Upsample(None, 2, "nearest")),                                         # 14
        ([-1, 3], lambda: M.Concat(1)),                                                       # 15
        (-1, lambda: csp(168, 128, 1, False)),                                                # 16 (P3/8)
        (-1, lambda: M.RFCBAMConv(128, 128, 3, 2)),                                           # 17
        ([-1, 13], lambda: M.Concat(1)),                                                      # 18
        (-1, lambda: csp(256, 256, 1, False)),                                                # 19 (P4/16)
        (-1, lambda: M.RFCBAMConv(256, 256, 3, 2)),                                           # 20
        ([-1, 9], lambda: M.Concat(1)),                                                       # 21
        (-1, lambda: csp(512, 512, 1, False)),                                                # 22 (P5/32)
        ([16, 19, 22], lambda: M.Detect(nc, [list(a) for a in ANCHORS], (128, 256, 512))),    # 23
    ]


class LeadYolo(nn.Module):
    """Same layout as the reference DetectionModel (models/yolo.py:253-305): `self.model` is an nn.Sequential
    whose children carry `.i` / `.f`, `self.save` lists the layer outputs kept for later layers."""

    def __init__(self, variant: str = "lead", nc: int = 1):
        super().__init__()
        assert variant in ("lead", "c3")
        layers, save = [], []
        for i, (f, ctor) in enumerate(_layer_specs(variant, nc)):
            m = ctor()
            m.i, m.f = i, f
            m.type = type(m).__name__
            m.np = sum(p.numel() for p in m.parameters())
            save.extend(x % i for x in ([f] if isinstance(f, int) else f) if x != -1)
            layers.append(m)
        self.model = nn.Sequential(*layers)
        self.save = sorted(set(save))
        self.nc = nc
        self.names = [str(i) for i in range(nc)]
        det = self.model[-1]
        det.stride = torch.tensor([8.0, 16.0, 32.0])     # yolo.py:287-289 (P3/P4/P5 of this graph)
        self.stride = det.stride
        det.anchors /= det.stride.view(-1, 1, 1)          # yolo.py:291
        self._initialize_biases()
        # utils/torch_utils.py:212-221: BatchNorm eps/momentum overrides (exact-type test as the reference)
        for m in self.modules():
            if type(m) is nn.BatchNorm2d:
                m.eps = 1e-3
                m.momentum = 0.03
            elif type(m) in (nn.ReLU, nn.ReLU6, nn.SiLU, nn.Hardswish, nn.LeakyReLU):
                m.inplace = True

    def _initialize_biases(self):
        """Detect bias prior — models/yolo.py:352-359."""
        det = self.model[-1]
        for mi, s in zip(det.m, det.stride):
            b = mi.bias.view(det.na, -1)
            b.data[:, 4] += math.log(8 / (640 / float(s)) ** 2)
            b.data[:, 5:5 + det.nc] += math.log(0.6 / (det.nc - 0.99999))
            mi.bias = nn.Parameter(b.view(-1), requires_grad=True)

    def forward(self, x):
        y = []
        lazy_cat = (not torch.is_grad_enabled()) and (not self.training)
        for m in self.model:
            if m.f != -1:
                x = y[m.f] if isinstance(m.f, int) else [x if j == -1 else y[j] for j in m.f]
            if lazy_cat and isinstance(m, M.Concat) and m.i not in self.save and m.i + 1 < len(self.model) and \
                    isinstance(self.model[m.i + 1], (M.C3, M.C3_CA)) and self.model[m.i + 1].f == -1:
                # inference: the only consumer is a CSP block whose 1x1 convs read the pieces as K segments of one
                # GEMM, so the concat is never materialised
                y.append(None)
                continue
            x = m(x)
            y.append(x if m.i in self.save else None)
        return x

    def fuse(self):
        """Reference BaseModel.fuse (models/yolo.py:213-233): fold BN into Conv / patch convolutions."""
        for m in self.model.modules():
            if isinstance(m, M.Conv) and hasattr(m, "bn"):
                m.conv = fuse_conv_and_bn(m.conv, m.bn)
                delattr(m, "bn")
                m.forward = m.forward_fuse
            if type(m) is M.PatchEmbed_FasterNet and hasattr(m, "norm"):
                m.proj = fuse_conv_and_bn(m.proj, m.norm)
                delattr(m, "norm")
                m.forward = m.fuseforward
            if type(m) is M.PatchMerging_FasterNet and hasattr(m, "norm"):
                m.reduction = fuse_conv_and_bn(m.reduction, m.norm)
                delattr(m, "norm")
                m.forward = m.fuseforward
        return self


def fuse_conv_and_bn(conv: nn.Conv2d, bn: nn.BatchNorm2d) -> nn.Conv2d:
    """W' = diag(gamma/sqrt(var+eps)) W, b' = beta + (b - mean) * gamma/sqrt(var+eps)
    (same algebra as utils/torch_utils.py:248-269)."""
    fused = nn.Conv2d(conv.in_channels, conv.out_channels, conv.kernel_size, conv.stride, conv.padding,
                      conv.dilation, conv.groups, bias=True).requires_grad_(False).to(conv.weight.device)
    with torch.no_grad():
        g = bn.weight / torch.sqrt(bn.running_var + bn.eps)
        fused.weight.copy_((conv.weight.view(conv.out_channels, -1) * g.view(-1, 1)).view(fused.weight.shape))
        b = torch.zeros(conv.out_channels, device=conv.weight.device) if conv.bias is None else conv.bias
        fused.bias.copy_(bn.bias + (b - bn.running_mean) * g)
    return fused


def build_lead_yolo(variant: str = "lead", nc: int = 1, device="cuda", seed: int | None = 0) -> LeadYolo:
    if seed is not None:
        torch.manual_seed(seed)
    return LeadYolo(variant, nc).to(device)
