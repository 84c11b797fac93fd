"""Drop-in replacements for the LEAD-YOLO hot-path nn.Modules.

Class names, constructor signatures, sub-module names (hence state_dict keys) and the attributes the reference
relies on (`fuse()`, pickling of bound `forward` attributes, `.add`, `.kernel_size`, ...) are those of the
reference (models/common.py, models/rfa.py, models/yolo.py — cited per class); only `forward` differs: it runs
the sm_100a kernels of libleadyolo.so through the autograd Functions in ops.py.  There is no CPU or PyTorch
fallback: tensors must live on a B200.

`forward` may rely only on attributes that exist on an un-pickled reference instance.
"""
from __future__ import annotations

import math
import warnings

import torch
import torch.nn as nn

from . import lowlevel as ll
from . import ops
from ._lib import LeadYoloError
from .packing import pack_conv_weight

CL = torch.channels_last


def _fast(mod: nn.Module) -> bool:
    """Inference fast path (folded BN, cached packed weights, no autograd graph)."""
    return (not torch.is_grad_enabled()) and (not mod.training)


def _act_name(act) -> str:
    if act is None or isinstance(act, nn.Identity):
        return "none"
    if isinstance(act, nn.SiLU):
        return "silu"
    if isinstance(act, nn.ReLU):
        return "relu"
    raise LeadYoloError(f"activation {type(act).__name__} has no sm_100a kernel in lead_yolo_b200")


def _conv_geometry(conv: nn.Conv2d):
    k, s = conv.kernel_size[0], conv.stride[0]
    ok = (conv.kernel_size[0] == conv.kernel_size[1] and conv.stride[0] == conv.stride[1] and conv.groups == 1
          and conv.dilation == (1, 1))
    if ok and s == 1 and k in (1, 3) and conv.padding == (k // 2, k // 2):
        return k, s
    if ok and s == k and k in (2, 4) and conv.padding == (0, 0):
        return k, s
    raise LeadYoloError(f"convolution {conv} has no sm_100a kernel in lead_yolo_b200")


def _identity_affine(c, device):
    key = ("ident", c, str(device))
    hit = ops._cache.get(key)
    if hit is None:
        hit = (torch.ones(c, device=device), torch.zeros(c, device=device))
        ops._cache[key] = hit
    return hit


def act_only(z, act: str):
    """Activation with autograd (used after fuse(), when a Conv has no BatchNorm left)."""
    if act == "none":
        return z
    one, zero = _identity_affine(z.shape[1], z.device)
    return ops.BnActFn.apply(z, one, zero, zero, one, 0.0, 0.0, act, False)


def conv_bn_act(conv: nn.Conv2d, bn, act: str, srcs, fast: bool, residual=None, out=None):
    """act(BN(conv(cat(srcs)))) with the reference's semantics; srcs is a list (virtual concat along C)."""
    k, s = _conv_geometry(conv)
    cout = conv.out_channels
    if fast:
        chans = tuple(int(x.shape[1]) for x in srcs)
        wp = ops.cached([conv.weight], ("wp", chans),
                        lambda: pack_conv_weight(conv.weight, list(chans), patch=(s == k and k > 1)))
        if bn is not None:
            scale, shift = ops.bn_fold(bn, conv.bias)
        else:
            scale = None
            shift = None if conv.bias is None else ops.cached([conv.bias], "b32",
                                                              lambda: conv.bias.detach().float().contiguous())
        return ops.conv_infer(srcs, wp, cout, k=k, s=s, scale=scale, shift=shift, act=act, residual=residual, out=out)
    assert out is None
    if bn is None:
        if residual is not None:
            assert act == "none"
            return ops.conv(srcs, conv.weight, conv.bias, k=k, s=s, residual=residual)
        z = ops.conv(srcs, conv.weight, conv.bias, k=k, s=s)
        return act_only(z, act)
    assert residual is None
    z = ops.conv(srcs, conv.weight, conv.bias, k=k, s=s)
    return ops.bn_act(z, bn, act)


def autopad(k, p=None, d=1):
    """'same' padding (reference models/common.py:1880-1888)."""
    if d > 1:
        k = d * (k - 1) + 1 if isinstance(k, int) else [d * (x - 1) + 1 for x in k]
    if p is None:
        p = k // 2 if isinstance(k, int) else [x // 2 for x in k]
    return p


class Conv(nn.Module):
    """Conv2d + BatchNorm2d + SiLU — reference models/common.py:1890-1910 (live re-definition of 82-101)."""
    default_act = nn.SiLU()

    def __init__(self, c1, c2, k=1, s=1, p=None, g=1, d=1, act=True):
        super().__init__()
        self.conv = nn.Conv2d(c1, c2, k, s, autopad(k, p, d), groups=g, dilation=d, bias=False)
        self.bn = nn.BatchNorm2d(c2)
        self.act = self.default_act if act is True else act if isinstance(act, nn.Module) else nn.Identity()

    def forward(self, x):
        return conv_bn_act(self.conv, self.bn, _act_name(self.act), [x], _fast(self))

    def forward_fuse(self, x):
        return conv_bn_act(self.conv, None, _act_name(self.act), [x], _fast(self))

    def run(self, srcs, out=None):
        """Library-internal: list input (virtual concat) and optional destination slice (fast path only)."""
        bn = getattr(self, "bn", None)
        return conv_bn_act(self.conv, bn, _act_name(self.act), srcs, _fast(self), out=out)


class Bottleneck(nn.Module):
    """reference models/common.py:195-207"""

    def __init__(self, c1, c2, shortcut=True, g=1, e=0.5):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c_, c2, 3, 1, g=g)
        self.add = shortcut and c1 == c2

    def forward(self, x):
        y = self.cv2(self.cv1(x))
        return x + y if self.add else y


class C3(nn.Module):
    """CSP bottleneck with 3 convolutions — reference models/common.py:254-268"""

    def __init__(self, c1, c2, n=1, shortcut=True, g=1, e=0.5):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c1, c_, 1, 1)
        self.cv3 = Conv(2 * c_, c2, 1)
        self.m = nn.Sequential(*(Bottleneck(c_, c_, shortcut, g, e=1.0) for _ in range(n)))

    def forward(self, x):
        return _csp_forward(self, x)


def _csp_forward(mod, x):
    """cv3(cat(m(cv1 x), cv2 x)) without materialising the concat: cv3 reads both tensors as K segments.
    Inference: cv1 and cv2 share their input, so they run as ONE GEMM with N = 2*c_ into one buffer."""
    if _fast(mod) and hasattr(mod.cv1, "bn") and hasattr(mod.cv2, "bn") and \
            type(mod.cv1.act) is type(mod.cv2.act):
        c_ = mod.cv1.conv.out_channels
        # a list input is a not-yet-materialised channel concat (see LeadYolo.forward): read it as K segments
        srcs = [ops.to_act(t) for t in x] if isinstance(x, (list, tuple)) else [ops.to_act(x)]
        chans = tuple(int(t.shape[1]) for t in srcs)
        keys = [mod.cv1.conv.weight, mod.cv2.conv.weight]

        def build():
            w = torch.cat([mod.cv1.conv.weight.detach().float(), mod.cv2.conv.weight.detach().float()], 0)
            return pack_conv_weight(w, list(chans))

        wp = ops.cached(keys, ("csp12", chans), build)
        s1, h1 = ops.bn_fold(mod.cv1.bn)
        s2, h2 = ops.bn_fold(mod.cv2.bn)
        ss = ops.cached([s1, s2, h1, h2], "csp12ss", lambda: (torch.cat([s1, s2]), torch.cat([h1, h2])))
        buf = ops.conv_infer(srcs, wp, 2 * c_, k=1, s=1, scale=ss[0], shift=ss[1], act=_act_name(mod.cv1.act))
        a = mod.m(buf[:, :c_])
        return mod.cv3.run([a, buf[:, c_:]])
    if isinstance(x, (list, tuple)):
        x = ops.UpsampleCatFn.apply(tuple(False for _ in x), *x)
    a = mod.m(mod.cv1(x))
    b = mod.cv2(x)
    return mod.cv3.run([a, b])


class SPPF(nn.Module):
    """Spatial pyramid pooling (fast) — reference models/common.py:348-367"""

    def __init__(self, c1, c2, k=5):
        super().__init__()
        c_ = c1 // 2
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c_ * 4, c2, 1, 1)
        self.m = nn.MaxPool2d(kernel_size=k, stride=1, padding=k // 2)

    def forward(self, x):
        if self.m.kernel_size != 5:
            raise LeadYoloError("SPPF: only the 5x5 pooling of LEAD-YOLO has a kernel")
        if _fast(self):
            x = ops.to_act(x)
            n, _, h, w = x.shape
            c_ = self.cv1.conv.out_channels
            buf = ll.empty_cl(n, 4 * c_, h, w, device=x.device)
            self.cv1.run([x], out=buf[:, :c_])
            if h * w <= 1600:
                # the three cascaded pools in one launch, straight into their concat slices
                ll.call("ly_sppf_pool3", buf[:, :c_], 4 * c_, n, h, w, c_, buf[:, c_:2 * c_], 4 * c_, c_)
            else:
                for i in range(3):
                    ops.maxpool5_into(buf[:, i * c_:(i + 1) * c_], buf[:, (i + 1) * c_:(i + 2) * c_])
            return self.cv2.run([buf])
        a = self.cv1(x)
        y1 = ops.MaxPool5Fn.apply(a)
        y2 = ops.MaxPool5Fn.apply(y1)
        y3 = ops.MaxPool5Fn.apply(y2)
        return self.cv2.run([a, y1, y2, y3])


class Concat(nn.Module):
    """Channel concatenation — reference models/common.py:531-538"""

    def __init__(self, dimension=1):
        super().__init__()
        self.d = dimension

    def forward(self, x):
        if self.d != 1:
            raise LeadYoloError("Concat: only dimension=1 (channels) has a kernel")
        return ops.UpsampleCatFn.apply(tuple(False for _ in x), *x)


class Upsample(nn.Module):
    """nn.Upsample(None, 2, 'nearest') of LEAD-YOLO.yaml:29,33 as a library kernel (used by our own builder;
    the reference's parse_model instantiates the stock torch module for this yaml row)."""

    def __init__(self, size=None, scale_factor=2, mode="nearest"):
        super().__init__()
        assert size is None and int(scale_factor) == 2 and mode == "nearest"
        self.scale_factor = scale_factor
        self.mode = mode

    def forward(self, x):
        return ops.UpsampleCatFn.apply((True,), x)


# ------------------------------------------------------------------------------------------------------------
# FasterNet — reference models/common.py:1411-1561
# ------------------------------------------------------------------------------------------------------------
# The partial conv runs on the implicit-GEMM kernel (tensor cores, 8-aligned channel slice, zero weight rows for the
# untouched channels inside the slice).


class Partial_conv3(nn.Module):
    """3x3 convolution over the first dim//n_div channels, rest untouched — reference common.py:1411-1437.
    Inside MLPBlock the partial conv is folded into the block's kernels; this standalone forward exists for
    API parity (the bound method is stored in the instance dict and pickled by name, common.py:1419-1421)."""

    def __init__(self, dim, n_div, forward):
        super().__init__()
        self.dim_conv3 = dim // n_div
        self.dim_untouched = dim - self.dim_conv3
        self.partial_conv3 = nn.Conv2d(self.dim_conv3, self.dim_conv3, 3, 1, 1, bias=False)
        if forward == "slicing":
            self.forward = self.forward_slicing
        elif forward == "split_cat":
            self.forward = self.forward_split_cat
        else:
            raise NotImplementedError

    def _t1(self, x, fast):
        """conv3x3 over channels [0, cp) read as an 8-aligned slice; returns (N, cpp, H, W), channels >= cp zero."""
        cp = self.dim_conv3
        cpp = (cp + 7) // 8 * 8
        w = self.partial_conv3.weight
        if x.shape[1] < cpp:
            raise LeadYoloError("Partial_conv3: channel count too small for the 8-aligned slice")
        if fast:
            wp = ops.cached([w], ("pconv", cpp), lambda: pack_conv_weight(
                nn.functional.pad(w.detach().float(), (0, 0, 0, 0, 0, cpp - cp, 0, cpp - cp)), [cpp]))
            return ops.conv_infer([x[:, :cpp]], wp, cpp, k=3, s=1)
        wpad = nn.functional.pad(w, (0, 0, 0, 0, 0, cpp - cp, 0, cpp - cp))
        return ops.conv([x[:, :cpp]], wpad, None, k=3, s=1)

    def forward_split_cat(self, x):
        x = ops.to_act(x)
        t1 = self._t1(x, _fast(self))
        return torch.cat((t1[:, :self.dim_conv3], x[:, self.dim_conv3:]), 1)

    def forward_slicing(self, x):
        return self.forward_split_cat(x)


class MLPBlock(nn.Module):
    """x + W2 relu(BN(W1 [pconv3x3(x[:, :C/4]) | x[:, C/4:]])) — reference common.py:1441-1489"""

    def __init__(self, dim, n_div, mlp_ratio, drop_path, layer_scale_init_value, act_layer, norm_layer,
                 pconv_fw_type):
        super().__init__()
        self.dim = dim
        self.mlp_ratio = mlp_ratio
        if drop_path > 0.:
            raise LeadYoloError("MLPBlock: stochastic depth (drop_path > 0) has no kernel; LEAD-YOLO uses 0")
        self.drop_path = nn.Identity()
        self.n_div = n_div
        hidden = int(dim * mlp_ratio)
        self.mlp = nn.Sequential(nn.Conv2d(dim, hidden, 1, bias=False), norm_layer(hidden), act_layer(),
                                 nn.Conv2d(hidden, dim, 1, bias=False))
        self.spatial_mixing = Partial_conv3(dim, n_div, pconv_fw_type)
        if layer_scale_init_value > 0:
            self.layer_scale = nn.Parameter(layer_scale_init_value * torch.ones((dim)), requires_grad=True)
            self.forward = self.forward_layer_scale
        else:
            self.forward = self.forward

    def forward(self, x):
        x = ops.to_act(x)
        fast = _fast(self)
        pc = self.spatial_mixing
        cp = pc.dim_conv3
        cpp = (cp + 7) // 8 * 8
        w1, bn, act, w2 = self.mlp[0], self.mlp[1], _act_name(self.mlp[2]), self.mlp[3]
        c = x.shape[1]
        if fast and c in (24, 40, 80) and act == "relu" and ll.pixel_ld(x) == c and x.data_ptr() % 16 == 0 and \
                (c != 80 or cp == 20):
            # whole block in one kernel (csrc/fasternet_block.cu)
            from .packing import fasternet_weight_image
            img = ops.cached([w1.weight, w2.weight, pc.partial_conv3.weight], ("fnb", c), lambda: fasternet_weight_image(
                w1.weight, w2.weight, pc.partial_conv3.weight))
            scale, shift = ops.bn_fold(bn)
            n, _, h, wd = x.shape
            y = ll.empty_cl(n, c, h, wd, device=x.device)
            ll.call("ly_fasternet_block_infer", x, y, n, h, wd, c, img[0], img[0].numel() * 2, img[1], img[2], scale,
                    shift)
            return y
        # u = W1[:, :cp] t1 + W1[:, cp:] x[:, cp:]  ==  [W1[:, :cp] | 0] t1(8-aligned)  +  [0 | W1[:, cp:]] x
        if fast:
            t1 = pc._t1(x, fast)
            def build():
                w = w1.weight.detach().float()
                wa = nn.functional.pad(w[:, :cp], (0, 0, 0, 0, 0, cpp - cp))
                wb = w.clone()
                wb[:, :cp] = 0
                return pack_conv_weight(torch.cat([wa, wb], 1), [cpp, c])

            wp1 = ops.cached([w1.weight], ("mlp1", cpp, c), build)
            scale, shift = ops.bn_fold(bn)
            v = ops.conv_infer([t1, x], wp1, w1.out_channels, scale=scale, shift=shift, act=act)
            wp2 = ops.cached([w2.weight], ("wp", (w2.in_channels,)), lambda: pack_conv_weight(w2.weight, None))
            return ops.conv_infer([v], wp2, c, residual=x)
        w = w1.weight
        if isinstance(bn, nn.BatchNorm2d) and bn.training and bn.running_mean is not None and act == "relu" and \
                bn.momentum is not None and pc.partial_conv3.weight.dtype == torch.float32:
            # training: the whole block as ONE autograd node (ops.MlpBlockFn)
            y = ops.MlpBlockFn.apply(x, pc.partial_conv3.weight, w, w2.weight, bn.weight, bn.bias, bn.running_mean,
                                     bn.running_var, float(bn.eps), float(bn.momentum), cp, cpp)
            if bn.num_batches_tracked is not None and not ops.NBT_EXTERNAL:
                bn.num_batches_tracked += 1
            return y
        wa = nn.functional.pad(w[:, :cp], (0, 0, 0, 0, 0, cpp - cp))
        mask = torch.ones(c, device=w.device, dtype=w.dtype)
        mask[:cp] = 0
        wcat = torch.cat([wa, w * mask.view(1, -1, 1, 1)], 1)
        t1 = pc._t1(x, fast)
        u = ops.conv([t1, x], wcat, None)
        v = ops.bn_act(u, bn, act)
        return ops.conv([v], w2.weight, None, residual=x)

    def forward_layer_scale(self, x):
        raise LeadYoloError("MLPBlock: layer_scale > 0 has no kernel; LEAD-YOLO uses layer_scale_init_value=0")


class BasicStage(nn.Module):
    """reference common.py:1493-1525"""

    def __init__(self, dim, depth=1, n_div=4, mlp_ratio=2, layer_scale_init_value=0, norm_layer=nn.BatchNorm2d,
                 act_layer=nn.ReLU, pconv_fw_type="split_cat"):
        super().__init__()
        self.blocks = nn.Sequential(*[
            MLPBlock(dim=dim, n_div=n_div, mlp_ratio=mlp_ratio, drop_path=0.0,
                     layer_scale_init_value=layer_scale_init_value, norm_layer=norm_layer, act_layer=act_layer,
                     pconv_fw_type=pconv_fw_type) for _ in range(depth)])

    def forward(self, x):
        return self.blocks(x)


def _patch_forward(conv: nn.Conv2d, norm, x, fast):
    has_bn = isinstance(norm, nn.BatchNorm2d)
    k = conv.kernel_size[0]
    if conv.in_channels == 3 and k == 4 and conv.stride[0] == 4 and x.shape[1] == 3:
        # stem: straight from the NCHW image
        if fast:
            if has_bn:
                scale, shift = ops.bn_fold(norm, conv.bias)
            else:
                scale, shift = _identity_affine(conv.out_channels, x.device)
                if conv.bias is not None:
                    shift = ops.cached([conv.bias], "b32", lambda: conv.bias.detach().float().contiguous())
            return ops.StemFn.apply(x, conv.weight, scale, shift)
        z = ops.StemFn.apply(x, conv.weight, None, None)
        if has_bn:
            return ops.bn_act(z, norm, "none")
        if conv.bias is not None:
            raise LeadYoloError("stem with bias but no BatchNorm is only supported on the inference path")
        return z
    return conv_bn_act(conv, norm if has_bn else None, "none", [x], fast)


class PatchEmbed_FasterNet(nn.Module):
    """reference common.py:1528-1543"""

    def __init__(self, in_chans, embed_dim, patch_size, patch_stride, norm_layer=nn.BatchNorm2d):
        super().__init__()
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_size, stride=patch_stride, bias=False)
        self.norm = norm_layer(embed_dim) if norm_layer is not None else nn.Identity()

    def forward(self, x):
        return _patch_forward(self.proj, self.norm, x, _fast(self))

    def fuseforward(self, x):
        return _patch_forward(self.proj, None, x, _fast(self))


class PatchMerging_FasterNet(nn.Module):
    """reference common.py:1546-1561"""

    def __init__(self, dim, out_dim, k, patch_stride2, norm_layer=nn.BatchNorm2d):
        super().__init__()
        self.reduction = nn.Conv2d(dim, out_dim, kernel_size=k, stride=patch_stride2, bias=False)
        self.norm = norm_layer(out_dim) if norm_layer is not None else nn.Identity()

    def forward(self, x):
        return _patch_forward(self.reduction, self.norm, x, _fast(self))

    def fuseforward(self, x):
        return _patch_forward(self.reduction, None, x, _fast(self))


# ------------------------------------------------------------------------------------------------------------
# Coordinate attention — reference models/common.py:1565-1637
# ------------------------------------------------------------------------------------------------------------
class h_sigmoid(nn.Module):
    """relu6(x+3)/6 — reference common.py:1565-1571 (parameter-free holder; the math is fused in the kernels)."""

    def __init__(self, inplace=True):
        super().__init__()
        self.relu = nn.ReLU6(inplace=inplace)

    def forward(self, x):
        raise LeadYoloError("h_sigmoid is evaluated inside the fused CoordAtt kernels, not standalone")


class h_swish(nn.Module):
    """x * h_sigmoid(x) — reference common.py:1574-1580"""

    def __init__(self, inplace=True):
        super().__init__()
        self.sigmoid = h_sigmoid(inplace=inplace)

    def forward(self, x):
        return act_only(ops.to_act(x), "hswish")


class CoordAtt(nn.Module):
    """reference common.py:1583-1609.  Pooling and gating are two streaming kernels over x; the 1x1
    convolutions / BatchNorm / h_swish / sigmoid between them run on the tiny pooled matrix."""

    def __init__(self, inp, oup, reduction=32):
        super().__init__()
        self.pool_h = nn.AdaptiveAvgPool2d((None, 1))
        self.pool_w = nn.AdaptiveAvgPool2d((1, None))
        mip = max(8, inp // reduction)
        self.conv1 = nn.Conv2d(inp, mip, kernel_size=1, stride=1, padding=0)
        self.bn1 = nn.BatchNorm2d(mip)
        self.act = h_swish()
        self.conv_h = nn.Conv2d(mip, oup, kernel_size=1, stride=1, padding=0)
        self.conv_w = nn.Conv2d(mip, oup, kernel_size=1, stride=1, padding=0)

    def forward(self, x):
        x = ops.to_act(x)
        n, c, h, w = x.shape
        if self.conv_h.out_channels != c:
            raise LeadYoloError("CoordAtt: oup must equal the input channels")
        fast = _fast(self)
        p = ops.CaPoolFn.apply(x)                                   # (1, c, n*h + n*w, 1)
        nh = n * h
        if fast and self.conv1.out_channels <= 32:
            f32 = lambda t, tag: ops.cached([t], tag, lambda: t.detach().float().reshape(t.shape[0], -1).contiguous())  # noqa: E731
            scale, shift = ops.bn_fold(self.bn1, self.conv1.bias)
            gates = ll.empty_cl(1, c, n * (h + w), 1, device=x.device)
            ll.call("ly_ca_mid_infer", p, c, n * (h + w), nh, c, self.conv1.out_channels,
                    f32(self.conv1.weight, "caw1"), scale, shift, f32(self.conv_h.weight, "cawh"),
                    f32(self.conv_h.bias, "cabh"), f32(self.conv_w.weight, "caww"), f32(self.conv_w.bias, "cabw"),
                    gates, c)
            return ops.CaApplyFn.apply(x, gates)
        y = conv_bn_act(self.conv1, self.bn1, "hswish", [p], fast)   # (1, mip, rows, 1)
        if fast:
            gates = ll.empty_cl(1, c, n * (h + w), 1, device=x.device)
            conv_bn_act(self.conv_h, None, "sigmoid", [y[:, :, :nh]], True, out=gates[:, :, :nh])
            conv_bn_act(self.conv_w, None, "sigmoid", [y[:, :, nh:]], True, out=gates[:, :, nh:])
        else:
            ah = ops.conv([y[:, :, :nh]], self.conv_h.weight, self.conv_h.bias, act="sigmoid")
            aw = ops.conv([y[:, :, nh:]], self.conv_w.weight, self.conv_w.bias, act="sigmoid")
            gates = torch.cat([ah, aw], 2)
        return ops.CaApplyFn.apply(x, gates)


class CA_Bottleneck(nn.Module):
    """reference common.py:1612-1623"""

    def __init__(self, c1, c2, shortcut=True, g=1, e=0.5):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c_, c2, 3, 1, g=g)
        self.ca = CoordAtt(c2, c2, 32)
        self.add = shortcut and c1 == c2

    def forward(self, x):
        y = self.ca(self.cv2(self.cv1(x)))
        return x + y if self.add else y


class C3_CA(nn.Module):
    """CSP bottleneck with 3 convolutions and coordinate attention — reference common.py:1626-1637"""

    def __init__(self, c1, c2, n=1, shortcut=True, g=1, e=0.5):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c1, c_, 1, 1)
        self.cv3 = Conv(2 * c_, c2, 1)
        self.m = nn.Sequential(*(CA_Bottleneck(c_, c_, shortcut, g, e=1.0) for _ in range(n)))

    def forward(self, x):
        return _csp_forward(self, x)


# ------------------------------------------------------------------------------------------------------------
# RFCBAMConv — reference models/rfa.py:77-129
# ------------------------------------------------------------------------------------------------------------
# A/B knob: LY_RF_UNFUSED=1 keeps the k = 3 RFCBAMConv on the round-1 path (gated matrix written to HBM, separate GEMM)
_RF_UNFUSED = bool(__import__('os').environ.get('LY_RF_UNFUSED'))


class SE(nn.Module):
    """rfa.SE — reference models/rfa.py:77-92 (hidden width is the literal `ratio`)."""

    def __init__(self, in_channel, ratio=16):
        super().__init__()
        self.gap = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Sequential(nn.Linear(in_channel, ratio, bias=False), nn.ReLU(),
                                nn.Linear(ratio, in_channel, bias=False), nn.Sigmoid())

    def forward(self, x):
        return self.from_pooled(ops.GapFn.apply(x))                # (n, c, 1, 1)

    def from_pooled(self, g):
        """The two Linear layers + sigmoid on an already pooled (n, c, 1, 1) tensor."""
        fast = _fast(self)
        w1, w2 = self.fc[0].weight, self.fc[2].weight
        if fast:
            wp1 = ops.cached([w1], "se1", lambda: pack_conv_weight(w1.detach()[:, :, None, None], None))
            wp2 = ops.cached([w2], "se2", lambda: pack_conv_weight(w2.detach()[:, :, None, None], None))
            hid = ops.conv_infer([g], wp1, w1.shape[0], act="relu")
            return ops.conv_infer([hid], wp2, w2.shape[0], act="sigmoid")
        hid = ops.conv([g], w1[:, :, None, None], None, act="relu")
        return ops.conv([hid], w2[:, :, None, None], None, act="sigmoid")


class RFCBAMConv(nn.Module):
    """reference models/rfa.py:95-129.  The stride is not stored on the reference module: it is read from
    generate[0].stride (un-pickle path)."""

    def __init__(self, in_channel, out_channel, kernel_size=3, stride=1, dilation=1):
        super().__init__()
        self.kernel_size = kernel_size
        self.generate = nn.Sequential(
            nn.Conv2d(in_channel, in_channel * (kernel_size ** 2), kernel_size, padding=kernel_size // 2,
                      stride=stride, groups=in_channel, bias=False),
            nn.BatchNorm2d(in_channel * (kernel_size ** 2)), nn.ReLU())
        self.get_weight = nn.Sequential(nn.Conv2d(2, 1, kernel_size=3, padding=1, bias=False), nn.Sigmoid())
        self.se = SE(in_channel)
        self.conv = nn.Sequential(nn.Conv2d(in_channel, out_channel, kernel_size, stride=kernel_size),
                                  nn.BatchNorm2d(out_channel), nn.ReLU())

    def forward(self, x):
        x = ops.to_act(x)
        fast = _fast(self)
        k = self.kernel_size
        gen, gbn = self.generate[0], self.generate[1]
        s = gen.stride[0]
        n, c, h, w = x.shape
        conv, cbn = self.conv[0], self.conv[1]
        cout = conv.out_channels
        perm = ops.rf_channel_perm(c, k, x.device) if k > 1 else None
        if fast and c <= 256 and (gen.stride[0] == 1 or (h % 2 == 0 and w % 2 == 0)):
            # two passes over x, no receptive-field matrix in HBM (ly_rfcbam_front_infer)
            gscale, gshift = ops.bn_fold(gbn, None, perm=perm)
            if ops.rfcbam_fused_ok(k, c, cout) and not _RF_UNFUSED:
                scale, shift = ops.bn_fold(cbn, conv.bias)
                return ops.rfcbam_infer_fused(x, gen.weight, s, gscale, gshift, self.se.fc[0].weight,
                                              self.se.fc[2].weight, self.get_weight[0].weight, conv.weight, scale, shift)
            a = ops.rfcbam_front_infer(x, gen.weight, k, s, gscale, gshift, self.se.fc[0].weight,
                                       self.se.fc[2].weight, self.get_weight[0].weight)
            tcl = ops.rf_tc_layout(k, c)
            wp = ops.cached([conv.weight], ("rfconv", c, k, tcl), lambda: pack_conv_weight(
                ops.rf_conv_weight_matrix(conv.weight, k, c, tcl), None))
            scale, shift = ops.bn_fold(cbn, conv.bias)
            return ops.conv_infer([a], wp, cout, scale=scale, shift=shift, act="relu")
        if not fast:
            z, gap = ops.RfFrontFn.apply(x, gen.weight, k, s)
            se = self.se.from_pooled(gap)
        else:
            se = self.se(x)
        if fast:
            gscale, gshift = ops.bn_fold(gbn, None, perm=perm)
            g = ops.RfGenerateFn.apply(x, gen.weight, k, s, gscale, gshift, "relu")
            a = ops.rf_gate_att_infer(g, se, self.get_weight[0].weight, k)
            wp = ops.cached([conv.weight], ("rfconv", c, k), lambda: pack_conv_weight(
                conv.weight.detach().float().permute(0, 2, 3, 1).reshape(cout, k * k * c, 1, 1), None))
            scale, shift = ops.bn_fold(cbn, conv.bias)
            return ops.conv_infer([a], wp, cout, scale=scale, shift=shift, act="relu")
        if perm is None:
            g = ops.bn_act(z, gbn, "relu")
        else:
            inv = ops.rf_channel_perm_inv(c, k, x.device)
            rm_p, rv_p = gbn.running_mean.index_select(0, perm), gbn.running_var.index_select(0, perm)
            g = ops.bn_act(z, gbn, "relu", gamma=ops.PermuteFn.apply(gbn.weight, perm, inv),
                           beta=ops.PermuteFn.apply(gbn.bias, perm, inv), rmean=rm_p, rvar=rv_p)
            if gbn.training:
                with torch.no_grad():
                    gbn.running_mean.index_copy_(0, perm, rm_p)
                    gbn.running_var.index_copy_(0, perm, rv_p)
        a = ops.RfGateAttFn.apply(g, se, self.get_weight[0].weight, k)
        wmat = conv.weight.permute(0, 2, 3, 1).reshape(cout, k * k * c, 1, 1)
        y = ops.conv([a], wmat, conv.bias)
        return ops.bn_act(y, cbn, "relu")


# ------------------------------------------------------------------------------------------------------------
# Detect head — reference models/yolo.py:39-153 (conv on the kernels; sigmoid/grid decode stays fp32 torch code,
# SURVEY §7 H8)
# ------------------------------------------------------------------------------------------------------------
class Detect(nn.Module):
    stride = None
    dynamic = False
    export = False

    def __init__(self, nc=80, anchors=(), ch=(), inplace=True):
        super().__init__()
        self.nc = nc
        self.no = nc + 5
        self.nl = len(anchors)
        self.na = len(anchors[0]) // 2
        self.grid = [torch.empty(0) for _ in range(self.nl)]
        self.anchor_grid = [torch.empty(0) for _ in range(self.nl)]
        self.register_buffer("anchors", torch.tensor(anchors).float().view(self.nl, -1, 2))
        self.m = nn.ModuleList(nn.Conv2d(x, self.no * self.na, 1) for x in ch)
        self.inplace = inplace

    def forward(self, x):
        z = []
        x = list(x)
        fast = _fast(self)
        if fast and not self.export and torch.is_tensor(self.stride):
            # inference: one tensor-core kernel per level: 1x1 head conv + bias, and its epilogue writes both the permuted
            # raw map and this level's decoded rows straight into the concatenated (bs, sum na*ny*nx, no) output
            shapes = [(int(t.shape[2]), int(t.shape[3])) for t in x]
            bs = int(x[0].shape[0])
            rows = sum(self.na * ny * nx for ny, nx in shapes)
            zcat = torch.empty((bs, rows, self.no), dtype=torch.float32, device=x[0].device)
            off = 0
            # strides as Python floats, read once per stride tensor: under the reference's DetectionModel._apply
            # (models/yolo.py:239-249) `stride` lives on the GPU, and a per-forward float() would be a device sync
            # (and an error inside CUDA-graph capture)
            strides = ops.cached([self.stride], "strides", lambda: [float(s) for s in self.stride.detach().cpu()])
            for i in range(self.nl):
                m = self.m[i]
                ny, nx = shapes[i]
                wp = ops.cached([m.weight], ("wp", (m.in_channels,)), lambda: pack_conv_weight(m.weight, None))
                b = ops.cached([m.bias], "b32", lambda: m.bias.detach().float().contiguous())
                raw = torch.empty((bs, self.na, ny, nx, self.no), dtype=torch.float32, device=x[i].device)
                anc = ops.cached([self.anchors], ("anc", i), lambda: self.anchors[i].detach().float().contiguous())
                xi = ops.to_act(x[i])
                if self.na * self.no <= 32 and self.na <= 4:
                    # head conv + bias + decode in one kernel (ly_detect_head_infer)
                    xi, ldx = ll.as_nhwc(xi)
                    ll.call("ly_detect_head_infer", xi, ldx, bs, ny, nx, m.in_channels, wp, b, self.na, self.no,
                            strides[i], anc, raw, zcat, rows, off)
                else:
                    y = ops.conv_infer([xi], wp, m.out_channels, shift=b, out_fp32=True)
                    ll.call("ly_detect_decode", y, bs, ny, nx, self.na, self.no, strides[i], anc, raw, zcat,
                            rows, off)
                x[i] = raw
                off += self.na * ny * nx
            return zcat, x
        for i in range(self.nl):
            m = self.m[i]
            if fast:
                wp = ops.cached([m.weight], ("wp", (m.in_channels,)), lambda: pack_conv_weight(m.weight, None))
                b = ops.cached([m.bias], "b32", lambda: m.bias.detach().float().contiguous())
                y = ops.conv_infer([x[i]], wp, m.out_channels, shift=b, out_fp32=True)
            else:
                y = ops.conv([x[i]], m.weight, m.bias, out_fp32=True)
            bs, _, ny, nx = y.shape
            # y is NHWC in memory: (bs, ny, nx, na*no) -> (bs, na, ny, nx, no); in training the maps stay strided views
            # of the convolution output (the reference makes a contiguous copy, models/yolo.py:95): the loss kernels
            # take strides, and the gradient then flows back into the head's GEMMs without a permute copy
            x[i] = y.permute(0, 2, 3, 1).reshape(bs, ny, nx, self.na, self.no).permute(0, 3, 1, 2, 4)
            if not self.training:
                x[i] = x[i].contiguous()
            if not self.training:
                if self.dynamic or self.grid[i].shape[2:4] != x[i].shape[2:4] or self.grid[i].device != y.device:
                    self.grid[i], self.anchor_grid[i] = self._make_grid(nx, ny, i)
                sg = x[i].sigmoid()
                xy = (sg[..., 0:2] * 2 + self.grid[i]) * self.stride[i]
                wh = (sg[..., 2:4] * 2) ** 2 * self.anchor_grid[i]
                z.append(torch.cat((xy, wh, sg[..., 4:]), 4).view(bs, self.na * nx * ny, self.no))
        return x if self.training else (torch.cat(z, 1),) if self.export else (torch.cat(z, 1), x)

    def _make_grid(self, nx=20, ny=20, i=0):
        d = self.anchors[i].device
        t = torch.float32
        yv, xv = torch.meshgrid(torch.arange(ny, device=d, dtype=t), torch.arange(nx, device=d, dtype=t),
                                indexing="ij")
        grid = torch.stack((xv, yv), 2).expand(1, self.na, ny, nx, 2) - 0.5
        anchor_grid = (self.anchors[i].float() * self.stride[i]).view(1, self.na, 1, 1, 2).expand(
            1, self.na, ny, nx, 2)
        return grid, anchor_grid
