"""Primitive operators of the LEAD-YOLO hot path as torch.autograd.Functions over the C-ABI kernels.

Every tensor that crosses these functions is a logical-NCHW, channels_last (NHWC in memory) bf16 tensor.
Parameters stay fp32 `nn.Parameter`s owned by stock submodules; packed bf16 copies for the tensor
cores are derived per call (training) or cached per parameter version (inference).
"""
from __future__ import annotations

import weakref

import os

import torch

from . import lowlevel as ll
from .packing import pack_conv_weight

BF16 = torch.bfloat16
CL = torch.channels_last

# ---------------------------------------------------------------------------------------------------------
# small caches (never stored on nn.Module instances: modules stay picklable / deepcopy-able)
# ---------------------------------------------------------------------------------------------------------
_cache = {}

# Gradient sink (set by train.TrainEngine while a step runs): data_ptr of a parameter -> its slice of the engine's flat
# gradient buffer, ZEROED at the start of the step.  A backward that owns the whole gradient of a leaf parameter writes
# it there and returns that tensor, so the engine has nothing to copy (one small copy kernel per parameter otherwise).
GRAD_SINK = {}
_SINK_USED = set()
NBT_EXTERNAL = False   # True: the caller advances every BatchNorm's num_batches_tracked itself (one add for all of them)


def grad_sink(param):
    """The flat-gradient slice of `param` if this backward may write there (leaf fp32 parameter of the running
    TrainEngine step, first and only use), else None."""
    if not GRAD_SINK or param is None:
        return None
    key = param.data_ptr()
    v = GRAD_SINK.get(key)
    if v is None or key in _SINK_USED or not param.is_leaf or param.dtype != torch.float32 or v.shape != param.shape:
        return None
    _SINK_USED.add(key)
    return v
# A captured training step replays the SAME launches while the parameters change underneath through raw pointers, so
# everything derived from a parameter has to be recomputed inside the step: train.TrainEngine sets this flag around
# its forward/backward and `cached` then rebuilds whatever is keyed on a tensor that requires grad.
BYPASS_PARAM_CACHE = False


def cached(key_tensors, tag, builder):
    """Memoise `builder()` on the identity + version of `key_tensors` (inference-time weight packing)."""
    if BYPASS_PARAM_CACHE and any(t.requires_grad for t in key_tensors):
        return builder()
    key = (tag,) + tuple(id(t) for t in key_tensors)
    ver = tuple((t._version, t.data_ptr()) for t in key_tensors)
    hit = _cache.get(key)
    if hit is not None and hit[0] == ver and all(r() is t for r, t in zip(hit[1], key_tensors)):
        return hit[2]
    val = builder()
    if len(_cache) > 4096:
        _cache.clear()
    _cache[key] = (ver, [weakref.ref(t) for t in key_tensors], val)
    return val


# ---------------------------------------------------------------------------------------------------------
# Zero-initialised fp32 accumulators.  Many kernels add into small fp32 buffers (BatchNorm sums, weight-gradient
# partials, attention gradients): one torch.zeros launch each.  A training engine can open an arena instead: ONE
# buffer zeroed once per step, handed out in 256-byte aligned slices (train.TrainEngine).  Outside an arena this
# is plain torch.zeros.
# ---------------------------------------------------------------------------------------------------------
class _Arena:
    buf = None
    off = 0
    need = 0
    active = False


def arena_begin(device, min_floats=0):
    """Start a step: (re)allocate to the high-water mark of the previous step and zero the whole arena."""
    want = max(int(_Arena.need * 1.05) + 1024, min_floats)
    if _Arena.buf is None or _Arena.buf.numel() < want or _Arena.buf.device != torch.device(device):
        _Arena.buf = torch.empty(want, dtype=torch.float32, device=device)
    _Arena.buf.zero_()
    _Arena.off = 0
    _Arena.need = 0
    _Arena.active = True


def arena_end():
    _Arena.active = False


def zeros_f32(n, device):
    """n zeroed floats: an arena slice when a training engine has one open, else torch.zeros."""
    n = int(n)
    padded = (n + 63) // 64 * 64
    _Arena.need += padded
    if _Arena.active and _Arena.buf is not None and _Arena.off + padded <= _Arena.buf.numel() and \
            _Arena.buf.device == torch.device(device):
        out = _Arena.buf[_Arena.off:_Arena.off + n]
        _Arena.off += padded
        return out
    return torch.zeros(n, dtype=torch.float32, device=device)


def to_act(x: torch.Tensor) -> torch.Tensor:
    """Any float NCHW tensor -> bf16 channels_last activation."""
    if x.dtype == BF16 and ll.pixel_ld(x) is not None:
        return x
    return x.to(dtype=BF16, memory_format=CL)


def _rows(x):
    n, c, h, w = x.shape
    return n * h * w


# ---------------------------------------------------------------------------------------------------------
# Convolution (implicit GEMM) with autograd
# ---------------------------------------------------------------------------------------------------------
def _geom(k, s):
    patch = (s == k and k > 1)
    pad = 0 if (patch or k == 1) else k // 2
    return patch, pad


def _w32(weight):
    w = weight.detach()
    return w if (w.dtype == torch.float32 and w.is_contiguous()) else w.float().contiguous()


# Batched weight packing (train.TrainEngine).  While PACK_RECORD is a list, every pack of a leaf fp32 parameter of the
# running step is remembered as (key, packed tensor, job fields); the engine turns the list into ONE ly_pack_batch launch at
# the start of its captured step and sets PACK_PLAN = {key: packed tensor}: the Functions then find their operand ready.
PACK_RECORD = None
PACK_PLAN = None


def _pack_key(kind, weight, *geom):
    return (kind, weight.data_ptr()) + tuple(geom)


def _plannable(weight):
    return weight.grad_fn is None and weight.is_leaf and weight.dtype == torch.float32 and weight.is_contiguous() and \
        weight.data_ptr() in GRAD_SINK


def packed_weight(weight, chans, k, patch):
    """bf16 [Cout][Kp] forward operand of `weight` for K-concatenated sources `chans`: one kernel launch, memoised on
    the parameter's version (training: once per optimiser step; inference: once)."""
    chans = tuple(int(c) for c in chans)
    key = None
    if (PACK_PLAN is not None or PACK_RECORD is not None) and _plannable(weight):
        key = _pack_key("f", weight, chans, k, patch)
        if PACK_PLAN is not None and key in PACK_PLAN:
            return PACK_PLAN[key]

    def build():
        cout = weight.shape[0]
        from .packing import packed_k
        kp = packed_k(list(chans), k, k, patch)
        wp = torch.empty((cout, kp), dtype=BF16, device=weight.device)
        c4 = list(chans) + [0] * (4 - len(chans))
        ll.call("ly_pack_weight", _w32(weight), cout, k, k, 1 if patch else 0, c4[0], c4[1], c4[2], c4[3], wp)
        if key is not None and PACK_RECORD is not None:
            PACK_RECORD.append((key, wp, dict(kind=0, w=weight.data_ptr(), Cout=cout, Cin=sum(chans), kh=k, kw=k,
                                              patch=1 if patch else 0, c=c4, coff=0, cs=0, P=0, Kp=kp, total=cout * kp)))
        return wp

    if weight.grad_fn is not None:          # derived tensor (e.g. a masked / permuted view): nothing stable to key on
        return build()
    return cached([weight], ("pw", chans, k, patch), build)


def packed_weight_dgrad(weight, cout, cin, k, patch, coff, cs):
    """bf16 weights of the GEMM that maps dZ to the gradient of source channels [coff, coff+cs) (one launch)."""
    key = None
    if (PACK_PLAN is not None or PACK_RECORD is not None) and _plannable(weight):
        key = _pack_key("d", weight, cout, cin, k, patch, coff, cs)
        if PACK_PLAN is not None and key in PACK_PLAN:
            return PACK_PLAN[key]

    def build():
        p64 = ((cout + 7) // 8 * 8 + 63) // 64 * 64
        rows, kp = (k * k * cs, p64) if patch else (cs, k * k * p64)
        wd = torch.empty((rows, kp), dtype=BF16, device=weight.device)
        ll.call("ly_pack_weight_dgrad", _w32(weight), cout, cin, k, k, 1 if patch else 0, coff, cs, wd)
        if key is not None and PACK_RECORD is not None:
            PACK_RECORD.append((key, wd, dict(kind=1, w=weight.data_ptr(), Cout=cout, Cin=cin, kh=k, kw=k,
                                              patch=1 if patch else 0, c=[0, 0, 0, 0], coff=coff, cs=cs, P=p64, Kp=kp,
                                              total=rows * kp)))
        return wd

    if weight.grad_fn is not None:
        return build()
    return cached([weight], ("pwd", k, patch, coff, cs), build)


def pack_batch_table(records, device):
    """(device table, n jobs, total blocks, plan dict) for ly_pack_batch from PACK_RECORD entries (first of each key)."""
    import ctypes as C
    from ._lib import PackJob
    seen, jobs = {}, []
    for key, buf, f in records:
        if key not in seen:
            seen[key] = buf
            jobs.append((buf, f))
    arr = (PackJob * len(jobs))()
    blocks = 0
    for i, (buf, f) in enumerate(jobs):
        j = arr[i]
        j.w, j.dst = f["w"], buf.data_ptr()
        j.kind, j.Cout, j.Cin, j.kh, j.kw, j.patch = f["kind"], f["Cout"], f["Cin"], f["kh"], f["kw"], f["patch"]
        for q in range(4):
            j.c[q] = f["c"][q]
        j.coff, j.cs, j.P, j.Kp = f["coff"], f["cs"], f["P"], f["Kp"]
        j.total, j.first_block = f["total"], blocks
        blocks += (f["total"] + 255) // 256
    raw = bytes(arr)
    table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(device)
    return table, len(jobs), blocks, seen


class ConvFn(torch.autograd.Function):
    """y = act(conv(cat(srcs), weight) + bias) (+ residual).  act in {none, relu, sigmoid} (derivative from y)."""

    @staticmethod
    def forward(ctx, weight, bias, residual, k, s, act, out_fp32, *srcs):
        patch, pad = _geom(k, s)
        srcs = [to_act(x) for x in srcs]
        chans = [x.shape[1] for x in srcs]
        cout = weight.shape[0]
        wp = packed_weight(weight, chans, k, patch)
        shift = bias.detach().float().contiguous() if bias is not None else None
        res = to_act(residual) if residual is not None else None
        y = ll.conv_gemm(srcs, wp, cout, k=k, stride=s, pad=pad, shift=shift, act=act, residual=res,
                         out_fp32=out_fp32)
        ctx.cfg = (k, s, act, patch, pad, chans, out_fp32, bias is not None, residual is not None)
        ctx.bias_ref = bias
        need_y = act != "none"
        ctx.save_for_backward(weight, y if need_y else None, *srcs)
        return y

    @staticmethod
    def backward(ctx, dy):
        k, s, act, patch, pad, chans, out_fp32, has_bias, has_res = ctx.cfg
        weight, y, *srcs = ctx.saved_tensors
        cout = weight.shape[0]
        dy = to_act(dy)
        d_res = dy if (has_res and ctx.needs_input_grad[2]) else None
        if act != "none":
            assert not has_res, "activation + residual backward is not needed on this path"
            dz = torch.empty_like(dy, memory_format=CL)
            ll.call("ly_act_bwd_from_out", dy, ll.pixel_ld(dy), y, ll.pixel_ld(y), _rows(dy), cout, ll.ACT[act],
                    dz, ll.pixel_ld(dz))
        else:
            dz = dy
        # channel counts that are not a multiple of 8 (Detect: 18) are zero-padded for the tensor-core kernels
        cpad = (cout + 7) // 8 * 8
        if cpad != cout:
            dzp = dz.new_zeros((dz.shape[0], cpad, dz.shape[2], dz.shape[3])).contiguous(memory_format=CL)
            dzp[:, :cout] = dz
            dz = dzp
        d_bias = None
        if has_bias and ctx.needs_input_grad[1]:
            sink = grad_sink(ctx.bias_ref) if cpad == cout else None
            sums = sink if sink is not None else zeros_f32(cpad, dz.device)
            ll.call("ly_col_stats", dz, ll.pixel_ld(dz), _rows(dz), cpad, sums, None)
            d_bias = sums[:cout] if sink is None else sink
        cin = sum(chans)
        d_srcs = []
        off = 0
        for i, x in enumerate(srcs):
            cs = chans[i]
            if ctx.needs_input_grad[7 + i]:
                wdp = packed_weight_dgrad(weight, cout, cin, k, patch, off, cs)
                if patch:
                    dx = ll.conv_gemm([dz], wdp, k * k * cs, k=1, patch_scatter=(k, cs))
                else:
                    dx = ll.conv_gemm([dz], wdp, cs, k=k, stride=1, pad=pad)
                d_srcs.append(dx)
            else:
                d_srcs.append(None)
            off += cs
        d_w = None
        if ctx.needs_input_grad[0]:
            d_w = _wgrad_oihw(srcs, dz, weight, chans, k, s, pad, patch)
        return (d_w, d_bias, d_res, None, None, None, None, *d_srcs)


def conv(srcs, weight, bias=None, k=1, s=1, act="none", residual=None, out_fp32=False):
    return ConvFn.apply(weight, bias, residual, k, s, act, out_fp32, *srcs)


def conv_infer(srcs, wp, cout, k=1, s=1, scale=None, shift=None, act="none", residual=None, out=None,
               out_fp32=False):
    """Inference-only convolution with pre-packed weights and folded BN scale/shift (no autograd)."""
    patch, pad = _geom(k, s)
    return ll.conv_gemm([to_act(x) for x in srcs], wp, cout, k=k, stride=s, pad=pad, scale=scale, shift=shift,
                        act=act, residual=None if residual is None else to_act(residual), out=out,
                        out_fp32=out_fp32)


# ---------------------------------------------------------------------------------------------------------
# BatchNorm (+ activation)
# ---------------------------------------------------------------------------------------------------------
class BnActFn(torch.autograd.Function):
    """y = act(BN(z)); batch statistics + running-stat update when `training`, running statistics otherwise."""

    @staticmethod
    def forward(ctx, z, gamma, beta, running_mean, running_var, eps, momentum, act, training):
        z = to_act(z)
        c = z.shape[1]
        m = _rows(z)
        dev = z.device
        stats = torch.empty((4, c), dtype=torch.float32, device=dev)  # mean, rstd, scale, shift
        g32, b32 = gamma.detach().float(), beta.detach().float()
        if training:
            y = torch.empty_like(z, memory_format=CL)
            acc = zeros_f32(5 * c, dev)
            ll.call("ly_bn_stats", z, ll.pixel_ld(z), m, c, acc)
            ll.call("ly_bn_apply_train", z, ll.pixel_ld(z), m, c, acc, g32, b32, eps, momentum, running_mean,
                    running_var, stats, ll.ACT[act], y, ll.pixel_ld(y))
            # the kernel updated the running statistics through raw pointers: bump their version counters so that
            # caches keyed on them (bn_fold and everything derived from it) see the change
            for t in (running_mean, running_var):
                if t is not None:
                    torch.autograd.graph.increment_version(t)
        else:
            ll.call("ly_bn_fold", c, eps, g32, b32, running_mean, running_var, None, stats[0], stats[1], stats[2],
                    stats[3])
            y = torch.empty_like(z, memory_format=CL)
            ll.call("ly_scale_shift_act", z, ll.pixel_ld(z), m, c, stats[2], stats[3], ll.ACT[act], y,
                    ll.pixel_ld(y))
        ctx.cfg = (act, training)
        ctx.affine = (gamma, beta)
        ctx.save_for_backward(z, stats)
        return y

    @staticmethod
    def backward(ctx, dy):
        act, training = ctx.cfg
        z, stats = ctx.saved_tensors
        dy = to_act(dy)
        c = z.shape[1]
        m = _rows(z)
        d_beta, d_gamma = _bn_grad_buffers(ctx.affine[0], ctx.affine[1], c, z.device)
        parts = torch.empty(ll.reduce_blocks(m, c) * 2 * c, dtype=torch.float32, device=z.device)
        dz = torch.empty_like(z, memory_format=CL)
        ll.call("ly_bn_act_bwd", dy, ll.pixel_ld(dy), z, ll.pixel_ld(z), m, c, stats[2], stats[3], stats[0],
                stats[1], ll.ACT[act], 1 if training else 0, d_beta, d_gamma, parts, zeros_f32(2, z.device), dz,
                ll.pixel_ld(dz))
        return dz, d_gamma, d_beta, None, None, None, None, None, None


def _bn_grad_buffers(gamma, beta, c, device):
    """(d_beta, d_gamma) fp32 [c] for ly_bn_act_bwd: the parameters' gradient sinks when there are any."""
    sb, sg = grad_sink(beta), grad_sink(gamma)
    if sb is None or sg is None:
        s12 = torch.empty((2, c), dtype=torch.float32, device=device)
        sb = s12[0] if sb is None else sb
        sg = s12[1] if sg is None else sg
    return sb, sg


def bn_act(z, bn, act="none", gamma=None, beta=None, rmean=None, rvar=None):
    """Apply nn.BatchNorm2d `bn` (+activation) with reference semantics, reading eps/momentum from the module."""
    training = bn.training or bn.running_mean is None
    gamma = bn.weight if gamma is None else gamma
    beta = bn.bias if beta is None else beta
    rmean = bn.running_mean if rmean is None else rmean
    rvar = bn.running_var if rvar is None else rvar
    if bn.momentum is None:
        # nn.BatchNorm2d(momentum=None) = cumulative moving average with factor 1/num_batches_tracked, which would
        # need a host read of the counter per step; LEAD-YOLO sets momentum=0.03 (utils/torch_utils.py:212-221)
        raise ll.LeadYoloError("BatchNorm2d(momentum=None) (cumulative average) has no kernel in lead_yolo_b200")
    mom = bn.momentum
    y = BnActFn.apply(z, gamma, beta, rmean, rvar, float(bn.eps), float(mom), act, training)
    if training and bn.num_batches_tracked is not None and not NBT_EXTERNAL:
        bn.num_batches_tracked += 1
    return y


def bn_fold(bn, conv_bias=None, perm=None):
    """(scale, shift) of an eval-mode BatchNorm folded behind a convolution (cached per parameter version)."""
    keys = [bn.weight, bn.bias, bn.running_mean, bn.running_var] + ([conv_bias] if conv_bias is not None else [])
    if bn.num_batches_tracked is not None:
        keys.append(bn.num_batches_tracked)     # bumped by every training-mode forward (bn_act)

    def build():
        c = bn.weight.numel()
        g, b, rm, rv = (t.detach().float() for t in (bn.weight, bn.bias, bn.running_mean, bn.running_var))
        cb = conv_bias.detach().float() if conv_bias is not None else None
        if perm is not None:
            g, b, rm, rv = g[perm].contiguous(), b[perm].contiguous(), rm[perm].contiguous(), rv[perm].contiguous()
        out = torch.empty((2, c), dtype=torch.float32, device=g.device)
        ll.call("ly_bn_fold", c, float(bn.eps), g, b, rm, rv, cb, None, None, out[0], out[1])
        return out[0], out[1]

    return cached(keys, ("bnfold", perm is not None), build)


# ---------------------------------------------------------------------------------------------------------
# Stem
# ---------------------------------------------------------------------------------------------------------
STEM_WGRAD_CUDA_CORES = bool(os.environ.get("LY_STEM_WGRAD_CUDA"))   # dev knob: fp32 CUDA-core stem weight gradient


class StemFn(torch.autograd.Function):
    """4x4/s4 patch embedding straight from the NCHW image (fp32, or uint8 with the /255 folded in)."""

    @staticmethod
    def forward(ctx, img, weight, scale, shift):
        assert img.dim() == 4 and img.shape[1] == 3 and tuple(weight.shape[1:]) == (3, 4, 4)
        is_u8 = img.dtype == torch.uint8
        if not is_u8 and img.dtype != torch.float32:
            img = img.float()
        img = img.contiguous()
        n, _, h, w = img.shape
        cout = weight.shape[0]
        # [48][Cout] fp32 copy of the weight: memoised on the weight's version, not rebuilt (one copy kernel) per step
        wk = cached([weight], "stem_wk", lambda: weight.detach().float().reshape(cout, 48).t().contiguous())
        out = ll.empty_cl(n, cout, h // 4, w // 4, device=img.device)
        in_scale = 1.0 / 255.0 if is_u8 else 1.0
        ll.call("ly_stem_fwd", img, 1 if is_u8 else 0, n, h, w, cout, wk, scale, shift, in_scale, out, cout)
        ctx.save_for_backward(img)
        ctx.weight_ref = weight
        ctx.cfg = (is_u8, in_scale, cout, weight.dtype)
        return out

    @staticmethod
    def backward(ctx, dz):
        (img,) = ctx.saved_tensors
        is_u8, in_scale, cout, wdt = ctx.cfg
        dz = to_act(dz)
        n, _, h, w = img.shape
        if cout % 8 == 0 and not STEM_WGRAD_CUDA_CORES:
            # patch matrix (bf16, exact for uint8 images) + the 1x1 weight-gradient GEMM on the tensor cores
            pm = ll.empty_cl(n, 64, h // 4, w // 4, device=dz.device)
            ll.call("ly_stem_im2col", img, 1 if is_u8 else 0, n, h, w, pm)
            dwp = ll.conv_wgrad([pm], dz, cout, k=1, zeros=zeros_f32)            # [cout][64], column c*16 + ky*4 + kx
            sink = grad_sink(ctx.weight_ref)
            d_w = sink if sink is not None else torch.empty((cout, 3, 4, 4), dtype=torch.float32, device=dz.device)
            torch.mul(dwp[:, :48].reshape(cout, 3, 4, 4), in_scale, out=d_w)
            return None, (d_w if sink is not None else d_w.to(wdt)), None, None
        dwk = zeros_f32(48 * cout, dz.device).view(48, cout)
        ll.call("ly_stem_wgrad", img, 1 if is_u8 else 0, n, h, w, cout, dz, ll.pixel_ld(dz), in_scale, dwk)
        return None, dwk.t().reshape(cout, 3, 4, 4).to(wdt), None, None


# ---------------------------------------------------------------------------------------------------------
# Pooling / resampling
# ---------------------------------------------------------------------------------------------------------
class MaxPool5Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = to_act(x)
        n, c, h, w = x.shape
        y = torch.empty_like(x, memory_format=CL)
        need = x.requires_grad and torch.is_grad_enabled()
        idx = torch.empty((n * h * w, c), dtype=torch.uint8, device=x.device)
        ll.call("ly_maxpool5_fwd", x, ll.pixel_ld(x), n, h, w, c, y, c, idx)
        ctx.save_for_backward(idx)
        ctx.shape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        n, c, h, w = ctx.shape
        dy = to_act(dy)
        dx = ll.empty_cl(n, c, h, w, device=dy.device)
        ll.call("ly_maxpool5_bwd", dy, ll.pixel_ld(dy), idx, n, h, w, c, dx, c, 0)
        return dx


def maxpool5_into(x, out):
    """Inference: 5x5 max pool of x written into `out` (a channel slice of a wider NHWC buffer)."""
    n, c, h, w = x.shape
    ll.call("ly_maxpool5_fwd", x, ll.pixel_ld(x), n, h, w, c, out, ll.pixel_ld(out), None)
    return out


class UpsampleCatFn(torch.autograd.Function):
    """cat([up2x?(a), b, ...], dim=1) materialised by slice copies; `ups` marks the tensors to upsample 2x."""

    @staticmethod
    def forward(ctx, ups, *xs):
        xs = [to_act(x) for x in xs]
        n = xs[0].shape[0]
        h = xs[0].shape[2] * (2 if ups[0] else 1)
        w = xs[0].shape[3] * (2 if ups[0] else 1)
        ctot = sum(x.shape[1] for x in xs)
        out = ll.empty_cl(n, ctot, h, w, device=xs[0].device)
        off = 0
        for x, u in zip(xs, ups):
            c = x.shape[1]
            ll.call("ly_copy_upsample", x, ll.pixel_ld(x), n, x.shape[2], x.shape[3], c, 2 if u else 1,
                    out[:, off:off + c], ctot)
            off += c
        ctx.cfg = (ups, [x.shape for x in xs])
        return out

    @staticmethod
    def backward(ctx, dy):
        ups, shapes = ctx.cfg
        dy = to_act(dy)
        grads, off = [], 0
        for u, shp in zip(ups, shapes):
            c = shp[1]
            sl = dy[:, off:off + c]
            if u:
                dx = ll.empty_cl(*shp, device=dy.device)
                ll.call("ly_downsum2", sl, ll.pixel_ld(sl), shp[0], shp[2], shp[3], c, dx, c)
                grads.append(dx)
            else:
                grads.append(sl)
            off += c
        return (None, *grads)


GAP_PARTS = 16   # LY_GAP_PARTS of include/leadyolo.h


def _gap_mean(x):
    """(N, C, 1, 1) bf16 mean over H*W of an NHWC activation."""
    n, c, h, w = x.shape
    if c <= 512:
        parts = torch.empty((GAP_PARTS, n, c), dtype=torch.float32, device=x.device)
        out = ll.empty_cl(n, c, 1, 1, device=x.device)
        ll.call("ly_gap_mean", x, ll.pixel_ld(x), n, h * w, c, parts, out)   # ordered partial sums: deterministic
        return out
    out = ll.empty_cl(n, c, 1, 1, device=x.device)
    ll.call("ly_gap", x, ll.pixel_ld(x), n, h * w, c, out, c)
    return out


class GapFn(torch.autograd.Function):
    """Global average pool -> (N, C, 1, 1) bf16."""

    @staticmethod
    def forward(ctx, x):
        x = to_act(x)
        n, c, h, w = x.shape
        ctx.shape = (n, c, h, w)
        return _gap_mean(x)

    @staticmethod
    def backward(ctx, d):
        n, c, h, w = ctx.shape
        return (d.float() / float(h * w)).to(BF16).expand(n, c, h, w).contiguous(memory_format=CL)


# ---------------------------------------------------------------------------------------------------------
# CoordAtt pieces
# ---------------------------------------------------------------------------------------------------------
class CaPoolFn(torch.autograd.Function):
    """x (N,C,H,W) -> pooled "HW rows" matrix as a (1, C, N*H + N*W, 1) activation."""

    @staticmethod
    def forward(ctx, x):
        x = to_act(x)
        n, c, h, w = x.shape
        p = ll.empty_cl(1, c, n * (h + w), 1, device=x.device)
        ll.call("ly_ca_pool", x, ll.pixel_ld(x), n, h, w, c, p, c)
        ctx.shape = (n, c, h, w)
        return p

    @staticmethod
    def backward(ctx, dp):
        n, c, h, w = ctx.shape
        dp = to_act(dp)
        dx = ll.empty_cl(n, c, h, w, device=dp.device)
        ll.call("ly_ca_pool_bwd", dp, ll.pixel_ld(dp), n, h, w, c, dx, c)
        return dx


class CaApplyFn(torch.autograd.Function):
    """out = x * a_h * a_w with the gates in "HW rows" layout."""

    @staticmethod
    def forward(ctx, x, gates):
        x, gates = to_act(x), to_act(gates)
        n, c, h, w = x.shape
        out = torch.empty_like(x, memory_format=CL)
        ll.call("ly_ca_apply", x, ll.pixel_ld(x), gates, ll.pixel_ld(gates), n, h, w, c, out, c)
        ctx.save_for_backward(x, gates)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, gates = ctx.saved_tensors
        dy = to_act(dy)
        n, c, h, w = x.shape
        dx = torch.empty_like(x, memory_format=CL)
        ll.call("ly_ca_apply", dy, ll.pixel_ld(dy), gates, ll.pixel_ld(gates), n, h, w, c, dx, c)
        dg = torch.empty_like(gates, memory_format=CL)
        ll.call("ly_ca_gate_grad", x, ll.pixel_ld(x), dy, ll.pixel_ld(dy), gates, ll.pixel_ld(gates), n, h, w, c,
                dg, c)
        return dx, dg


# ---------------------------------------------------------------------------------------------------------
# RFCBAM pieces
# ---------------------------------------------------------------------------------------------------------
def rf_out_size(h, w, k, s):
    p = k // 2
    return (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1


def rf_pack_generate_weight(wg, c, k):
    """generate.0.weight [c*k*k, 1, k, k] -> fp32 [k*k (t)][k*k (ij)][c]."""
    return wg.detach().float().reshape(c, k * k, k * k).permute(1, 2, 0).contiguous()


def rf_channel_perm(c, k, device):
    """index p with G-column (t*c + ch) <- reference channel (ch*k*k + t) (a constant per (c, k, device))."""
    key = ("rfperm", c, k, str(device))
    hit = _cache.get(key)
    if hit is None:
        t = torch.arange(k * k, device=device).view(-1, 1)
        ch = torch.arange(c, device=device).view(1, -1)
        hit = _cache[key] = (ch * (k * k) + t).reshape(-1)
    return hit


def rf_channel_perm_inv(c, k, device):
    """inverse of rf_channel_perm: reference channel (ch*k*k + t) <- G-column (t*c + ch)."""
    key = ("rfperminv", c, k, str(device))
    hit = _cache.get(key)
    if hit is None:
        hit = _cache[key] = torch.argsort(rf_channel_perm(c, k, device))
    return hit


class PermuteFn(torch.autograd.Function):
    """y = x[perm] for a PERMUTATION of dim 0; the gradient is a gather with the inverse permutation (torch's index
    backward is an index_put with accumulation: a sort and an atomic scatter per call)."""

    @staticmethod
    def forward(ctx, x, perm, inv):
        ctx.inv = inv
        return x.index_select(0, perm)

    @staticmethod
    def backward(ctx, dy):
        return dy.index_select(0, ctx.inv), None, None


RF_GENERATE_CUDA_CORES = bool(os.environ.get("LY_RF_GENERATE_CUDA"))   # dev knob: fp32 CUDA-core feature generation


def _rf_generate_raw(x, wgk, k, s, ho, wo):
    """z = the grouped k x k convolution of rfa.py:95-100 alone, (N, k*k*C, ho, wo), channel t*C + c.  k = 3 with
    C % 16 == 0 runs on the tensor cores (ly_rf_tc_generate, one tcgen05.mma per tap and 16-channel group, bf16 weights);
    everything else on ly_rf_generate (fp32 weights)."""
    n, c, h, w = x.shape
    g = ll.empty_cl(n, k * k * c, ho, wo, device=x.device)
    if k == 3 and c % 16 == 0 and not RF_GENERATE_CUDA_CORES:
        wc = wgk.view(9, 9, c // 16, 16).permute(2, 1, 0, 3).reshape(c // 16, 1296).to(BF16).contiguous()
        ll.call("ly_rf_tc_generate", x, ll.pixel_ld(x), n, h, w, c, s, ho, wo, wc, _rf_zero_shift(c, x.device), g)
    else:
        ll.call("ly_rf_generate", x, ll.pixel_ld(x), n, h, w, c, k, s, ho, wo, wgk, None, None, ll.ACT["none"], g)
    return g


def _rf_zero_shift(c, device):
    key = ("rfzero", c, device)
    zero = _cache.get(key)
    if zero is None:
        zero = _cache[key] = torch.zeros((c // 16, 144), dtype=torch.int32, device=device)
    return zero


def _rf_generate_bwd(dz, x, wgk, k, s, ho, wo, extra, need_dx, need_dw):
    """(dx, dwgk) of the receptive-field generation.  k = 3 with C % 16 == 0: the input gradient runs on the tensor cores
    (ly_rf_tc_taps: patch gradients, ly_rf_col2im: fold), the weight gradient on ly_rf_generate_bwd; otherwise both there."""
    n, c, h, w = x.shape
    dev = x.device
    dx = ll.empty_cl(n, c, h, w, device=dev) if need_dx else None
    tc = need_dx and k == 3 and c % 16 == 0 and not RF_GENERATE_CUDA_CORES
    if tc:
        wct = wgk.view(9, 9, c // 16, 16).permute(2, 0, 1, 3).reshape(c // 16, 1296).to(BF16).contiguous()
        r = ll.empty_cl(n, 9 * c, ho, wo, device=dev)
        ll.call("ly_rf_tc_taps", dz, n, ho, wo, c, wct, _rf_zero_shift(c, dev), r)
        ll.call("ly_rf_col2im", r, n, h, w, c, s, ho, wo, extra, dx, c)
    dwgk = zeros_f32(wgk.numel(), dev).view(wgk.shape) if need_dw else None
    if dwgk is not None or (need_dx and not tc):
        scr = ll.rf_wgrad_scratch(c, k, dev) if dwgk is not None else None
        ll.call("ly_rf_generate_bwd", dz, x, ll.pixel_ld(x), n, h, w, c, k, s, ho, wo, wgk, extra, None if tc else dx, c,
                dwgk, scr, 0 if scr is None else scr.numel())
    return dx, dwgk


class RfGenerateFn(torch.autograd.Function):
    """Depthwise receptive-field features z (N, k*k*C, ho, wo), channel t*C+c; optional fused affine+act."""

    @staticmethod
    def forward(ctx, x, wg, k, s, scale, shift, act):
        x = to_act(x)
        n, c, h, w = x.shape
        ho, wo = rf_out_size(h, w, k, s)
        wgk = rf_pack_generate_weight(wg, c, k)
        if scale is None:
            g = _rf_generate_raw(x, wgk, k, s, ho, wo)
        else:
            g = ll.empty_cl(n, k * k * c, ho, wo, device=x.device)
            ll.call("ly_rf_generate", x, ll.pixel_ld(x), n, h, w, c, k, s, ho, wo, wgk, scale, shift, ll.ACT[act], g)
        ctx.save_for_backward(x, wgk)
        ctx.cfg = (k, s, ho, wo, wg.dtype, scale is not None)
        return g

    @staticmethod
    def backward(ctx, dz):
        x, wgk = ctx.saved_tensors
        k, s, ho, wo, wdt, fused = ctx.cfg
        assert not fused, "fused affine epilogue is inference-only"
        dz = to_act(dz).contiguous(memory_format=CL)
        n, c, h, w = x.shape
        dx, dwgk = _rf_generate_bwd(dz, x, wgk, k, s, ho, wo, None, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        dwg = None
        if dwgk is not None:
            dwg = dwgk.permute(2, 0, 1).reshape(c * k * k, 1, k, k).to(wdt)
        return dx, dwg, None, None, None, None, None


class RfFrontFn(torch.autograd.Function):
    """(z, gap) = (receptive-field features of x, SE average pool of x) — rfa.py:113-115 and 80,90 — as ONE autograd
    node, so that x receives ONE gradient: the pooling gradient rides into the feature kernel's input-gradient pass
    (`extra`) instead of being expanded to a full tensor and added by autograd."""

    @staticmethod
    def forward(ctx, x, wg, k, s):
        x = to_act(x)
        n, c, h, w = x.shape
        ho, wo = rf_out_size(h, w, k, s)
        wgk = rf_pack_generate_weight(wg, c, k)
        g = _rf_generate_raw(x, wgk, k, s, ho, wo)
        gap = _gap_mean(x)
        ctx.save_for_backward(x, wgk)
        ctx.cfg = (k, s, ho, wo, wg.dtype)
        return g, gap

    @staticmethod
    def backward(ctx, dz, dgap):
        x, wgk = ctx.saved_tensors
        k, s, ho, wo, wdt = ctx.cfg
        dz = to_act(dz).contiguous(memory_format=CL)
        n, c, h, w = x.shape
        extra = None
        if dgap is not None:
            extra = (dgap.float() * (1.0 / (h * w))).reshape(n, c).contiguous()
        dx, dwgk = _rf_generate_bwd(dz, x, wgk, k, s, ho, wo, extra, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        dwg = None
        if dwgk is not None:
            dwg = dwgk.permute(2, 0, 1).reshape(c * k * k, 1, k, k).to(wdt)
        return dx, dwg, None, None


class MlpBlockFn(torch.autograd.Function):
    """FasterNet block body (models/common.py:1432-1437, 1478-1482) as ONE autograd node:
        t1 = conv3x3(x[:, :cpp], Wp);  u = W1cat [t1 | x];  v = relu(BN(u));  y = x + W2 v
    Wp = the partial conv's weight zero-padded to the 8-aligned slice (cp -> cpp channels), W1cat = [W1[:, :cp] padded |
    W1 with its first cp columns zeroed]; both are built here from the raw parameters and their gradients are folded
    back here, so autograd sees neither the padding nor the slicing (a zero fill, a copy and an add per slice otherwise).
    x gets ONE gradient, written in NHWC — the slice, the two convolutions reading x and the residual would otherwise be
    summed by autograd through a zero-filled NCHW tensor and two extra adds."""

    @staticmethod
    def forward(ctx, x, wpc, w1, w2, gamma, beta, rmean, rvar, eps, momentum, cp, cpp):
        x = to_act(x)
        n, c, h, w = x.shape
        m = n * h * w
        hid = w1.shape[0]
        with torch.no_grad():
            wp = torch.nn.functional.pad(wpc, (0, 0, 0, 0, 0, cpp - cp, 0, cpp - cp))
            wcat = torch.zeros((hid, cpp + c, 1, 1), dtype=w1.dtype, device=w1.device)
            wcat[:, :cp] = w1[:, :cp]
            wcat[:, cpp + cp:] = w1[:, cp:]
        xs = x[:, :cpp]
        t1 = ll.conv_gemm([xs], packed_weight(wp, [cpp], 3, False), cpp, k=3, stride=1, pad=1)
        u = ll.conv_gemm([t1, x], packed_weight(wcat, [cpp, c], 1, False), hid)
        v = torch.empty_like(u, memory_format=CL)
        stats = torch.empty((4, hid), dtype=torch.float32, device=x.device)
        acc = zeros_f32(5 * hid, x.device)
        ll.call("ly_bn_stats", u, ll.pixel_ld(u), m, hid, acc)
        ll.call("ly_bn_apply_train", u, ll.pixel_ld(u), m, hid, acc, gamma.detach().float(), beta.detach().float(), eps,
                momentum, rmean, rvar, stats, ll.ACT["relu"], v, ll.pixel_ld(v))
        for t in (rmean, rvar):
            torch.autograd.graph.increment_version(t)
        y = ll.conv_gemm([v], packed_weight(w2, [hid], 1, False), c, residual=x)
        ctx.save_for_backward(x, t1, u, v, stats, wp, wcat, w2)
        ctx.cfg = (cp, cpp)
        ctx.params = (wpc, w1, gamma, beta)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, t1, u, v, stats, wp, wcat, w2 = ctx.saved_tensors
        cp, cpp = ctx.cfg
        wpc, w1, gamma, beta = ctx.params
        dy = to_act(dy)
        n, c, h, w = x.shape
        m = n * h * w
        hid = wcat.shape[0]
        dev = x.device
        # mlp.3: dv, dW2
        dv = ll.conv_gemm([dy], packed_weight_dgrad(w2, c, hid, 1, False, 0, hid), hid)
        d_w2 = _wgrad_oihw([v], dy, w2, [hid], 1, 1, 0, False)
        # BatchNorm + ReLU
        d_beta, d_gamma = _bn_grad_buffers(gamma, beta, hid, dev)
        parts = torch.empty(ll.reduce_blocks(m, hid) * 2 * hid, dtype=torch.float32, device=dev)
        du = torch.empty_like(u, memory_format=CL)
        ll.call("ly_bn_act_bwd", dv, ll.pixel_ld(dv), u, ll.pixel_ld(u), m, hid, stats[2], stats[3], stats[0], stats[1],
                ll.ACT["relu"], 1, d_beta, d_gamma, parts, zeros_f32(2, dev), du, ll.pixel_ld(du))
        # mlp.0: gradient of the x source lands on top of the residual gradient; t1 source separately
        dx = ll.conv_gemm([du], packed_weight_dgrad(wcat, hid, cpp + c, 1, False, cpp, c), c, residual=dy)
        dt1 = ll.conv_gemm([du], packed_weight_dgrad(wcat, hid, cpp + c, 1, False, 0, cpp), cpp)
        d_wcat = _wgrad_oihw([t1, x], du, wcat, [cpp, c], 1, 1, 0, False)
        # W1 = [columns 0..cp of the t1 half | columns cp.. of the x half]
        d_w1 = grad_sink(w1)
        if d_w1 is None:
            d_w1 = torch.empty(w1.shape, dtype=d_wcat.dtype, device=dev)
        torch.cat([d_wcat[:, :cp], d_wcat[:, cpp + cp:]], 1, out=d_w1)
        # partial conv: accumulate into the first cpp channels of dx in place
        dxs = dx[:, :cpp]
        ll.conv_gemm([dt1], packed_weight_dgrad(wp, cpp, cpp, 3, False, 0, cpp), cpp, k=3, stride=1, pad=1, residual=dxs,
                     out=dxs)
        d_wp = _wgrad_oihw([x[:, :cpp]], dt1, wp, [cpp], 3, 1, 1, False)
        d_wpc = grad_sink(wpc)
        if d_wpc is None:
            d_wpc = d_wp[:cp, :cp].contiguous()
        else:
            d_wpc.copy_(d_wp[:cp, :cp])
        return dx, d_wpc, d_w1, d_w2, d_gamma, d_beta, None, None, None, None, None, None


def _wgrad_oihw(srcs, dz, weight, chans, k, s, pad, patch):
    """fp32 OIHW weight gradient of conv(cat(srcs), weight) given dz (the tail of ConvFn.backward); lands in the
    gradient sink of `weight` when there is one."""
    cout = weight.shape[0]
    cpad = (cout + 7) // 8 * 8
    sink = grad_sink(weight)
    if len(chans) == 1 and k == 1 and chans[0] % 64 == 0 and cpad == cout:
        # packed == OIHW: the kernel accumulates straight into the (zeroed) sink, nothing to unpack
        dwp = ll.conv_wgrad(srcs, dz, cpad, k=k, stride=s, pad=pad, zeros=zeros_f32, out=sink)
        return dwp.view(weight.shape) if sink is not None else dwp.view(weight.shape).to(weight.dtype)
    dwp = ll.conv_wgrad(srcs, dz, cpad, k=k, stride=s, pad=pad, zeros=zeros_f32)
    d_w = sink if sink is not None else torch.empty(weight.shape, dtype=torch.float32, device=dz.device)
    c4 = list(chans) + [0] * (4 - len(chans))
    ll.call("ly_unpack_wgrad", dwp, cout, k, k, 1 if patch else 0, c4[0], c4[1], c4[2], c4[3], d_w, 0)
    return d_w if sink is not None else d_w.to(weight.dtype)


RF_GATE_PARTS = 16   # LY_RF_GATE_PARTS of include/leadyolo.h


class RfGateAttFn(torch.autograd.Function):
    """A = G * se[n,c] * sigmoid(conv3x3([max_c G; mean_c G])) — rfa.py:124-128 on the unfolded grid."""

    @staticmethod
    def forward(ctx, g, se, wa, k):
        g, se = to_act(g), to_act(se)
        n, kkc, ho, wo = g.shape
        c = kkc // (k * k)
        dev = g.device
        maps = torch.empty((n, ho * k, wo * k, 2), dtype=torch.float32, device=dev)
        amax = torch.empty((n, ho * k, wo * k), dtype=torch.int16, device=dev)
        att = torch.empty((n, ho * k, wo * k), dtype=torch.float32, device=dev)
        wa32 = wa.detach().float().reshape(18).contiguous()
        ll.call("ly_rf_attention", g, n, ho, wo, c, k, wa32, maps, amax, att)
        a = torch.empty_like(g, memory_format=CL)
        ll.call("ly_rf_gate", g, n, ho, wo, c, k, se, ll.pixel_ld(se), att, a)
        ctx.save_for_backward(g, se, wa32, maps, amax, att)
        ctx.cfg = (k, c, wa.dtype)
        return a

    @staticmethod
    def backward(ctx, da):
        g, se, wa32, maps, amax, att = ctx.saved_tensors
        k, c, wdt = ctx.cfg
        da = to_act(da).contiguous(memory_format=CL)
        n, kkc, ho, wo = g.shape
        dev = g.device
        dg = torch.empty_like(g, memory_format=CL)
        dse_parts = torch.empty((RF_GATE_PARTS, n, c), dtype=torch.float32, device=dev)
        datt = torch.empty_like(att)
        ll.call("ly_rf_gate_bwd", da, g, n, ho, wo, c, k, se, ll.pixel_ld(se), att, dg, dse_parts, datt)
        dse = dse_parts.sum(0)                                  # fixed order: deterministic
        dmaps = torch.empty_like(maps)
        dwa = zeros_f32(18, dev)
        ll.call("ly_rf_attention_bwd", datt, att, maps, amax, n, ho, wo, c, k, wa32, dmaps, dwa, dg)
        return dg, dse.to(BF16).view(n, c, 1, 1), dwa.view(1, 2, 3, 3).to(wdt), None


def _rf_tc_weights(wg, c, gscale, gshift):
    """Diagonal generate weights [C/16][tap][t][16] (BatchNorm scale folded in) and shift [C/16][t][16] as (bf16 hi | bf16 lo << 16) words for ly_rf_tc_*."""
    wgk = rf_pack_generate_weight(wg, c, 3) * gscale.view(9, 1, c)          # [t][tap][c]
    wc = wgk.view(9, 9, c // 16, 16).permute(2, 1, 0, 3).reshape(c // 16, 1296)
    sh = gshift.float().view(9, c // 16, 16).permute(1, 0, 2).reshape(c // 16, 144)
    hi = sh.to(BF16)
    lo = (sh - hi.float()).to(BF16)
    packed = (hi.view(torch.int16).to(torch.int32) & 0xFFFF) | (lo.view(torch.int16).to(torch.int32) << 16)
    return wc.to(BF16).contiguous(), packed.contiguous()


def rf_tc_layout(k, c):
    """True when rfcbam_front_infer takes the tensor-core path, whose A columns are channel-group-major:
    column (ch // 16) * 144 + t * 16 + ch % 16 instead of t * C + ch."""
    return k == 3 and c % 16 == 0


def rf_conv_weight_matrix(w, k, c, tc_layout):
    """RFCBAMConv.conv weight [Cout, C, k, k] -> [Cout, k*k*C, 1, 1] in the column order of the gated operand A."""
    cout = w.shape[0]
    m = w.detach().float().permute(0, 2, 3, 1).reshape(cout, k * k, c)            # [cout][t][ch]
    if tc_layout:
        m = m.reshape(cout, k * k, c // 16, 16).permute(0, 2, 1, 3)
    return m.reshape(cout, k * k * c, 1, 1)


def rfcbam_front_infer(x, wg, k, s, gscale, gshift, w1, w2, wa):
    """Inference: x -> gated GEMM operand A (N, k*k*C, ho, wo) in two passes over x (ly_rfcbam_front_infer, or the
    tensor-core pair ly_rf_tc_stats / ly_rf_tc_apply when rf_tc_layout(k, C))."""
    x = to_act(x)
    n, c, h, w = x.shape
    ho, wo = rf_out_size(h, w, k, s)
    dev = x.device
    if rf_tc_layout(k, c):
        # feature generation on the tensor cores, twice: channel statistics first, then the gated GEMM operand
        wc, sh = cached([wg, gscale, gshift], ("rftc", c), lambda: _rf_tc_weights(wg, c, gscale, gshift))
        w1f = cached([w1], "sew1", lambda: w1.detach().float().contiguous())
        w2f = cached([w2], "sew2", lambda: w2.detach().float().contiguous())
        wa32 = cached([wa], "wa32", lambda: wa.detach().float().reshape(18).contiguous())
        nm = n * ho * 3 * wo * 3
        nmp = (nm + 3) // 4 * 4
        ws = torch.empty(3 * nmp + 2 * n * c, dtype=torch.float32, device=dev)
        maps, att = ws[:2 * nm], ws[2 * nmp:2 * nmp + nm]
        gap, se = ws[3 * nmp:3 * nmp + n * c], ws[3 * nmp + n * c:]
        a = ll.empty_cl(n, 9 * c, ho, wo, device=dev)
        ld = ll.pixel_ld(x)
        ll.call("ly_rf_gap_sum", x, ld, n, h * w, c, gap)
        ll.call("ly_rf_tc_stats", x, ld, n, h, w, c, s, ho, wo, wc, sh, maps)
        ll.call("ly_rf_se_att", gap, n, c, h * w, w1f, w2f, int(w1.shape[0]), maps, ho * 3, wo * 3, wa32, se, att)
        ll.call("ly_rf_tc_apply", x, ld, n, h, w, c, s, ho, wo, wc, sh, se, att, a)
        return a
    wgk = cached([wg], ("wgk", c, k), lambda: rf_pack_generate_weight(wg, c, k))
    w1f = cached([w1], "sew1", lambda: w1.detach().float().contiguous())
    w2f = cached([w2], "sew2", lambda: w2.detach().float().contiguous())
    wa32 = cached([wa], "wa32", lambda: wa.detach().float().reshape(18).contiguous())
    nm = n * ho * k * wo * k
    nmp = (nm + 3) // 4 * 4          # keep every segment 16-byte aligned
    ws = torch.empty(3 * nmp + 2 * n * c, dtype=torch.float32, device=dev)
    maps, att = ws[:2 * nm], ws[2 * nmp:2 * nmp + nm]
    gap, se = ws[3 * nmp:3 * nmp + n * c], ws[3 * nmp + n * c:]
    a = ll.empty_cl(n, k * k * c, ho, wo, device=dev)
    ll.call("ly_rfcbam_front_infer", x, ll.pixel_ld(x), n, h, w, c, k, s, ho, wo, wgk, gscale, gshift, w1f, w2f,
            int(w1.shape[0]), wa32, maps, gap, se, att, a)
    return a


def rfcbam_fused_ok(k, c, cout):
    """True when the whole k = 3 module runs as two tensor-core passes with the final convolution fused in
    (ly_rfcbam_fused_infer): the gated receptive-field matrix never reaches HBM."""
    return rf_tc_layout(k, c) and cout in (128, 256)


def rfcbam_infer_fused(x, wg, s, gscale, gshift, w1, w2, wa, conv_weight, scale2, shift2):
    """Inference, k = 3: x -> RFCBAMConv output (N, Cout, ho, wo).  Pass 1 (ly_rf_tc_stats) gives the channel max / mean
    maps, ly_rf_se_att the SE gate and the spatial attention, pass 2 (ly_rfcbam_fused_infer) regenerates the features,
    gates them and multiplies by the stride-k convolution's weights inside the same kernel, with BatchNorm + ReLU."""
    from .packing import rf_fused_weight_image
    x = to_act(x)
    n, c, h, w = x.shape
    ho, wo = rf_out_size(h, w, 3, s)
    dev = x.device
    cout = conv_weight.shape[0]
    wc, sh = cached([wg, gscale, gshift], ("rftc", c), lambda: _rf_tc_weights(wg, c, gscale, gshift))
    w1f = cached([w1], "sew1", lambda: w1.detach().float().contiguous())
    w2f = cached([w2], "sew2", lambda: w2.detach().float().contiguous())
    wa32 = cached([wa], "wa32", lambda: wa.detach().float().reshape(18).contiguous())
    w2img = cached([conv_weight], ("rffw", c), lambda: rf_fused_weight_image(conv_weight, c))
    nm = n * ho * 3 * wo * 3
    nmp = (nm + 3) // 4 * 4
    ws = torch.empty(3 * nmp + 2 * n * c, dtype=torch.float32, device=dev)
    maps, att = ws[:2 * nm], ws[2 * nmp:2 * nmp + nm]
    gap, se = ws[3 * nmp:3 * nmp + n * c], ws[3 * nmp + n * c:]
    y = ll.empty_cl(n, cout, ho, wo, device=dev)
    ld = ll.pixel_ld(x)
    ll.call("ly_rf_gap_sum", x, ld, n, h * w, c, gap)
    ll.call("ly_rf_tc_stats", x, ld, n, h, w, c, s, ho, wo, wc, sh, maps)
    ll.call("ly_rf_se_att", gap, n, c, h * w, w1f, w2f, int(w1.shape[0]), maps, ho * 3, wo * 3, wa32, se, att)
    ll.call("ly_rfcbam_fused_infer", x, ld, n, h, w, c, s, ho, wo, wc, sh, se, att, w2img, scale2, shift2, cout, y, cout)
    return y


def rf_gate_att_infer(g, se, wa, k):
    n, kkc, ho, wo = g.shape
    c = kkc // (k * k)
    dev = g.device
    maps = torch.empty((n, ho * k, wo * k, 2), dtype=torch.float32, device=dev)
    att = torch.empty((n, ho * k, wo * k), dtype=torch.float32, device=dev)
    wa32 = cached([wa], "wa32", lambda: wa.detach().float().reshape(18).contiguous())
    ll.call("ly_rf_attention", g, n, ho, wo, c, k, wa32, maps, None, att)
    a = torch.empty_like(g, memory_format=CL)
    ll.call("ly_rf_gate", g, n, ho, wo, c, k, se, ll.pixel_ld(se), att, a)
    return a
