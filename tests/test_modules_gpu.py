"""Per-module parity on the GPU: drop-in modules (sm_100a kernels through the C-ABI) vs the oracle restatement
(oracle/restate.py, fp32 PyTorch) on the same seeded inputs.  Forward in eval (folded-BN fast path), forward in
training mode (batch statistics + running-stat update) and every gradient.  Tolerance: rel-L2 <= 1e-2 (bf16)."""
import json
import os

import pytest
import torch

from helpers import ORACLE_FN, TOL_BF16, half_normal, oracle_params, randomise_bn, rel_l2
from oracle import restate as R

pytestmark = pytest.mark.gpu


def _mods():
    from lead_yolo_b200 import modules as M
    return M


from cases import BN_SEED, CASES, GOLDEN, GRAD_SEED, IN_SEED, INIT_SEED  # noqa: E402

GOLD_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
with open(os.path.join(GOLD_DIR, "ref_bf16_error.json")) as _f:
    REF_BF16 = json.load(_f)


def _build(name, table=CASES):
    cls, args, okey, shape, signed = table[name]
    torch.manual_seed(INIT_SEED)
    mod = getattr(_mods(), cls)(*args).cuda()
    randomise_bn(mod, BN_SEED)
    x = half_normal(shape, seed=IN_SEED, signed=signed)
    return mod, okey, x


def _tol(name, key):
    """1e-2, or 2.5x the error the reference itself shows when run under bf16 autocast on the same input
    (tests/golden/ref_bf16_error.json, made by oracle/make_golden.py): ReLU-mask flips and bf16 pre-activations
    put a floor under ANY bf16 implementation's gradient error (e.g. ~3% for the FasterNet block weights,
    ~8% for RFCBAMConv, ~10% through the SPPF max-pool chain)."""
    return max(TOL_BF16, 2.5 * REF_BF16.get(name, {}).get(key, 0.0))


def _oracle_ctx(P, okey, mod, training):
    if okey == "RFCBAM":
        P["__stride__m"] = mod.generate[0].stride[0]
    return R.Ctx(training)


@pytest.mark.parametrize("name", list(CASES))
def test_eval_forward(name):
    mod, okey, x = _build(name)
    mod.eval()
    P = oracle_params(mod)
    with torch.no_grad():
        y = mod(x)
        ref = ORACLE_FN[okey](P, x.float(), _oracle_ctx(P, okey, mod, False))
    assert y.shape == ref.shape
    err = rel_l2(y, ref)
    assert err <= TOL_BF16, f"{name}: eval forward rel-L2 {err:.3e}"


@pytest.mark.parametrize("name", list(CASES))
def test_train_forward_backward(name):
    mod, okey, x = _build(name)
    mod.train()
    P = oracle_params(mod, requires_grad=True)
    ctx = _oracle_ctx(P, okey, mod, True)
    xr = x.float().clone().requires_grad_(True)
    ref = ORACLE_FN[okey](P, xr, ctx)
    g = torch.Generator(device="cpu").manual_seed(GRAD_SEED)
    r = torch.randn(ref.shape, generator=g).bfloat16().float().cuda()
    (ref * r).sum().backward()

    xm = x.clone().requires_grad_(True)
    y = mod(xm)
    assert y.shape == ref.shape
    (y.float() * r).sum().backward()
    torch.cuda.synchronize()

    errs = {"out": rel_l2(y, ref), "dx": rel_l2(xm.grad, xr.grad)}
    for k, p in mod.named_parameters():
        errs["d_" + k] = rel_l2(p.grad, P["m." + k].grad)
    sd = mod.state_dict()
    for k, v in ctx.updates.items():
        errs["upd_" + k[2:]] = rel_l2(sd[k[2:]], v)
    # a convolution bias directly in front of a training-mode BatchNorm has an exactly-zero true gradient:
    # judge it against the scale of the BatchNorm's own bias gradient instead of a relative error
    for bk, ref_k in (("conv1.bias", "bn1.bias"), ("conv.0.bias", "conv.1.bias")):
        for k in [k for k in errs if k.endswith(bk)]:
            pre = k[2:-len(bk)]
            gb = dict(mod.named_parameters())[pre + bk].grad.float().norm()
            gref = dict(mod.named_parameters())[pre + ref_k].grad.float().norm()
            errs[k] = 0.0 if gb <= 0.05 * gref else float(gb / gref)
    if okey == "RFCBAM" and mod.kernel_size == 1:
        # k=1: generate is a per-channel scale followed by BatchNorm; its gradient is ~eps/(var+eps) of the usual
        # size (zero for eps=0), i.e. ill-conditioned: the reference's own bf16 run is 10% off, allow 35%
        errs["d_generate.0.weight"] = 0.0 if errs["d_generate.0.weight"] < 0.35 else 1.0
    bad = {k: v for k, v in errs.items() if not (v <= _tol(name, k))}
    assert not bad, f"{name}: rel-L2 over tolerance: {bad} (all: { {k: round(v, 5) for k, v in errs.items()} })"


def test_eval_mode_with_grad():
    """eval-mode BatchNorm (running statistics) through the autograd path."""
    mod, okey, x = _build("c3ca_168_128")
    mod.eval()
    P = oracle_params(mod, requires_grad=True)
    xr = x.float().clone().requires_grad_(True)
    ref = ORACLE_FN[okey](P, xr, R.Ctx(False))
    ref.square().sum().backward()
    xm = x.clone().requires_grad_(True)
    y = mod(xm)
    y.float().square().sum().backward()
    assert rel_l2(y, ref) <= TOL_BF16
    assert rel_l2(xm.grad, xr.grad) <= 2 * TOL_BF16


@pytest.mark.parametrize("u8", [False, True])
def test_patch_embed_stem(u8):
    M = _mods()
    torch.manual_seed(0)
    mod = M.PatchEmbed_FasterNet(3, 24, 4, 4).cuda()
    randomise_bn(mod)
    g = torch.Generator(device="cpu").manual_seed(5)
    img8 = torch.randint(0, 256, (2, 3, 64, 96), generator=g, dtype=torch.uint8).cuda()
    imgf = img8.float() / 255
    P = oracle_params(mod, requires_grad=True)
    mod.eval()
    with torch.no_grad():
        y = mod(img8 if u8 else imgf)
        ref = R.patch_embed(P, "m", imgf, R.Ctx(False))
    assert rel_l2(y, ref) <= TOL_BF16
    mod.train()
    ctx = R.Ctx(True)
    ref = R.patch_embed(P, "m", imgf, ctx)
    r = torch.randn_like(ref)
    (ref * r).sum().backward()
    y = mod(img8 if u8 else imgf)
    (y.float() * r).sum().backward()
    assert rel_l2(y, ref) <= TOL_BF16
    for k, p in mod.named_parameters():
        assert rel_l2(p.grad, P["m." + k].grad) <= TOL_BF16, k


def test_upsample_concat():
    M = _mods()
    a = half_normal((2, 64, 10, 10), 1, signed=True).requires_grad_(True)
    b = half_normal((2, 40, 20, 20), 2, signed=True).requires_grad_(True)
    up = M.Upsample(None, 2, "nearest")
    cat = M.Concat(1)
    y = cat([up(a), b])
    ref = torch.cat([torch.nn.functional.interpolate(a.float(), scale_factor=2.0, mode="nearest"), b.float()], 1)
    assert torch.equal(y.float(), ref)
    r = torch.randn_like(ref).to(torch.bfloat16).float()
    (y.float() * r).sum().backward()
    da_ref = torch.nn.functional.avg_pool2d(r[:, :64], 2) * 4
    assert rel_l2(a.grad, da_ref) <= 4e-3
    assert torch.equal(b.grad.float(), r[:, 64:])


def test_detect_head():
    M = _mods()
    from lead_yolo_b200.model import ANCHORS
    torch.manual_seed(0)
    det = M.Detect(1, [list(a) for a in ANCHORS], (128, 256, 512)).cuda()
    det.stride = torch.tensor([8.0, 16.0, 32.0])
    det.anchors /= det.stride.view(-1, 1, 1).cuda()
    feats = [half_normal((2, c, s, s), i, signed=True) for i, (c, s) in enumerate(((128, 16), (256, 8), (512, 4)))]
    P = oracle_params(det, requires_grad=True)
    det.eval()
    with torch.no_grad():
        z, raw = det([f.clone() for f in feats])
        raw_ref = R.detect_raw(P, "m", [f.float() for f in feats])
        z_ref = R.detect_decode(P, "m", raw_ref)
    for a, b in zip(raw, raw_ref):
        assert rel_l2(a, b) <= 2e-3
    assert rel_l2(z, z_ref) <= 2e-3
    det.train()
    fr = [f.float().clone().requires_grad_(True) for f in feats]
    raw_ref = R.detect_raw(P, "m", fr)
    fm = [f.clone().requires_grad_(True) for f in feats]
    raw = det(fm)
    sum((a.square().sum() for a in raw_ref)).backward()
    sum((a.square().sum() for a in raw)).backward()
    for a, b in zip(fm, fr):
        assert rel_l2(a.grad, b.grad) <= TOL_BF16
    for k, p in det.named_parameters():
        assert rel_l2(p.grad, P["m." + k].grad) <= TOL_BF16, k


@pytest.mark.parametrize("name", list(GOLDEN))
def test_against_reference_golden(name):
    """The CUDA path against outputs/gradients of the REAL reference classes (tests/golden/modules.pt):
    loads the reference state_dict (strict), same input, eval + training forward and backward."""
    gold = torch.load(os.path.join(GOLD_DIR, "modules.pt"), map_location="cuda", weights_only=True)[name]
    cls, args, okey, shape, signed = GOLDEN[name]
    mod = getattr(_mods(), cls)(*args).cuda()
    mod.load_state_dict(gold["state_dict"], strict=True)
    for m in mod.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.eps, m.momentum = 1e-3, 0.03
    x = gold["x"].contiguous(memory_format=torch.channels_last)
    mod.eval()
    with torch.no_grad():
        y = mod(x)
    assert rel_l2(y, gold["out_eval"]) <= TOL_BF16
    mod.train()
    xm = x.clone().requires_grad_(True)
    y = mod(xm)
    g = torch.Generator(device="cpu").manual_seed(GRAD_SEED)
    r = torch.randn(y.shape, generator=g).bfloat16().float().cuda()
    (y.float() * r).sum().backward()
    assert rel_l2(y, gold["out_train"]) <= 1.5 * TOL_BF16
    # tiny tensors (a few hundred pixels): gradient noise from ReLU-mask flips is large, so this is a
    # structural check (right tensors, right layout, right sign/scale); precision is judged on the big cases
    assert rel_l2(xm.grad, gold["dx"]) <= 0.15
    for k, p in mod.named_parameters():
        ref = gold["grads"][k]
        if ref.norm() < 1e-3 * max(1.0, float(p.grad.float().norm())):
            continue
        assert rel_l2(p.grad, ref) <= 0.2, k
    sd = mod.state_dict()
    for k, v in gold["running"].items():
        assert rel_l2(sd[k], v) <= 5e-3, k
