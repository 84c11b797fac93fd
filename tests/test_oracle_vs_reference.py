"""CPU tests that pin the oracle: oracle/restate.py against (a) the committed fixtures generated from the REAL
reference (tests/golden/, oracle/make_golden.py) and (b) the imported reference itself when /root/reference is present
(build container only)."""
import os

import pytest
import torch

from cases import GOLDEN, GRAD_SEED, rfcbam_stride
from helpers import ORACLE_FN
from helpers_cpu import rel_l2
from oracle import ref_import, restate as R

GOLD_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def gold():
    return torch.load(os.path.join(GOLD_DIR, "modules.pt"), map_location="cpu", weights_only=True)


@pytest.mark.parametrize("name", list(GOLDEN))
def test_restatement_matches_reference_fixture(gold, name):
    g = gold[name]
    cls, args, okey, shape, signed = GOLDEN[name]
    x = g["x"].float()
    P = {"m." + k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running" not in k else v.clone())
         for k, v in g["state_dict"].items()}
    if okey == "RFCBAM":
        P["__stride__m"] = rfcbam_stride(args)
    with torch.no_grad():
        y = ORACLE_FN[okey](P, x, R.Ctx(False))
    assert rel_l2(y, g["out_eval"]) < 1e-5
    ctx = R.Ctx(True)
    xr = x.clone().requires_grad_(True)
    y = ORACLE_FN[okey](P, xr, ctx)
    gen = torch.Generator().manual_seed(GRAD_SEED)
    r = torch.randn(y.shape, generator=gen).bfloat16().float()
    (y * r).sum().backward()
    assert rel_l2(y, g["out_train"]) < 1e-5
    assert rel_l2(xr.grad, g["dx"]) < 1e-4
    gmax = max(float(v.norm()) for v in g["grads"].values())
    for k, ref in g["grads"].items():
        got = P["m." + k].grad
        if ref.norm() < 1e-4 * gmax:  # analytically zero gradients (conv bias in front of a training-mode BN)
            assert got.norm() < 1e-3 * gmax
        else:
            assert rel_l2(got, ref) < 2e-4, k
    for k, ref in g["running"].items():
        assert rel_l2(ctx.updates["m." + k], ref) < 1e-5, k


def test_whole_model_restatement_matches_fixture():
    """decoded boxes / raw heads of the real reference on the 64x64 fixture image vs the restatement, using the
    same seed-0 initialisation (parameters are bit-identical by construction order) and BN randomisation."""
    from cases import BN_SEED, INIT_SEED
    from helpers_cpu import randomise_bn
    from lead_yolo_b200.model import LeadYolo
    g = torch.load(os.path.join(GOLD_DIR, "model_small.pt"), map_location="cpu", weights_only=True)
    torch.manual_seed(INIT_SEED)
    m = LeadYolo("lead", 1)
    randomise_bn(m, BN_SEED)
    P = {k: v.detach().float() for k, v in m.state_dict().items()}
    x = g["img"].float() / 255
    with torch.no_grad():
        dec, raw, _ = R.model_forward(P, x, R.Ctx(False))
        _, raw_tr, _ = R.model_forward(P, x, R.Ctx(True), decode=False)
    assert rel_l2(dec, g["decoded"]) < 1e-5
    for a, b in zip(raw, g["raw"]):
        assert rel_l2(a, b) < 1e-5
    for a, b in zip(raw_tr, g["raw_train"]):
        assert rel_l2(a, b) < 1e-4


@pytest.mark.skipif(not ref_import.available(), reason="reference tree not present (GPU box)")
def test_restatement_matches_live_reference():
    model = ref_import.build_model(seed=0)
    from helpers_cpu import randomise_bn
    randomise_bn(model, 1)
    g = torch.Generator().manual_seed(1)
    x = torch.rand(2, 3, 128, 128, generator=g)
    P0 = {k: v.clone() for k, v in model.state_dict().items()}
    model.eval()
    with torch.no_grad():
        yref = model(x)
        dec, raw, _ = R.model_forward(P0, x, R.Ctx(False))
    assert (dec - yref[0]).abs().max() < 5e-4
    for a, b in zip(raw, yref[1]):
        assert (a - b).abs().max() < 1e-4
    model.train()
    out_ref = model(x)
    ctx = R.Ctx(True)
    _, raw, _ = R.model_forward(P0, x, ctx, decode=False)
    for a, b in zip(raw, out_ref):
        assert (a - b).abs().max() < 1e-4
    sd = model.state_dict()
    assert max((ctx.updates[k] - sd[k]).abs().max().item() for k in ctx.updates) < 1e-5
    # ablation yaml (config 4): plain C3 instead of C3_CA
    m2 = ref_import.build_model("models/yolov5s-FasterNet+RFCBAMConv.yaml", seed=0)
    m2.eval()
    with torch.no_grad():
        y2 = m2(x)
        d2, _, _ = R.model_forward({k: v for k, v in m2.state_dict().items()}, x, R.Ctx(False), variant="c3")
    assert (d2 - y2[0]).abs().max() < 5e-4
