"""TEST INFRASTRUCTURE ONLY — generates the committed fixtures under tests/golden/ from the REAL reference
(imported unmodified from /root/reference through oracle/ref_import.py).  Run in the build container:

    python oracle/make_golden.py

Outputs
  tests/golden/modules.pt        per-module state_dict, input, eval output, train output, input/parameter
                                 gradients and BatchNorm running-stat updates of the reference classes (fp32 math)
  tests/golden/model_small.pt    whole LEAD-YOLO (seed 0 init, randomised BN) on a 1x3x64x64 input: decoded
                                 boxes + raw heads (eval) and raw heads of a training-mode forward
  tests/golden/ref_bf16_error.json
                                 rel-L2 of the reference's OWN bf16-autocast run against its fp32 run, per case
                                 and per tensor (output, dx, every parameter gradient) — the precision floor of
                                 any bf16 implementation (ReLU-mask flips, bf16 pre-activations), used as the
                                 gradient tolerance reference by tests/test_modules_gpu.py
"""
from __future__ import annotations

import copy
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from cases import BN_SEED, CASES, GOLDEN, GRAD_SEED, IN_SEED, INIT_SEED  # noqa: E402
from helpers_cpu import make_input, randomise_bn, rel_l2  # noqa: E402
from oracle import ref_import  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def ref_class(ref, name):
    if name in ("SE", "RFCBAMConv"):
        return getattr(ref.rfa, name)
    return getattr(ref.common, name)


def build(ref, cls, args):
    torch.manual_seed(INIT_SEED)
    m = ref_class(ref, cls)(*args)
    randomise_bn(m, BN_SEED)
    for sub in m.modules():  # in-place activations would corrupt saved inputs of this harness
        if hasattr(sub, "inplace"):
            sub.inplace = False
    return m


def run_train(m, x, autocast=False):
    m = copy.deepcopy(m).train()
    xx = x.clone().requires_grad_(True)
    if autocast:
        with torch.autocast("cpu", dtype=torch.bfloat16):
            y = m(xx)
    else:
        y = m(xx)
    g = torch.Generator().manual_seed(GRAD_SEED)
    r = torch.randn(y.shape, generator=g).bfloat16().float()
    (y.float() * r).sum().backward()
    grads = {k: p.grad.detach().clone() for k, p in m.named_parameters()}
    return y.detach().float(), xx.grad.detach().clone(), grads, m.state_dict()


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = ref_import.load()
    torch.set_grad_enabled(True)

    # ---- reference bf16-autocast vs fp32 error per case
    errs = {}
    for name, (cls, args, okey, shape, signed) in CASES.items():
        m = build(ref, cls, args)
        x = make_input(shape, IN_SEED, signed)
        y32, dx32, g32, _ = run_train(m, x, False)
        y16, dx16, g16, _ = run_train(m, x, True)
        e = {"out": rel_l2(y16, y32), "dx": rel_l2(dx16, dx32)}
        for k in g32:
            e["d_" + k] = rel_l2(g16[k], g32[k])
        errs[name] = e
        print(name, {k: round(v, 4) for k, v in e.items()}, flush=True)
    with open(os.path.join(OUT, "ref_bf16_error.json"), "w") as f:
        json.dump(errs, f, indent=1, sort_keys=True)

    # ---- small per-module goldens
    gold = {}
    for name, (cls, args, okey, shape, signed) in GOLDEN.items():
        m = build(ref, cls, args)
        x = make_input(shape, IN_SEED, signed)
        sd0 = {k: v.clone() for k, v in m.state_dict().items()}
        m.eval()
        with torch.no_grad():
            y_eval = m(x).float()
        y_tr, dx, grads, sd1 = run_train(m, x, False)
        upd = {k: v.clone() for k, v in sd1.items() if "running_" in k}
        gold[name] = {"state_dict": sd0, "x": x.bfloat16(), "out_eval": y_eval, "out_train": y_tr, "dx": dx,
                      "grads": grads, "running": upd}
        print("golden", name, tuple(y_eval.shape), flush=True)
    torch.save(gold, os.path.join(OUT, "modules.pt"))

    # ---- whole model, small input
    model = ref_import.build_model(seed=INIT_SEED)
    randomise_bn(model, BN_SEED)
    g = torch.Generator().manual_seed(IN_SEED)
    img = torch.randint(0, 256, (1, 3, 64, 64), generator=g, dtype=torch.uint8)
    x = img.float() / 255
    model.eval()
    with torch.no_grad():
        dec, raw = model(x)
    m2 = copy.deepcopy(model).train()
    with torch.no_grad():
        raw_tr = m2(x)
    torch.save({"img": img, "decoded": dec.float(), "raw": [r.float() for r in raw],
                "raw_train": [r.float() for r in raw_tr],
                "n_params": sum(p.numel() for p in model.parameters())},
               os.path.join(OUT, "model_small.pt"))
    print("model_small", tuple(dec.shape))


if __name__ == "__main__":
    main()
