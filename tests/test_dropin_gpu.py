"""GPU: the drop-in proven through the REFERENCE'S OWN code.  With this library's classes installed
in the reference's namespaces, `models/yolo.py` builds `LEAD-YOLO.yaml` on the B200 (its constructor runs a
training-mode 1x3x256x256 forward, yolo.py:284-289), `fuse()`, `_apply`, `attempt_load` of a checkpoint pickled from
the reference classes (train.py:387-403 format) and the training branch all run on the sm_100a kernels, and the
results match the oracle.  Needs the git-ignored copy of the reference that tools/install_ref.py makes in
baseline/_ref (it travels to the GPU box); skipped when absent."""
import os

import pytest
import torch

from helpers import rel_l2
from helpers_cpu import randomise_bn
from oracle import ref_import
from oracle import restate as R

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref_import.available(), reason="no reference copy (baseline/_ref)")]
DEV = "cuda"


@pytest.fixture(scope="module")
def ref():
    from lead_yolo_b200 import model as LM
    r = ref_import.load()
    saved = LM.install(r.common, r.rfa, r.yolo)
    yield r
    LM.uninstall(r.common, r.rfa, r.yolo, saved)


def _img(n, s, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (n, 3, s, s), generator=g, dtype=torch.uint8)


@pytest.mark.parametrize("cfg,csp", [("models/LEAD-YOLO.yaml", "C3_CA"), ("models/yolov5s-FasterNet+RFCBAMConv.yaml", "C3")])
def test_reference_model_builds_and_runs_on_gpu(ref, cfg, csp):
    from lead_yolo_b200 import modules as M
    from lead_yolo_b200.engine import InferenceEngine
    from lead_yolo_b200.loss import ComputeLoss
    with torch.device(DEV):
        torch.manual_seed(0)
        model = ref.yolo.Model(os.path.join(ref.root, cfg), ch=3, nc=1)       # reference code, our classes, on the GPU
    assert type(model.model[12]) is getattr(M, csp) and type(model.model[9]) is M.RFCBAMConv
    assert type(model.model[-1]) is M.Detect and model.stride.tolist() == [8.0, 16.0, 32.0]
    assert next(model.parameters()).is_cuda
    randomise_bn(model, 1)
    variant = "lead" if csp == "C3_CA" else "c3"
    x8 = _img(2, 128)
    x = (x8.float() / 255).to(DEV)
    P = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}
    ref_dec, _, _ = R.model_forward(P, x.cpu(), R.Ctx(False), variant=variant)
    model.eval()
    with torch.no_grad():
        dec, raw = model(x)
    assert rel_l2(dec.float().cpu(), ref_dec) < 3e-2
    # training branch through the reference's _forward_once + this library's ComputeLoss
    model.train()
    model.hyp = R.scaled_hyp(3, 1, 128)
    targets = torch.tensor([[0, 0, 0.5, 0.5, 0.2, 0.3], [1, 0, 0.3, 0.6, 0.1, 0.15]], device=DEV)
    loss, items = ComputeLoss(model)(model(x), targets)
    loss.backward()
    assert torch.isfinite(loss).all() and all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())
    # reference fuse() (yolo.py:213-233) + eval: same predictions
    model.zero_grad(set_to_none=True)
    P2 = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}   # BN buffers moved in the train forward
    ref_dec2, _, _ = R.model_forward(P2, x.cpu(), R.Ctx(False), variant=variant)
    model.eval().fuse()
    with torch.no_grad():
        dec2, _ = model(x)
    assert rel_l2(dec2.float().cpu(), ref_dec2) < 3e-2
    # the serving engine captures a reference-built model (Detect.stride lives on the GPU there)
    eng = InferenceEngine(model, 2, 128, slots=2, device=torch.device(DEV))
    out = eng(x8.pin_memory())
    assert rel_l2(out, ref_dec2) < 3e-2


def test_reference_saved_checkpoint_unpickles_and_runs(ref):
    import sys
    from lead_yolo_b200 import modules as M
    ckpt_path = os.path.join(ref.root, "lead_yolo_refsaved.pt")
    if not os.path.exists(ckpt_path):
        pytest.skip("no reference-saved checkpoint (tools/install_ref.py writes it)")
    if ref.root not in sys.path:
        sys.path.insert(0, ref.root)
    from models.experimental import attempt_load
    model = attempt_load(ckpt_path, device=torch.device(DEV))                 # un-pickle -> .float() -> fuse() -> eval()
    assert type(model.model[1].blocks[0]) is M.MLPBlock and type(model.model[17]) is M.RFCBAMConv
    assert not hasattr(model.model[8].cv1, "bn")                              # fused
    raw_ckpt = torch.load(ckpt_path, map_location="cpu", weights_only=False)["model"].float()
    P = {k: v.detach().float() for k, v in raw_ckpt.state_dict().items()}
    x8 = _img(2, 160, seed=4)
    x = x8.float() / 255
    ref_dec, _, _ = R.model_forward(P, x, R.Ctx(False))
    with torch.no_grad():
        dec, _ = model(x.to(DEV))
    # the checkpoint stores fp16 weights (train.py:391): the oracle runs on the same rounded values
    assert rel_l2(dec.float().cpu(), ref_dec) < 3e-2
