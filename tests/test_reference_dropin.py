"""CPU tests of the module-level drop-in contract: class names, constructor signatures, state_dict
layout, object lifecycle (deepcopy / pickle / half / fuse) and — when the reference tree is present — that the
reference's own parse_model() builds LEAD-YOLO.yaml out of the replacement classes."""
import copy
import inspect
import io
import os

import pytest
import torch
import torch.nn as nn

from oracle import ref_import

GOLD_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _model():
    from lead_yolo_b200.model import LeadYolo
    torch.manual_seed(0)
    return LeadYolo("lead", 1)


def test_layout_counts():
    m = _model()
    sd = m.state_dict()
    assert len(sd) == 319
    assert sum(p.numel() for p in m.parameters()) == 3_135_478
    assert sum(isinstance(x, nn.BatchNorm2d) for x in m.modules()) == 44
    assert all(x.eps == 1e-3 and x.momentum == 0.03 for x in m.modules() if type(x) is nn.BatchNorm2d)
    for k, shape in {"model.1.blocks.0.spatial_mixing.partial_conv3.weight": (6, 6, 3, 3),
                     "model.9.get_weight.0.weight": (1, 2, 3, 3), "model.9.se.fc.0.weight": (16, 160),
                     "model.9.se.fc.2.weight": (160, 16), "model.17.generate.0.weight": (1152, 1, 3, 3),
                     "model.17.conv.0.bias": (128,), "model.12.m.0.ca.conv1.weight": (8, 128, 1, 1),
                     "model.23.anchors": (3, 3, 2), "model.23.m.2.weight": (18, 512, 1, 1)}.items():
        assert tuple(sd[k].shape) == shape, k
    c3 = __import__("lead_yolo_b200.model", fromlist=["LeadYolo"]).LeadYolo("c3", 1)
    assert "model.12.m.0.ca.conv1.weight" not in c3.state_dict()


def test_golden_state_dicts_load_strictly():
    from cases import GOLDEN
    from lead_yolo_b200 import modules as M
    gold = torch.load(os.path.join(GOLD_DIR, "modules.pt"), map_location="cpu", weights_only=True)
    for name, (cls, args, *_rest) in GOLDEN.items():
        mod = getattr(M, cls)(*args)
        mod.load_state_dict(gold[name]["state_dict"], strict=True)


def test_lifecycle_deepcopy_pickle_half_fuse():
    m = _model()
    m2 = copy.deepcopy(m)
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), m2.state_dict().values()))
    buf = io.BytesIO()
    torch.save(m, buf)                                    # whole-module pickle, as train.py:387-403
    buf.seek(0)
    m3 = torch.load(buf, weights_only=False)
    blk = m3.model[1].blocks[0]
    assert type(blk).__name__ == "MLPBlock" and blk.forward.__self__ is blk
    assert blk.spatial_mixing.forward.__name__ == "forward_split_cat"
    m3.half().float()
    n_before = len(m3.state_dict())
    m3.eval().fuse()
    assert len(m3.state_dict()) < n_before
    assert not hasattr(m3.model[8].cv1, "bn") and not hasattr(m3.model[0], "norm")
    assert sum(isinstance(x, nn.BatchNorm2d) for x in m3.modules()) == 18    # 18 BNs survive fusion
    assert sum(p.numel() for p in m3.parameters()) == 3_131_478              # authors' fused count (val.txt:6)


@pytest.mark.skipif(not ref_import.available(), reason="reference tree not present (GPU box)")
def test_signatures_and_state_dict_match_reference():
    ref = ref_import.load()
    from lead_yolo_b200 import modules as M
    from lead_yolo_b200.model import DROP_IN, DROP_IN_RFA
    for name in DROP_IN + DROP_IN_RFA:
        rcls = getattr(ref.rfa if name in DROP_IN_RFA else ref.common, name)
        ours = getattr(M, name)
        assert list(inspect.signature(ours.__init__).parameters) == list(inspect.signature(rcls.__init__).parameters), name
    assert list(inspect.signature(M.Detect.__init__).parameters) == \
        list(inspect.signature(ref.yolo.Detect.__init__).parameters)
    rm = ref_import.build_model(seed=0)
    mine = _model()
    a, b = rm.state_dict(), mine.state_dict()
    assert list(a) == list(b)
    for k in a:
        assert a[k].shape == b[k].shape, k
        if "running" not in k and "num_batches" not in k:   # the reference's build-time forward touches BN buffers
            assert torch.equal(a[k].float(), b[k].float()), k
    mine.load_state_dict(a, strict=True)
    rm.load_state_dict(b, strict=True)


@pytest.mark.skipif(not ref_import.available(), reason="reference tree not present (GPU box)")
def test_reference_parse_model_builds_from_replacement_classes():
    import yaml
    ref = ref_import.load()
    from lead_yolo_b200 import model as LM
    from lead_yolo_b200 import modules as M
    saved = LM.install(ref.common, ref.rfa, ref.yolo)
    try:
        for cfg, csp in (("models/LEAD-YOLO.yaml", M.C3_CA), ("models/yolov5s-FasterNet+RFCBAMConv.yaml", M.C3)):
            with open(os.path.join(ref.root, cfg)) as f:
                d = yaml.safe_load(f)
            torch.manual_seed(0)
            seq, save = ref.yolo.parse_model(copy.deepcopy(d), ch=[3])      # reference code, our classes
            assert type(seq[0]) is M.PatchEmbed_FasterNet and type(seq[9]) is M.RFCBAMConv
            assert type(seq[12]) is csp and type(seq[23]) is M.Detect
            assert save == [3, 5, 9, 13, 16, 19, 22]
            mine = LM.LeadYolo("lead" if csp is M.C3_CA else "c3", 1)
            assert [k for k in seq.state_dict()] == [k[len("model."):] for k in mine.state_dict()]
    finally:
        LM.uninstall(ref.common, ref.rfa, ref.yolo, saved)
    assert ref.common.Conv is not M.Conv
