"""CPU: the oracle's loss / NMS restatement (oracle/restate.py) against the fixtures generated from the REAL
reference (utils/loss.py ComputeLoss, utils/general.py non_max_suppression + torchvision.ops.nms) by
oracle/make_golden_post.py."""
import os

import pytest
import torch

from oracle import restate as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = torch.load(os.path.join(ROOT, "tests", "golden", "loss_nms.pt"))


@pytest.mark.parametrize("name", sorted(GOLD["loss"]))
def test_compute_loss_restatement_matches_reference(name):
    c = GOLD["loss"][name]
    p = [t.clone().requires_grad_(True) for t in c["p"]]
    loss, items = R.compute_loss(p, c["targets"], c["anchors"], c["hyp"], c["nc"])
    loss.backward()
    assert torch.allclose(loss, c["loss"], rtol=1e-6, atol=1e-7)
    assert torch.allclose(items, c["items"], rtol=1e-6, atol=1e-7)
    for a, b in zip(p, c["dp"]):
        assert torch.allclose(a.grad, b, rtol=1e-5, atol=1e-8)


def test_compute_loss_ignores_padding_rows():
    c = GOLD["loss"]["nc1_128"]
    pad = torch.full((5, 6), -1.0)
    t = torch.cat([c["targets"][:3], pad, c["targets"][3:]])
    loss, _ = R.compute_loss(c["p"], t, c["anchors"], c["hyp"], c["nc"])
    assert torch.allclose(loss, c["loss"], rtol=1e-6)


@pytest.mark.parametrize("name", sorted(GOLD["nms"]))
def test_nms_restatement_matches_reference(name):
    c = GOLD["nms"][name]
    out = R.non_max_suppression(c["pred"], 0.25, 0.45)
    for a, b in zip(out, c["out"]):
        assert a.shape == b.shape
        assert torch.equal(a, b)
