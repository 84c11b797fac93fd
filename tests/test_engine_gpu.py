"""InferenceEngine (the end-to-end serving path bench.py times): the pipelined submit / result protocol with slot reuse
must hand back, for every batch, exactly what a plain forward of the same model gives for that batch."""
import collections

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("slots", [2, 3])
def test_engine_pipelined_results_match_forward(slots):
    from lead_yolo_b200.engine import InferenceEngine
    from lead_yolo_b200.model import build_lead_yolo

    bs, img, nbatch = 4, 320, 7
    model = build_lead_yolo("lead", 1, "cuda", seed=3).eval()
    g = torch.Generator().manual_seed(5)
    host = [torch.randint(0, 256, (bs, 3, img, img), generator=g, dtype=torch.uint8).pin_memory() for _ in range(nbatch)]
    with torch.no_grad():
        want = [model(h.cuda())[0].float().cpu() for h in host]

    eng = InferenceEngine(model, bs, img, slots=slots, device="cuda")
    got = [None] * nbatch
    pending = collections.deque()
    for i in range(nbatch):
        pending.append((i, eng.submit(host[i])))
        if len(pending) >= eng.slots:            # consume the oldest before its slot is submitted into again
            j, s = pending.popleft()
            got[j] = eng.result(s).clone()
    while pending:
        j, s = pending.popleft()
        got[j] = eng.result(s).clone()
    eng.drain()
    for i in range(nbatch):
        assert got[i].shape == want[i].shape
        # same kernels on the same inputs: graph replay against eager launch is bit-identical
        assert torch.equal(got[i], want[i]), f"batch {i}: max diff {(got[i] - want[i]).abs().max().item()}"


def test_engine_call_is_submit_plus_result():
    from lead_yolo_b200.engine import InferenceEngine
    from lead_yolo_b200.model import build_lead_yolo

    model = build_lead_yolo("lead", 1, "cuda", seed=3).eval()
    x = torch.randint(0, 256, (2, 3, 320, 320), dtype=torch.uint8).pin_memory()
    eng = InferenceEngine(model, 2, 320, slots=2, device="cuda")
    a = eng(x).clone()
    b = eng(x).clone()      # other slot, same input
    with torch.no_grad():
        ref = model(x.cuda())[0].float().cpu()
    assert torch.equal(a, ref) and torch.equal(b, ref)
