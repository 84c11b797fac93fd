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


def test_engine_nms_mode_matches_postprocess_of_dense_output():
    """nms=(conf, iou): the captured graph's packed detections equal non_max_suppression of the dense output."""
    from lead_yolo_b200.engine import InferenceEngine
    from lead_yolo_b200.model import build_lead_yolo
    from lead_yolo_b200.postprocess import non_max_suppression

    model = build_lead_yolo("lead", 1, "cuda", seed=3).eval()
    with torch.no_grad():
        for mi in model.model[-1].m:
            mi.bias.view(3, -1)[:, 4] += 4.0              # enough objectness for a few hundred candidates
    x = torch.randint(0, 256, (2, 3, 320, 320), dtype=torch.uint8).pin_memory()
    with torch.no_grad():
        want = non_max_suppression(model(x.cuda())[0], 0.25, 0.45)
    eng = InferenceEngine(model, 2, 320, slots=2, device="cuda", nms=(0.25, 0.45))
    for _ in range(3):                                      # slot reuse
        got = InferenceEngine.unpack(eng(x))
        assert sum(w.shape[0] for w in want) > 0
        for a, b in zip(got, want):
            assert torch.equal(a, b.cpu())


def _train_pair(graph_a, graph_b, steps=3):
    from lead_yolo_b200.model import build_lead_yolo
    from lead_yolo_b200.train import TrainEngine
    g = torch.Generator().manual_seed(9)
    imgs = [torch.randint(0, 256, (4, 3, 128, 128), generator=g, dtype=torch.uint8).cuda() for _ in range(steps)]
    tg = torch.tensor([[i, 0, 0.3 + 0.1 * i, 0.5, 0.2, 0.25] for i in range(4)], device="cuda")
    outs = []
    for graph in (graph_a, graph_b):
        model = build_lead_yolo("lead", 1, "cuda", seed=1)
        eng = TrainEngine(model, 4, 128, graph=graph, device="cuda")
        start = eng.flat.clone()
        losses = [float(eng.step(im, tg)) for im in imgs]
        outs.append((losses, eng.flat - start, eng))
    return outs


def test_train_engine_graph_replay_equals_eager_step():
    (la, da, ea), (lb, db, _) = _train_pair(True, False)
    assert ea.graphs is not None and ea.launches_per_step > 100
    # the forward pass is bit-reproducible (fixed-point BatchNorm statistics, ordered pooling sums): the first loss is
    # identical, later ones differ only through the fp32 atomics of the weight-gradient kernels
    assert la[0] == lb[0], (la, lb)
    for a, b in zip(la, lb):
        assert abs(a - b) <= 1e-3 * abs(b), (la, lb)
    assert ((da - db).norm() / db.norm()).item() < 1e-2
    # a model whose weights moved through the engine must not serve stale packed weights at inference
    m = ea.model.eval()
    x = torch.randint(0, 256, (2, 3, 128, 128), dtype=torch.uint8).cuda()
    with torch.no_grad():
        y1 = m(x)[0].clone()
        ea.model.train()
        ea.step(x.repeat(2, 1, 1, 1), None)
        y2 = ea.model.eval()(x)[0]
    assert not torch.equal(y1, y2)


def test_train_engine_matches_torch_sgd_reference_loop():
    """Same step written with stock pieces (this library's modules + ComputeLoss, torch clip_grad_norm_, torch.optim.SGD
    with the reference's three groups, utils/torch_utils.py:318-346): parameter updates agree."""
    from lead_yolo_b200.loss import ComputeLoss
    from lead_yolo_b200.model import build_lead_yolo
    from lead_yolo_b200.train import HYP_DEFAULT, TrainEngine, param_groups, scale_hyp
    g = torch.Generator().manual_seed(9)
    imgs = [torch.randint(0, 256, (4, 3, 128, 128), generator=g, dtype=torch.uint8).cuda() for _ in range(2)]
    tg = torch.tensor([[i, 0, 0.3 + 0.1 * i, 0.5, 0.2, 0.25] for i in range(4)], device="cuda")
    model = build_lead_yolo("lead", 1, "cuda", seed=1).train()
    model.hyp = scale_hyp(HYP_DEFAULT, 3, 1, 128)
    g0, g1, g2 = param_groups(model)
    assert len(g0) + len(g1) + len(g2) == len(list(model.parameters()))
    opt = torch.optim.SGD(g1 + g2, lr=0.01, momentum=0.937, nesterov=True)
    opt.add_param_group({"params": g0, "weight_decay": 5e-4})
    closs = ComputeLoss(model)
    before = torch.cat([p.detach().flatten() for p in g0 + g1 + g2])
    for im in imgs:
        loss, _ = closs(model(im), tg)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=10.0)
        opt.step()
        opt.zero_grad(set_to_none=True)
    d_ref = torch.cat([p.detach().flatten() for p in g0 + g1 + g2]) - before

    model2 = build_lead_yolo("lead", 1, "cuda", seed=1)
    eng = TrainEngine(model2, 4, 128, graph=True, device="cuda")
    before2 = torch.cat([p.detach().flatten() for p in eng.params]).clone()
    for im in imgs:
        eng.step(im, tg)
    d_eng = torch.cat([p.detach().flatten() for p in eng.params]) - before2
    assert ((d_eng - d_ref).norm() / d_ref.norm()).item() < 1e-2
    # EMA follows ModelEMA.update: d = decay * (1 - exp(-updates / tau))
    ema = eng.ema_model()
    assert not any(p.requires_grad for p in ema.parameters()) and not ema.training
