"""GPU, 2 ranks over NCCL (skipped with fewer than 2 GPUs): the data-parallel training semantics of the reference —
`smart_DDP` = DistributedDataParallel(static_graph=True) (utils/torch_utils.py:55-63), `loss *= WORLD_SIZE`
(train.py:321-322), local (non-synchronised) BatchNorm statistics — on this library's modules.

  * stock DDP: the all-reduced gradients equal the sum over ranks of the single-GPU gradients of each rank's shard
    (each shard normalised with its own batch statistics), rel-L2 <= 1e-3 per parameter tensor set;
  * TrainEngine (flat-gradient all-reduce inside the replayed step): same gradients, and both ranks hold identical
    parameters after the optimiser step."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _shard(rank, n=4, s=128):
    g = torch.Generator().manual_seed(50 + rank)
    img = torch.randint(0, 256, (n, 3, s, s), generator=g, dtype=torch.uint8)
    t = torch.tensor([[i, 0, 0.3 + 0.1 * rank, 0.4 + 0.05 * i, 0.15, 0.2] for i in range(n)], dtype=torch.float32)
    return img, t


def _setup(rank, world, port):
    import datetime
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=dev, timeout=datetime.timedelta(seconds=120))
    return dist, dev


def _single_gpu_sum(model, closs, world, dev, sd0):
    """Every shard on its own (own batch statistics) on ONE GPU, gradients summed: what DDP must reproduce."""
    g_sum = None
    for r in range(world):
        model.load_state_dict(sd0)              # BatchNorm buffers back to the start for every shard
        im_r, tg_r = _shard(r)
        l_r, _ = closs(model(im_r.to(dev)), tg_r.to(dev))
        gr = torch.autograd.grad(l_r, list(model.parameters()))
        g_sum = list(gr) if g_sum is None else [a + b for a, b in zip(g_sum, gr)]
    return g_sum


def _rel(a, b):
    num = sum(float((x - y).double().pow(2).sum()) for x, y in zip(a, b))
    den = sum(float(y.double().pow(2).sum()) for y in b)
    return (num / den) ** 0.5


def _worker_ddp(rank, world, port, results):
    """Reference semantics through stock DistributedDataParallel."""
    from lead_yolo_b200.loss import ComputeLoss
    from lead_yolo_b200.model import build_lead_yolo
    from lead_yolo_b200.train import HYP_DEFAULT, scale_hyp
    dist, dev = _setup(rank, world, port)
    try:
        model = build_lead_yolo("lead", 1, dev, seed=0).train()
        model.hyp = scale_hyp(HYP_DEFAULT, 3, 1, 128)
        sd0 = {k: v.clone() for k, v in model.state_dict().items()}
        img, tg = _shard(rank)
        img, tg = img.to(dev), tg.to(dev)
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[rank], output_device=rank, static_graph=True)
        closs = ComputeLoss(ddp)
        loss, _ = closs(ddp(img), tg)
        (loss * world).backward()
        g_ddp = [p.grad.detach().clone() for p in model.parameters()]
        model.zero_grad(set_to_none=True)
        g_sum = _single_gpu_sum(model, closs, world, dev, sd0)
        if rank == 0:
            results.put((_rel(g_ddp, g_sum), float(loss)))
    except Exception:  # noqa: BLE001
        import traceback
        results.put(("error", rank, traceback.format_exc()))
        raise
    finally:
        dist.destroy_process_group()


def _worker_engine(rank, world, port, results):
    """TrainEngine: flat-gradient all-reduce inside the replayed step, fused optimiser."""
    from lead_yolo_b200.loss import ComputeLoss
    from lead_yolo_b200.model import build_lead_yolo
    from lead_yolo_b200.train import HYP_DEFAULT, TrainEngine, scale_hyp
    dist, dev = _setup(rank, world, port)
    try:
        model = build_lead_yolo("lead", 1, dev, seed=0).train()
        sd0 = {k: v.clone() for k, v in model.state_dict().items()}
        img, tg = _shard(rank)
        img, tg = img.to(dev), tg.to(dev)
        eng = TrainEngine(model, 4, 128, graph=True, device=dev, max_norm=0.0)
        eng.set_lr(0.0)                          # first step: expose the gradients, leave the weights
        flat0 = eng.flat.detach().clone()
        loss = eng.step(img, tg)
        # engine gradients are the all-reduced SUM of d(loss*world); the optimiser applies hp[7] = 1/world
        order = {id(p): i for i, p in enumerate(model.parameters())}
        g_eng = [None] * len(eng.params)
        for p, gv in zip(eng.params, eng.grad_views):
            g_eng[order[id(p)]] = gv.detach().clone() / world
        eng.set_lr(0.01)
        eng.step(img, tg)
        flat = eng.flat.detach().clone()
        other = flat.clone()
        dist.broadcast(other, 0)
        same = bool(torch.equal(flat, other))
        moved = not torch.equal(flat, flat0)
        # single-GPU reference on a fresh copy of the model
        ref = build_lead_yolo("lead", 1, dev, seed=0).train()
        ref.hyp = scale_hyp(HYP_DEFAULT, 3, 1, 128)
        g_sum = _single_gpu_sum(ref, ComputeLoss(ref), world, dev, sd0)
        if rank == 0:
            results.put((_rel(g_eng, g_sum), same, moved, float(loss)))
    except Exception:  # noqa: BLE001
        import traceback
        results.put(("error", rank, traceback.format_exc()))
        raise
    finally:
        dist.destroy_process_group()


def _run(worker):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    for p in procs:
        if p.is_alive():
            p.kill()
    res = q.get() if not q.empty() else None
    assert res is not None and res[0] != "error", res
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    return res


def test_stock_ddp_gradients_match_single_gpu_sum():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    err, loss = _run(_worker_ddp)
    assert loss == loss and loss > 0
    assert err <= 1e-3, f"stock DDP gradients vs single-GPU sum: rel-L2 {err:.2e}"


def test_train_engine_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    err, same, moved, loss = _run(_worker_engine)
    assert loss == loss and loss > 0
    assert err <= 1e-3, f"TrainEngine gradients vs single-GPU sum: rel-L2 {err:.2e}"
    assert same and moved, "ranks hold different (or unchanged) parameters after the engine's step"
