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


def _worker(rank, world, port, results):
    import torch.distributed as dist
    from lead_yolo_b200.loss import ComputeLoss
    from lead_yolo_b200.model import build_lead_yolo
    from lead_yolo_b200.train import HYP_DEFAULT, TrainEngine, scale_hyp
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=dev)
    try:
        model = build_lead_yolo("lead", 1, dev, seed=0).train()
        model.hyp = scale_hyp(HYP_DEFAULT, 3, 1, 128)
        sd0 = {k: v.clone() for k, v in model.state_dict().items()}
        img, tg = _shard(rank)
        img, tg = img.to(dev), tg.to(dev)
        # ---- reference semantics through stock DDP
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[rank], output_device=rank, static_graph=True)
        closs = ComputeLoss(ddp)
        loss, _ = closs(ddp(img), tg)
        (loss * world).backward()
        g_ddp = [p.grad.detach().clone() for p in model.parameters()]
        # ---- the same thing on one GPU: every shard on its own (own batch statistics), gradients summed
        model.load_state_dict(sd0)
        model.zero_grad(set_to_none=True)
        g_sum = None
        for r in range(world):
            model.load_state_dict(sd0)              # BatchNorm buffers back to the start for every shard
            im_r, tg_r = _shard(r)
            l_r, _ = closs(model(im_r.to(dev)), tg_r.to(dev))
            gr = torch.autograd.grad(l_r, list(model.parameters()))
            g_sum = list(gr) if g_sum is None else [a + b for a, b in zip(g_sum, gr)]
        num = sum(float((a - b).double().pow(2).sum()) for a, b in zip(g_ddp, g_sum))
        den = sum(float(b.double().pow(2).sum()) for b in g_sum)
        err_ddp = (num / den) ** 0.5
        # ---- TrainEngine: flat-gradient all-reduce, fused optimiser; lr 0 on the first step exposes the gradients
        model.load_state_dict(sd0)
        eng = TrainEngine(model, 4, 128, graph=True, device=dev, max_norm=0.0)
        eng.step(img, tg)
        g_eng = [v.detach().clone() for v in eng.grad_views]
        # engine gradients are the all-reduced SUM; the optimiser applies hp[7] = 1/world (DDP's mean of loss*world)
        order = {id(p): i for i, p in enumerate(model.parameters())}
        g_eng_model_order = [None] * len(g_eng)
        for p, gv in zip(eng.params, g_eng):
            g_eng_model_order[order[id(p)]] = gv / world
        num = sum(float((a - b).double().pow(2).sum()) for a, b in zip(g_eng_model_order, g_sum))
        err_eng = (num / den) ** 0.5
        eng.step(img, tg)
        flat = eng.flat.detach().clone()
        other = flat.clone()
        dist.broadcast(other, 0)
        same = bool(torch.equal(flat, other))
        if rank == 0:
            results.put((err_ddp, err_eng, same, float(loss)))
    finally:
        dist.destroy_process_group()


def test_ddp_gradients_match_single_gpu_sum():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    err_ddp, err_eng, same, loss = q.get()
    assert loss == loss and loss > 0
    assert err_ddp <= 1e-3, f"stock DDP gradients vs single-GPU sum: rel-L2 {err_ddp:.2e}"
    assert err_eng <= 1e-3, f"TrainEngine gradients vs single-GPU sum: rel-L2 {err_eng:.2e}"
    assert same, "ranks hold different parameters after the engine's step"
