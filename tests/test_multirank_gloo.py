"""world_size-2 gloo test (CPU) of the host-side multi-process logic: per-rank synthetic shards differ, DDP-style
parameter broadcast makes replicas identical, gradient averaging and the loss*WORLD_SIZE convention
(train.py:321-322) give the single-process gradient, and bench.py's max-over-ranks timing reduction."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import bench
        from lead_yolo_b200.model import LeadYolo
        # per-rank shard of the synthetic workload (seed 100*rank + i as in bench.py)
        shard = bench.sar_like_batch(2, 64, 100 * rank)
        gathered = [torch.empty_like(shard) for _ in range(world)]
        dist.all_gather(gathered, shard)
        assert not torch.equal(gathered[0], gathered[1])
        # replicas: different local init, identical after the rank-0 broadcast DDP performs at wrap time
        torch.manual_seed(rank)
        m = LeadYolo("lead", 1)
        for p in list(m.parameters()) + list(m.buffers()):
            dist.broadcast(p.data, 0)
        flat = torch.cat([p.detach().flatten() for p in m.parameters()])
        ref = flat.clone()
        dist.broadcast(ref, 0)
        assert torch.equal(flat, ref)
        # gradient all-reduce(mean) with loss * WORLD_SIZE == sum of per-rank gradients
        g = torch.full((1000,), float(rank + 1)) * world
        dist.all_reduce(g)
        g /= world
        assert torch.allclose(g, torch.full((1000,), 3.0))
        # bench.py timing: max over ranks
        t = torch.tensor([10.0 + rank, 20.0 - rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.tolist() == [11.0, 20.0]
        out.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        out.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_two_rank_host_logic():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
