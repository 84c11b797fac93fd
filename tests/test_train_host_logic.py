"""CPU tests of the host logic behind `train.TrainEngine` that needs no kernel: the gradient-sink rules (ops.grad_sink),
the job table of the batched weight packing (ops.pack_batch_table / ly_pack_job) and the optimiser parameter grouping
(train.param_groups against the reference's smart_optimizer rule, utils/torch_utils.py:318-329)."""
import ctypes as C

import torch
import torch.nn as nn


def test_grad_sink_rules():
    from lead_yolo_b200 import ops
    flat = torch.zeros(32)
    w = nn.Parameter(torch.randn(4, 4))
    b = nn.Parameter(torch.randn(4))
    h = nn.Parameter(torch.randn(4).half())
    views = {w.data_ptr(): flat[:16].view(4, 4), b.data_ptr(): flat[16:20], h.data_ptr(): flat[20:24]}
    assert ops.grad_sink(w) is None                      # no engine step is running
    ops.GRAD_SINK = views
    ops._SINK_USED.clear()
    try:
        s = ops.grad_sink(w)
        assert s is not None and s.data_ptr() == flat.data_ptr() and s.shape == w.shape
        assert ops.grad_sink(w) is None                  # second use of the same parameter: autograd has to add
        assert ops.grad_sink(w.t()) is None              # a view / non-leaf never writes into the parameter's slice
        assert ops.grad_sink(w * 2) is None
        assert ops.grad_sink(h) is None                  # not fp32
        assert ops.grad_sink(None) is None
        assert ops.grad_sink(nn.Parameter(torch.randn(4))) is None       # not a parameter of the engine
        sb = ops.grad_sink(b)
        assert sb is not None and sb.data_ptr() == flat[16:].data_ptr()
    finally:
        ops.GRAD_SINK = {}
        ops._SINK_USED.clear()


def test_pack_batch_table_layout():
    from lead_yolo_b200 import ops
    from lead_yolo_b200._lib import PackJob
    bufs = [torch.empty(10, 64, dtype=torch.bfloat16), torch.empty(64, 576, dtype=torch.bfloat16)]
    rec = [
        (("f", 1), bufs[0], dict(kind=0, w=4096, Cout=10, Cin=48, kh=1, kw=1, patch=0, c=[48, 0, 0, 0], coff=0, cs=0, P=0,
                                 Kp=64, total=640)),
        (("f", 1), bufs[0], dict(kind=0, w=4096, Cout=10, Cin=48, kh=1, kw=1, patch=0, c=[48, 0, 0, 0], coff=0, cs=0, P=0,
                                 Kp=64, total=640)),                      # same key twice: one job
        (("d", 2), bufs[1], dict(kind=1, w=8192, Cout=64, Cin=64, kh=3, kw=3, patch=0, c=[0, 0, 0, 0], coff=0, cs=64, P=64,
                                 Kp=576, total=64 * 576)),
    ]
    table, njobs, blocks, plan = ops.pack_batch_table(rec, "cpu")
    assert njobs == 2 and set(plan) == {("f", 1), ("d", 2)}
    assert blocks == 3 + 144                              # ceil(640 / 256) + 36864 / 256
    assert table.numel() == 2 * C.sizeof(PackJob)
    jobs = (PackJob * 2).from_buffer_copy(bytes(table.numpy().tobytes()))
    assert (jobs[0].w, jobs[0].dst, jobs[0].kind, jobs[0].total, jobs[0].first_block) == (4096, bufs[0].data_ptr(), 0, 640, 0)
    assert (jobs[1].w, jobs[1].dst, jobs[1].kind, jobs[1].total, jobs[1].first_block) == (8192, bufs[1].data_ptr(), 1,
                                                                                     64 * 576, 3)
    assert list(jobs[0].c) == [48, 0, 0, 0] and (jobs[1].cs, jobs[1].P, jobs[1].Kp) == (64, 64, 576)


def test_param_groups_follow_smart_optimizer():
    """weights with decay | normalisation weights | biases — every parameter in exactly one group."""
    from lead_yolo_b200.train import param_groups

    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv = nn.Conv2d(3, 8, 3, bias=True)
            self.bn = nn.BatchNorm2d(8)
            self.fc = nn.Linear(8, 2, bias=False)
            self.scale = nn.Parameter(torch.ones(1))
            self.frozen = nn.Parameter(torch.ones(1), requires_grad=False)

    m = M()
    g0, g1, g2 = param_groups(m)
    assert {id(p) for p in g0} == {id(m.conv.weight), id(m.fc.weight), id(m.scale)}
    assert {id(p) for p in g1} == {id(m.bn.weight)}
    assert {id(p) for p in g2} == {id(m.conv.bias), id(m.bn.bias)}
    assert all(p is not m.frozen for g in (g0, g1, g2) for p in g)
