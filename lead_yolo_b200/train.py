"""One training step of LEAD-YOLO as a replayable device program.

`TrainEngine.step(images, targets)` is the body of the reference's training loop (train.py:291-352) for one batch:
uint8 images -> forward (batch-statistics BatchNorm) -> ComputeLoss (utils/loss.py:121-191) -> `loss *= WORLD_SIZE`
(train.py:321-322) -> backward -> gradient mean over ranks (DDP, utils/torch_utils.py:55-63) -> clip_grad_norm_(10)
-> SGD-Nesterov step over the reference's three parameter groups (utils/torch_utils.py:318-346) -> ModelEMA.update
(utils/torch_utils.py:404-432).

What differs from the reference is only how it is executed:
  * forward + loss + backward are captured ONCE into a CUDA graph over static buffers and replayed (the eager step is
    bound by ~3000 Python-side launches, not by the GPU);
  * every parameter lives in ONE flat fp32 buffer (the nn.Parameters become views of it, so state_dict / checkpoints /
    `model.parameters()` are unchanged), gradients are gathered into one flat buffer: the data-parallel exchange is a
    single NCCL all-reduce per step and the optimiser + EMA is one kernel (csrc/optim.cu);
  * all small fp32 accumulators of a step come out of one arena that is zeroed once (ops.arena_begin).
The modules stay ordinary nn.Modules: the same model also trains eagerly under stock DistributedDataParallel
(tests/test_ddp_gpu.py); this engine is the fast path for `bench.py` and for a training loop that wants it.
"""
from __future__ import annotations

import copy
import math

import torch
import torch.nn as nn

from . import lowlevel as ll
from . import ops
from .loss import ComputeLoss

HYP_DEFAULT = {  # data/hyps/hyp.scratch-low.yaml of the reference (optimiser + loss part)
    "lr0": 0.01, "momentum": 0.937, "weight_decay": 0.0005, "box": 0.05, "cls": 0.5, "cls_pw": 1.0, "obj": 1.0,
    "obj_pw": 1.0, "anchor_t": 4.0, "fl_gamma": 0.0, "label_smoothing": 0.0}


def scale_hyp(hyp, nl, nc, imgsz):
    """train.py:238-242."""
    h = dict(hyp)
    h["box"] *= 3 / nl
    h["cls"] *= nc / 80 * 3 / nl
    h["obj"] *= (imgsz / 640) ** 2 * 3 / nl
    h["label_smoothing"] = h.get("label_smoothing", 0.0)
    return h


def param_groups(model: nn.Module):
    """smart_optimizer's grouping (utils/torch_utils.py:318-329): weights with decay | norm weights | biases."""
    norm = tuple(v for k, v in nn.__dict__.items() if "Norm" in k)
    g = [], [], []
    seen = set()
    for v in model.modules():
        for name, p in v.named_parameters(recurse=False):
            if id(p) in seen or not p.requires_grad:
                continue
            seen.add(id(p))
            if name == "bias":
                g[2].append(p)
            elif name == "weight" and isinstance(v, norm):
                g[1].append(p)
            else:
                g[0].append(p)
    return g


def _flatten(tensors, align=4):
    """Move `tensors` into one flat fp32 buffer; each becomes a view of it (16-byte aligned). Returns (flat, views)."""
    sizes = [(t.numel() + align - 1) // align * align for t in tensors]
    dev = tensors[0].device
    flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
    views, off = [], 0
    for t, sz in zip(tensors, sizes):
        v = flat[off:off + t.numel()].view(t.shape)
        v.copy_(t.detach())
        t.data = v
        views.append(v)
        off += sz
    return flat, views


class TrainEngine:
    def __init__(self, model: nn.Module, batch: int, img: int = 640, hyp: dict | None = None, max_targets: int | None = None,
                 ema_decay: float = 0.9999, ema_tau: float = 2000.0, max_norm: float = 10.0, graph: bool = True,
                 broadcast_buffers: bool = True, device=None):
        import torch.distributed as dist
        self.dist = dist if (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1) else None
        self.world = self.dist.get_world_size() if self.dist else 1
        self.model = model.train()
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        det = model.model[-1]
        self.hyp = scale_hyp(HYP_DEFAULT if hyp is None else hyp, det.nl, det.nc, img)
        model.hyp, model.nc = self.hyp, det.nc
        self.compute_loss = ComputeLoss(model)
        self.batch, self.img = batch, img
        self.max_targets = max_targets or 32 * batch
        self.ema_decay, self.ema_tau, self.updates = ema_decay, ema_tau, 0
        self.broadcast_buffers = broadcast_buffers

        # ---- flat parameter / buffer storage (group order = optimiser groups)
        g0, g1, g2 = param_groups(model)
        self.params = g0 + g1 + g2
        missing = [n for n, p in model.named_parameters() if p.requires_grad and all(p is not q for q in self.params)]
        assert not missing, missing
        self.flat, _ = _flatten(self.params)
        al = lambda ts: sum((t.numel() + 3) // 4 * 4 for t in ts)  # noqa: E731
        self.n0, self.n1, self.n = al(g0), al(g0) + al(g1), self.flat.numel()
        self.flat_grad = torch.zeros_like(self.flat)
        self.grad_views, off = [], 0
        for p in self.params:
            self.grad_views.append(self.flat_grad[off:off + p.numel()].view(p.shape))
            off += (p.numel() + 3) // 4 * 4
        self.momentum_buf = torch.zeros_like(self.flat)
        self.ema = self.flat.clone()
        self.bufs = [b for b in model.buffers() if b.is_floating_point() and b.numel() > 0 and
                     not any(b is a for a in [det.anchors])]
        self.flat_buf, _ = _flatten(self.bufs) if self.bufs else (None, [])
        self.ema_buf = self.flat_buf.clone() if self.flat_buf is not None else None
        for p, gv in zip(self.params, self.grad_views):
            p.grad = gv
        # backward kernels write parameter gradients straight into their flat slice where they can (ops.grad_sink)
        self._sink = {p.data_ptr(): gv for p, gv in zip(self.params, self.grad_views)}
        # every BatchNorm's num_batches_tracked in one int64 buffer: one add per step instead of one per layer
        self.nbt = [b for b in model.buffers() if b.dtype == torch.int64 and b.numel() == 1]
        if self.nbt:
            self.nbt_flat = torch.zeros(len(self.nbt), dtype=torch.int64, device=self.device)
            for i, b in enumerate(self.nbt):
                self.nbt_flat[i] = b.reshape(())
                b.data = self.nbt_flat[i].view(b.shape)

        # ---- device-resident hyper-parameters and static inputs
        h = self.hyp
        self.hp_host = [h["lr0"], h["lr0"], h["lr0"], h["momentum"], h["weight_decay"], max_norm, 0.0, 1.0 / self.world]
        self.hp = torch.tensor(self.hp_host, dtype=torch.float32, device=self.device)
        self.gsq = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.images = torch.zeros((batch, 3, img, img), dtype=torch.uint8, device=self.device)
        self.targets = torch.full((self.max_targets, 6), -1.0, dtype=torch.float32, device=self.device)
        self.loss_out = None
        self._pack = None                 # (device job table, jobs, blocks, {key: packed tensor}) of the captured step
        self.graphs = None
        self.use_graph = graph
        self.launches_per_step = 0

    # -- hyper-parameters (picked up by the next step without re-capture)
    def set_lr(self, lr_weights, lr_norm=None, lr_bias=None, momentum=None):
        self.hp_host[0] = lr_weights
        self.hp_host[1] = lr_weights if lr_norm is None else lr_norm
        self.hp_host[2] = lr_weights if lr_bias is None else lr_bias
        if momentum is not None:
            self.hp_host[3] = momentum

    # -- the step, as enqueued work
    def _forward_backward(self):
        ops.arena_begin(self.device)
        self.flat_grad.zero_()            # gradient sinks accumulate (weight-gradient atomics land there directly)
        if self._pack is not None:        # every parameter in every packed layout of the step: one launch
            ll.call("ly_pack_batch", self._pack[0], self._pack[1], self._pack[2])
            ops.PACK_PLAN = self._pack[3]
        ops.BYPASS_PARAM_CACHE = True     # replayed launches must re-derive everything that depends on a parameter
        ops.GRAD_SINK = self._sink
        ops._SINK_USED.clear()
        ops.NBT_EXTERNAL = bool(self.nbt)
        try:
            raw = self.model(self.images)
            if self.nbt:
                self.nbt_flat.add_(1)
            loss, _ = self.compute_loss(raw, self.targets)
            total = loss * float(self.world)                      # train.py:321-322
            grads = torch.autograd.grad(total, self.params, allow_unused=False)
        finally:
            ops.BYPASS_PARAM_CACHE = False
            ops.GRAD_SINK = {}
            ops.PACK_PLAN = None
            ops.NBT_EXTERNAL = False
            ops.arena_end()
        # whatever did not go through a sink is copied into its slice
        dst, src = [], []
        for gv, g in zip(self.grad_views, grads):
            if g.data_ptr() != gv.data_ptr():
                dst.append(gv)
                src.append(g)
        if dst:
            torch._foreach_copy_(dst, src)
        self.sunk = len(self.grad_views) - len(dst)
        return loss.detach()

    def _optimizer(self):
        self.gsq.zero_()
        ll.call("ly_sq_sum", self.flat_grad, self.n, self.gsq)
        ll.call("ly_sgd_ema_step", self.flat, self.flat_grad, self.momentum_buf, self.ema, self.n, self.n0, self.n1,
                self.hp, self.gsq)
        if self.flat_buf is not None:
            ll.call("ly_ema_update", self.ema_buf, self.flat_buf, self.flat_buf.numel(), self.hp)

    def _capture(self):
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(3):           # warm-up: caches, arena high-water mark, cuBLAS-free but allocator-heavy
                self._forward_backward()
            # one more pass that records which parameter is packed into which layout: the captured step packs them all
            # with ONE launch (ops.PACK_PLAN) instead of ~150 small ones
            ops.PACK_RECORD = []
            try:
                self._forward_backward()
            finally:
                rec, ops.PACK_RECORD = ops.PACK_RECORD, None
            if rec:
                self._pack = ops.pack_batch_table(rec, self.device)
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        g1, g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        ll.COUNT[0] = 0
        # thread_local: other threads of the process (NCCL's watchdog polls events, a DDP reducer may be alive) must not
        # invalidate the capture; the kernels the autograd worker thread launches still land in the capturing stream
        with torch.cuda.graph(g1, capture_error_mode="thread_local"):
            self.loss_out = self._forward_backward()
        with torch.cuda.graph(g2, capture_error_mode="thread_local"):
            self._optimizer()
        self.launches_per_step = ll.COUNT[0]
        self.graphs = (g1, g2)

    def step(self, images: torch.Tensor, targets: torch.Tensor | None = None) -> torch.Tensor:
        """Enqueue one training step.  images: uint8 (B,3,S,S) on the host (pinned) or the device; targets: (nt, 6)
        `(image, class, x, y, w, h)` normalised (utils/dataloaders.py:886-890), nt <= max_targets.  Returns the
        device scalar `(lbox + lobj + lcls) * bs` of this rank (before the WORLD_SIZE scaling)."""
        if images is not self.images:
            self.images.copy_(images, non_blocking=True)
        if targets is not None and targets is not self.targets:
            nt = int(targets.shape[0])
            if nt > self.max_targets:
                raise ll.LeadYoloError(f"TrainEngine: {nt} targets exceed max_targets={self.max_targets}")
            self.targets.fill_(-1.0)
            if nt:
                self.targets[:nt].copy_(targets, non_blocking=True)
        self.updates += 1
        self.hp_host[6] = self.ema_decay * (1 - math.exp(-self.updates / self.ema_tau))   # torch_utils.py:414,421
        self.hp.copy_(torch.tensor(self.hp_host, dtype=torch.float32))
        if self.dist and self.broadcast_buffers and self.flat_buf is not None:
            self.dist.broadcast(self.flat_buf, 0)             # DDP broadcast_buffers=True: rank 0's BN statistics
        if self.use_graph:
            if self.graphs is None:
                self._capture()
            self.graphs[0].replay()
            loss = self.loss_out
        else:
            ll.COUNT[0] = 0
            loss = self._forward_backward()
        if self.dist:
            self.dist.all_reduce(self.flat_grad)              # sum; hp[7] = 1/world turns it into DDP's mean
        if self.use_graph:
            self.graphs[1].replay()
        else:
            self._optimizer()
            self.launches_per_step = ll.COUNT[0]
        # parameters and BatchNorm buffers changed through raw pointers: invalidate version-keyed caches
        torch.autograd.graph.increment_version(self.params)
        return loss

    # -- EMA access (ModelEMA.ema of the reference: a detached eval copy of the model)
    def ema_model(self) -> nn.Module:
        m = copy.deepcopy(self.model).eval()
        src = dict(zip((id(p) for p in self.params), range(len(self.params))))
        off = [0]
        for p in self.params:
            off.append(off[-1] + (p.numel() + 3) // 4 * 4)
        with torch.no_grad():
            for (_, p), (_, q) in zip(self.model.named_parameters(), m.named_parameters()):
                i = src[id(p)]
                q.copy_(self.ema[off[i]:off[i] + p.numel()].view(p.shape))
            if self.flat_buf is not None:
                boff = 0
                mb = {n: b for n, b in m.named_buffers()}
                for (n, b) in self.model.named_buffers():
                    if any(b is x for x in self.bufs):
                        mb[n].copy_(self.ema_buf[boff:boff + b.numel()].view(b.shape))
                        boff += (b.numel() + 3) // 4 * 4
        for q in m.parameters():
            q.requires_grad_(False)
        return m
