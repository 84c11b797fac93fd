#!/usr/bin/env python
"""bench.py — LEAD-YOLO images/s at 640x640 on N B200s (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode infer|train]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

ours / infer (default; BASELINE.json configs[1]): batched inference bs=64/GPU, 640x640, bf16 activations, random-init
weights, synthetic SAR-like uint8 images.  A step = one forward of the whole network (stem .. Detect decode) on one
batch.  `value` = device-resident throughput (CUDA-graph replay, inputs in HBM); `e2e` = same metric through the public
call `model(images)` with pinned HOST uint8 images copied in and the decoded predictions copied out every step.
reference: the reference's CPU PyTorch path for the same forward, restated in oracle/restate.py (the reference tree
is not present on the GPU box), fp32 on all host threads, on a bounded sample of the workload.
"""
from __future__ import annotations

import argparse
import collections
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

BS = 64
IMG = 640
FWD_GFLOP_PER_IMG = 5.935      # 2*MAC over every nn.Conv2d at 640x640
FWD_MB_PER_IMG = 36.44         # bf16 activations at module boundaries + weights once per batch


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, "fallback"


def sar_like_batch(n, s, seed):
    """uint8 (n,3,s,s): single-look-SAR-like clutter, Exponential(mean 40) clipped, gray replicated, with a few bright
    ship-like rectangles (statistics measured on data/SSDD)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.empty(n, 1, s, s).exponential_(1 / 40.0, generator=g).clamp_(0, 255)
    nb = torch.poisson(torch.full((n,), 2.2), generator=g).clamp_(1, 29).int().tolist()
    for i in range(n):
        for _ in range(nb[i]):
            w = int(max(0.01, min(0.5, torch.normal(0.070, 0.070, (1,), generator=g).item())) * s)
            h = int(max(0.01, min(0.6, torch.normal(0.116, 0.097, (1,), generator=g).item())) * s)
            cx = int(torch.randint(w // 2 + 1, s - w // 2 - 1, (1,), generator=g))
            cy = int(torch.randint(h // 2 + 1, s - h // 2 - 1, (1,), generator=g))
            x[i, 0, cy - h // 2:cy + h // 2 + 1, cx - w // 2:cx + w // 2 + 1] = \
                float(torch.empty(1).uniform_(180, 255, generator=g))
    return x.expand(n, 3, s, s).round().to(torch.uint8).contiguous()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [t.strip() for t in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nme, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(n):
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not dist.is_initialized():
        torch.cuda.set_device(local)
        import datetime
        # a short timeout: a collective mismatch should fail the run in minutes, not hold N GPUs for ten
        dist.init_process_group("nccl", device_id=torch.device("cuda", local),
                                timeout=datetime.timedelta(seconds=180))
    return rank, local, world


def reference_available():
    from oracle import ref_import
    return ref_import.available()


def cpu_reference_run(steps, warmup, sample_imgs, detail=False):
    """The reference's CPU path, eval forward incl. Detect decode, fp32 on all host threads.  With the git-ignored copy
    of the reference present (baseline/_ref, tools/install_ref.py) this is the UNMODIFIED reference
    `Model('models/LEAD-YOLO.yaml')` after `.fuse().eval()` as `attempt_load` leaves it (kind "reference"); otherwise
    the restatement in oracle/restate.py (kind "port").  Returns (img/s, ms/step, cores, kind, extra)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    x = sar_like_batch(sample_imgs, IMG, 0).float() / 255
    extra = {}
    if reference_available():
        from oracle import ref_import
        model = ref_import.build_model(seed=0).eval()
        kind = "reference"

        def timed(m, inp, k, w):
            with torch.no_grad():
                for _ in range(w):
                    m(inp)
                t0 = time.perf_counter()
                for _ in range(k):
                    m(inp)
                return (time.perf_counter() - t0) / k
        if detail:
            # bs=1 unfused and fused; bs=8 forward + ComputeLoss + backward
            extra["bs1_unfused_ms"] = round(timed(model, x[:1], 5, 2) * 1e3, 2)
        model.fuse()
        if detail:
            extra["bs1_fused_ms"] = round(timed(model, x[:1], 5, 2) * 1e3, 2)
        dt = timed(model, x, steps, warmup)
        if detail:
            try:
                extra.update(_cpu_train_reference(cores))
            except Exception as e:  # noqa: BLE001
                extra["train_bs8"] = f"failed: {e}"
        return sample_imgs / dt, dt * 1e3, cores, kind, extra
    from oracle import restate as R
    from lead_yolo_b200.model import LeadYolo
    torch.manual_seed(0)
    model = LeadYolo("lead", 1).eval()
    P = {k: v.detach().float() for k, v in model.state_dict().items()}
    with torch.no_grad():
        for _ in range(warmup):
            R.model_forward(P, x, R.Ctx(False))
        t0 = time.perf_counter()
        for _ in range(steps):
            R.model_forward(P, x, R.Ctx(False))
        dt = (time.perf_counter() - t0) / steps
    return sample_imgs / dt, dt * 1e3, cores, "port", extra


def ssdd_like_targets(n, seed):
    """(nt, 6) labels [image, 0, x, y, w, h] with SSDD statistics."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    rows = []
    nb = torch.poisson(torch.full((n,), 2.2), generator=g).clamp_(1, 29).int().tolist()
    for i in range(n):
        for _ in range(nb[i]):
            w = float(torch.normal(0.070, 0.070, (1,), generator=g).clamp(0.01, 0.5))
            h = float(torch.normal(0.116, 0.097, (1,), generator=g).clamp(0.01, 0.6))
            cx = float(torch.empty(1).uniform_(w / 2, 1 - w / 2, generator=g))
            cy = float(torch.empty(1).uniform_(h / 2, 1 - h / 2, generator=g))
            rows.append([i, 0, cx, cy, w, h])
    return torch.tensor(rows, dtype=torch.float32)


def _ref_train_setup(device):
    """Reference model + ComputeLoss + SGD exactly as train.py builds them (SURVEY §8d recipe)."""
    import yaml
    from oracle import ref_import
    ref = ref_import.load()
    if ref.root not in sys.path:
        sys.path.insert(0, ref.root)
    from utils.loss import ComputeLoss
    model = ref_import.build_model(seed=0).to(device).train()
    hyp = yaml.safe_load(open(os.path.join(ref.root, "data", "hyps", "hyp.scratch-low.yaml")))
    nl = model.model[-1].nl
    hyp["box"] *= 3 / nl
    hyp["cls"] *= 1 / 80 * 3 / nl
    hyp["obj"] *= (IMG / 640) ** 2 * 3 / nl
    hyp["label_smoothing"] = 0.0
    model.nc, model.hyp = 1, hyp
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.937, nesterov=True)
    return model, ComputeLoss(model), opt


def _cpu_train_reference(cores):
    model, closs, opt = _ref_train_setup("cpu")
    x = sar_like_batch(8, IMG, 1).float() / 255
    t = ssdd_like_targets(8, 1)
    times = []
    for i in range(3):
        t0 = time.perf_counter()
        loss, _ = closs(model(x), t)
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        times.append(time.perf_counter() - t0)
    dt = min(times[1:])
    return {"train_bs8_fwd_bwd_ms": round(dt * 1e3, 1), "train_bs8_img_s": round(8 / dt, 2)}


def eager_gpu_reference(dev, batch, steps=4):
    """The practical bar: the UNMODIFIED reference, PyTorch eager on this B200 — inference and
    forward + ComputeLoss + backward + SGD, fp32 and bf16 autocast, channels_last.  None when there is no copy."""
    if not reference_available():
        return None
    out = {}
    x8 = sar_like_batch(batch, IMG, 2).to(dev)
    tg = ssdd_like_targets(batch, 2).to(dev)
    for tag, amp in (("fp32", False), ("bf16_autocast", True)):
        try:
            model, closs, opt = _ref_train_setup(dev)
            model = model.to(memory_format=torch.channels_last)

            def train_step():
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                    pred = model((x8.float() / 255).contiguous(memory_format=torch.channels_last))
                    loss, _ = closs([p.float() for p in pred], tg)
                loss.backward()
                opt.step()
                opt.zero_grad(set_to_none=True)
            for _ in range(2):
                train_step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                train_step()
            e1.record()
            torch.cuda.synchronize()
            out[f"train_{tag}_img_s"] = round(batch * steps / (e0.elapsed_time(e1) / 1e3), 1)
            model.eval()
            with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                xin = (x8.float() / 255).contiguous(memory_format=torch.channels_last)
                for _ in range(2):
                    model(xin)
                e0.record()
                for _ in range(steps):
                    model(xin)
                e1.record()
                torch.cuda.synchronize()
            out[f"infer_{tag}_img_s"] = round(batch * steps / (e0.elapsed_time(e1) / 1e3), 1)
            del model, closs, opt
        except Exception as e:  # noqa: BLE001
            out[f"{tag}_error"] = str(e)[:200]
    torch.cuda.empty_cache()
    return out


def train_measure(batch, steps, warm, rank, local, world, dev, profile=True, variant="lead", eager_bar=False):
    """configs[2]: the reference's training step (train.py:291-352) at bs=64/GPU through lead_yolo_b200.train.TrainEngine:
    uint8 images -> forward -> ComputeLoss (EIoU) -> loss * WORLD_SIZE -> backward -> NCCL gradient mean -> clip ->
    SGD-Nesterov -> EMA.  Returns the dict that rides in the bench line as `train`."""
    import torch.distributed as dist
    from lead_yolo_b200 import lowlevel as ll
    from lead_yolo_b200.model import build_lead_yolo
    from lead_yolo_b200.train import TrainEngine
    model = build_lead_yolo(variant, 1, dev, seed=0)
    eng = TrainEngine(model, batch, IMG, device=dev, graph=not os.environ.get("LY_TRAIN_NO_GRAPH"))
    nbuf = 3
    host = [sar_like_batch(batch, IMG, 100 * rank + i).pin_memory() for i in range(nbuf)]
    tgt = [ssdd_like_targets(batch, 100 * rank + i).to(dev) for i in range(nbuf)]
    dimg = [h.to(dev, non_blocking=True) for h in host]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(warm):
        eng.step(dimg[i % nbuf], tgt[i % nbuf])
    barrier()
    if os.environ.get("LY_PROFILE_STEP"):     # ncu --profile-from-start off: exactly ONE training step is profiled
        torch.cuda.profiler.start()
        eng.step(dimg[0], tgt[0])
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ar = []
    e0.record()
    for i in range(steps):
        loss = eng.step(dimg[i % nbuf], tgt[i % nbuf])
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    loss_val = float(loss)
    # end to end: pinned host uint8 batch + labels in, loss scalar out, every step
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()
    tgt_host = [t.cpu().pin_memory() for t in tgt]
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(steps):
        loss_host.copy_(eng.step(host[i % nbuf], tgt_host[i % nbuf]), non_blocking=True)
    t1.record()
    barrier()
    ms_e2e = t0.elapsed_time(t1)
    # exposed all-reduce: the same steps with the collective's duration bracketed on the stream
    ar_ms = 0.0
    if world > 1:
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ea.record()
        for _ in range(5):
            dist.all_reduce(eng.flat_grad)
        eb.record()
        barrier()
        ar_ms = ea.elapsed_time(eb) / 5
    t = torch.tensor([ms_total, ms_e2e, ar_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, ar_ms = t.tolist()
    out = {"value": batch * world * steps / (ms_total / 1e3), "unit": "images/s", "ms_per_step": ms_total / steps,
           "steps": steps, "warmup": warm + 3,
           "workload": f"LEAD-YOLO.yaml fwd + ComputeLoss(EIoU) + bwd + clip + SGD-Nesterov + EMA, bs={batch}/GPU 640x640 "
                       "bf16 (configs[2]), SSDD-shaped synthetic labels",
           "parallelism": f"DDP dp{world}: one NCCL all-reduce of the flat fp32 gradient (12.5 MB) per step" if world > 1
                          else "single GPU",
           "cuda_graph": eng.graphs is not None, "gpu_launches_per_step": eng.launches_per_step,
           "frac_of_roofline_40.6k": batch * steps / (ms_total / 1e3) / 40600.0,
           "allreduce_ms_exposed": ar_ms, "loss_last_step": loss_val,
           "e2e": {"value": batch * world * steps / (ms_e2e / 1e3), "unit": "images/s",
                   "h2d_bytes_per_step": int(host[0].numel() + tgt_host[0].numel() * 4), "d2h_bytes_per_step": 4}}
    if profile:
        # per-entry-point device time of ONE eager step (events around every C-ABI call, device kept busy ahead).
        # Every rank takes the step (it contains the collectives); rank 0's table is reported.
        eng.use_graph = False
        ll.PROFILE = []
        torch.cuda._sleep(int(2e8))
        eng.step(dimg[0], tgt[0])
        torch.cuda.synchronize()
        prof = {}
        for name, a, b, nbytes, flops in ll.PROFILE:
            d = prof.setdefault(name, [0.0, 0])
            d[0] += a.elapsed_time(b)
            d[1] += 1
        ll.PROFILE = None
        eng.use_graph = True
        out["kernel_ms_per_step"] = {k: round(v[0], 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])[:10]}
        out["kernel_ms_total"] = round(sum(v[0] for v in prof.values()), 3)
    del eng
    if eager_bar and rank == 0 and world == 1:
        bar = eager_gpu_reference(dev, batch)
        if bar is not None:
            out["eager_torch_reference"] = bar
    return out


def calibrate_head(model, dimg, per_image=100, conf=0.25):
    """Random-init objectness sits at ~0.02, so nothing would reach the NMS.  Shift the Detect objectness biases (one
    constant for the whole head) until about `per_image` cells per image pass conf_thres — the load a trained
    detector hands to the NMS on a scene with a few ships and clutter."""
    det = model.model[-1]
    with torch.no_grad():
        dec, _ = model(dimg)
        obj = dec[..., 4].flatten().double().clamp(1e-9, 1 - 1e-9)
        o = torch.log(obj / (1 - obj))                      # objectness logits
        c = dec[..., 5:].max(-1)[0].flatten().double()      # class score
        want = per_image * dec.shape[0]
        lo, hi = -20.0, 20.0
        for _ in range(40):                                 # bisection on the shift: count is monotone in it
            d = 0.5 * (lo + hi)
            n = int(((torch.sigmoid(o + d) * c) > conf).sum())
            lo, hi = (d, hi) if n < want else (lo, d)
        d = 0.5 * (lo + hi)
        for mi in det.m:
            mi.bias.view(det.na, -1)[:, 4] += d
        torch.autograd.graph.increment_version(list(det.m.parameters()))
    return d


def main():
    global _REAL_STDOUT
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="infer", choices=["infer", "train"],
                    help="infer = BASELINE configs[1] (default, the driver's line; it also carries a `train` object "
                         "for configs[2]); train = configs[2] as the line itself")
    ap.add_argument("--batch", type=int, default=BS)
    ap.add_argument("--variant", default="lead", choices=["lead", "c3"], help="c3 = the ablation yaml (configs[3])")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the configs[2] training measurement in the line")
    ap.add_argument("--no-extras", action="store_true", help="skip the bs=1 latency / 1024x1024 / eager-PyTorch extras")
    ap.add_argument("--profile-kernels", action="store_true", help="print the per-entry-point time table to stderr")
    args = ap.parse_args()
    warm = max(args.warmup, 3)

    if args.impl == "reference":
        rank = int(os.environ.get("RANK", "0"))
        if rank != 0:
            return
        sample = 4
        ips, ms, cores, kind, extra = cpu_reference_run(args.steps, min(warm, 3), sample, detail=True)
        what = "unmodified reference Model('models/LEAD-YOLO.yaml').fuse().eval() from baseline/_ref" \
            if kind == "reference" else "oracle/restate.py"
        _emit(json.dumps({
            "impl": "reference", "metric": "LEAD-YOLO images/sec @640", "value": ips, "unit": "images/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": min(warm, 3), "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "LEAD-YOLO.yaml batched inference 640x640 (forward + Detect decode)",
                       "sample": f"{sample} images per step"},
            "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": kind,
                             "sample": f"{sample} images x {args.steps} steps, fp32, {what}", "detail": extra},
            "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    rank, local, world = dist_setup(args.gpus)
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from lead_yolo_b200 import lowlevel as ll
    from lead_yolo_b200.model import build_lead_yolo
    hbm, tf, which = peaks()

    if args.mode == "train":
        sampler = ClockSampler(local)
        sampler.start()
        tr = train_measure(args.batch, args.steps, warm, rank, local, world, dev, variant=args.variant,
                           eager_bar=not args.no_extras)
        clocks = sampler.stop()
        if rank == 0:
            _emit(json.dumps({
                "metric": "LEAD-YOLO images/sec @640", "value": tr["value"], "unit": "images/s", "n_gpus": world,
                "steps": args.steps, "warmup": tr["warmup"], "ms_per_step": tr["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": tr["workload"], "batch_per_gpu": args.batch, "parallelism": tr["parallelism"],
                           "variant": args.variant, "l2": "inputs rotate over 3 batches; activations >> L2",
                           "cuda_graph": tr["cuda_graph"]},
                "clocks": clocks, "e2e": tr["e2e"], "gpu_launches": int(tr["gpu_launches_per_step"] * args.steps),
                "train": tr}))
        if world > 1:
            dist.destroy_process_group()
        return

    model = build_lead_yolo(args.variant, 1, dev, seed=0).eval()
    nbuf = 3   # rotate 3 input batches (3 x 78.6 MB > 126 MB L2)
    host = [sar_like_batch(args.batch, IMG, 100 * rank + i).pin_memory() for i in range(nbuf)]
    dimg = [h.to(dev, non_blocking=True) for h in host]
    torch.cuda.synchronize()
    bias_shift = calibrate_head(model, dimg[0])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.no_grad():
        for i in range(warm):
            out = model(dimg[i % nbuf])
        torch.cuda.synchronize()
        ll.COUNT[0] = 0
        model(dimg[0])
        launches_per_step = ll.COUNT[0]

        graphs = None
        if not args.no_graph:
            try:
                graphs, gouts = [], []
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    model(dimg[0])
                torch.cuda.current_stream().wait_stream(side)
                for i in range(nbuf):
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        o = model(dimg[i])
                    graphs.append(g)
                    gouts.append(o)
                for g in graphs:
                    g.replay()
                torch.cuda.synchronize()
            except Exception as e:  # noqa: BLE001
                sys.stderr.write(f"[bench] CUDA graph capture failed, timing eager launches: {e}\n")
                graphs = None

        def step(i):
            if graphs is not None:
                graphs[i % nbuf].replay()
            else:
                model(dimg[i % nbuf])

        for i in range(warm):
            step(i)
        sampler = ClockSampler(local)
        barrier()
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            step(i)
        e1.record()
        barrier()
        clocks = sampler.stop()
        ms_total = e0.elapsed_time(e1)

        # ---- end to end through the public serving call (lead_yolo_b200.engine.InferenceEngine): pinned host uint8
        # batch in; out comes what the reference's val.py / detect.py consume after the forward — the detections that
        # survive conf 0.25 / IoU 0.45 NMS (utils/general.py:884-994), here computed on the device inside the same
        # graph, (B, 301, 6) fp32 back on the host every step.  `e2e_dense` is the same loop returning the dense
        # (B, 25200, 6) decoded tensor instead (round-1 definition of e2e).
        from lead_yolo_b200.engine import InferenceEngine
        slots = int(os.environ.get("LY_ENGINE_SLOTS", "3"))

        def e2e_measure(eng):
            def loop(k):
                # serving loop: the host keeps slots-1 batches in flight and consumes the oldest result before it
                # submits the next batch into that result's slot
                pending = collections.deque()
                for i in range(k):
                    pending.append(eng.submit(host[i % nbuf]))
                    if len(pending) >= eng.slots:
                        eng.result(pending.popleft())
                while pending:
                    eng.result(pending.popleft())
            loop(3)
            barrier()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record(eng.s_in)                    # all streams idle after the barrier: first H2D starts here
            loop(args.steps)
            t1.record(eng.s_out)                   # after the last D2H
            barrier()
            return t0.elapsed_time(t1)

        eng = InferenceEngine(model, args.batch, IMG, slots=slots, device=dev, nms=(0.25, 0.45))
        ms_e2e = e2e_measure(eng)
        d2h_bytes = int(eng.host_out[0].numel() * 4)
        last = eng.host_out[(eng._n - 1) % eng.slots]
        det_per_img = float(last[:, -1, 0].mean())
        cand_per_img = float(last[:, -1, 1].mean())
        del eng
        eng = InferenceEngine(model, args.batch, IMG, slots=slots, device=dev)
        ms_e2e_dense = e2e_measure(eng)
        d2h_dense = int(eng.host_out[0].numel() * 4)
        del eng

        # ---- per-entry-point device time (one extra, untimed step with events around every C-ABI call)
        # The device first sleeps ~25 ms so that the host enqueues the whole step ahead of it: each event pair then
        # brackets the kernel alone, not the Python / ctypes launch latency in front of it.
        ll.PROFILE = []
        torch.cuda._sleep(int(5e7))
        model(dimg[0])
        torch.cuda.synchronize()
        prof = {}
        bound_ms = 0.0
        for name, a, b, nbytes, flops in ll.PROFILE:
            d = prof.setdefault(name, [0.0, 0, 0.0, 0.0, 0.0])
            d[0] += a.elapsed_time(b)
            d[1] += 1
            d[2] += nbytes
            d[3] += flops
            lb = max(nbytes / (hbm * 1e9), flops / (tf * 1e12)) * 1e3      # this launch's own roofline time
            d[4] += lb
            bound_ms += lb
        ll.PROFILE = None

    extras = {}
    if not args.no_extras and rank == 0:
        extras = infer_extras(dev, args.variant)

    train = None
    if not args.no_train:
        del graphs
        torch.cuda.empty_cache()
        try:
            train = train_measure(args.batch, max(5, args.steps // 2), 3, rank, local, world, dev,
                                  variant=args.variant, eager_bar=not args.no_extras)
        except Exception as e:  # noqa: BLE001  (the inference line must still come out)
            import traceback
            traceback.print_exc()
            train = {"error": str(e)[:300]}

    t = torch.tensor([ms_total, ms_e2e, ms_e2e_dense], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, ms_e2e_dense = t.tolist()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    imgs = args.batch * world * args.steps
    value = imgs / (ms_total / 1e3)
    tot_ms = sum(v[0] for v in prof.values())
    if args.profile_kernels:
        for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0]):
            sys.stderr.write(f"[bench] {k:28s} {v[1]:4d} launches {v[0]:9.3f} ms {100 * v[0] / tot_ms:5.1f}%"
                             f"  {v[2] / 1e6:10.1f} MB {v[3] / 1e9:9.2f} GFLOP  bound {v[4]:7.3f} ms\n")
    top = max(prof.items(), key=lambda kv: kv[1][0])
    tname, (tms, tn, tbytes, tflops, tbound) = top
    # the dominant kernel mixes HBM-bound and tensor-bound launches: `frac` = Σ per-launch roofline time / measured time;
    # `achieved`/`peak` are quoted in the unit of the side that carries more of that bound time
    hbm_time = tbytes / (hbm * 1e9)
    tc_time = tflops / (tf * 1e12)
    if tc_time > hbm_time:
        roof = {"bound": "tensor", "achieved": tflops / (tms / 1e3) / 1e12, "peak": tf, "unit": "TFLOP/s"}
    else:
        roof = {"bound": "hbm", "achieved": tbytes / (tms / 1e3) / 1e9, "peak": hbm, "unit": "GB/s"}
    roof["frac"] = roof["achieved"] / roof["peak"]
    traffic = None
    for tp in ("r02_conv_traffic.json", "r01_conv_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", tp)
        if tname == "ly_conv_fwd" and os.path.exists(tpath) and args.batch == BS:
            tj = json.load(open(tpath))   # DRAM bytes of all launches of this kernel in one step (ncu --set full)
            if tj.get("launches_per_step") == tn:
                traffic = tj["dram_bytes_read_per_step"] + tj["dram_bytes_write_per_step"]
                break
    roof.update({"traffic": traffic, "kernel": tname, "launches_per_step": tn, "share_of_step": tms / tot_ms,
                 "peak_source": which, "algorithmic_bytes_per_step": tbytes, "algorithmic_flops_per_step": tflops,
                 "per_launch_bound_ms": tbound, "measured_ms": tms, "frac_per_launch_bound": tbound / tms})

    line = {
        "metric": "LEAD-YOLO images/sec @640", "value": value, "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": warm, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "LEAD-YOLO.yaml batched inference bs=64/GPU 640x640 bf16 (configs[1]), forward + "
                               "Detect decode, random init; e2e adds conf 0.25 / IoU 0.45 NMS on the device",
                   "variant": args.variant,
                   "batch_per_gpu": args.batch, "parallelism": f"replicas x{world} (no collective)",
                   "l2": f"inputs rotate over {nbuf} batches ({nbuf}x78.6 MB > 126 MB L2); per-step activation "
                         "traffic >> L2", "cuda_graph": graphs is not None if args.no_train else True,
                   "head": f"objectness bias shifted by {bias_shift:+.2f} so that ~100 cells/image pass conf 0.25 "
                           f"(measured: {cand_per_img:.0f} candidates, {det_per_img:.0f} detections per image)"},
        "clocks": clocks,
        "e2e": {"value": imgs / (ms_e2e / 1e3), "unit": "images/s", "h2d_bytes_per_step": int(host[0].numel()),
                "d2h_bytes_per_step": d2h_bytes, "ms_per_step": ms_e2e / args.steps,
                "result": "post-NMS detections (B, 301, 6) fp32"},
        "e2e_dense": {"value": imgs / (ms_e2e_dense / 1e3), "unit": "images/s",
                      "h2d_bytes_per_step": int(host[0].numel()), "d2h_bytes_per_step": d2h_dense,
                      "result": "dense decoded predictions (B, 25200, 6) fp32"},
        "gpu_launches": int(launches_per_step * args.steps),
        "roofline": roof,
    }
    # whole-model bounds: §8(d) module-boundary budget (132.7k img/s per GPU) and Σ per-launch roofline times
    line["model_roofline"] = {"images_per_s_bound_per_gpu": 132700.0, "frac": value / world / 132700.0,
                              "per_launch_bound_ms_per_step": bound_ms, "kernel_ms_per_step": tot_ms,
                              "frac_of_per_launch_bound": bound_ms / tot_ms}
    if train is not None:
        line["train"] = train
    if extras:
        line["extras"] = extras
    if not args.no_cpu_baseline and world >= 1:
        try:
            ips, ms, cores, kind, extra = cpu_reference_run(5, 2, 4, detail=True)
            line["cpu_baseline"] = {"value": ips, "unit": "images/s", "cores": cores, "kind": kind,
                                    "sample": "4 images x 5 steps (+2 warm-up), fp32 eval forward + decode on all host "
                                              "threads, " + ("unmodified reference from baseline/_ref, fused as "
                                                             "attempt_load leaves it" if kind == "reference" else
                                                             "oracle/restate.py"), "detail": extra}
        except Exception as e:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": "images/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"failed: {e}"}
    _emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def infer_extras(dev, variant):
    """configs[0] / configs[4] as measurements: bs=1 latency (graph replay, device-resident and through the engine
    with host copies) and 1024x1024 bs=4 (the per-GPU share of bs=32 on 8 GPUs) throughput."""
    from lead_yolo_b200.engine import InferenceEngine
    from lead_yolo_b200.model import build_lead_yolo
    out = {}
    try:
        model = build_lead_yolo(variant, 1, dev, seed=0).eval()
        for tag, bs, img, iters in (("bs1_640", 1, 640, 200), ("bs4_1024", 4, 1024, 50)):
            eng = InferenceEngine(model, bs, img, slots=2, device=dev)
            x = sar_like_batch(bs, img, 7).pin_memory()
            for _ in range(3):
                eng(x)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(eng.s_run):
                e0.record(eng.s_run)
                for _ in range(iters):
                    eng.graphs[0].replay()
                e1.record(eng.s_run)
            torch.cuda.synchronize()
            dev_ms = e0.elapsed_time(e1) / iters
            t0 = time.perf_counter()
            for _ in range(iters):
                eng(x)
            lat_ms = (time.perf_counter() - t0) / iters * 1e3
            out[tag] = {"device_ms_per_batch": round(dev_ms, 4), "device_img_s": round(bs / dev_ms * 1e3, 1),
                        "host_to_host_latency_ms": round(lat_ms, 4)}
            del eng
        out["note"] = ("bs1_640 = configs[0] geometry on the GPU (graph replay; latency = pinned uint8 in -> decoded fp32 "
                       "out, synchronous); bs4_1024 = configs[4] per-GPU share (32 images over 8 GPUs), "
                       "roofline 51.9k img/s")
    except Exception as e:  # noqa: BLE001
        out["error"] = str(e)[:300]
    torch.cuda.empty_cache()
    return out


def _emit(*a, **k):
    """print() onto the ORIGINAL stdout: everything else a rank writes to fd 1 (NCCL prints its version there) goes to
    stderr, so stdout carries exactly the one JSON line the contract asks for."""
    sys.stdout.flush()
    if _REAL_STDOUT != 1:
        os.dup2(_REAL_STDOUT, 1)
    print(*a, **k)
    sys.stdout.flush()
    if _REAL_STDOUT != 1:
        os.dup2(2, 1)


_REAL_STDOUT = 1

if __name__ == "__main__":
    main()
