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
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, local, world


def cpu_reference_run(steps, warmup, sample_imgs):
    """The reference's CPU path (fp32, all host threads) restated in oracle/restate.py, eval forward incl. decode."""
    from oracle import restate as R
    from lead_yolo_b200.model import LeadYolo
    torch.manual_seed(0)
    model = LeadYolo("lead", 1).eval()
    P = {k: v.detach().float() for k, v in model.state_dict().items()}
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    x = sar_like_batch(sample_imgs, IMG, 0).float() / 255
    with torch.no_grad():
        for _ in range(warmup):
            R.model_forward(P, x, R.Ctx(False))
        t0 = time.perf_counter()
        for _ in range(steps):
            R.model_forward(P, x, R.Ctx(False))
        dt = time.perf_counter() - t0
    return sample_imgs * steps / dt, dt / steps * 1e3, cores


def train_bench(args, warm, rank, local, world, dev):
    """configs[2]: training forward + backward (+ SGD step) at bs=64/GPU, DDP gradient all-reduce over NCCL.
    The loss is a surrogate (mean square of the three raw heads, scaled by WORLD_SIZE as train.py:321-322): the
    reference's ComputeLoss is outside the hot path."""
    import torch.distributed as dist
    from lead_yolo_b200 import lowlevel as ll
    from lead_yolo_b200.model import build_lead_yolo
    model = build_lead_yolo("lead", 1, dev, seed=0).train()
    net = model
    if world > 1:
        net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], output_device=local,
                                                        static_graph=True)
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.937, nesterov=True)
    nbuf = 3
    host = [sar_like_batch(args.batch, IMG, 100 * rank + i).pin_memory() for i in range(nbuf)]
    dimg = [h.to(dev, non_blocking=True) for h in host]

    def step(img):
        raw = net(img)
        loss = sum(r.float().square().mean() for r in raw) * world
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(warm):
        step(dimg[i % nbuf])
    torch.cuda.synchronize()
    ll.COUNT[0] = 0
    step(dimg[0])
    launches_per_step = ll.COUNT[0]
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(dimg[i % nbuf])
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_total = e0.elapsed_time(e1)
    # end to end: pinned host uint8 batch in, loss scalar out, every step
    stage = torch.empty_like(dimg[0])
    loss_host = torch.empty((), dtype=torch.float32).pin_memory()
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(args.steps):
        stage.copy_(host[i % nbuf], non_blocking=True)
        loss_host.copy_(step(stage).detach(), non_blocking=True)
    t1.record()
    barrier()
    ms_e2e = t0.elapsed_time(t1)
    ll.PROFILE = []
    step(dimg[0])
    torch.cuda.synchronize()
    prof = {}
    for name, a, b, nbytes, flops in ll.PROFILE:
        d = prof.setdefault(name, [0.0, 0, 0.0, 0.0])
        d[0] += a.elapsed_time(b); d[1] += 1; d[2] += nbytes; d[3] += flops
    ll.PROFILE = None
    t = torch.tensor([ms_total, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = t.tolist()
    if rank == 0:
        imgs = args.batch * world * args.steps
        tot_ms = sum(v[0] for v in prof.values())
        if args.profile_kernels:
            for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0]):
                sys.stderr.write(f"[bench] {k:28s} {v[1]:4d} calls {v[0]:9.3f} ms {100 * v[0] / tot_ms:5.1f}%\n")
        _emit(json.dumps({
            "metric": "LEAD-YOLO images/sec @640", "value": imgs / (ms_total / 1e3), "unit": "images/s",
            "n_gpus": world, "steps": args.steps, "warmup": warm, "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "LEAD-YOLO.yaml training fwd+bwd+SGD bs=64/GPU 640x640 (configs[2]), surrogate "
                                   "loss on the raw heads", "batch_per_gpu": args.batch,
                       "parallelism": f"DDP dp{world} (NCCL gradient all-reduce)" if world > 1 else "single GPU",
                       "l2": "inputs rotate over 3 batches; activations >> L2", "cuda_graph": False},
            "clocks": clocks,
            "e2e": {"value": imgs / (ms_e2e / 1e3), "unit": "images/s", "h2d_bytes_per_step": int(host[0].numel()),
                    "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches_per_step * args.steps),
            "kernel_ms_per_step": {k: round(v[0], 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])[:8]},
            "train_roofline_images_per_s_per_gpu": 40600.0}))
    if world > 1:
        dist.destroy_process_group()


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
                    help="infer = BASELINE configs[1] (default, the driver's line); train = configs[2]: forward + "
                         "backward + SGD step, DDP over NCCL when --gpus > 1")
    ap.add_argument("--batch", type=int, default=BS)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-kernels", action="store_true", help="print the per-entry-point time table to stderr")
    args = ap.parse_args()
    warm = max(args.warmup, 3)

    if args.impl == "reference":
        rank = int(os.environ.get("RANK", "0"))
        if rank != 0:
            return
        sample = 4
        ips, ms, cores = cpu_reference_run(args.steps, min(warm, 3), sample)
        _emit(json.dumps({
            "impl": "reference", "metric": "LEAD-YOLO images/sec @640", "value": ips, "unit": "images/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": min(warm, 3), "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "LEAD-YOLO.yaml batched inference 640x640 (forward + Detect decode)",
                       "sample": f"{sample} images per step"},
            "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": "port",
                             "sample": f"{sample} images x {args.steps} steps, fp32, oracle/restate.py"},
            "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    rank, local, world = dist_setup(args.gpus)
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from lead_yolo_b200 import lowlevel as ll
    from lead_yolo_b200.model import build_lead_yolo
    if args.mode == "train":
        return train_bench(args, warm, rank, local, world, dev)

    model = build_lead_yolo("lead", 1, dev, seed=0).eval()
    nbuf = 3   # rotate 3 input batches (3 x 78.6 MB > 126 MB L2)
    host = [sar_like_batch(args.batch, IMG, 100 * rank + i).pin_memory() for i in range(nbuf)]
    dimg = [h.to(dev, non_blocking=True) for h in host]
    torch.cuda.synchronize()

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
        # batch in, decoded predictions (B,25200,6) fp32 back on the host, every step; copies overlap the replay
        from lead_yolo_b200.engine import InferenceEngine
        eng = InferenceEngine(model, args.batch, IMG, slots=int(os.environ.get("LY_ENGINE_SLOTS", "3")), device=dev)
        out_host = eng.host_out[0]

        def e2e_loop(k):
            # serving loop: the host keeps slots-1 batches in flight and consumes the oldest result before it submits the
            # next batch into that result's slot.  (With one batch in flight a step costs download + upload back to back,
            # 0.7 + 1.4 ms of PCIe time; with two the copies hide behind the replay — tests/dev/e2e_parts.py.)
            pending = collections.deque()
            for i in range(k):
                pending.append(eng.submit(host[i % nbuf]))
                if len(pending) >= eng.slots:
                    eng.result(pending.popleft())
            while pending:
                eng.result(pending.popleft())

        e2e_loop(3)
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(eng.s_in)                    # all streams idle after the barrier: first H2D starts here
        e2e_loop(args.steps)
        t1.record(eng.s_out)                   # after the last D2H
        barrier()
        ms_e2e = t0.elapsed_time(t1)

        # ---- per-entry-point device time (one extra, untimed step with events around every C-ABI call)
        # The device first sleeps ~25 ms so that the host enqueues the whole step ahead of it: each event pair then
        # brackets the kernel alone, not the Python / ctypes launch latency in front of it.
        ll.PROFILE = []
        torch.cuda._sleep(int(5e7))
        model(dimg[0])
        torch.cuda.synchronize()
        prof = {}
        for name, a, b, nbytes, flops in ll.PROFILE:
            d = prof.setdefault(name, [0.0, 0, 0.0, 0.0])
            d[0] += a.elapsed_time(b)
            d[1] += 1
            d[2] += nbytes
            d[3] += flops
        ll.PROFILE = None

    t = torch.tensor([ms_total, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = t.tolist()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    imgs = args.batch * world * args.steps
    value = imgs / (ms_total / 1e3)
    e2e = imgs / (ms_e2e / 1e3)
    hbm, tf, which = peaks()
    tot_ms = sum(v[0] for v in prof.values())
    if args.profile_kernels:
        for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0]):
            sys.stderr.write(f"[bench] {k:28s} {v[1]:4d} launches {v[0]:9.3f} ms {100 * v[0] / tot_ms:5.1f}%"
                             f"  {v[2] / 1e6:10.1f} MB {v[3] / 1e9:9.2f} GFLOP\n")
    top = max(prof.items(), key=lambda kv: kv[1][0])
    tname, (tms, tn, tbytes, tflops) = top
    hbm_time = tbytes / (hbm * 1e9)
    tc_time = tflops / (tf * 1e12)
    if tc_time > hbm_time:
        roof = {"bound": "tensor", "achieved": tflops / (tms / 1e3) / 1e12, "peak": tf, "unit": "TFLOP/s"}
    else:
        roof = {"bound": "hbm", "achieved": tbytes / (tms / 1e3) / 1e9, "peak": hbm, "unit": "GB/s"}
    roof["frac"] = roof["achieved"] / roof["peak"]
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r01_conv_traffic.json")
    if tname == "ly_conv_fwd" and os.path.exists(tpath) and args.batch == BS:
        tj = json.load(open(tpath))   # DRAM bytes of all launches of this kernel in one step (ncu --set full capture)
        if tj.get("launches_per_step") == tn:
            traffic = tj["dram_bytes_read_per_step"] + tj["dram_bytes_write_per_step"]
    roof.update({"traffic": traffic, "kernel": tname, "launches_per_step": tn, "share_of_step": tms / tot_ms,
                 "peak_source": which, "algorithmic_bytes_per_step": tbytes, "algorithmic_flops_per_step": tflops})

    line = {
        "metric": "LEAD-YOLO images/sec @640", "value": value, "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": warm, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "LEAD-YOLO.yaml batched inference bs=64/GPU 640x640 bf16 (configs[1]), forward + "
                               "Detect decode, random init",
                   "batch_per_gpu": args.batch, "parallelism": f"replicas x{world} (no collective)",
                   "l2": f"inputs rotate over {nbuf} batches ({nbuf}x78.6 MB > 126 MB L2); per-step activation "
                         "traffic >> L2", "cuda_graph": graphs is not None},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": int(host[0].numel()),
                "d2h_bytes_per_step": int(out_host.numel() * 4), "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches_per_step * args.steps),
        "roofline": roof,
        "model_roofline": {"images_per_s_bound": 1.0 / max(FWD_MB_PER_IMG * 1e6 / (hbm * 1e9),
                                                            FWD_GFLOP_PER_IMG * 1e9 / (tf * 1e12)),
                           "frac": None},
    }
    # whole-model bound: per-module max(hbm, tensor) summed (132.7k img/s per GPU)
    line["model_roofline"] = {"images_per_s_bound_per_gpu": 132700.0, "frac": value / world / 132700.0}
    if not args.no_cpu_baseline and world >= 1:
        try:
            ips, ms, cores = cpu_reference_run(3, 1, 2)
            line["cpu_baseline"] = {"value": ips, "unit": "images/s", "cores": cores, "kind": "port",
                                    "sample": "2 images x 3 steps (+1 warm-up), fp32 eval forward + decode, "
                                              "oracle/restate.py on all host threads"}
        except Exception as e:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": "images/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"failed: {e}"}
    _emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


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
