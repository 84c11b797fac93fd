"""Which leg of the upload / run / download pipeline costs the end-to-end step its overlap?  python tests/dev/e2e_parts.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from bench import sar_like_batch
from lead_yolo_b200.model import build_lead_yolo
from lead_yolo_b200.engine import InferenceEngine

model = build_lead_yolo("lead", 1, "cuda", seed=0).eval()
eng = InferenceEngine(model, 64, 640, slots=2, device="cuda")
host = [sar_like_batch(64, 640, i).pin_memory() for i in range(3)]


def loop(up, down, steps=30):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for st in (eng.s_in, eng.s_run, eng.s_out):
        st.wait_event(e0)
    for n in range(steps):
        s = n % 2
        if up:
            with torch.cuda.stream(eng.s_in):
                if n >= 2:
                    eng.s_in.wait_event(eng.ev_run[s])
                eng.inp[s].copy_(host[n % 3], non_blocking=True)
                eng.ev_in[s].record(eng.s_in)
        with torch.cuda.stream(eng.s_run):
            if up:
                eng.s_run.wait_event(eng.ev_in[s])
            if down and n >= 2:
                eng.s_run.wait_event(eng.ev_out[s])
            eng.graphs[s].replay()
            eng.ev_run[s].record(eng.s_run)
        if down:
            with torch.cuda.stream(eng.s_out):
                eng.s_out.wait_event(eng.ev_run[s])
                eng.host_out[s].copy_(eng.out[s], non_blocking=True)
                eng.ev_out[s].record(eng.s_out)
    cur = torch.cuda.current_stream()
    for st in (eng.s_in, eng.s_run, eng.s_out):
        cur.wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


for name, u, d in (("run only", 0, 0), ("upload + run", 1, 0), ("run + download", 0, 1), ("all", 1, 1)):
    loop(u, d, 6)
    print(f"{name:16s} {loop(u, d):.3f} ms/step")
