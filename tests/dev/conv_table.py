"""Per-call table of ly_conv_fwd in the bs=64 640x640 forward: shape, time, achieved GB/s and TFLOP/s.
python tests/dev/conv_table.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from bench import sar_like_batch
from lead_yolo_b200 import lowlevel as ll
from lead_yolo_b200.model import build_lead_yolo

model = build_lead_yolo("lead", 1, "cuda", seed=0).eval()
img = sar_like_batch(64, 640, 0).cuda()
rec = []
orig = ll.conv_gemm


def wrapped(srcs, w_packed, cout, k=1, stride=1, pad=0, **kw):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = orig(srcs, w_packed, cout, k=k, stride=stride, pad=pad, **kw)
    e1.record()
    n, _, h, w = srcs[0].shape
    cin = [int(s.shape[1]) for s in srcs]
    ho, wo = out.shape[2], out.shape[3]
    nbytes = sum(s.numel() * 2 for s in srcs) + out.numel() * out.element_size() + w_packed.numel() * 2
    if kw.get("residual") is not None:
        nbytes += kw["residual"].numel() * 2
    flops = 2.0 * n * ho * wo * sum(cin) * k * k * cout if kw.get("patch_scatter") is None else 2.0 * n * h * w * sum(cin) * cout
    rec.append((e0, e1, f"{n}x{h}x{w} cin={cin} k={k}s{stride} cout={cout} act={kw.get('act','none')}"
                f"{' res' if kw.get('residual') is not None else ''}{' f32' if kw.get('out_fp32') else ''}", nbytes, flops))
    return out


with torch.no_grad():
    for _ in range(3):
        model(img)
    ll.conv_gemm = wrapped
    import lead_yolo_b200.ops as ops
    ops.ll.conv_gemm = wrapped
    torch.cuda._sleep(int(5e7))   # the device sleeps while the host enqueues: events then time kernels only
    model(img)
torch.cuda.synchronize()
tot = 0.0
rows = []
for e0, e1, desc, nb, fl in rec:
    ms = e0.elapsed_time(e1)
    tot += ms
    rows.append((ms, desc, nb, fl))
print(f"{len(rows)} conv calls, {tot:.3f} ms")
for i, (ms, desc, nb, fl) in enumerate(rows):
    print(f"{i:2d} {ms*1e3:7.1f} us  {nb/ms/1e6:7.0f} GB/s  {fl/ms/1e9:6.0f} TF/s  {nb/1e6:7.1f} MB  {desc}")
