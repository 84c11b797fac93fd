"""Dev: time ly_rf_generate_bwd pieces (dx / wgrad) at the LEAD-YOLO layer shapes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from lead_yolo_b200 import lowlevel as ll, ops

def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

for (N, C, H, k, s) in ((64, 128, 80, 3, 2), (64, 256, 40, 3, 2), (64, 160, 20, 1, 1), (64, 256, 40, 1, 1)):
    ho, wo = ops.rf_out_size(H, H, k, s)
    x = torch.randn(N, C, H, H, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dz = torch.randn(N, k * k * C, ho, wo, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wgk = torch.randn(k * k, k * k, C, device="cuda")
    dx = torch.empty_like(x)
    dw = torch.zeros_like(wgk)
    scr = ll.rf_wgrad_scratch(C, k, "cuda")
    a = t(lambda: ll.call("ly_rf_generate_bwd", dz, x, C, N, H, H, C, k, s, ho, wo, wgk, None, dx, C, None, None, 0))
    b = t(lambda: ll.call("ly_rf_generate_bwd", dz, x, C, N, H, H, C, k, s, ho, wo, wgk, None, None, C, dw, scr, scr.numel()))
    print(f"C={C} {H}x{H} k={k} s={s}: dx {a:7.1f} us   wgrad {b:7.1f} us   (dZ {dz.numel()*2/1e6:.0f} MB)")
