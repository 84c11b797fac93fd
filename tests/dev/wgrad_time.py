"""Dev: time ly_conv_wgrad at the LEAD-YOLO layer shapes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lead_yolo_b200 import lowlevel as ll   # noqa: E402


def t(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for (n, cin, cout, h, k) in ((64, 64, 64, 80, 3), (64, 64, 64, 80, 1), (64, 128, 128, 40, 3), (64, 128, 128, 40, 1), (64, 256, 256, 20, 3),
                             (64, 128, 64, 80, 1), (64, 64, 128, 80, 1), (64, 8, 8, 160, 3), (64, 336, 128, 40, 1)):
    x = torch.randn(n, cin, h, h, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dz = torch.randn(n, cout, h, h, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    us = t(lambda: ll.conv_wgrad([x], dz, cout, k=k, stride=1, pad=k // 2))
    fl = 2.0 * n * h * h * cin * k * k * cout
    print(f"wgrad {cin:4d}->{cout:4d} k={k} @{h}x{h}: {us:7.1f} us  ({fl / us / 1e6:6.0f} TFLOP/s)")
